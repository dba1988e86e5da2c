"""Instructions executed / stall samples per CUDA source line of one kernel in an .ncu-rep (needs -lineinfo and
--import-source on):  python tools/ncu_lines.py rep.ncu-rep kernel_regex [top]"""
import collections
import csv
import subprocess
import sys


def main():
    rep, kern = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv",
                          "--kernel-name", "regex:" + kern], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    agg = collections.OrderedDict()
    tot = tots = 0
    hdr = None
    for r in rows:
        if "Instructions Executed" in r and "Line No" in r:
            hdr = r   # one section per source file, each with its own header
            ci, si, li, src = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Line No"), hdr.index("Source")
            continue
        if hdr is None or len(r) <= ci or not r[li].strip():
            continue  # SASS rows have no line number
        try:
            n, sm = int(r[ci]), int(r[si])
        except ValueError:
            continue
        key = (r[li], r[src].strip()[:100])
        a = agg.setdefault(key, [0, 0])
        a[0] += n
        a[1] += sm
        tot += n
        tots += sm
    print("total warp-instructions %d, samples %d" % (tot, tots))
    for (ln, text), (n, sm) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print("%5.1f%% inst %5.1f%% smp  L%-4s %s" % (100 * n / max(tot, 1), 100 * sm / max(tots, 1), ln, text))


if __name__ == "__main__":
    main()
