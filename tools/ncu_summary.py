"""Summarise an .ncu-rep: per kernel headline metrics, aggregated stall reasons, and the hottest SASS lines.
   python tools/ncu_summary.py gpurun_out/x.ncu-rep [--kernel regex] [--top 25]"""
import argparse
import csv
import io
import subprocess


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    return rows[0], rows[1], rows[2:]


KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers", "smsp__inst_executed.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "lts__t_sector_hit_rate.pct"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rep")
    ap.add_argument("--kernel", default="")
    ap.add_argument("--top", type=int, default=25)
    ap.add_argument("--no-source", action="store_true")
    a = ap.parse_args()
    hdr, units, data = raw(a.rep)
    seen = set()
    for r in data:
        name = r[hdr.index("Kernel Name")]
        short = name.split("(")[0][-60:]
        if short in seen or (a.kernel and a.kernel not in name):
            continue
        seen.add(short)
        print("==", short)
        for k in KEYS:
            if k in hdr:
                print("   %-62s %s %s" % (k, r[hdr.index(k)], units[hdr.index(k)]))
        st = []
        for i, h in enumerate(hdr):
            if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("per_issue_active.ratio"):
                try:
                    st.append((float(r[i]), h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]))
                except ValueError:
                    pass
        print("   stalls (warps per issue):", ", ".join("%s %.2f" % (n, v) for v, n in sorted(st, reverse=True)[:6]))
    if a.no_source:
        return
    for short in seen:
        kn = short.split("::")[-1].split("<")[0].replace("void ", "").strip()
        out = subprocess.run(["ncu", "-i", a.rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kn],
                             capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(out)))
        h = None
        recs = []
        for row in rows:
            if row and row[0] == "Address":
                if h is not None:
                    break  # first launch only
                h = row
                continue
            if h and len(row) == len(h):
                recs.append(row)
        if not h:
            continue
        si, ii, src = h.index("# Samples"), h.index("Instructions Executed"), h.index("Source")
        tot = sum(int(r[si] or 0) for r in recs) or 1
        print("== hottest SASS of", kn, "(samples, % of kernel, warp-instr, sass)")
        for r in sorted(recs, key=lambda r: -int(r[si] or 0))[:a.top]:
            print("   %6s %5.1f%% %10s  %s" % (r[si], 100.0 * int(r[si] or 0) / tot, r[ii], r[src].strip()[:90]))


if __name__ == "__main__":
    main()
