"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name: launches, total and share.
   python tools/launch_summary.py gpurun_out/launches.csv [--skip N] > profiles/rNN_launches.md"""
import csv
import sys
from collections import defaultdict


def main():
    path = sys.argv[1]
    skip = int(sys.argv[sys.argv.index("--skip") + 1]) if "--skip" in sys.argv else 0
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    rd = csv.DictReader(lines)
    for r in rd:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        if v != v:
            continue  # ncu reports nan for a few graph nodes (memset nodes)
        unit = r.get("Metric Unit", "ns")
        scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3}.get(unit, 1e-3)
        rows.append((int(r["ID"]), r["Kernel Name"], v * scale))
    rows = [r for r in rows if r[0] >= skip]
    agg = defaultdict(lambda: [0, 0.0])
    for _, name, us in rows:
        short = name.split("(")[0]
        short = short if len(short) < 90 else short[:40] + "…" + short[-45:]
        agg[short][0] += 1
        agg[short][1] += us
    tot = sum(v[1] for v in agg.values())
    print("| kernel | launches | total us | share | avg us |")
    print("|---|---:|---:|---:|---:|")
    for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("| `%s` | %d | %.1f | %.1f%% | %.1f |" % (k, n, us, 100 * us / tot, us / n))
    print("| total | %d | %.1f | 100%% | |" % (len(rows), tot))


if __name__ == "__main__":
    main()
