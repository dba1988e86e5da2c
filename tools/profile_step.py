"""Kernel-time breakdown of one bench step with torch.profiler (CUPTI): python tools/profile_step.py [--workload W] [--batch B]
A quick orientation tool; the judged launch lists come from ncu (profiles/)."""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ml-autofocusformermod_b200"))
import bench_workloads as W  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="")
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--rows", type=int, default=45)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    wl = W.get(a.workload)
    run = wl.build(a.batch or wl.default_batch, dev, 1)
    batch = run.to_device(run.make_host_batch(0))
    for _ in range(3):
        run.step(batch)
    torch.cuda.synchronize()
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        run.step(batch)
        torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=a.rows, max_name_column_width=90))


if __name__ == "__main__":
    main()
