"""aten-level attribution of the kernels inside the graphed training step: runs GraphedTrainStep._fwd_bwd() eagerly (same
code the graph captured: flat parameter shadows, gradient sinks) under torch.profiler and prints the stock (non-affb200)
ops by shape.   python tools/profile_graph_ops.py [--rows 60]"""
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
    ap.add_argument("--rows", type=int, default=60)
    ap.add_argument("--workload", default="", help="e.g. aff_tiny_1_5th_infer: profiles the eager inference forward")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    wl = W.get(a.workload)
    run = wl.build(wl.default_batch, dev, 1)
    batch = run.to_device(run.make_host_batch(0))
    for _ in range(2):
        run.step(batch)
    torch.cuda.synchronize()
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU], record_shapes=True) as prof:
        if getattr(run, "infer", False):
            with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
                run.model(batch[0])
        else:
            run.gstep._fwd_bwd()
        torch.cuda.synchronize()
    rows = []
    for e in prof.key_averages(group_by_input_shape=True):
        if e.key.startswith("aten::") and e.self_device_time_total > 0:
            rows.append((e.self_device_time_total, e.count, e.key, str(e.input_shapes)[:110]))
    rows.sort(reverse=True)
    tot = sum(r[0] for r in rows)
    print("stock aten ops with device time: %.2f ms in %d calls" % (tot / 1e3, sum(r[1] for r in rows)))
    for us, n, k, sh in rows[:a.rows]:
        print("%8.0f us %4d  %-28s %s" % (us, n, k, sh))


if __name__ == "__main__":
    main()
