"""BASELINE config 1 ("mandatory"): the reference's pure-PyTorch gather expressions
(clusten/test_qk_kernel.py:39-40, test_av_kernel.py:39-40, test_wf_kernel.py:40-41 — restated in oracle/clusten_ref.py)
timed on this box's host cores, beside the B200 operators at the same shapes: B=2, N=3136, heads=2, nbhd=48, per-head
dim 16 (AFF-Mini stage 1) and 32 (the kernel-test head dim); WF 784 <- 3136, C=32, 4 inner channels.  Forward + backward
(`out.mean().backward()` as the reference's tests do), fp32 on the CPU; fp32 and bf16 on the GPU.

  python tests/config1_table.py --out profiles/r02_config1_table
"""
import argparse
import json
import os
import statistics
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ml-autofocusformermod_b200"))
sys.path.insert(0, ROOT)

from oracle import clusten_ref as R  # noqa: E402
from clusten import CLUSTENQKFunction, CLUSTENAVFunction, CLUSTENWFFunction  # noqa: E402
from tests.opbench import make_idx, flush_l2, peak_gbs  # noqa: E402

QKAV = {"c16": dict(B=2, H=2, N=3136, c=16, M=48), "c32": dict(B=2, H=2, N=3136, c=32, M=48)}
WF = dict(B=2, Nq=784, Nk=3136, M=48, IC=4, C=32)


def operands(op, S, dtype, dev, idx_mode, seed):
    g = torch.Generator().manual_seed(seed)
    if op == "wf":
        a = torch.randn(S["B"], S["Nq"], S["M"], S["IC"], generator=g)
        b = torch.randn(S["B"], S["Nk"], S["C"], generator=g)
        idx = make_idx(S["B"], S["Nq"], S["Nk"], S["M"], idx_mode, torch.device("cpu"))
    else:
        B, H, N, c, M = S["B"], S["H"], S["N"], S["c"], S["M"]
        a = torch.randn(B, H, N, c, generator=g) if op == "qk" else torch.softmax(torch.randn(B, H, N, M, generator=g), -1)
        b = torch.randn(B, H, N, c, generator=g)
        idx = make_idx(B, N, N, M, idx_mode, torch.device("cpu"))
    return a.to(dtype).to(dev).requires_grad_(True), b.to(dtype).to(dev).requires_grad_(True), idx.to(dev)


def alg_bytes(op, S, s):
    if op == "wf":
        B, Nq, Nk, M, IC, C = S["B"], S["Nq"], S["Nk"], S["M"], S["IC"], S["C"]
        return (s * (B * Nq * M * IC + B * Nk * C + B * Nq * IC * C) + 8 * B * Nq * M
                + s * (B * Nq * IC * C + 2 * B * Nq * M * IC + 2 * B * Nk * C) + 8 * B * Nq * M)
    B, H, N, c, M = S["B"], S["H"], S["N"], S["c"], S["M"]
    if op == "qk":
        return s * (2 * B * H * N * c + B * H * N * M) + s * (B * H * N * M + 4 * B * H * N * c) + 16 * B * N * M
    return s * (B * H * N * M + 2 * B * H * N * c) + s * (2 * B * H * N * M + 4 * B * H * N * c) + 16 * B * N * M


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    peak, how = peak_gbs()
    cpu_fn = {"qk": R.qk_ref, "av": R.av_ref, "wf": R.wf_ref}
    gpu_fn = {"qk": CLUSTENQKFunction.apply, "av": CLUSTENAVFunction.apply, "wf": CLUSTENWFFunction.apply}
    rows = []
    for op, name, S in [("qk", "c16", QKAV["c16"]), ("av", "c16", QKAV["c16"]), ("qk", "c32", QKAV["c32"]),
                        ("av", "c32", QKAV["c32"]), ("wf", "784<-3136", WF)]:
        for mode in ("S", "U"):
            x, y, idx = operands(op, S, torch.float32, "cpu", mode, 1001)

            def cpu_step():
                x.grad = y.grad = None
                cpu_fn[op](x, y, idx).mean().backward()
            for _ in range(2):
                cpu_step()
            ts = []
            for _ in range(5):
                t0 = time.perf_counter()
                cpu_step()
                ts.append((time.perf_counter() - t0) * 1e3)
            cpu_ms = statistics.median(ts)
            row = dict(op=op, shape=name, idx=mode, cpu_ms=round(cpu_ms, 2), cpu_cores=cores, dims=S)
            for dt, dn in ((torch.float32, "f32"), (torch.bfloat16, "bf16")):
                gx, gy, gidx = operands(op, S, dt, "cuda", mode, 1001)
                fn = gpu_fn[op]

                def gpu_step():
                    gx.grad = gy.grad = None
                    fn(gx, gy, gidx).mean().backward()
                for _ in range(5):
                    gpu_step()
                torch.cuda.synchronize()
                gts = []
                for _ in range(20):
                    flush_l2()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    gpu_step()
                    e1.record()
                    e1.synchronize()
                    gts.append(e0.elapsed_time(e1) * 1e3)
                us = statistics.median(gts)
                nb = alg_bytes(op, S, 4 if dt == torch.float32 else 2)
                row["gpu_%s_us" % dn] = round(us, 1)
                row["gpu_%s_GBps" % dn] = round(nb / us / 1e3, 1)
                row["gpu_%s_alg_MB" % dn] = round(nb / 1e6, 2)
            rows.append(row)
            print(json.dumps(row), flush=True)
    md = ["| op | shape | idx | CPU fwd+bwd ms (fp32, %d threads) | B200 fp32 us | GB/s | B200 bf16 us | GB/s |" % cores,
          "|---|---|---|---:|---:|---:|---:|---:|"]
    for r in rows:
        md.append("| %s | %s | %s | %.2f | %.1f | %.1f | %.1f | %.1f |" % (
            r["op"], r["shape"], r["idx"], r["cpu_ms"], r["gpu_f32_us"], r["gpu_f32_GBps"], r["gpu_bf16_us"], r["gpu_bf16_GBps"]))
    md.append("")
    md.append("GPU times are CUDA-event medians of 20 forward+backward calls through the autograd.Function surface (index plans "
              "cached across calls as in a model stage), L2 flushed between calls; they include the host-side launch path, "
              "which dominates at this size (the working set is 6-13 MB: ~2 us of HBM time at the measured %.0f GB/s, %s). "
              "GB/s = SURVEY 8(d) algorithmic bytes of forward + backward / time." % (peak, how))
    print("\n".join(md))
    if a.out:
        open(a.out + ".md", "w").write("\n".join(md) + "\n")
        json.dump(dict(rows=rows, torch=torch.__version__, cores=cores, peak_gbs=peak), open(a.out + ".json", "w"), indent=1)


if __name__ == "__main__":
    main()
