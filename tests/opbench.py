"""Op-level microbenchmark of the CLUSTEN kernels through the C ABI: CUDA-event time per launch group, achieved
GB/s on the algorithmic bytes of SURVEY.md §8(d), fraction of the measured HBM peak.  Working sets are larger than
L2 (126 MB) at the training shapes; smaller shapes get an L2 flush between iterations.

  python tests/opbench.py [--shapes tiny15_s1,...] [--dtype bf16] [--idx S|U] [--iters 20]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ml-autofocusformermod_b200"))
sys.path.insert(0, ROOT)

from affb200 import _native as N  # noqa: E402
from affb200.plan import get_plan  # noqa: E402

# (B, H, N, c, M) local-attention stages and WF shapes (B, Nq, Nk, M, IC, C)
SHAPES = {
    "cfg1_c16": dict(B=2, H=2, N=3136, c=16, M=48),
    "cfg1_c32": dict(B=2, H=2, N=3136, c=32, M=48),
    "reftest": dict(B=256, H=4, N=196, c=32, M=48),
    "tiny15_s1": dict(B=256, H=2, N=3136, c=32, M=48),
    "tiny15_s2": dict(B=256, H=4, N=627, c=32, M=48),
    "tiny15_s3": dict(B=256, H=8, N=125, c=32, M=48),
    "base384_s1": dict(B=16, H=4, N=9216, c=32, M=144),
    "mixres_j3": dict(B=1, H=2, N=30720, c=32, M=48),
}
WF_SHAPES = {
    "cfg1_wf": dict(B=2, Nq=784, Nk=3136, M=48, IC=4, C=32),
    "tiny15_wf1": dict(B=256, Nq=627, Nk=3136, M=48, IC=4, C=64),
    "tiny15_wf2": dict(B=256, Nq=125, Nk=627, M=48, IC=4, C=128),
    "base384_wf1": dict(B=16, Nq=2304, Nk=9216, M=144, IC=4, C=128),
}
DT = {"f32": torch.float32, "bf16": torch.bfloat16, "f16": torch.float16}


def peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p))["hbm_gbs"], "measured"
    return 6650.0, "fallback"


def make_idx(B, Nq, Nk, M, mode, dev, m=8):
    if mode == "U":
        return torch.randint(Nk, (B, Nq, M), device=dev)
    # S: block-structured like member_idx.gather(nearest_cluster) (aff_transformer.py:480): nnc = M/m clusters of
    # m consecutive rows around the token's own cluster
    nnc = max(1, M // m)
    k = (Nk + m - 1) // m
    own = (torch.arange(Nq, device=dev) * Nk // max(Nq, 1)) // m
    cl = (own[:, None] + torch.arange(nnc, device=dev)[None, :] - nnc // 2).clamp_(0, k - 1)
    idx = (cl[:, :, None] * m + torch.arange(m, device=dev)[None, None, :]).reshape(Nq, nnc * m)
    idx = idx.clamp_(max=Nk - 1)
    if idx.shape[1] < M:
        idx = torch.cat([idx, idx[:, : M - idx.shape[1]]], 1)
    return idx[None].expand(B, -1, -1).contiguous()


_flush = None
NO_DENSE = False
REF = False
WARM = 5


def flush_l2():
    global _flush
    if _flush is None:
        _flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    _flush.zero_()


def time_call(fn, iters, warm, flush):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush:
            flush_l2()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def bench_qkav(name, S, dtype, mode, iters):
    B, H, Nn, c, M = S["B"], S["H"], S["N"], S["c"], S["M"]
    dev = torch.device("cuda")
    s = torch.empty((), dtype=dtype).element_size()
    q = torch.randn(B, H, Nn, c, device=dev).to(dtype)
    k = torch.randn(B, H, Nn, c, device=dev).to(dtype)
    attn = torch.softmax(torch.randn(B, H, Nn, M, device=dev), -1).to(dtype)
    idx = make_idx(B, Nn, Nn, M, mode, dev, m=8 if M % 8 == 0 and M != 144 else 24)
    o_attn, o_feat = torch.empty_like(attn), torch.empty_like(q)
    d_q, d_k = torch.empty_like(q), torch.empty_like(k)
    pl = get_plan(idx)
    dense = dtype in (torch.float16, torch.bfloat16) and c in (16, 32) and not NO_DENSE
    pf, pb = pl.struct(idx, Nn, dense, False), pl.struct(idx, Nn, dense, True)
    print(json.dumps(dict(shape=name, idx=mode, plan=pl.stats)), flush=True)
    st = N.stream_ptr(dev)
    de = N.dtype_enum(q)
    A = (B, H, Nn, Nn, M, c, de, st)
    ops = {
        "qk_fwd": (lambda: N.check(N.LIB.affb200_qk_fwd(N.p(q), N.p(k), N.p(idx), pf, N.p(o_attn), *A), "qk"),
                   s * (2 * B * H * Nn * c + B * H * Nn * M) + 8 * B * Nn * M),
        "qk_bwd": (lambda: N.check(N.LIB.affb200_qk_bwd(N.p(attn), N.p(q), N.p(k), N.p(idx), pb,
                                                         N.p(d_q), N.p(d_k), *A), "qkb"),
                   s * (B * H * Nn * M + 4 * B * H * Nn * c) + 8 * B * Nn * M),
        "av_fwd": (lambda: N.check(N.LIB.affb200_av_fwd(N.p(attn), N.p(k), N.p(idx), pf, N.p(o_feat), *A), "av"),
                   s * (B * H * Nn * M + 2 * B * H * Nn * c) + 8 * B * Nn * M),
        "av_bwd": (lambda: N.check(N.LIB.affb200_av_bwd(N.p(q), N.p(attn), N.p(k), N.p(idx), pb,
                                                         N.p(o_attn), N.p(d_k), *A), "avb"),
                   s * (2 * B * H * Nn * M + 4 * B * H * Nn * c) + 8 * B * Nn * M),
    }
    if REF and dtype in (torch.float32, torch.float16):
        # the reference's own CUDA kernels (oracle/_ref) on the same tensors, called as clusten/clusten.py calls them
        from oracle import ref_cuda as RC
        nb = {k_: v_[1] for k_, v_ in ops.items()}
        ops.update({
            "ref_qk_fwd": (lambda: RC.forward("qk", q, k, idx), nb["qk_fwd"]),
            "ref_qk_bwd": (lambda: RC.backward("qk", attn, q, k, idx), nb["qk_bwd"]),
            "ref_av_fwd": (lambda: RC.forward("av", attn, k, idx), nb["av_fwd"]),
            "ref_av_bwd": (lambda: RC.backward("av", q, attn, k, idx), nb["av_bwd"]),
        })
    return run_ops(name, ops, dtype, mode, iters, S)


def bench_wf(name, S, dtype, mode, iters):
    B, Nq, Nk, M, IC, C = S["B"], S["Nq"], S["Nk"], S["M"], S["IC"], S["C"]
    dev = torch.device("cuda")
    s = torch.empty((), dtype=dtype).element_size()
    w = torch.randn(B, Nq, M, IC, device=dev).to(dtype)
    f = torch.randn(B, Nk, C, device=dev).to(dtype)
    idx = make_idx(B, Nq, Nk, M, mode, dev, m=8 if M != 144 else 24)
    out = torch.empty(B, Nq, IC, C, device=dev, dtype=dtype)
    go = torch.randn_like(out)
    d_w, d_f = torch.empty_like(w), torch.empty_like(f)
    pb = get_plan(idx).struct(idx, Nk, False, True)
    st = N.stream_ptr(dev)
    A = (B, Nq, Nk, M, IC, C, N.dtype_enum(w), st)
    ops = {
        "wf_fwd": (lambda: N.check(N.LIB.affb200_wf_fwd(N.p(w), N.p(f), N.p(idx), None, N.p(out), *A), "wf"),
                   s * (B * Nq * M * IC + B * Nk * C + B * Nq * IC * C) + 8 * B * Nq * M),
        "wf_bwd": (lambda: N.check(N.LIB.affb200_wf_bwd(N.p(go), N.p(w), N.p(f), N.p(idx), pb,
                                                         N.p(d_w), N.p(d_f), *A), "wfb"),
                   s * (B * Nq * IC * C + 2 * B * Nq * M * IC + 2 * B * Nk * C) + 8 * B * Nq * M),
    }
    if REF and dtype in (torch.float32, torch.float16):
        from oracle import ref_cuda as RC
        nb = {k_: v_[1] for k_, v_ in ops.items()}
        ops.update({
            "ref_wf_fwd": (lambda: RC.forward("wf", w, f, idx), nb["wf_fwd"]),
            "ref_wf_bwd": (lambda: RC.backward("wf", go, w, f, idx), nb["wf_bwd"]),
        })
    return run_ops(name, ops, dtype, mode, iters, S)


def run_ops(name, ops, dtype, mode, iters, S):
    peak, how = peak_gbs()
    rows = []
    for op, (fn, nbytes) in ops.items():
        flush = nbytes < (512 << 20)
        med, best = time_call(fn, iters, WARM, flush)
        gbs = nbytes / (med * 1e-3) / 1e9
        row = dict(shape=name, op=op, dtype=str(dtype).replace("torch.", ""), idx=mode, ms=round(med, 4),
                   ms_best=round(best, 4), alg_MB=round(nbytes / 1e6, 2), GBps=round(gbs, 1),
                   frac=round(gbs / peak, 3), peak=how, l2_flush=flush, dims=S)
        print(json.dumps(row), flush=True)
        rows.append(row)
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", default="cfg1_c16,tiny15_s1,tiny15_s2,tiny15_s3,cfg1_wf,tiny15_wf1,tiny15_wf2")
    ap.add_argument("--dtype", default="bf16")
    ap.add_argument("--idx", default="S,U")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--out", default="")
    ap.add_argument("--warm", type=int, default=5)
    ap.add_argument("--no-dense", action="store_true", help="force the generic (row-gather) kernels")
    ap.add_argument("--ref", action="store_true",
                    help="also time the reference's own CUDA kernels from oracle/_ref (f32/f16 only)")
    a = ap.parse_args()
    global NO_DENSE, WARM, REF
    NO_DENSE = a.no_dense
    REF = a.ref
    WARM = a.warm
    rows = []
    for dt in a.dtype.split(","):
        for mode in a.idx.split(","):
            for nm in a.shapes.split(","):
                if nm in SHAPES:
                    rows += bench_qkav(nm, SHAPES[nm], DT[dt], mode, a.iters)
                else:
                    rows += bench_wf(nm, WF_SHAPES[nm], DT[dt], mode, a.iters)
                torch.cuda.empty_cache()
    if a.out:
        with open(a.out, "w") as f:
            for r in rows:
                f.write(json.dumps(r) + "\n")


if __name__ == "__main__":
    main()
