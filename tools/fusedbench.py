"""Fused neighbourhood attention at the AFF-Tiny-1/5 stage shapes with the real index pipeline:
CUDA-event time of forward and backward per stage.   python tools/fusedbench.py [--stages 1,2,3] [--batch 256] [--iters 10]"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ml-autofocusformermod_b200"))
sys.path.insert(0, ROOT)
from affb200.point_utils import space_filling_cluster, build_nbhd  # noqa: E402
from affb200.fused import fused_cluster_attention  # noqa: E402
from affb200.plan import get_plan  # noqa: E402

STAGES = {1: (3136, 2, 32), 2: (627, 4, 32), 3: (125, 8, 32)}


def positions(n, B, seed):
    g = torch.Generator().manual_seed(seed)
    ys, xs = torch.meshgrid(torch.arange(56), torch.arange(56), indexing="ij")
    allp = torch.stack([xs, ys], 2).reshape(-1, 2).float()
    if n == 3136:
        return allp.unsqueeze(0).expand(B, -1, -1).contiguous()
    return torch.stack([allp[torch.randperm(3136, generator=g)[:n]] for _ in range(B)])


def ev_time(fn, iters, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--stages", default="1,2,3")
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--iters", type=int, default=10)
    a = ap.parse_args()
    B = a.batch
    for s in [int(x) for x in a.stages.split(",")]:
        n, H, c = STAGES[s]
        hc = H * c
        pos = positions(n, B, s).cuda()
        p_s, mean, _, _, _ = space_filling_cluster(pos, 8, 56, 56)
        nbhd, mask, pe_idx, _ = build_nbhd(p_s, mean, 8, 48, 55)
        q = torch.randn(B, n, hc, device="cuda").bfloat16().requires_grad_(True)
        kv = torch.randn(B, n, 2 * hc, device="cuda").bfloat16().requires_grad_(True)
        table = torch.randn(111 * 111, H, device="cuda").bfloat16().requires_grad_(True)
        bk = torch.randn(hc, device="cuda").requires_grad_(True)
        bv = torch.randn(hc, device="cuda").requires_grad_(True)
        go = torch.randn(B, n, hc, device="cuda").bfloat16()
        out = fused_cluster_attention(q, kv, H, c ** -0.5, nbhd, pe_idx, mask, table, bk, bv)
        st = get_plan(nbhd).stats
        t_f = ev_time(lambda: fused_cluster_attention(q, kv, H, c ** -0.5, nbhd, pe_idx, mask, table, bk, bv), a.iters)
        t_b = ev_time(lambda: out.backward(go, retain_graph=True), a.iters)
        M = 48
        s_b = 2
        fwd_bytes = s_b * (4 * B * n * hc) + 8 * B * n * M * 2  # q,k,v,out + nbhd_idx + pe_idx (API types)
        print(json.dumps(dict(stage=s, B=B, n=n, H=H, c=c, fwd_ms=round(t_f, 4), bwd_ms=round(t_b, 4), plan=st,
                              fwd_alg_MB=round(fwd_bytes / 1e6, 1), fwd_GBps=round(fwd_bytes / t_f / 1e6, 1))), flush=True)


if __name__ == "__main__":
    main()
