"""The drop-in CLUSTEN QK/AV operators at the AFF-Tiny-1/5 stage shapes with neighbourhoods from the real index
pipeline (grid/subsampled positions -> clustering -> kNN -> nbhd build): CUDA-event time, GB/s on SURVEY §8(d) bytes.
   python tools/densebench.py [--stages 1,2,3] [--batch 256] [--iters 10] [--dtype bf16]"""
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
from affb200.point_utils import space_filling_cluster, build_nbhd  # noqa: E402
from tools.fusedbench import STAGES, positions, ev_time  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--stages", default="1,2,3")
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--dtype", default="bf16")
    a = ap.parse_args()
    dt = {"bf16": torch.bfloat16, "f16": torch.float16, "f32": torch.float32}[a.dtype]
    peak = 6539.5
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = json.load(open(p))["hbm_gbs"]
    B = a.batch
    for s in [int(x) for x in a.stages.split(",")]:
        n, H, c = STAGES[s]
        M = 48
        pos = positions(n, B, s).cuda()
        p_s, mean, _, _, _ = space_filling_cluster(pos, 8, 56, 56)
        idx, mask, pe_idx, _ = build_nbhd(p_s, mean, 8, 48, 55)
        q = torch.randn(B, H, n, c, device="cuda").to(dt)
        k = torch.randn(B, H, n, c, device="cuda").to(dt)
        attn = torch.softmax(torch.randn(B, H, n, M, device="cuda"), -1).to(dt)
        o_attn, o_feat, d_q, d_k = torch.empty_like(attn), torch.empty_like(q), torch.empty_like(q), torch.empty_like(k)
        pl = get_plan(idx)
        dense = dt != torch.float32
        pf, pb = pl.struct(idx, n, dense, False), pl.struct(idx, n, dense, True)
        A = (B, H, n, n, M, c, N.dtype_enum(q), N.stream_ptr(q.device))
        P = N.p
        sz = q.element_size()
        ops = {
            "qk_fwd": (lambda: N.LIB.affb200_qk_fwd(P(q), P(k), P(idx), pf, P(o_attn), *A),
                       sz * (2 * B * H * n * c + B * H * n * M) + 8 * B * n * M),
            "av_fwd": (lambda: N.LIB.affb200_av_fwd(P(attn), P(k), P(idx), pf, P(o_feat), *A),
                       sz * (B * H * n * M + 2 * B * H * n * c) + 8 * B * n * M),
            "qk_bwd": (lambda: N.LIB.affb200_qk_bwd(P(attn), P(q), P(k), P(idx), pb, P(d_q), P(d_k), *A),
                       sz * (B * H * n * M + 4 * B * H * n * c) + 8 * B * n * M),
            "av_bwd": (lambda: N.LIB.affb200_av_bwd(P(q), P(attn), P(k), P(idx), pb, P(o_attn), P(d_k), *A),
                       sz * (2 * B * H * n * M + 4 * B * H * n * c) + 8 * B * n * M),
        }
        for nm, (fn, nbytes) in ops.items():
            ms = ev_time(fn, a.iters)
            print(json.dumps(dict(stage=s, op=nm, ms=round(ms, 4), alg_MB=round(nbytes / 1e6, 1),
                                  GBps=round(nbytes / ms / 1e6, 1), frac=round(nbytes / ms / 1e6 / peak, 3),
                                  plan=pl.stats)), flush=True)


if __name__ == "__main__":
    main()
