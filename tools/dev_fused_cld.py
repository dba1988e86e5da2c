"""Dev harness: kernel-only timings of the resident-key dense kernels (csrc/fused_cld.cu) at the AFF-Tiny-1/5 stage-3 shape
beside the round-1 slot kernels, and of the bias-generator gradient that follows the backward.
Run on the GPU box:  python tools/dev_fused_cld.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ml-autofocusformermod_b200"))
import torch
from affb200 import fused, fused_cl
from affb200.aff_transformer import build_pe_lookup
from affb200.point_utils import space_filling_cluster, build_nbhd
from tools.dev_fused_cl import positions, timeit


def main():
    h = w = 56
    n, B, H, c, m, M, img = 125, 256, 8, 32, 8, 48, 224
    rpw = img // 4 - 1
    tw = 2 * rpw + 1
    pos = positions(h, w, n, B, seed=n).cuda()
    p_s, mean, _, _, _ = space_filling_cluster(pos, m, h, w)
    nbhd, mask, pe_idx, _ = build_nbhd(p_s, mean, m, M, rpw)
    fused_cl.hint_geometry(nbhd, p_s, rpw, tw)
    g = torch.Generator().manual_seed(1)
    dt = torch.bfloat16
    q = torch.randn(B, n, H * c, generator=g).to(dt).cuda()
    kv = torch.randn(B, n, 2 * H * c, generator=g).to(dt).cuda()
    table = (0.5 * torch.randn(tw * tw, H, generator=g)).to(dt).cuda()
    bk, bv = torch.randn(H * c, generator=g).cuda(), torch.randn(H * c, generator=g).cuda()
    pre_table = build_pe_lookup(img)[2].cuda()
    for name, env in (("dense", "1"), ("slots", "0")):
        os.environ["AFFB200_FUSED_CLD"] = env
        calls = fused.raw_calls(q, kv, H, c ** -0.5, nbhd, pe_idx, mask, table, bk, bv, pre_table=pre_table)
        fwd, bwd = calls
        print("%-6s %-20s fwd %.1f us   bwd group %.1f us" % (name, getattr(fwd, "family", "resident-key slots"), timeit(fwd) * 1e3, timeit(bwd) * 1e3))
    clp = fused_cl.cl_plan(nbhd, mask, pe_idx, tw * tw)
    print("plan: sub_r", clp.sub_r, "stats", clp.stats.tolist())
    ds = torch.randn(B, H, n, M, generator=g).to(dt).cuda()
    keep = []
    pg = fused._raw_pe_grad(pre_table, ds, clp.q_pe, B, H, n * M, keep)
    print("pe_feat_grad alone %.1f us" % (timeit(pg) * 1e3))


if __name__ == "__main__":
    main()
