"""Dev harness: the raw fused-attention calls (what bench.py times for the roofline) at one AFF-Tiny-1/5 stage shape,
run a few times — for ncu captures of single kernels.
    python tools/dev_raw_stage.py <stage 1|2|3|all> [fwd|bwd|both] [reps] [--time]"""
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

SHAPES = {1: (3136, 2), 2: (627, 4), 3: (125, 8)}


def run_stage(stage, which, reps, time_it):
    n, H = SHAPES[stage]
    B, c, m, M, img = 256, 32, 8, 48, 224
    rpw = img // 4 - 1
    tw = 2 * rpw + 1
    pos = positions(56, 56, n, B, seed=n).cuda()
    p_s, mean, _, _, _ = space_filling_cluster(pos, m, 56, 56)
    nbhd, mask, pe_idx, _ = build_nbhd(p_s, mean, m, M, rpw)
    fused_cl.hint_geometry(nbhd, p_s, rpw, tw)
    g = torch.Generator().manual_seed(1)
    dt = torch.bfloat16
    q = torch.randn(B, n, H * c, generator=g).to(dt).cuda()
    kv = torch.randn(B, n, 2 * H * c, generator=g).to(dt).cuda()
    table = (0.5 * torch.randn(tw * tw, H, generator=g)).to(dt).cuda()
    bk, bv = torch.randn(H * c, generator=g).cuda(), torch.randn(H * c, generator=g).cuda()
    pre_table = build_pe_lookup(img)[2].cuda()
    fwd, bwd = fused.raw_calls(q, kv, H, c ** -0.5, nbhd, pe_idx, mask, table, bk, bv, pre_table=pre_table)
    torch.cuda.synchronize()
    torch.cuda.nvtx.range_push("stage%d" % stage)
    for _ in range(reps):
        if which in ("fwd", "both"):
            fwd()
        if which in ("bwd", "both"):
            bwd()
    torch.cuda.synchronize()
    torch.cuda.nvtx.range_pop()
    if time_it:
        print("stage %d: fwd %.1f us  bwd group %.1f us" % (stage, timeit(fwd) * 1e3, timeit(bwd) * 1e3))


def main():
    stage = sys.argv[1] if len(sys.argv) > 1 else "1"
    which = sys.argv[2] if len(sys.argv) > 2 else "both"
    reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
    for st in ([1, 2, 3] if stage == "all" else [int(stage)]):
        run_stage(st, which, reps, "--time" in sys.argv)


if __name__ == "__main__":
    main()
