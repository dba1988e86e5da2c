"""Development check of the block-structured fused attention (forward + backward through autograd) against the
row-union kernels and the fp64 oracle chain; optional timings at the AFF-Tiny-1/5 stage shapes."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ml-autofocusformermod_b200"))
import torch
from affb200 import fused, fused_cl
from affb200.point_utils import space_filling_cluster, build_nbhd
from tools.dev_fused_cl import positions, timeit
from tests.test_fused_attn import _reference


def run(args, H, scale, nbhd, pe_idx, mask, go, cl):
    os.environ["AFFB200_FUSED_CL"] = "1" if cl else "0"
    os.environ["AFFB200_FUSED_IMG"] = "0"
    a = [x.detach().clone().requires_grad_(True) for x in args]
    out = fused.fused_cluster_attention(a[0], a[1], H, scale, nbhd, pe_idx, mask, a[2], a[3], a[4])
    out.backward(go)
    return [out.detach()] + [x.grad for x in a]


def case(h, w, n, B, H, c, m, nbhd_size, img, dtype=torch.bfloat16, check_ref=False, time_it=False):
    rpw = img // 4 - 1
    tw = 2 * rpw + 1
    P = tw * tw
    pos = positions(h, w, n, B, seed=n).cuda()
    p_s, mean, _, _, _ = space_filling_cluster(pos, m, h, w)
    nbhd, mask, pe_idx, _ = build_nbhd(p_s, mean, m, nbhd_size, rpw)
    fused_cl.hint_geometry(nbhd, p_s, rpw, tw)
    g = torch.Generator().manual_seed(7 + n)
    hc = H * c
    args = [torch.randn(B, n, hc, generator=g).to(dtype).cuda(), torch.randn(B, n, 2 * hc, generator=g).to(dtype).cuda(),
            (0.5 * torch.randn(P, H, generator=g)).to(dtype).cuda(), torch.randn(hc, generator=g).cuda(),
            torch.randn(hc, generator=g).cuda()]
    go = torch.randn(B, n, hc, generator=g).to(dtype).cuda()
    scale = c ** -0.5
    new = run(args, H, scale, nbhd, pe_idx, mask, go, True)
    clp = fused_cl.cl_plan(nbhd, mask, pe_idx, P)
    print("case", (h, w, n, B, H, c, m, nbhd_size, str(dtype)), "plan", None if clp is None else (clp.ucap, clp.kbcap, clp.stats.tolist()))
    old = run(args, H, scale, nbhd, pe_idx, mask, go, False)
    names = ["out", "d_q", "d_kv", "d_table", "d_blank_k", "d_blank_v"]
    for nm, a, b in zip(names, new, old):
        e = ((a.double() - b.double()).norm() / b.double().norm()).item()
        print("   %-10s new vs old %.3e  nan %d" % (nm, e, torch.isnan(a.float()).sum().item()))
    if check_ref:
        ref = _reference(args[0], args[1], H, scale, nbhd, pe_idx, mask, args[2], args[3].to(dtype), args[4].to(dtype), go)
        for nm, a, b, r in zip(names, new, old, ref):
            en = ((a.double().cpu() - r).norm() / r.norm()).item()
            eo = ((b.double().cpu() - r).norm() / r.norm()).item()
            print("   %-10s vs fp64 oracle: new %.3e  old %.3e" % (nm, en, eo))
    if time_it:
        for cl in (False, True):
            os.environ["AFFB200_FUSED_CL"] = "1" if cl else "0"
            a = [x.detach().clone().requires_grad_(True) for x in args]
            def fb():
                out = fused.fused_cluster_attention(a[0], a[1], H, scale, nbhd, pe_idx, mask, a[2], a[3], a[4])
                out.backward(go)
            print("   fwd+bwd through autograd, cl=%s: %.4f ms" % (cl, timeit(fb, 10)))


if __name__ == "__main__":
    case(56, 56, 3136, 2, 2, 32, 8, 48, 224, check_ref=True)
    case(56, 56, 627, 3, 4, 32, 8, 48, 224, check_ref=True)
    case(56, 56, 784, 2, 4, 16, 8, 48, 224, check_ref=True)
    case(56, 56, 627, 3, 4, 32, 8, 48, 224, dtype=torch.float16, check_ref=True)
    if len(sys.argv) > 1:
        case(56, 56, 3136, 256, 2, 32, 8, 48, 224, time_it=True)
        case(56, 56, 627, 256, 4, 32, 8, 48, 224, time_it=True)
