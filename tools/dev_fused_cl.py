"""Development check of the block-structured fused attention forward against the row-union kernels and the fp64
oracle chain; prints timings of both at the AFF-Tiny-1/5 stage shapes.  Run on the GPU box."""
import ctypes
import sys
import os
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ml-autofocusformermod_b200"))
import torch
from affb200 import _native as N
from affb200 import fused, fused_cl
from affb200.fused import _view
from affb200.point_utils import space_filling_cluster, build_nbhd


def positions(h, w, n, B, seed):
    g = torch.Generator().manual_seed(seed)
    ys, xs = torch.meshgrid(torch.arange(h), torch.arange(w), indexing="ij")
    allp = torch.stack([xs, ys], 2).reshape(-1, 2).float()
    if n == h * w:
        return allp.unsqueeze(0).expand(B, -1, -1).contiguous()
    return torch.stack([allp[torch.randperm(h * w, generator=g)[:n]] for _ in range(B)])


def cl_fwd(q_lin, kv_lin, H, scale, pl, table, bk, bv, w, tw):
    B, n, hc = q_lin.shape
    c = hc // H
    dev = q_lin.device
    out = torch.empty_like(q_lin)
    lse = torch.empty((B, H, n), dtype=torch.float32, device=dev)
    guard = torch.zeros(1, dtype=torch.int32, device=dev)

    def run():
        N.check(N.LIB.affb200_fused_cl_fwd(
            _view(q_lin, n * hc, hc, c), _view(kv_lin, n * 2 * hc, 2 * hc, 2 * c), _view(kv_lin, n * 2 * hc, 2 * hc, 2 * c, c),
            _view(out, n * hc, hc, c), N.p(table), N.p(bk), N.p(bv), pl.byref(), N.p(lse), N.p(guard), B, H, n, c, w, tw,
            float(scale), N.dtype_enum(q_lin), N.stream_ptr(dev)), "fused_cl_fwd")
    run()
    return out, lse, guard, run


def timeit(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def case(h, w, n, B, H, c, m, nbhd_size, img, dtype=torch.bfloat16, check_ref=False, time_it=False):
    rpw = img // 4 - 1
    tw = 2 * rpw + 1
    P = tw * tw
    pos = positions(h, w, n, B, seed=n).cuda()
    p_s, mean, _, _, _ = space_filling_cluster(pos, m, h, w)
    nbhd, mask, pe_idx, _ = build_nbhd(p_s, mean, m, nbhd_size, rpw)
    fused_cl.hint_geometry(nbhd, p_s, rpw, tw)
    pl = fused_cl.cl_plan(nbhd, mask, pe_idx, P)
    print("case", (h, w, n, B, H, c, m, nbhd_size), "plan:", None if pl is None else (pl.ucap, pl.kbcap, pl.stats.tolist()))
    if pl is None:
        return
    g = torch.Generator().manual_seed(7 + n)
    hc = H * c
    q_lin = torch.randn(B, n, hc, generator=g).to(dtype).cuda()
    kv_lin = torch.randn(B, n, 2 * hc, generator=g).to(dtype).cuda()
    table = (0.5 * torch.randn(P, H, generator=g)).to(dtype).cuda()
    bk = torch.randn(hc, generator=g).cuda()
    bv = torch.randn(hc, generator=g).cuda()
    scale = c ** -0.5
    tab32 = table.float().contiguous()
    out, lse, guard, run = cl_fwd(q_lin, kv_lin, H, scale, pl, tab32, bk, bv, rpw, tw)
    torch.cuda.synchronize()
    os.environ["AFFB200_FUSED_IMG"] = "0"
    old = fused.fused_cluster_attention(q_lin, kv_lin, H, scale, nbhd, pe_idx, mask, table, bk, bv)
    err = (out.float() - old.float()).norm() / old.float().norm()
    print("   guard", guard.item(), "rel err vs row-union kernel %.3e" % err.item(), "nan", torch.isnan(out.float()).sum().item())
    if check_ref:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from tests.test_fused_attn import _reference
        go = torch.zeros_like(q_lin)
        ref = _reference(q_lin, kv_lin, H, scale, nbhd, pe_idx, mask, table, bk.to(dtype), bv.to(dtype), go)[0]
        e = (out.double().cpu() - ref).norm() / ref.norm()
        eo = (old.double().cpu() - ref).norm() / ref.norm()
        print("   rel err vs fp64 oracle: new %.3e  old %.3e" % (e.item(), eo.item()))
    if time_it:
        raw = fused.raw_calls(q_lin, kv_lin, H, scale, nbhd, pe_idx, mask, table, bk, bv)
        t_old = timeit(raw[0])
        t_new = timeit(run)
        print("   forward ms: row-union %.4f   block-structured %.4f" % (t_old, t_new))


if __name__ == "__main__":
    case(56, 56, 3136, 2, 2, 32, 8, 48, 224, check_ref=True)
    case(56, 56, 627, 3, 4, 32, 8, 48, 224, check_ref=True)
    case(56, 56, 784, 2, 4, 16, 8, 48, 224, check_ref=True)
    case(56, 56, 627, 3, 4, 32, 8, 48, 224, dtype=torch.float16, check_ref=True)
    if len(sys.argv) > 1:
        case(56, 56, 3136, 256, 2, 32, 8, 48, 224, time_it=True)
        case(56, 56, 627, 256, 4, 32, 8, 48, 224, time_it=True)
