"""The MixResNeighbour stage glue (affb200/mixres_neighbour.py) against the golden produced by the REFERENCE's own
`BasicLayer` / `divide_tokens_to_split_and_keep` (tests/golden/mixres_stage.pt, generator make_mixres_golden.py):
  * CPU: glue + the oracle's operator set reproduces indices, positions, features and gradients;
  * GPU: glue + the CUDA operator set does (integer tensors bit-exact; fp32 1e-5 on outputs, 1e-4 on gradients),
    with the reference's full-table indexing, with the compacted table, and through the fused attention operator;
    the token split order is bit-exact."""
import os

import pytest
import torch

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "mixres_stage.pt")
CASES = ["ragged", "grid", "global"]


def rel(a, b):
    return ((a.float().cpu() - b).norm() / b.norm()).item()


def _counting(fn, hits):
    """wrap the fused operator: count the calls it accepted (it returns None when it declines a shape)"""
    def f(*a, **kw):
        out = fn(*a, **kw)
        hits.append(out is not None)
        return out
    return f


def _run(layer, case, dev, capture, autocast=False):
    def pre(mod, args):
        if not capture:
            capture.update(member_idx=args[1], cluster_mask=args[2], pe_idx=args[3], global_attn=bool(args[4]))
    hk = layer.blocks[0].register_forward_pre_hook(pre)
    for p in layer.parameters():
        p.grad = None
    feat = case["feat"].to(dev).clone().requires_grad_(True)
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
        pos, out = layer(case["pos"].to(dev).clone(), feat, case["h"], case["w"], case["on_grid"])
    hk.remove()
    w = torch.linspace(-1, 1, out.numel(), device=dev).reshape(out.shape)
    (out.float() * w).sum().backward()
    grads = {k: p.grad.detach() for k, p in layer.named_parameters()}
    return pos.detach(), out.detach(), feat.grad.detach(), grads


def _check(g, name, layer, dev, index_exact, tol_out, tol_grad, autocast=False):
    c = g["cases"][name]
    cap = {}
    pos, out, dx, grads = _run(layer, c["case"], dev, cap, autocast)
    assert cap["global_attn"] == c["global_attn"]
    if index_exact:
        for k, v in c["index"].items():
            assert (v is None) == (cap[k] is None), (name, k)
            if v is not None:
                assert cap[k].dtype == torch.int64 and torch.equal(cap[k].cpu().to(torch.int32), v), (name, k)
    assert torch.equal(pos.cpu(), c["pos"]), name
    assert rel(out, c["out"]) < tol_out, (name, rel(out, c["out"]))
    assert rel(dx, c["dx"]) < tol_grad, (name, rel(dx, c["dx"]))
    for k, v in c["grads"].items():
        assert rel(grads[k], v) < tol_grad, (name, k, rel(grads[k], v))


@pytest.mark.parametrize("name", CASES)
def test_glue_with_oracle_ops_reproduces_reference_stage(name):
    from affb200.mixres_neighbour import BasicLayer
    from oracle.aff_ops_ref import OracleOps, fill_parameters_by_name
    g = torch.load(GOLD)
    layer = BasicLayer(ops=OracleOps(), **g["layer"])
    fill_parameters_by_name(layer, seed=g["seed"])
    layer.train()
    layer.compact = False
    _check(g, name, layer, "cpu", True, 1e-6, 1e-5)
    layer.compact = True
    _check(g, name, layer, "cpu", False, 1e-6, 1e-5)


def test_split_with_stable_sort_reproduces_reference():
    from affb200.mixres_neighbour import divide_tokens_to_split_and_keep
    s = torch.load(GOLD)["split"]
    res = divide_tokens_to_split_and_keep(s["feat"], s["pos"], s["scores"], s["upscale_ratio"],
                                          argsort=lambda x: x.sort(dim=1, stable=True)[1])
    for a, b in zip(res, s["result"]):
        assert torch.equal(a, b)


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["full_table", "compact", "fused"])
@pytest.mark.parametrize("name", CASES)
def test_cuda_stage_matches_reference_golden(name, mode):
    from affb200.aff_transformer import CudaOps
    from affb200.mixres_neighbour import BasicLayer
    from oracle.aff_ops_ref import fill_parameters_by_name
    torch.backends.cuda.matmul.allow_tf32 = False
    g = torch.load(GOLD)
    ops = CudaOps()
    hits = []
    if mode == "fused":
        from affb200 import fused as fa
        ops.fused_attention = _counting(fa.fused_cluster_attention, hits)
    else:
        ops.fused_attention = None
    layer = BasicLayer(ops=ops, **g["layer"])
    fill_parameters_by_name(layer, seed=g["seed"])
    layer.cuda().train()
    layer.compact = mode != "full_table"
    if mode == "fused":   # the fused operator computes in bf16/f16 only: autocast, bf16 tolerance (1e-2 per operator)
        _check(g, name, layer, "cuda", False, 2e-2, 6e-2, autocast=True)
    else:
        _check(g, name, layer, "cuda", mode == "full_table", 1e-5, 1e-4)
    if mode == "fused" and name != "global":
        assert hits and all(hits), "the fused operator declined the compacted-table stage"


@pytest.mark.gpu
def test_cuda_split_is_bit_exact():
    from affb200.mixres_neighbour import divide_tokens_to_split_and_keep
    s = torch.load(GOLD)["split"]
    res = divide_tokens_to_split_and_keep(s["feat"].cuda(), s["pos"].cuda(), s["scores"].cuda(), s["upscale_ratio"])
    for a, b in zip(res, s["result"]):
        assert torch.equal(a.cpu(), b)


@pytest.mark.gpu
def test_cuda_stage_config5_shape_bf16_runs_fused():
    """maskfiner_up_down_small stage j=3 shape (N=30720, C=64, 2 heads, canvas 256 x 512) at depth 1,
    bf16 autocast: the compacted table fits the fused operator (<= 32768 rows) and the result is finite and close to
    the unfused operators on the same weights."""
    from affb200.aff_transformer import CudaOps
    from affb200.mixres_neighbour import BasicLayer, compact_table
    from affb200 import fused as fa
    from oracle.aff_ops_ref import fill_parameters_by_name
    gen = torch.Generator().manual_seed(3)
    h, w, n = 256, 512, 30720
    cell = torch.randperm(h * w, generator=gen)[:n]
    pos = torch.stack([torch.zeros(n), (cell % w).float(), (cell // w).float()], 1).unsqueeze(0).cuda()
    feat = torch.randn(1, n, 64, generator=gen).cuda()
    outs, hits = [], []
    for fused in (False, True):
        ops = CudaOps()
        ops.fused_attention = _counting(fa.fused_cluster_attention, hits) if fused else None
        layer = BasicLayer(dim=64, cluster_size=8, nbhd_size=48, depth=1, num_heads=2, mlp_ratio=2., ops=ops)
        fill_parameters_by_name(layer, seed=1)
        layer.cuda().eval()
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
            p, o = layer(pos, feat, h, w, False)
        outs.append((p, o.float()))
    assert hits == [True]
    assert torch.equal(outs[0][0], outs[1][0])
    assert torch.isfinite(outs[1][1]).all()
    assert rel(outs[1][1], outs[0][1].cpu()) < 2e-2
    # under no_grad the stage kept the full 1023 x 1023 table and the block-structured forward evaluated pos_embed on
    # the central rows its offsets reach; with gradients enabled the stage compacts the table instead (row-union
    # kernels): the same numbers up to the bf16 rounding of two different kernels
    from affb200 import fused_cl
    obs = fused_cl.observed((1, n, 48))
    assert obs is not None and obs["ok"] and 0 < obs["maxoff"] <= 255, obs
    with torch.autocast("cuda", dtype=torch.bfloat16):
        p2, o2 = layer(pos, feat, h, w, False)
    assert torch.equal(p2, outs[1][0])
    assert rel(o2.float().detach(), outs[1][1].cpu()) < 1e-2, rel(o2.float().detach(), outs[1][1].cpu())


def _mixed_scale_positions(seed, h=256, w=512):
    """token positions of a 4-scale split as the U-shaped model produces them (coarse tokens on an 8-unit grid, half of
    each scale split into four children, parents kept): 2048 + 4096 + 8192 + 16384 = 30720 tokens"""
    g = torch.Generator().manual_seed(seed)
    ys, xs = torch.meshgrid(torch.arange(0, h, 8), torch.arange(0, w, 8), indexing="ij")
    cur = torch.stack([xs.reshape(-1), ys.reshape(-1)], 1)
    allp, alls = [], []
    for s in range(4):
        allp.append(cur)
        alls.append(torch.full((cur.shape[0],), s))
        if s == 3:
            break
        pick = torch.randn(cur.shape[0], generator=g).argsort()[cur.shape[0] // 2:]
        par, half = cur[pick], 2 ** (2 - s)
        cur = torch.stack([par, par + torch.tensor([half, 0]), par + torch.tensor([0, half]), par + torch.tensor([half, half])],
                          1).reshape(-1, 2)
    xy, sc = torch.cat(allp).float(), torch.cat(alls).float()
    return torch.cat([sc[:, None], xy], 1).unsqueeze(0)


@pytest.mark.gpu
def test_cuda_stage_split_tiles_match_unfused():
    """Mixed-resolution token sets put more than 64 key blocks behind some tiles of 64 tokens; those tiles carry one
    union per half (AFFB200_CL_SPLIT) and the block-structured forward still serves the stage.  Checked against the
    unfused operators on the same weights, on a draw that has split tiles."""
    from affb200.aff_transformer import CudaOps
    from affb200.mixres_neighbour import BasicLayer
    from affb200 import fused as fa, fused_cl
    from oracle.aff_ops_ref import fill_parameters_by_name
    seen_split = 0
    for seed in range(8):
        pos = _mixed_scale_positions(seed).cuda()
        n = pos.shape[1]
        feat = torch.randn(1, n, 64, generator=torch.Generator().manual_seed(seed)).cuda()
        outs, hits = [], []
        for fused in (False, True):
            ops = CudaOps()
            ops.fused_attention = _counting(fa.fused_cluster_attention, hits) if fused else None
            layer = BasicLayer(dim=64, cluster_size=8, nbhd_size=48, depth=1, num_heads=2, mlp_ratio=2., ops=ops)
            fill_parameters_by_name(layer, seed=1)
            layer.cuda().eval()
            with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
                p, o = layer(pos, feat, 256, 512, False)
            outs.append((p, o.float()))
        st = fused_cl.last_eager_stats
        assert hits == [True], (seed, st)
        assert st[0] == 0 and st[1] == 0 and st[4] <= 64, st
        assert torch.equal(outs[0][0], outs[1][0])
        err = (outs[1][1] - outs[0][1]).norm(dim=2) / outs[0][1].norm(dim=2).clamp_min(1e-6)      # per token
        assert rel(outs[1][1], outs[0][1].cpu()) < 2e-2 and err.max().item() < 0.06, (seed, st, err.max().item())
        seen_split += int(st[3] > 0)
        if seen_split >= 2:
            break
    assert seen_split >= 1, "no draw produced a tile with more than 64 blocks"
