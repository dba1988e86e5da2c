"""GPU parity of the fused merge operator (csrc/merge.cu) with the reference's expressions
(aff_transformer.py:343-361: table gather, lp gather, mask, CLUSTENWF) evaluated in float64 through the oracle."""
import pytest
import torch

from tests.util import assert_close

pytestmark = pytest.mark.gpu


def _reference(table, pe_idx, lp, mask, feat, nbhd):
    from oracle.aff_ops_ref import OracleOps
    b, n_new, m = nbhd.shape
    w = OracleOps.table_gather(table, pe_idx)
    if lp is not None:
        l = lp.gather(1, nbhd.reshape(b, -1, 1)).reshape(b, n_new, m, 1)
        if mask is not None:
            l = l * mask.unsqueeze(3)
        w = w * l
    elif mask is not None:
        w = w * mask.unsqueeze(3)
    return OracleOps.wf(w, feat, nbhd).reshape(b, n_new, -1)


@pytest.mark.parametrize("fdt", [torch.bfloat16, torch.float32, torch.float16])
@pytest.mark.parametrize("case", [(3, 157, 627, 48, 64, True, True), (2, 31, 125, 48, 128, True, True),
                                  (2, 6, 25, 25, 256, False, True), (1, 40, 200, 7, 32, True, False),
                                  (2, 33, 64, 49, 512, False, True), (1, 3, 10, 144, 128, True, True)])
def test_merge_wf_matches_reference(case, fdt):
    from affb200.merge_ops import merge_wf
    B, Nq, Nk, M, C, use_mask, use_lp = case
    g = torch.Generator().manual_seed(B * Nq + M + C)
    P = 111 * 111
    table = torch.randn(P, 4, generator=g)
    pe_idx = torch.randint(0, P, (B, Nq, M), generator=g)
    nbhd = torch.randint(0, Nk, (B, Nq, M), generator=g)
    nbhd[:, :, 1] = nbhd[:, :, 0]                      # duplicates inside a neighbourhood are legal
    mask = (torch.rand(B, Nq, M, generator=g) > 0.2).long() if use_mask else None
    lp = torch.rand(B, Nk, 1, generator=g).to(fdt) if use_lp else None
    feat = torch.randn(B, Nk, C, generator=g).to(fdt)
    go = torch.randn(B, Nq, 4 * C, generator=g)

    t1 = table.cuda().requires_grad_(True)
    f1 = feat.cuda().requires_grad_(True)
    l1 = lp.cuda().requires_grad_(True) if use_lp else None
    out = merge_wf(t1, pe_idx.cuda(), l1, None if mask is None else mask.cuda(), f1, nbhd.cuda())
    assert out is not None and out.dtype == torch.float32
    out.backward(go.cuda())

    t2 = table.double().requires_grad_(True)
    f2 = feat.double().requires_grad_(True)
    l2 = lp.double().requires_grad_(True) if use_lp else None
    ref = _reference(t2, pe_idx, l2, mask, f2, nbhd)
    ref.backward(go.double())

    assert_close(out, ref, torch.float32, "merge fwd")
    assert_close(t1.grad, t2.grad, torch.float32, "d_table")
    assert f1.grad.dtype == fdt
    assert_close(f1.grad, f2.grad, fdt, "d_feat")
    if use_lp:
        assert l1.grad.shape == lp.shape and l1.grad.dtype == fdt
        assert_close(l1.grad, l2.grad, fdt, "d_lp")


def test_merge_wf_declines_unsupported_cases():
    from affb200.merge_ops import merge_wf
    t = torch.randn(50, 4, device="cuda")
    pe = torch.zeros(1, 2, 3, dtype=torch.long, device="cuda")
    nb = torch.zeros(1, 2, 3, dtype=torch.long, device="cuda")
    assert merge_wf(t, pe, None, None, torch.randn(1, 5, 48, device="cuda"), nb) is None          # C = 48
    assert merge_wf(t.half(), pe, None, None, torch.randn(1, 5, 64, device="cuda"), nb) is None   # table not fp32
    assert merge_wf(torch.randn(50, 3, device="cuda"), pe, None, None, torch.randn(1, 5, 64, device="cuda"), nb) is None
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        merge_wf(t, pe, None, None, torch.randn(1, 5, 64), nb)


@pytest.mark.parametrize("autocast", [False, True])
@pytest.mark.parametrize("P", [12321, 633, 1])
def test_weight_net_table_matches_the_module(P, autocast):
    """weight_net = Sequential(Linear(5, 4), LayerNorm(4), GELU) on the relative-position feature table (reference
    aff_transformer.py:262-264,343) as one kernel per direction: output and the four parameter gradients of the stock
    modules, fp32 at 1e-5, under bf16 autocast (Linear in bf16, LayerNorm / GELU in fp32) at 1e-2."""
    from affb200.merge_ops import weight_net_table
    from affb200.aff_transformer import build_pe_lookup
    g = torch.Generator().manual_seed(P)
    _, _, table = build_pe_lookup(224)
    table = table[torch.randperm(table.shape[0], generator=g)[:P]].cuda()
    go = torch.randn(P, 4, generator=g).cuda()
    res = []
    for ours in (True, False):
        torch.manual_seed(7)
        net = torch.nn.Sequential(torch.nn.Linear(5, 4), torch.nn.LayerNorm(4), torch.nn.GELU()).cuda()
        with torch.no_grad():
            net[1].weight.normal_(1, 0.2)
            net[1].bias.normal_(0, 0.2)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
            out = weight_net_table(net, table) if ours else net(table)
        assert out.dtype == torch.float32 and out.shape == (P, 4)
        out.backward(go)
        res.append([out.detach()] + [p.grad for p in net.parameters()])
    tol = 1e-2 if autocast else 1e-5
    for a, b, what in zip(res[0], res[1], ("out", "d_w", "d_b", "d_gamma", "d_beta")):
        assert a.shape == b.shape and a.dtype == b.dtype, what
        err = float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-12))
        assert err < tol, (what, err)
