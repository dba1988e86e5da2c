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
