"""GPU parity of the relative-position bias / table-gather operators (csrc/attn_glue.cu) against the reference's
expressions (aff_transformer.py:131-137, 347-349) restated in oracle/aff_ops_ref.py."""
import pytest
import torch

from tests.util import assert_close

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("shape", [(3, 2, 300, 48, 111), (2, 8, 125, 48, 111), (4, 16, 25, 25, 111), (1, 4, 500, 144, 191),
                                   (2, 2, 64, 48, 1023)])
def test_attn_bias_add(shape, dtype):
    from affb200.attn_ops import attn_bias_add
    from oracle.aff_ops_ref import OracleOps
    B, H, n, m, tw = shape
    g = torch.Generator().manual_seed(B * n + m)
    P = tw * tw
    attn = torch.randn(B, H, n, m, generator=g).to(dtype)
    table = torch.randn(P, H, generator=g).to(dtype)
    centre = P // 2
    pe = (centre + torch.randint(-5 * tw - 5, 5 * tw + 5, (B, n, m), generator=g)).clamp(0, P - 1)
    mask = (torch.rand(B, n, m, generator=g) > 0.1).long() if m == 48 else None
    go = torch.randn(B, H, n, m, generator=g).to(dtype)

    a1 = attn.cuda().clone().requires_grad_(True)
    t1 = table.cuda().clone().requires_grad_(True)
    out = attn_bias_add(a1 * 1, t1, pe.cuda(), None if mask is None else mask.cuda())
    out.backward(go.cuda())
    a2 = attn.double().requires_grad_(True)
    t2 = table.double().requires_grad_(True)
    ref = OracleOps.attn_bias_add(a2, t2, pe, mask)
    ref.backward(go.double())
    assert out.dtype == dtype
    assert_close(out, ref, dtype, "bias fwd")
    assert_close(a1.grad, a2.grad, dtype, "d_attn")
    assert_close(t1.grad, t2.grad, dtype, "d_table")
    if dtype == torch.float32 and mask is None:
        # same rounding sequence as the reference's adds in the tensor dtype
        ref32 = OracleOps.attn_bias_add(attn, table, pe, mask)
        assert torch.equal(out.detach().cpu(), ref32)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_table_gather(dtype):
    from affb200.attn_ops import table_gather
    from oracle.aff_ops_ref import OracleOps
    g = torch.Generator().manual_seed(5)
    P, IC, B, n, m = 111 * 111, 4, 3, 200, 48
    table = torch.randn(P, IC, generator=g).to(dtype)
    pe = torch.randint(P // 2 - 600, P // 2 + 600, (B, n, m), generator=g)
    go = torch.randn(B, n, m, IC, generator=g).to(dtype)
    t1 = table.cuda().clone().requires_grad_(True)
    out = table_gather(t1, pe.cuda())
    out.backward(go.cuda())
    t2 = table.double().requires_grad_(True)
    ref = OracleOps.table_gather(t2, pe)
    ref.backward(go.double())
    assert torch.equal(out.detach().cpu().double(), ref.detach())
    assert_close(t1.grad, t2.grad, dtype, "d_table")
