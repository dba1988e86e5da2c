"""GPU parity of the fused residual + stochastic-depth pass (csrc/residual.cu) with the reference expression
`shortcut + drop_path(x)` (aff_transformer.py:224-229, timm DropPath) under the same random draw."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("shape", [(16, 300, 64), (7, 125, 256), (3, 8), (5, 33, 24)])
def test_residual_add_matches_reference_expression(shape, dtype):
    from affb200.residual_ops import residual_add
    from oracle.aff_ops_ref import OracleOps
    g = torch.Generator().manual_seed(sum(shape))
    feat = torch.randn(shape, generator=g).to(dtype).cuda()
    x = torch.randn(shape, generator=g).to(dtype).cuda()
    go = torch.randn(shape, generator=g).to(dtype).cuda()
    res = []
    for fn in (residual_add, OracleOps.residual_add):
        f, xx = feat.clone().requires_grad_(True), x.clone().requires_grad_(True)
        torch.manual_seed(11)           # same bernoulli draw: both draw B values from the device generator
        out = fn(f, xx, 0.3, True)
        out.backward(go)
        res.append((out, f.grad, xx.grad))
    for a, b, what in zip(res[0], res[1], ("out", "d_feat", "d_x")):
        assert a.dtype == b.dtype
        assert torch.equal(a, b), what   # same roundings as the two-operation reference: bit-exact
    # some samples dropped, some kept and scaled
    s = (res[0][2] / go.float().clamp_min(1e-3).to(dtype)).float().flatten(1).median(1).values
    assert (s == 0).any() or shape[0] < 5


def test_residual_add_eval_and_p0_are_plain_adds():
    from affb200.residual_ops import residual_add
    a, b = torch.randn(4, 10, 16, device="cuda"), torch.randn(4, 10, 16, device="cuda")
    assert torch.equal(residual_add(a, b, 0.5, False), a + b)
    assert torch.equal(residual_add(a, b, 0.0, True), a + b)
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        residual_add(a.cpu(), b.cpu(), 0.1, True)
