"""GPU parity of the lookup GELU (csrc/gelu.cu) against nn.GELU as the reference's Mlp applies it
(aff_transformer.py:35-51): bit-exact forward over every 2-byte input, backward against the fp64 derivative and the
stock bf16 operator, and the fused fc1 + GELU Function with its bias-gradient column sums."""
import pytest
import torch

from tests.util import assert_close

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_gelu_forward_is_bit_exact_over_all_inputs(dtype):
    from affb200.linear_ops import gelu
    codes = torch.arange(65536, dtype=torch.int32).to(torch.int16).cuda()
    x = codes.view(dtype).reshape(64, 1024)
    y = gelu(x)
    ref = torch.nn.functional.gelu(x)
    assert y.dtype == dtype
    assert torch.equal(y.view(torch.int16)[~ref.isnan()], ref.view(torch.int16)[~ref.isnan()])
    assert torch.equal(y.isnan(), ref.isnan())


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("shape", [(3, 125, 512), (2, 3136, 128), (1, 7, 8), (5, 24)])
def test_gelu_backward(shape, dtype):
    from affb200.linear_ops import gelu
    g = torch.Generator().manual_seed(shape[-1])
    x = (2.5 * torch.randn(shape, generator=g)).to(dtype)
    go = torch.randn(shape, generator=g).to(dtype)
    x1 = x.cuda().requires_grad_(True)
    gelu(x1).backward(go.cuda())
    x2 = x.double().requires_grad_(True)
    torch.nn.functional.gelu(x2).backward(go.double())
    assert_close(x1.grad, x2.grad, dtype, "gelu dx")
    x3 = x.cuda().requires_grad_(True)
    torch.nn.functional.gelu(x3).backward(go.cuda())
    # against the stock operator in the same dtype: the tables hold the same fp32 derivative (up to 1 - G for x < 0)
    diff = (x1.grad.float() - x3.grad.float()).abs()
    assert float((diff > 0).float().mean()) < 0.1
    assert float(diff.max()) <= 2.0 ** -7 * float(x3.grad.float().abs().max())


def test_gelu_every_derivative_entry():
    from affb200.linear_ops import gelu
    codes = torch.arange(65536, dtype=torch.int32).to(torch.int16).cuda()
    x = codes.view(torch.bfloat16).reshape(64, 1024)
    ok = x.isfinite()
    x1 = torch.where(ok, x, torch.zeros_like(x)).requires_grad_(True)
    gelu(x1).backward(torch.ones_like(x1))
    x2 = x1.detach().double().requires_grad_(True)
    torch.nn.functional.gelu(x2).backward(torch.ones_like(x2))
    err = (x1.grad.double() - x2.grad).abs()
    assert float(err.max()) <= 2.0 ** -8        # one bf16 rounding of a derivative in [-0.13, 1.13]


@pytest.mark.parametrize("sink", [False, True])
@pytest.mark.parametrize("dims", [(2, 300, 64, 128), (3, 125, 256, 512), (1, 25, 384, 768), (2, 10, 32, 24)])
def test_linear_gelu_matches_reference(dims, sink):
    from affb200.linear_ops import linear_gelu
    from oracle.aff_ops_ref import OracleOps
    B, n, cin, cout = dims
    torch.manual_seed(cin + cout)
    lin = torch.nn.Linear(cin, cout).cuda()
    x = torch.randn(B, n, cin, device="cuda")
    go = torch.randn(B, n, cout, device="cuda").bfloat16()
    if sink:
        for p in lin.parameters():
            p.grad = torch.full_like(p, 0.5)
            p._affb200_sink = True
    x1 = x.clone().requires_grad_(True)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        y = linear_gelu(x1, lin.weight, lin.bias)
    assert y.dtype == torch.bfloat16
    y.backward(go)
    # reference on the bf16-rounded operands, fp64 arithmetic
    x2 = x.bfloat16().double().cpu().requires_grad_(True)
    w2 = lin.weight.detach().bfloat16().double().cpu().requires_grad_(True)
    b2 = lin.bias.detach().bfloat16().double().cpu().requires_grad_(True)
    ref = OracleOps.linear_gelu(x2, w2, b2)
    ref.backward(go.double().cpu())
    off = 0.5 if sink else 0.0
    assert_close(y, ref, torch.bfloat16, "linear_gelu fwd")
    assert_close(x1.grad, x2.grad, torch.bfloat16, "linear_gelu dx")
    assert_close(lin.weight.grad - off, w2.grad, torch.bfloat16, "linear_gelu dW")
    assert_close(lin.bias.grad - off, b2.grad, torch.bfloat16, "linear_gelu db")


def test_linear_gelu_fp32_uses_the_stock_operators():
    from affb200.linear_ops import linear_gelu
    torch.manual_seed(0)
    lin = torch.nn.Linear(64, 128).cuda()
    x = torch.randn(4, 50, 64, device="cuda")
    assert torch.allclose(linear_gelu(x, lin.weight, lin.bias), torch.nn.functional.gelu(lin(x)), atol=1e-6)
