"""GPU parity of the fused BatchNorm + GELU of the stem (csrc/batch_norm.cu) against nn.BatchNorm2d -> nn.GELU as the
reference applies them (aff_transformer.py:524-527,541), evaluated in float64 on the same inputs.
Tolerances (BASELINE.json): 1e-5 relative for fp32, 1e-2 for bf16/fp16."""
import pytest
import torch

from tests.util import assert_close

pytestmark = pytest.mark.gpu


def _case(shape, dtype, seed, offset=0.5):
    g = torch.Generator().manual_seed(seed)
    B, C, H, W = shape
    x = (torch.randn(shape, generator=g) * 1.5 + offset).to(dtype)
    go = torch.randn(shape, generator=g).to(dtype)
    w = 1 + 0.2 * torch.randn(C, generator=g)
    b = 0.2 * torch.randn(C, generator=g)
    return x, go, w, b


def _bn(C, w, b, dev, dtype=torch.float32):
    bn = torch.nn.BatchNorm2d(C).to(dev).to(dtype)
    with torch.no_grad():
        bn.weight.copy_(w)
        bn.bias.copy_(b)
        bn.running_mean.normal_(0, 0.3)
        bn.running_var.uniform_(0.5, 1.5)
    return bn


@pytest.mark.parametrize("gelu", [True, False])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("shape", [(4, 32, 28, 28), (3, 16, 17, 9), (2, 8, 5, 7), (2, 64, 12, 12), (1, 128, 6, 6),
                                   (64, 32, 56, 56), (1, 32, 1, 2)])
def test_batch_norm_act_fwd_bwd(shape, dtype, gelu):
    from affb200.norm_ops import batch_norm_act
    B, C, H, W = shape
    x, go, w, b = _case(shape, dtype, C + H)
    bn1 = _bn(C, w, b, "cuda")
    bn2 = _bn(C, w, b, "cpu", torch.float64)
    bn2.running_mean.copy_(bn1.running_mean.double().cpu())
    bn2.running_var.copy_(bn1.running_var.double().cpu())
    x1 = x.cuda().contiguous(memory_format=torch.channels_last).requires_grad_(True)
    y = batch_norm_act(x1, bn1, gelu=gelu)
    assert y.dtype == dtype and y.is_contiguous(memory_format=torch.channels_last)
    y.backward(go.cuda())
    x2 = x.double().requires_grad_(True)
    z = bn2(x2)
    ref = torch.nn.functional.gelu(z) if gelu else z
    ref.backward(go.double())
    assert_close(y, ref, dtype, "bn fwd")
    if B * H * W > 2:
        # dx is a difference of O(1) terms scaled by gamma/std; with 2-byte storage g is rounded before the reduction
        assert_close(x1.grad, x2.grad, dtype, "bn dx")
        assert_close(bn1.weight.grad, bn2.weight.grad, torch.float32 if dtype == torch.float32 else dtype, "bn dgamma")
        assert_close(bn1.bias.grad, bn2.bias.grad, torch.float32 if dtype == torch.float32 else dtype, "bn dbeta")
    assert_close(bn1.running_mean, bn2.running_mean, torch.float32, "running_mean")
    assert_close(bn1.running_var, bn2.running_var, torch.float32, "running_var")
    assert int(bn1.num_batches_tracked) == 1


def test_batch_norm_statistics_survive_a_large_mean():
    """mean >> std: the shifted sums keep the variance (a plain E[x^2] - E[x]^2 in fp32 would lose it)."""
    from affb200.norm_ops import batch_norm_act
    shape = (8, 32, 40, 40)
    x, go, w, b = _case(shape, torch.float32, 5, offset=3000.0)
    bn1 = _bn(32, w, b, "cuda")
    x1 = x.cuda().contiguous(memory_format=torch.channels_last)
    y = batch_norm_act(x1, bn1, gelu=False)
    bn2 = _bn(32, w, b, "cpu", torch.float64)
    ref = bn2(x.double())
    assert (y.cpu().double() - ref).abs().max() < 2e-3   # x itself carries 2.4e-4 of fp32 spacing at 3000


def test_batch_norm_act_matches_the_stock_bf16_chain():
    """bf16 in / bf16 out, against the stock CUDA operators run one after the other in bf16 (what autocast runs):
    same rounding points, so the results agree to an ulp of bf16 on almost every element."""
    from affb200.norm_ops import batch_norm_act
    shape = (16, 32, 112, 112)
    x, go, w, b = _case(shape, torch.bfloat16, 11)
    bn1, bn2 = _bn(32, w, b, "cuda"), _bn(32, w, b, "cuda")
    x1 = x.cuda().contiguous(memory_format=torch.channels_last).requires_grad_(True)
    x2 = x.cuda().contiguous(memory_format=torch.channels_last).requires_grad_(True)
    y1 = batch_norm_act(x1, bn1)
    y2 = torch.nn.functional.gelu(bn2(x2))
    y1.backward(go.cuda().contiguous(memory_format=torch.channels_last))
    y2.backward(go.cuda().contiguous(memory_format=torch.channels_last))
    d = (y1.detach().float() - y2.detach().float()).abs()
    assert float((d > 0).float().mean()) < 0.01 and float(d.max()) <= 0.03125
    dg = (x1.grad.float() - x2.grad.float()).abs()
    assert float(dg.max()) <= 0.02 * float(x2.grad.float().abs().max())
    assert_close(bn1.weight.grad, bn2.weight.grad, torch.bfloat16, "dgamma")


def test_batch_norm_act_eval_mode_uses_running_statistics():
    from affb200.norm_ops import batch_norm_act
    x, go, w, b = _case((2, 32, 8, 8), torch.float32, 2)
    bn = _bn(32, w, b, "cuda").eval()
    x1 = x.cuda().contiguous(memory_format=torch.channels_last)
    assert torch.equal(batch_norm_act(x1, bn), torch.nn.functional.gelu(bn(x1)))
    assert int(bn.num_batches_tracked) == 0


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_conv_batch_norm_act_folds_the_conv_bias(dtype):
    """conv -> BatchNorm(train) -> GELU with the conv bias dropped from the activation and put back into running_mean:
    output, running statistics and gradients of the stock chain; the conv bias gets an exactly-zero gradient (the stock
    chain's is rounding noise around zero); evaluation mode keeps the bias in the activation."""
    from affb200.norm_ops import conv_batch_norm_act
    g = torch.Generator().manual_seed(11)
    x = torch.randn(8, 3, 32, 32, generator=g).cuda().contiguous(memory_format=torch.channels_last)
    go = torch.randn(8, 32, 16, 16, generator=g).cuda().contiguous(memory_format=torch.channels_last)
    mods = []
    for _ in range(2):
        torch.manual_seed(5)
        conv = torch.nn.Conv2d(3, 32, 3, 2, 1).cuda()
        with torch.no_grad():
            conv.bias.normal_(0, 0.5)
        _, _, w, b = _case((1, 32, 1, 1), torch.float32, 6)
        mods.append((conv, _bn(32, w, b, "cuda")))
    ac = dict(device_type="cuda", dtype=torch.bfloat16, enabled=dtype == torch.bfloat16)
    (c1, b1), (c2, b2) = mods
    x1, x2 = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    with torch.autocast(**ac):
        y1 = conv_batch_norm_act(x1, c1, b1)
        y2 = torch.nn.functional.gelu(b2(c2(x2)))
    assert y1.dtype == y2.dtype
    y1.backward(go.to(y1.dtype))
    y2.backward(go.to(y2.dtype))
    assert_close(y1, y2.double(), dtype, "fwd")
    assert_close(x1.grad, x2.grad.double(), dtype, "dx")
    assert_close(c1.weight.grad, c2.weight.grad.double(), dtype, "d conv.weight")
    assert_close(b1.weight.grad, b2.weight.grad.double(), dtype, "d gamma")
    assert_close(b1.bias.grad, b2.bias.grad.double(), dtype, "d beta")
    assert c1.bias.grad is not None and c1.bias.grad.abs().max().item() == 0
    assert c2.bias.grad.abs().max().item() < (1e-3 if dtype == torch.float32 else 0.5)
    assert_close(b1.running_mean, b2.running_mean.double(), dtype, "running_mean")
    assert_close(b1.running_var, b2.running_var.double(), dtype, "running_var")
    assert int(b1.num_batches_tracked) == 1
    # evaluation mode: bias, running statistics, affine and GELU as one pass over the bias-free convolution output
    b1.eval()
    with torch.no_grad():
        want = torch.nn.functional.gelu(b1(c1(x))).double()
        assert_close(conv_batch_norm_act(x, c1, b1), want, torch.float32, "eval fp32")
        xl = x.to(dtype)
        with torch.autocast(**ac):
            got = conv_batch_norm_act(xl, c1, b1)
        assert got.dtype == dtype
        assert_close(got, want, dtype, "eval")
    # with gradients wanted in evaluation mode the stock operators run
    xg = x.clone().requires_grad_(True)
    y = conv_batch_norm_act(xg, c1, b1)
    y.sum().backward()
    assert xg.grad is not None and torch.equal(y, torch.nn.functional.gelu(b1(c1(x))))


@pytest.mark.parametrize("sink", [False, True])
@pytest.mark.parametrize("autocast", [False, True])
def test_conv_bias_matches_the_stock_convolution(autocast, sink):
    """proj2 of the stem: conv without bias + the row bias kernel; output bit-identical to the stock conv with bias
    in bf16 (same rounding: conv result rounded, bias rounded, sum rounded), gradients of x / weight identical, bias
    gradient = fp32 column sums (stock: a bf16 reduction)."""
    from affb200.norm_ops import conv_bias
    g = torch.Generator().manual_seed(4)
    x = torch.randn(4, 32, 28, 28, generator=g).cuda().contiguous(memory_format=torch.channels_last)
    go = torch.randn(4, 64, 14, 14, generator=g).cuda().contiguous(memory_format=torch.channels_last)
    res = []
    for ours in (True, False):
        torch.manual_seed(9)
        conv = torch.nn.Conv2d(32, 64, 3, 2, 1).cuda()
        if sink and ours:
            conv.bias.grad = torch.zeros_like(conv.bias)
            conv.bias._affb200_sink = True
        xx = x.clone().requires_grad_(True)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
            y = conv_bias(xx, conv) if ours else conv(xx)
        y.backward(go.to(y.dtype))
        res.append((y, xx.grad, conv.weight.grad, conv.bias.grad))
    dt = torch.bfloat16 if autocast else torch.float32
    (y1, dx1, dw1, db1), (y2, dx2, dw2, db2) = res
    assert y1.dtype == y2.dtype and y1.is_contiguous(memory_format=torch.channels_last)
    assert_close(y1, y2.double(), dt, "conv_bias y")
    assert torch.equal(dx1, dx2) and torch.equal(dw1, dw2)
    assert_close(db1, go.double().sum((0, 2, 3)) if not autocast else go.bfloat16().double().sum((0, 2, 3)), torch.float32,
                 "conv_bias d_bias")
    assert_close(db1, db2.double(), dt, "conv_bias d_bias vs stock")


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("shape,cout", [((3, 3, 224, 224), 32), ((2, 3, 160, 160), 32), ((2, 3, 40, 24), 16),
                                        ((1, 3, 64, 64), 64), ((2, 3, 30, 56), 48)])
def test_stem_convolution_kernel_matches_the_library(shape, cout, dtype):
    """proj1 (3 -> cout, 3x3, stride 2, padding 1, no bias) on channels-last images: the kernel against the stock
    convolution evaluated in float64 on the same rounded operands (1e-2 for 2-byte outputs; the kernel accumulates in
    fp32 and rounds once), including the zero padding at the top / left edges and a row count that is not a multiple of
    the four rows a CTA owns; the weight gradient comes from the library."""
    from affb200.norm_ops import ConvStemFunction, _stem_conv_ok
    g = torch.Generator().manual_seed(sum(shape) + cout)
    x = torch.randn(shape, generator=g).to(dtype).cuda().contiguous(memory_format=torch.channels_last)
    conv = torch.nn.Conv2d(3, cout, 3, 2, 1).cuda()
    assert _stem_conv_ok(x, conv)
    y = ConvStemFunction.apply(x, conv.weight)
    assert y.shape == (shape[0], cout, shape[2] // 2, shape[3] // 2) and y.dtype == dtype
    assert y.is_contiguous(memory_format=torch.channels_last)
    want = torch.nn.functional.conv2d(x.double(), conv.weight.detach().to(dtype).double(), None, 2, 1)
    assert_close(y, want, dtype, "stem conv")
    go = torch.randn(y.shape, generator=g).to(dtype).cuda().contiguous(memory_format=torch.channels_last)
    y.backward(go)
    wantw = torch.nn.grad.conv2d_weight(x.double(), conv.weight.shape, go.double(), 2, 1)
    assert_close(conv.weight.grad, wantw, dtype, "stem conv d_weight")
    assert not _stem_conv_ok(x.float(), conv) and not _stem_conv_ok(x.clone().requires_grad_(True), conv)
