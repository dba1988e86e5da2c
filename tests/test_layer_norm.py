"""GPU parity of the fused LayerNorm (csrc/layer_norm.cu) against nn.LayerNorm as the reference applies it
(aff_transformer.py:199-200,284), restated in oracle/aff_ops_ref.py and evaluated in float64 on the same inputs.
Tolerances (BASELINE.json): 1e-5 relative for fp32, 1e-2 for bf16/fp16."""
import pytest
import torch

from tests.util import assert_close

pytestmark = pytest.mark.gpu

SHAPES = [((4, 300, 64), None), ((2, 627, 128), None), ((3, 125, 256), None), ((2, 25, 512), None), ((2, 25, 1024), None),
          ((5, 37, 96), None), ((1, 1, 320), None), ((7, 64), None), ((2, 2000, 64), None)]


@pytest.mark.parametrize("io", [(torch.float32, torch.float32), (torch.bfloat16, torch.bfloat16),
                                (torch.float16, torch.float16), (torch.bfloat16, torch.float32),
                                (torch.float32, torch.bfloat16)])
@pytest.mark.parametrize("shape", [s for s, _ in SHAPES])
def test_layer_norm_fwd_bwd(shape, io):
    from affb200.norm_ops import layer_norm
    from oracle.aff_ops_ref import OracleOps
    din, dout = io
    C = shape[-1]
    g = torch.Generator().manual_seed(C + len(shape))
    x = (torch.randn(shape, generator=g) * 2 + 0.5).to(din)
    w = 1 + 0.1 * torch.randn(C, generator=g)
    b = 0.1 * torch.randn(C, generator=g)
    go = torch.randn(shape, generator=g).to(dout)

    x1 = x.cuda().requires_grad_(True)
    w1 = w.cuda().requires_grad_(True)
    b1 = b.cuda().requires_grad_(True)
    y = layer_norm(x1, w1, b1, 1e-5, out_dtype=dout)
    y.backward(go.cuda())
    x2 = x.double().requires_grad_(True)
    w2 = w.double().requires_grad_(True)
    b2 = b.double().requires_grad_(True)
    ref = OracleOps.layer_norm(x2, w2, b2, 1e-5)
    ref.backward(go.double())
    assert y.dtype == dout and x1.grad.dtype == din and w1.grad.dtype == torch.float32
    assert_close(y, ref, dout, "ln fwd")
    assert_close(x1.grad, x2.grad, din, "ln dx")
    # parameter gradients are fp32 sums of products of (rounded) inputs
    assert_close(w1.grad, w2.grad, torch.float32 if din == torch.float32 else din, "ln dgamma")
    assert_close(b1.grad, b2.grad, torch.float32, "ln dbeta")


def test_layer_norm_follows_autocast_and_matches_the_reference_chain():
    """Under autocast the reference computes LayerNorm in fp32 and the next Linear rounds its input to bf16: the fused
    operator must hand the Linear the same bf16 values (up to fp32 summation order: at most 1 bf16 ulp on a few)."""
    from affb200.norm_ops import layer_norm
    torch.manual_seed(3)
    x = torch.randn(8, 627, 128, device="cuda").bfloat16()
    ln = torch.nn.LayerNorm(128).cuda()
    with torch.no_grad():
        ln.weight.normal_(1, 0.1)
        ln.bias.normal_(0, 0.1)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        ref = ln(x)                      # fp32 under autocast
        ours = layer_norm(x, ln.weight, ln.bias, ln.eps)
    assert ref.dtype == torch.float32 and ours.dtype == torch.bfloat16
    diff = (ours.float() - ref.bfloat16().float()).abs()
    ulp = ref.abs().clamp_min(2.0 ** -126) * 2.0 ** -7
    assert (diff <= ulp).all()
    assert (diff > 0).float().mean() < 0.01


def test_layer_norm_rejects_cpu_tensors():
    from affb200.norm_ops import layer_norm
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        layer_norm(torch.randn(4, 64), torch.ones(64), torch.zeros(64))


def test_layer_norm_empty_and_unsupported_widths_use_stock_cuda():
    from affb200.norm_ops import layer_norm
    x = torch.randn(3, 10, 50, device="cuda")       # 50 is not a multiple of 8
    w, b = torch.ones(50, device="cuda"), torch.zeros(50, device="cuda")
    assert torch.allclose(layer_norm(x, w, b), torch.nn.functional.layer_norm(x, (50,), w, b))
    e = torch.randn(0, 64, device="cuda")
    assert layer_norm(e, torch.ones(64, device="cuda"), torch.zeros(64, device="cuda")).shape == (0, 64)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_layer_norm_fork_adds_the_shortcut_gradient_in_the_backward_pass(dtype):
    """(x, LN(x)) from one node: d_x = d_shortcut + LN'(d_y), against the two-branch expression in fp64; with
    `_affb200_sink` the parameter gradients are added into the existing fp32 .grad."""
    from affb200.norm_ops import layer_norm_fork
    from oracle.aff_ops_ref import OracleOps
    g = torch.Generator().manual_seed(9)
    shape = (3, 125, 256)
    x = torch.randn(shape, generator=g).to(dtype)
    w = 1 + 0.1 * torch.randn(256, generator=g)
    b = 0.1 * torch.randn(256, generator=g)
    g1 = torch.randn(shape, generator=g).to(dtype)
    g2 = torch.randn(shape, generator=g).to(dtype)
    for sink in (False, True):
        x1 = x.cuda().requires_grad_(True)
        w1 = w.cuda().requires_grad_(True)
        b1 = b.cuda().requires_grad_(True)
        if sink:
            for p in (w1, b1):
                p.grad = torch.full_like(p, 2.0)
                p._affb200_sink = True
        s, y = layer_norm_fork(x1, w1, b1, 1e-5, out_dtype=dtype)
        assert s.data_ptr() == x1.data_ptr() and y.dtype == dtype
        (s * g1.cuda() + y * g2.cuda()).sum().backward()
        x2 = x.double().requires_grad_(True)
        w2 = w.double().requires_grad_(True)
        b2 = b.double().requires_grad_(True)
        s2, y2 = OracleOps.layer_norm_fork(x2, w2, b2, 1e-5)
        (s2 * g1.double() + y2 * g2.double()).sum().backward()
        off = 2.0 if sink else 0.0
        assert_close(y, y2, dtype, "fork y")
        assert_close(x1.grad, x2.grad, dtype, "fork dx")
        assert_close(w1.grad - off, w2.grad, torch.float32 if dtype == torch.float32 else dtype, "fork dgamma")
        assert_close(b1.grad - off, b2.grad, torch.float32 if dtype == torch.float32 else dtype, "fork dbeta")
    # only the shortcut is used downstream: the gradient passes through untouched
    x1 = x.cuda().requires_grad_(True)
    s, y = layer_norm_fork(x1, w.cuda().requires_grad_(True), b.cuda().requires_grad_(True), 1e-5, out_dtype=dtype)
    (s * g1.cuda()).sum().backward()
    assert torch.equal(x1.grad, g1.cuda())
