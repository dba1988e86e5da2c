"""GPU parity of the column-sum bias gradient (csrc/col_sum.cu) and of the Linear operator built on it against
nn.functional.linear evaluated in float64 on the same inputs (reference layers: aff_transformer.py:85-100,285)."""
import pytest
import torch

from tests.util import assert_close

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("shape", [(1, 64), (37, 64), (5000, 64), (100003, 128), (777, 96), (4096, 256), (300, 1024),
                                   (64, 2048), (9, 8), (32000, 768), (802816, 64), (700, 4096), (300, 4104)])
def test_col_sum(shape, dtype):
    from affb200.linear_ops import col_sum
    g = torch.Generator().manual_seed(shape[0] + shape[1])
    x = torch.randn(shape, generator=g).to(dtype)
    out = col_sum(x.cuda())
    assert out.dtype == torch.float32 and out.shape == (shape[1],)
    assert_close(out, x.double().sum(0), torch.float32, "col_sum")
    # fixed summation order: bit-reproducible
    assert torch.equal(out, col_sum(x.cuda()))


def test_col_sum_accumulates_and_is_exact_across_calls_and_streams():
    """`into`: sums are added in place; back-to-back calls of different shapes on one stream and calls on a second
    stream give the same bits."""
    from affb200.linear_ops import col_sum
    g = torch.Generator(device="cuda").manual_seed(3)
    xs = [torch.randn(r, c, device="cuda", generator=g).bfloat16() for r, c in
          [(32000, 256), (32000, 768), (6400, 512), (160512, 128), (32000, 256), (1, 4096), (33, 8)]]
    want = [x.double().sum(0) for x in xs]
    for _ in range(3):
        for x, w in zip(xs, want):
            assert_close(col_sum(x), w, torch.float32, "col_sum")
    base = torch.randn(256, device="cuda", generator=g)
    acc = base.clone()
    assert col_sum(xs[0], into=acc) is acc
    assert_close(acc, base.double() + want[0], torch.float32, "col_sum into")
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        outs = [col_sum(x) for x in xs]
    mains = [col_sum(x) for x in xs]
    torch.cuda.synchronize()
    for a, b in zip(outs, mains):
        assert torch.equal(a, b)


def test_col_sum_edge_cases():
    from affb200.linear_ops import col_sum
    assert torch.equal(col_sum(torch.zeros(0, 64, device="cuda")), torch.zeros(64, device="cuda"))
    x = torch.randn(50, 12, device="cuda")   # width not a multiple of 8 -> stock reduction
    assert torch.allclose(col_sum(x), x.sum(0))
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        col_sum(torch.zeros(4, 8))


@pytest.mark.parametrize("autocast", [False, True])
@pytest.mark.parametrize("shape", [(4, 300, 64, 192), (2, 125, 256, 512), (3, 25, 1024, 512), (8, 64, 8)])
def test_linear_matches_nn_linear(shape, autocast):
    from affb200.linear_ops import linear
    *lead, cin, cout = shape
    g = torch.Generator().manual_seed(cin + cout)
    x = torch.randn(*lead, cin, generator=g)
    w = torch.randn(cout, cin, generator=g) / cin ** 0.5
    b = torch.randn(cout, generator=g)
    go = torch.randn(*lead, cout, generator=g)

    def run(fn):
        xs, ws, bs = (t.cuda().clone().requires_grad_(True) for t in (x, w, b))
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
            y = fn(xs, ws, bs)
        y.backward(go.cuda().to(y.dtype))
        return y, xs.grad, ws.grad, bs.grad

    ours = run(linear)
    ref = run(torch.nn.functional.linear)
    dt = torch.bfloat16 if autocast else torch.float32
    assert ours[0].dtype == ref[0].dtype
    for a, r, what in zip(ours, ref, ("y", "dx", "dw", "db")):
        assert a.dtype == r.dtype, what
        assert_close(a, r.double(), dt, "linear " + what)


@pytest.mark.parametrize("sink", [False, True])
@pytest.mark.parametrize("autocast", [False, True])
def test_linear_pair_matches_two_linears(autocast, sink):
    """q / kv of the attention: two Linears on one input, one dx in the backward (second GEMM accumulates onto the
    first); with fp32 gradient sinks the parameter gradients land in .grad directly."""
    from affb200.linear_ops import linear_pair
    g = torch.Generator().manual_seed(2)
    x = torch.randn(3, 125, 256, generator=g)
    ws = [torch.randn(256, 256, generator=g) / 16, torch.randn(512, 256, generator=g) / 16]
    bs = [torch.randn(256, generator=g), torch.randn(512, generator=g)]
    gos = [torch.randn(3, 125, 256, generator=g), torch.randn(3, 125, 512, generator=g)]

    def run(pair):
        xs = x.cuda().clone().requires_grad_(True)
        ps = [t.cuda().clone().requires_grad_(True) for t in (ws[0], bs[0], ws[1], bs[1])]
        if sink and pair:
            for p_ in ps:
                p_.grad = torch.zeros_like(p_)
                p_._affb200_sink = True
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
            if pair:
                y1, y2 = linear_pair(xs, *ps)
            else:
                y1 = torch.nn.functional.linear(xs, ps[0], ps[1])
                y2 = torch.nn.functional.linear(xs, ps[2], ps[3])
        torch.autograd.backward([y1, y2], [gos[0].cuda().to(y1.dtype), gos[1].cuda().to(y2.dtype)])
        return [y1, y2, xs.grad] + [p_.grad for p_ in ps]

    ours, ref = run(True), run(False)
    dt = torch.bfloat16 if autocast else torch.float32
    for a, r, what in zip(ours, ref, ("y1", "y2", "dx", "dw1", "db1", "dw2", "db2")):
        assert a.dtype == r.dtype and a.shape == r.shape, what
        assert_close(a, r.double(), dt, "linear_pair " + what)
