"""GPU parity of the CLUSTEN operators (through the reference's autograd.Function surface → ctypes → C ABI →
sm_100a kernels) against the pure-PyTorch gather oracle (oracle/clusten_ref.py = the expressions of the
reference's clusten/test_{qk,av,wf}_kernel.py)."""
import pytest
import torch

from tests.util import TOL, assert_close, make_inputs, rel_err

pytestmark = pytest.mark.gpu

F = {}


def _fns():
    if not F:
        from clusten import CLUSTENQKFunction, CLUSTENAVFunction, CLUSTENWFFunction
        F.update(qk=CLUSTENQKFunction, av=CLUSTENAVFunction, wf=CLUSTENWFFunction)
    return F


def run_ours(op, a, b, idx, go=None):
    a = a.detach().clone().requires_grad_(True)
    b = b.detach().clone().requires_grad_(True)
    out = _fns()[op].apply(a, b, idx)
    if go is None:
        out.mean().backward()
    else:
        out.backward(go)
    return out.detach(), a.grad, b.grad


def run_oracle(op, a, b, idx, go=None, device="cpu"):
    """fp64 evaluation of the gather expressions on the (already rounded) operands."""
    from oracle.clusten_ref import fwd_bwd
    a64, b64 = a.detach().double().to(device), b.detach().double().to(device)
    go64 = None if go is None else go.detach().double().to(device)
    return fwd_bwd(op, a64, b64, idx.to(device), go64)


def check(op, dims, dtype, seed=0, idx_mode="uniform", mean_loss=False, oracle_device="cpu"):
    a, b, idx, go = make_inputs(op, dims, dtype, seed, device="cuda", idx_mode=idx_mode)
    if mean_loss:
        go = None
    out, ga, gb = run_ours(op, a, b, idx, go)
    ro, rga, rgb = run_oracle(op, a, b, idx, go, oracle_device)
    assert out.dtype == dtype and ga.dtype == dtype and gb.dtype == dtype
    assert_close(out, ro, dtype, "%s fwd %s" % (op, dims))
    assert_close(ga, rga, dtype, "%s grad_a %s" % (op, dims))
    assert_close(gb, rgb, dtype, "%s grad_b %s" % (op, dims))


# BASELINE config 1: AFF-Mini stage 1 (B=2, N=3136, heads 2, nbhd 48); c=16 is the model's per-head dim, c=32 the
# reference kernel tests' head dim.
CFG1 = {"qk": [(2, 2, 3136, 3136, 48, 16), (2, 2, 3136, 3136, 48, 32)],
        "av": [(2, 2, 3136, 3136, 48, 16), (2, 2, 3136, 3136, 48, 32)],
        "wf": [(2, 784, 3136, 48, 4, 32)]}


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16, torch.float64])
@pytest.mark.parametrize("op", ["qk", "av", "wf"])
def test_config1_shapes(op, dtype):
    for dims in CFG1[op]:
        check(op, dims, dtype, seed=1001)


@pytest.mark.parametrize("op", ["qk", "av", "wf"])
def test_reference_test_shape_mean_loss(op):
    """The reference's own test: b=256,h=4,n=196,m=48,c=32 (test_qk_kernel.py:14-18), loss = out.mean();
    WF: n_=64, ic=4 (test_wf_kernel.py:14-19).  Oracle evaluated on the GPU in fp64 (size)."""
    dims = (256, 4, 196, 196, 48, 32) if op != "wf" else (256, 64, 196, 48, 4, 32)
    check(op, dims, torch.float32, seed=7, mean_loss=True, oracle_device="cuda")
    check(op, dims, torch.bfloat16, seed=8, oracle_device="cuda")


DENSE_CASES = [
    # (B, H, Nq, Nk, M, c, m) — local-attention stage shapes of BASELINE configs 2-5 at reduced batch
    (4, 2, 3136, 3136, 48, 32, 8),    # AFF-Tiny stage 1
    (4, 2, 3136, 3136, 48, 16, 8),    # AFF-Mini stage 1 (head dim 16)
    (8, 4, 627, 627, 48, 32, 8),      # AFF-Tiny-1/5 stage 2: N % 8 != 0 -> padded last cluster, duplicate row 0
    (40, 8, 125, 125, 48, 32, 8),     # stage 3
    (2, 4, 2304, 2304, 144, 32, 24),  # AFF-Base 384 stage 2 (nbhd 144)
    (1, 2, 30720, 30720, 48, 32, 8),  # maskfiner j=3 (largest token count)
    (3, 2, 100, 333, 40, 32, 8),      # Nq != Nk, partial last tile
]


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("op", ["qk", "av"])
def test_dense_tile_path_on_clustered_indices(op, dtype):
    """Clustered neighbourhoods (the model's regime) take the dense-tile tensor-core kernels: the tile plans must
    build without overflow, and outputs/gradients must match the fp64 oracle within the bf16/fp16 tolerance."""
    from tests.util import clustered_idx
    from affb200.plan import get_plan
    for s, (B, H, Nq, Nk, M, c, m) in enumerate(DENSE_CASES):
        a, b, _, go = make_inputs(op, (B, H, Nq, Nk, M, c), dtype, 40 + s, device="cuda")
        idx = clustered_idx(B, Nq, Nk, M, m, 50 + s).cuda()
        out, ga, gb = run_ours(op, a, b, idx, go)
        st = get_plan(idx).stats
        assert st["q"]["overflow"] == 0 and st["k"]["overflow"] == 0, st
        if Nk % m:
            assert st["q"]["dup_tiles"] > 0  # padded slots all point at row 0
        ro, rga, rgb = run_oracle(op, a, b, idx, go, "cuda")
        assert_close(out, ro, dtype, "%s fwd %s" % (op, DENSE_CASES[s]))
        assert_close(ga, rga, dtype, "%s grad_a %s" % (op, DENSE_CASES[s]))
        assert_close(gb, rgb, dtype, "%s grad_b %s" % (op, DENSE_CASES[s]))


def test_dense_and_generic_paths_agree():
    """Same bf16 inputs through the dense-tile kernels and through the generic row-gather kernels (fp32 math on
    the same rounded operands): both round an fp32 sum once, so they agree to ~1 ulp of bf16."""
    from tests.util import clustered_idx
    from affb200 import _native as N
    from affb200.plan import get_plan
    B, H, Nn, M, c = 4, 2, 784, 48, 32
    q, k, _, _ = make_inputs("qk", (B, H, Nn, Nn, M, c), torch.bfloat16, 77, device="cuda")
    idx = clustered_idx(B, Nn, Nn, M, 8, 78).cuda()
    o1, o2 = torch.empty(B, H, Nn, M, device="cuda", dtype=torch.bfloat16), torch.empty(B, H, Nn, M, device="cuda", dtype=torch.bfloat16)
    pl = get_plan(idx)
    A = (B, H, Nn, Nn, M, c, N.dtype_enum(q), N.stream_ptr(q.device))
    N.check(N.LIB.affb200_qk_fwd(N.p(q), N.p(k), N.p(idx), pl.struct(idx, Nn, True, False), N.p(o1), *A), "dense")
    N.check(N.LIB.affb200_qk_fwd(N.p(q), N.p(k), N.p(idx), None, N.p(o2), *A), "generic")
    assert pl.stats["q"]["overflow"] == 0
    assert rel_err(o1, o2) < 4e-3
    assert (o1 == o2).float().mean() > 0.9


RAGGED = {
    "qk": [(1, 1, 1, 1, 1, 1), (3, 2, 17, 5, 7, 3), (2, 3, 33, 70, 13, 8), (1, 5, 64, 64, 48, 24),
           (2, 1, 100, 9, 144, 32), (1, 2, 40, 40, 50, 20), (2, 2, 9, 300, 6, 128), (1, 16, 25, 25, 25, 32)],
    "wf": [(1, 1, 1, 1, 1, 1), (3, 17, 5, 7, 3, 5), (2, 33, 70, 13, 4, 8), (1, 64, 64, 48, 4, 24),
           (2, 25, 125, 48, 4, 256), (1, 40, 40, 50, 2, 20), (2, 9, 300, 6, 4, 64), (1, 7, 30, 5, 6, 16)],
}
RAGGED["av"] = RAGGED["qk"]


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("op", ["qk", "av", "wf"])
def test_ragged_shapes(op, dtype):
    """Nq != Nk, M and c not multiples of the vector width (scalar fallback), single elements."""
    for s, dims in enumerate(RAGGED[op]):
        check(op, dims, dtype, seed=100 + s)


@pytest.mark.parametrize("mode", ["dup", "hot"])
@pytest.mark.parametrize("op", ["qk", "av", "wf"])
def test_duplicates_and_hot_rows(op, mode):
    """Duplicate neighbours accumulate (reference semantics); 'hot' makes three key rows receive
    every contribution, which exercises the long-row branch of the transposed index (rows > 1024 entries)."""
    dims = (2, 2, 200, 50, 48, 32) if op != "wf" else (2, 200, 50, 48, 4, 32)
    check(op, dims, torch.float32, seed=5, idx_mode=mode)
    check(op, dims, torch.bfloat16, seed=6, idx_mode=mode)


@pytest.mark.parametrize("op", ["qk", "av", "wf"])
def test_empty(op):
    """Zero-sized batch / token / neighbour dimensions do not launch and return correctly shaped tensors;
    M == 0 sums are zero."""
    for dims in ([(0, 2, 8, 8, 4, 8), (2, 2, 0, 8, 4, 8), (2, 2, 8, 8, 0, 8)] if op != "wf" else
                 [(0, 8, 8, 4, 4, 8), (2, 0, 8, 4, 4, 8), (2, 8, 8, 0, 4, 8)]):
        a, b, idx, go = make_inputs(op, dims, torch.float32, 0, device="cuda")
        out, ga, gb = run_ours(op, a, b, idx, go)
        ro, rga, rgb = run_oracle(op, a, b, idx, go)
        assert out.shape == ro.shape
        assert torch.equal(out.double().cpu(), ro) and torch.equal(gb.double().cpu(), rgb)


@pytest.mark.parametrize("op", ["qk", "av", "wf"])
def test_backward_is_deterministic(op):
    """Gradients are bit-identical run to run (the reference's atomics are not: clustenqk_cuda_kernel.cu:125)."""
    dims = (4, 4, 196, 196, 48, 32) if op != "wf" else (4, 64, 196, 48, 4, 32)
    a, b, idx, go = make_inputs(op, dims, torch.bfloat16, 3, device="cuda")
    r1 = run_ours(op, a, b, idx, go)
    from affb200.plan import clear_plans
    clear_plans()
    r2 = run_ours(op, a, b, idx.clone(), go)
    for x, y in zip(r1, r2):
        assert torch.equal(x, y)


def test_dtype_coercion_and_noncontiguous():
    """Second operand is cast to the first operand's dtype (clusten.py:25-26,52-53,78-79); non-contiguous
    operands are made contiguous (clusten.py:23-24)."""
    q, k, idx, go = make_inputs("qk", (2, 2, 64, 64, 48, 32), torch.float32, 11, device="cuda")
    fn = _fns()
    out = fn["qk"].apply(q.half(), k, idx)
    assert out.dtype == torch.float16
    ro = run_oracle("qk", q.half(), k.half(), idx)[0]
    assert_close(out, ro, torch.float16, "qk mixed dtype")
    # AV under AMP: attn fp32 (softmax output), v half -> runs and returns fp32
    a, v, idx, go = make_inputs("av", (2, 2, 64, 64, 48, 32), torch.float32, 12, device="cuda")
    out = fn["av"].apply(a, v.half(), idx)
    assert out.dtype == torch.float32
    assert_close(out, run_oracle("av", a, v.half().float(), idx)[0], torch.float32, "av mixed dtype")
    # non-contiguous q (a permuted view, as produced by the model's reshape/permute: aff_transformer.py:103)
    qt = q.permute(0, 2, 1, 3).contiguous().permute(0, 2, 1, 3)
    assert not qt.is_contiguous()
    assert torch.equal(fn["qk"].apply(qt, k, idx), fn["qk"].apply(q, k, idx))
    idx_nc = idx.transpose(1, 2).contiguous().transpose(1, 2)
    assert torch.equal(fn["qk"].apply(q, k, idx_nc), fn["qk"].apply(q, k, idx))


def test_grad_wrt_idx_is_none_and_errors():
    fn = _fns()
    q, k, idx, go = make_inputs("qk", (1, 1, 8, 8, 4, 8), torch.float32, 0, device="cuda")
    q.requires_grad_(True)
    out = fn["qk"].apply(q, k, idx)
    assert out.grad_fn is not None
    with pytest.raises(RuntimeError, match="must be a CUDA tensor"):
        fn["qk"].apply(q.cpu(), k.cpu(), idx.cpu())
    with pytest.raises(RuntimeError):
        fn["qk"].apply(q.to(torch.int32), k.to(torch.int32), idx)
