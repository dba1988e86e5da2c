"""GPU, BASELINE.json's full sizes (AFF-Tiny-1/5 stage 1 at the yaml batch: B=256, H=2, N=3136, M=48, c=32; the
stage-2 ragged shape N=627 with a padded last cluster): checks that need no oracle run at this size — exact scaling
by a power of two, one-hot attention = row gather, the adjoint identity <op(x), g> = <x, op^T(g)> that ties every backward
kernel to its forward, and structural properties of the index builders (rankings are permutations, kNN distances ascend,
the merge selection is duplicate-free and contains the reserved grid)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

B, H, N, M, C = 256, 2, 3136, 48, 32


def _idx_real(b, n, hw, m=8, nbhd=48, seed=0):
    """neighbourhoods from the real pipeline on an hw x hw canvas (grid when n == hw*hw, else a random subset)"""
    from affb200.point_utils import space_filling_cluster, build_nbhd
    g = torch.Generator().manual_seed(seed)
    ys, xs = torch.meshgrid(torch.arange(hw), torch.arange(hw), indexing="ij")
    allp = torch.stack([xs, ys], 2).reshape(-1, 2).float()
    if n == hw * hw:
        pos = allp.unsqueeze(0).expand(b, -1, -1).contiguous()
    else:
        pos = torch.stack([allp[torch.randperm(hw * hw, generator=g)[:n]] for _ in range(b)])
    ps, mean, _, _, rank = space_filling_cluster(pos.cuda(), m, hw, hw)
    nb, mask, pe, near = build_nbhd(ps, mean, m, nbhd, hw - 1)
    return pos, ps, mean, rank, nb, mask, pe, near


def _dot(a, b):
    return (a.double() * b.double()).sum().item()


def _adjoint_ok(lhs, rhs, out, go, dtype):
    """|<out, go> - <x, dx>| against the size of the sum: rounding noise grows like sqrt(#terms) * |term| * eps, far
    below eps_dtype * ||out|| * ||go||; a wrong backward kernel misses by O(||out|| * ||go|| / sqrt(#terms))."""
    scale = out.double().norm().item() * go.double().norm().item()
    eps = {torch.float32: 1e-9, torch.bfloat16: 2e-6}[dtype]
    return abs(lhs - rhs) <= eps * scale


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_qk_av_full_size_scaling_onehot_adjoint(dtype):
    from clusten import CLUSTENQKFunction as QK, CLUSTENAVFunction as AV
    _, _, _, _, nb, _, _, _ = _idx_real(B, N, 56)
    g = torch.Generator(device="cuda").manual_seed(1)
    q = torch.randn(B, H, N, C, device="cuda", generator=g).to(dtype).requires_grad_(True)
    k = torch.randn(B, H, N, C, device="cuda", generator=g).to(dtype).requires_grad_(True)
    a = QK.apply(q, k, nb)
    assert a.shape == (B, H, N, M) and a.dtype == dtype
    # scaling by 2 is exact in binary floating point: bit-identical results
    assert torch.equal(QK.apply((q * 2).detach(), k.detach(), nb), a.detach() * 2)
    # adjoint identity: <QK(q,k), g> = <q, dq> = <k, dk> (QK is bilinear)
    go = torch.randn(a.shape, device="cuda", generator=g).to(dtype)
    a.backward(go)
    lhs = _dot(a.detach(), go)
    for name, x in (("dq", q), ("dk", k)):
        rhs = _dot(x.detach(), x.grad)
        assert _adjoint_ok(lhs, rhs, a.detach(), go, dtype), (name, lhs, rhs)
    del a, go
    # AV: one-hot attention returns exactly the gathered value rows
    v = torch.randn(B, H, N, C, device="cuda", generator=g).to(dtype)
    j0 = 17
    att = torch.zeros(B, H, N, M, device="cuda", dtype=dtype)
    att[..., j0] = 1
    out = AV.apply(att, v, nb)
    want = v.gather(2, nb[:, None, :, j0, None].expand(-1, H, -1, C))
    assert torch.equal(out, want)
    # AV adjoint in both operands
    att = torch.softmax(torch.randn(B, H, N, M, device="cuda", generator=g), -1).to(dtype).requires_grad_(True)
    v = v.requires_grad_(True)
    out = AV.apply(att, v, nb)
    go = torch.randn(out.shape, device="cuda", generator=g).to(dtype)
    out.backward(go)
    lhs = _dot(out.detach(), go)
    for name, x in (("d_attn", att), ("d_v", v)):
        rhs = _dot(x.detach(), x.grad)
        assert _adjoint_ok(lhs, rhs, out.detach(), go, dtype), (name, lhs, rhs)


def test_wf_full_size_adjoint_and_onehot():
    from clusten import CLUSTENWFFunction as WF
    Nq, Nk, IC, Cf = 627, 3136, 4, 64
    g = torch.Generator(device="cuda").manual_seed(2)
    idx = torch.randint(Nk, (B, Nq, M), device="cuda", generator=g)
    w = torch.randn(B, Nq, M, IC, device="cuda", generator=g).requires_grad_(True)
    f = torch.randn(B, Nk, Cf, device="cuda", generator=g).requires_grad_(True)
    out = WF.apply(w, f, idx)
    assert out.shape == (B, Nq, IC, Cf)
    go = torch.randn(out.shape, device="cuda", generator=g)
    out.backward(go)
    lhs = _dot(out.detach(), go)
    for name, x in (("d_w", w), ("d_f", f)):
        rhs = _dot(x.detach(), x.grad)
        assert _adjoint_ok(lhs, rhs, out.detach(), go, torch.float32), (name, lhs, rhs)
    oh = torch.zeros(B, Nq, M, IC, device="cuda")
    oh[:, :, 5, 2] = 1
    got = WF.apply(oh, f.detach(), idx)
    assert torch.equal(got[:, :, 2], f.detach().gather(1, idx[:, :, 5, None].expand(-1, -1, Cf)))
    assert got[:, :, [0, 1, 3]].abs().max().item() == 0


@pytest.mark.parametrize("n,hw", [(3136, 56), (627, 56)])
def test_index_builders_full_batch_structure(n, hw):
    from affb200.point_utils import knn_keops, select_merge_centers
    pos, ps, mean, rank, nb, mask, pe, near = _idx_real(B, n, hw, seed=3)
    ar = torch.arange(n, device="cuda")
    # pos_ranking is a permutation per image and pos_sorted is pos gathered by it
    assert torch.equal(rank.squeeze(2).sort(1)[0], ar.expand(B, -1))
    assert torch.equal(ps, pos.cuda().gather(1, rank.expand(-1, -1, 2)))
    # neighbourhoods: whole clusters (8 consecutive rows, or row 0 in the padded tail), indices in range
    k = (n + 7) // 8
    assert int(nb.min()) >= 0 and int(nb.max()) < n and int(near.max()) < k
    first = nb[..., ::8]
    cl = nb.reshape(B, n, 6, 8)
    inside = (cl - first.unsqueeze(3) == torch.arange(8, device="cuda")) | (cl == 0)
    assert bool(inside.all())
    if n % 8:
        assert mask is not None and int(mask.sum()) == B * n * 48 - int((near == k - 1).sum()) * (k * 8 - n)
    else:
        assert mask is None
    # kNN distances ascend; the nearest cluster of a token is at least as close as any other centre
    idx2, d2 = knn_keops(ps, mean, 6, return_dist=True)
    assert torch.equal(idx2, near) and bool((d2[..., 1:] >= d2[..., :-1]).all())
    # merge selection: keep = int(n * 0.2) distinct tokens, the reserved coarse grid last and ascending
    stride = 2 if n == hw * hw else 4
    res = ((ps % (2 * stride)).sum(-1) == 0)
    reserve_num = int(res[0].sum())
    if bool((res.sum(1) == reserve_num).all()):
        lp = torch.rand(B, n, 1, device="cuda", generator=torch.Generator(device="cuda").manual_seed(4))
        sel, _ = select_merge_centers(ps, lp, stride, 4.0, 0.2, reserve_num, check=True)
        assert sel.shape == (B, int(n * 0.2))
        srt = sel.sort(1)[0]
        assert bool((srt[:, 1:] != srt[:, :-1]).all())
        tail = sel[:, sel.shape[1] - reserve_num:]
        assert bool(res.gather(1, tail).all()) and bool((tail[:, 1:] > tail[:, :-1]).all())
