"""GPU parity of the index builders (C ABI kernels behind affb200/point_utils.py) — bit-exact integer outputs against
(a) the golden vectors produced by the reference's own space_filling_cluster (tests/golden/point_sfc.pt) and
(b) the CPU oracle (oracle/point_ref.py) on seeded inputs at every BASELINE config's token counts."""
import math
import os

import pytest
import torch

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def grid_pos(h, w):
    ys, xs = torch.meshgrid(torch.arange(h), torch.arange(w), indexing="ij")
    return torch.stack([xs, ys], 2).reshape(-1, 2).float()


def subset(h, w, n, B, seed, dup=0):
    g = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(B):
        p = grid_pos(h, w)[torch.randperm(h * w, generator=g)[: n - dup]]
        if dup:
            p = torch.cat([p, p[torch.randint(p.shape[0], (dup,), generator=g)]])
            p = p[torch.randperm(n, generator=g)]
        out.append(p)
    return torch.stack(out)


def test_sfc_matches_reference_golden():
    from affb200.point_utils import space_filling_cluster
    gold = torch.load(os.path.join(GOLD, "point_sfc.pt"))
    for name, c in gold.items():
        pos = c["pos"].float().cuda()
        p_s, mean, member, mask, ranking = space_filling_cluster(pos, c["m"], c["h"], c["w"])
        assert ranking.dtype == torch.int64 and ranking.shape == (pos.shape[0], pos.shape[1], 1)
        assert torch.equal(ranking.squeeze(2).cpu().to(torch.int32), c["ranking"]), name
        if c["ordering_rule_imposed"]:   # integer positions: sums are exact in fp32, centres are bit-exact
            assert torch.equal(mean.cpu(), c["mean"]), name
        else:                            # fractional positions: fp32 centre within 1e-5 relative (summation order)
            assert torch.allclose(mean.cpu(), c["mean"], rtol=1e-5, atol=0), name
        assert (mask is not None) == c["has_mask"], name
        assert torch.equal(p_s, pos.gather(1, ranking.expand(-1, -1, 2))), name


SFC_CASES = [  # (h, w, n, B, m)
    (56, 56, 3136, 3, 8), (56, 56, 627, 5, 8), (56, 56, 125, 7, 8), (56, 56, 784, 2, 8), (56, 56, 196, 2, 8),
    (96, 96, 9216, 1, 24), (96, 96, 2304, 2, 24), (96, 96, 576, 2, 24),
    (256, 512, 30720, 1, 8), (256, 512, 6144, 1, 8), (100, 100, 2499, 2, 20), (8, 8, 24, 2, 8),
]


@pytest.mark.parametrize("case", SFC_CASES)
def test_sfc_matches_oracle(case):
    from affb200.point_utils import space_filling_cluster
    from oracle import point_ref as O
    h, w, n, B, m = case
    pos = subset(h, w, n, B, seed=n + B, dup=(n // 10 if h == 256 else 0))
    p_s, mean, member, mask, ranking = space_filling_cluster(pos.cuda(), m, h, w)
    rp, rmean, rmember, rmask, rrank = O.sfc_ref(pos, m, h, w)
    assert torch.equal(ranking.cpu(), rrank)
    assert torch.equal(p_s.cpu(), rp)
    assert torch.equal(member.cpu(), rmember)
    assert (mask is None) == (rmask is None) and (mask is None or torch.equal(mask.cpu(), rmask))
    assert torch.equal(mean.cpu(), rmean)


def test_sfc_batch_global_max_coupling():
    """The key uses the maximum ratio of the whole batch (point_utils.py:158): results are compared on the same
    batch composition, and a different composition may legitimately reorder an image."""
    from affb200.point_utils import space_filling_cluster
    from oracle import point_ref as O
    pos = subset(56, 56, 627, 4, seed=9)
    r_all = space_filling_cluster(pos.cuda(), 8, 56, 56)[4].cpu()
    assert torch.equal(r_all, O.sfc_ref(pos, 8, 56, 56)[4])
    r_one = space_filling_cluster(pos[:1].cuda(), 8, 56, 56)[4].cpu()
    assert torch.equal(r_one, O.sfc_ref(pos[:1], 8, 56, 56)[4])


KNN_CASES = [(3136, 392, 6, 2), (627, 79, 6, 4), (125, 16, 6, 8), (9216, 384, 6, 1), (627, 627, 2, 3),
             (125, 125, 2, 3), (2000, 250, 6, 2), (50, 7, 7, 2), (300, 40, 12, 2), (100, 100, 32, 1)]


@pytest.mark.parametrize("case", KNN_CASES)
def test_knn_matches_oracle(case):
    """Integer grid queries against cluster centres (multiples of 1/8): exact ties at the k-th boundary are common
    (SURVEY.md §4), so this exercises the lowest-index tie rule."""
    from affb200.point_utils import knn_keops
    from oracle import point_ref as O
    nq, nd, k, B = case
    q = subset(56, 56, min(nq, 3136), B, seed=nq) if nq <= 3136 else subset(96, 96, nq, B, seed=nq)
    if nd == nq:
        d = q
    else:
        g = torch.Generator().manual_seed(nd)
        d = torch.randint(0, 56 * 8, (B, nd, 2), generator=g).float() / 8
    idx, dist = knn_keops(q.cuda(), d.cuda(), k, return_dist=True)
    ridx, rdist = O.knn_ref(q, d, k, return_dist=True)
    assert torch.equal(idx.cpu(), ridx)
    bad = (dist.cpu() != rdist)
    assert not bad.any(), (int(bad.sum()), dist.cpu()[bad][:4].tolist(), rdist[bad][:4].tolist(),
                           (dist.cpu()[bad][:4] ** 2).tolist())
    assert torch.equal(knn_keops(q.cuda(), d.cuda(), k).cpu(), ridx)


@pytest.mark.parametrize("case", [(56, 56, 3136, 2, 8, 48, 55), (56, 56, 627, 3, 8, 48, 55), (56, 56, 125, 4, 8, 48, 55),
                                  (96, 96, 2304, 1, 24, 144, 95), (64, 128, 2000, 2, 8, 48, 31), (8, 8, 20, 2, 8, 48, 7)])
def test_pipeline_sfc_knn_nbhd_matches_oracle(case):
    from affb200.point_utils import space_filling_cluster, build_nbhd
    from oracle import point_ref as O
    h, w, n, B, m, nbhd_size, rpw = case
    pos = subset(h, w, n, B, seed=3 * n)
    p_s, mean, _, _, _ = space_filling_cluster(pos.cuda(), m, h, w)
    for clamp in (False, True):
        nbhd, mask, pe, near = build_nbhd(p_s, mean, m, nbhd_size, rpw, clamp=clamp)
        rp, rmean = O.sfc_ref(pos, m, h, w)[:2]
        rn, rm, rpe, rnear = O.nbhd_ref(rp, rmean, m, nbhd_size, rpw, clamp=clamp)
        assert torch.equal(near.cpu(), rnear)
        assert torch.equal(nbhd.cpu(), rn)
        assert (mask is None) == (rm is None) and (mask is None or torch.equal(mask.cpu(), rm))
        assert torch.equal(pe.cpu(), rpe)


SEL_CASES = [  # (h, w, n, B, stride, ds_rate, dtype, levels)
    (56, 56, 3136, 3, 2, 0.2, torch.float32, 0), (56, 56, 3136, 2, 2, 0.25, torch.bfloat16, 0),
    (56, 56, 627, 4, 4, 0.2, torch.float32, 0), (56, 56, 784, 3, 4, 0.25, torch.bfloat16, 0),
    (56, 56, 125, 5, 8, 0.2, torch.float16, 0), (56, 56, 196, 3, 8, 0.25, torch.float32, 0),
    (96, 96, 9216, 1, 2, 0.25, torch.float32, 0), (56, 56, 3136, 2, 2, 0.2, torch.float32, 3),
    (56, 56, 627, 2, 4, 0.2, torch.float32, 2),
]


@pytest.mark.parametrize("case", SEL_CASES)
def test_merge_selection_matches_oracle(case):
    """Set AND order of the kept tokens; `levels` > 0 quantises the scores (saturated sigmoid => ties at the k-th
    value), which exercises torch.topk's CUDA output order."""
    from affb200.point_utils import select_merge_centers
    from oracle import point_ref as O
    h, w, n, B, stride, ds, dtype, levels = case
    g = torch.Generator().manual_seed(n + stride)
    if stride == 2:
        pos = grid_pos(h, w).unsqueeze(0).expand(B, -1, -1).contiguous()
    else:
        # tokens that survived earlier merges: all reserved coarse-grid tokens + a random subset of the rest
        allp = grid_pos(h, w)
        res = ((allp % stride).sum(-1) == 0)
        pos = []
        for _ in range(B):
            rest = allp[~res][torch.randperm(int((~res).sum()), generator=g)[: n - int(res.sum())]]
            p = torch.cat([rest, allp[res]])
            pos.append(p[torch.randperm(n, generator=g)])
        pos = torch.stack(pos)
    lp = torch.rand(B, n, 1, generator=g)
    if levels:
        lp = (lp * levels).floor() / levels
    lp = lp.to(dtype)
    reserve_num = math.ceil(h / (stride * 2)) * math.ceil(w / (stride * 2))
    idx, final = select_merge_centers(pos.cuda(), lp.cuda(), stride, 4.0, ds, reserve_num, check=True)
    ridx, rfinal = O.select_ref(pos, lp, stride, 4.0, ds, reserve_num)
    assert torch.equal(final.cpu(), rfinal)
    assert torch.equal(idx.cpu(), ridx)


def test_argsort_and_gather_rows():
    from affb200.point_utils import argsort_stable, gather_rows
    g = torch.Generator().manual_seed(0)
    for (B, n) in [(2, 2048), (1, 8192), (3, 4096), (4, 77)]:
        x = torch.randn(B, n, generator=g)
        x[:, 0:n - 7:7] = x[:, 1:n - 6:7]  # ties
        for desc in (False, True):
            assert torch.equal(argsort_stable(x.cuda(), desc).cpu(), x.sort(dim=1, descending=desc, stable=True)[1])
    src = torch.randint(0, 1000, (3, 50, 48), generator=g)
    idx = torch.randint(0, 50, (3, 20), generator=g)
    assert torch.equal(gather_rows(src.cuda(), idx.cuda()).cpu(), src.gather(1, idx.unsqueeze(2).expand(-1, -1, 48)))


@pytest.mark.gpu
def test_sfc_variants_match_reference_golden():
    """use_anchor=False / no_reorder=True call forms through the CUDA path against outputs of the reference itself;
    integer outputs bit-exact, cluster means within 1e-5."""
    import os
    from affb200.point_utils import space_filling_cluster
    g = torch.load(os.path.join(os.path.dirname(__file__), "golden", "point_sfc_variants.pt"))
    for name, c in g.items():
        out = space_filling_cluster(c["pos"].float().cuda(), c["m"], c["h"], c["w"], no_reorder=c["no_reorder"],
                                    use_anchor=c["use_anchor"])
        assert len(out) == len(c["out"]), name
        for a, b in zip(out, c["out"]):
            if b is None:
                assert a is None, name
            elif b.dtype == torch.float32:
                assert torch.allclose(a.cpu(), b, rtol=1e-5, atol=0), name
            else:
                assert torch.equal(a.cpu().to(b.dtype), b), name


@pytest.mark.gpu
def test_curve_orders_match_reference_golden():
    """calculate_peano_order / calculate_hilbert_order through affb200_curve_key + the stable rank sort against the
    reference's own outputs (square, widened-first-level, non-power and fractional-position cases): bit-exact."""
    import os
    from affb200.point_utils import calculate_peano_order, calculate_hilbert_order
    g = torch.load(os.path.join(os.path.dirname(__file__), "golden", "point_sfc_curves.pt"))["orders"]
    assert len(g) == 9
    for name, c in g.items():
        fn = calculate_peano_order if c["curve"] == "peano" else calculate_hilbert_order
        rank, order = fn(c["h"], c["w"], c["pos"].cuda())
        assert rank.dtype == torch.int64 and order.dtype == torch.int64
        assert torch.equal(order.cpu().to(torch.int32), c["order"]), name
        assert torch.equal(rank.cpu().to(torch.int32), c["rank"]), name


@pytest.mark.gpu
def test_sfc_curve_variants_match_reference_golden():
    """space_filling_cluster(sf_type='peano'|'hilbert') with and without anchors / reordering against outputs of the
    reference itself; integer outputs bit-exact, cluster means within 1e-5."""
    import os
    from affb200.point_utils import space_filling_cluster
    g = torch.load(os.path.join(os.path.dirname(__file__), "golden", "point_sfc_curves.pt"))["clusters"]
    assert len(g) == 19
    for name, c in g.items():
        out = space_filling_cluster(c["pos"].float().cuda(), c["m"], c["h"], c["w"], no_reorder=c["no_reorder"],
                                    sf_type=c["sf_type"], use_anchor=c["use_anchor"])
        assert len(out) == len(c["out"]), name
        for a, b in zip(out, c["out"]):
            if b is None:
                assert a is None, name
            elif b.dtype == torch.float32:
                assert torch.allclose(a.cpu(), b, rtol=1e-5, atol=0), name
            else:
                assert torch.equal(a.cpu().to(b.dtype), b), name


@pytest.mark.gpu
def test_hilbert_on_a_wide_canvas_is_refused():
    """round(2*w/h) != 2: the reference indexes its 4-entry digit table out of range (point_utils.py:322-338)."""
    from affb200.point_utils import calculate_hilbert_order
    with pytest.raises(RuntimeError, match="Hilbert"):
        calculate_hilbert_order(16, 48, torch.zeros(1, 8, 2, device="cuda"))


@pytest.mark.gpu
@pytest.mark.parametrize("case", [(2, 2048, 64, 32), (1, 16384, 256, 128), (3, 500, 30, 40), (1, 1, 4, 4)])
def test_find_pos_org_order_matches_reference_expression(case):
    """argmin(cdist(p=1)) of maskfiner_up_down.py:150-154: a shuffled copy (exact matches), and a database with
    duplicates and missing points (ties -> first index)."""
    from affb200.point_utils import find_pos_org_order
    from oracle import point_ref as O
    B, n, w, h = case
    g = torch.Generator().manual_seed(n)
    ys, xs = torch.meshgrid(torch.arange(h), torch.arange(w), indexing="ij")
    allp = torch.stack([xs, ys], 2).reshape(-1, 2).float()
    org = torch.stack([allp[torch.randperm(h * w, generator=g)[:n]] for _ in range(B)])
    shuf = torch.stack([org[b][torch.randperm(n, generator=g)] for b in range(B)])
    out = find_pos_org_order(org.cuda(), shuf.cuda())
    assert out.dtype == torch.int64 and torch.equal(out.cpu(), O.find_pos_org_order_ref(org, shuf))
    assert torch.equal(shuf.gather(1, out.cpu().unsqueeze(2).expand(-1, -1, 2)), org)
    if n > 8:
        db = torch.cat([shuf[:, : n // 2], shuf[:, : n // 4]], 1)    # duplicates, and half of the points missing
        out = find_pos_org_order(org.cuda(), db.cuda())
        assert torch.equal(out.cpu(), O.find_pos_org_order_ref(org, db))


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_reorder_tokens_matches_indexing_and_its_backward(dtype):
    """feat[b, reorder] (aff_transformer.py:466,472) through affb200_permute_rows: forward bit-equal to stock indexing,
    backward (conflict-free scatter) bit-equal to autograd's scatter_add of the same permutation."""
    from affb200.point_utils import reorder_tokens
    g = torch.Generator().manual_seed(12)
    for (B, n, c) in ((3, 627, 64), (2, 3136, 32), (1, 5, 8), (2, 125, 256)):
        perm = torch.stack([torch.randperm(n, generator=g) for _ in range(B)]).cuda()
        x = torch.randn(B, n, c, generator=g).to(dtype).cuda().requires_grad_(True)
        y = reorder_tokens(x, perm.unsqueeze(2))
        ref = x.detach().clone().requires_grad_(True)
        yr = ref.gather(1, perm.unsqueeze(2).expand(-1, -1, c))
        assert torch.equal(y, yr)
        go = torch.randn(B, n, c, generator=g).to(dtype).cuda()
        y.backward(go)
        yr.backward(go)
        assert torch.equal(x.grad, ref.grad)
