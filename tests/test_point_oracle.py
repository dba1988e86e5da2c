"""CPU: the index-builder oracle (oracle/point_ref.py) reproduces the golden vectors produced by the reference's own
space_filling_cluster (tests/golden/point_sfc.pt, generator tests/golden/make_point_golden.py), and its other
restatements satisfy the properties the reference code implies."""
import os

import torch

from oracle import point_ref as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_sfc_oracle_matches_reference_golden():
    gold = torch.load(os.path.join(GOLD, "point_sfc.pt"))
    assert len(gold) >= 10
    for name, c in gold.items():
        pos = c["pos"].float()
        p_s, mean, member, mask, ranking = O.sfc_ref(pos, c["m"], c["h"], c["w"])
        assert torch.equal(ranking.squeeze(2).to(torch.int32), c["ranking"]), name
        assert torch.equal(mean, c["mean"]), name
        assert (mask is not None) == c["has_mask"], name
        n = pos.shape[1]
        # a permutation, and pos_sorted is the gather of it
        assert torch.equal(ranking.squeeze(2).sort(dim=1)[0], torch.arange(n).expand(pos.shape[0], n)), name
        assert torch.equal(p_s, pos.gather(1, ranking.expand(-1, -1, 2))), name


def test_knn_oracle_properties():
    g = torch.Generator().manual_seed(3)
    q = torch.randint(0, 20, (2, 50, 2), generator=g).float()
    x = torch.randint(0, 20, (2, 70, 2), generator=g).float()
    idx, d = O.knn_ref(q, x, 6, return_dist=True)
    full = ((q[:, :, None] - x[:, None]) ** 2).sum(-1).sqrt()
    assert torch.equal(d, full.gather(2, idx))
    assert (d[..., 1:] >= d[..., :-1]).all()
    # ties: equal distances come out in ascending database index
    same = d[..., 1:] == d[..., :-1]
    assert (idx[..., 1:][same] > idx[..., :-1][same]).all()
    # nothing outside the list is closer than the last listed neighbour, and equal ones have a higher index
    kth_d, kth_i = d[..., -1:], idx[..., -1:]
    listed = torch.zeros_like(full, dtype=torch.bool).scatter_(2, idx, True)
    ar = torch.arange(70).view(1, 1, -1)
    assert not ((full < kth_d) & ~listed).any()
    assert not ((full == kth_d) & (ar < kth_i) & ~listed).any()


def test_topk_order_rule_and_selection():
    g = torch.Generator().manual_seed(4)
    x = torch.randint(0, 6, (3, 40), generator=g).float()
    got = O.topk_unsorted_cuda_order(x, 11)
    for b in range(3):
        kth = x[b].sort(descending=True)[0][10]
        want = [i for i in range(40) if x[b, i] > kth] + [i for i in range(40) if x[b, i] == kth]
        assert got[b].tolist() == want[:11]
    # selection on a stride-2 grid: reserved tokens (multiples of 4) come last, ascending; sampled ones exclude them
    ys, xs = torch.meshgrid(torch.arange(8), torch.arange(8), indexing="ij")
    pos = torch.stack([xs, ys], 2).reshape(1, 64, 2).float()
    lp = torch.rand(1, 64, 1, generator=g)
    idx, final = O.select_ref(pos, lp, 2, 4.0, 0.25, 4)
    assert idx.shape == (1, 16)
    res = idx[0, -4:].tolist()
    assert res == sorted(res) and all(int(pos[0, i, 0]) % 4 == 0 and int(pos[0, i, 1]) % 4 == 0 for i in res)
    assert len(set(idx[0].tolist())) == 16


def test_sfc_variant_oracle_matches_reference_golden():
    """use_anchor=False / no_reorder=True call forms (SURVEY §8 a4'): the oracle restatement against outputs of the
    reference's own space_filling_cluster (tests/golden/make_point_variants_golden.py)."""
    import os
    import torch
    from oracle import point_ref as O
    g = torch.load(os.path.join(os.path.dirname(__file__), "golden", "point_sfc_variants.pt"))
    assert len(g) == 9
    for name, c in g.items():
        mine = O.sfc_variant_ref(c["pos"].float(), c["m"], c["h"], c["w"], no_reorder=c["no_reorder"],
                                 use_anchor=c["use_anchor"])
        assert len(mine) == len(c["out"]) == (3 if c["no_reorder"] else 5)
        for a, b in zip(mine, c["out"]):
            if b is None:
                assert a is None, name
            else:
                assert torch.equal(a.to(b.dtype), b), name


def test_curve_oracle_matches_reference_golden():
    """Peano / Hilbert restatement (oracle/point_ref.py::curve_order_ref, sfc_curve_ref) against outputs of the
    reference's own calculate_*_order and space_filling_cluster(sf_type=...) (tests/golden/make_point_curves_golden.py)."""
    import os
    import torch
    from oracle import point_ref as O
    g = torch.load(os.path.join(os.path.dirname(__file__), "golden", "point_sfc_curves.pt"))
    assert len(g["orders"]) == 9 and len(g["clusters"]) == 19
    for name, c in g["orders"].items():
        rank, order = O.curve_order_ref(c["h"], c["w"], c["pos"], c["curve"])
        assert torch.equal(rank.to(torch.int32), c["rank"]) and torch.equal(order.to(torch.int32), c["order"]), name
    for name, c in g["clusters"].items():
        mine = O.sfc_curve_ref(c["pos"].float(), c["m"], c["h"], c["w"], c["sf_type"], no_reorder=c["no_reorder"],
                               use_anchor=c["use_anchor"])
        assert len(mine) == len(c["out"])
        for a, b in zip(mine, c["out"]):
            if b is None:
                assert a is None, name
            else:
                assert torch.equal(a.to(b.dtype), b), name
