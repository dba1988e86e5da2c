"""Generates tests/golden/point_sfc_curves.pt: the REFERENCE's calculate_peano_order / calculate_hilbert_order and
space_filling_cluster(sf_type='peano'|'hilbert') (/root/reference/models/point_utils.py:98-122,162-165,211-349,
imported unmodified, `sort` made stable as torch-CUDA sorts are) on seeded token sets (SURVEY §8 a4'), and checks the
oracle restatement (oracle/point_ref.py::curve_key_ref, sfc_curve_ref) against them.
Run:  python tests/golden/make_point_curves_golden.py"""
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from ref_stubs import load_reference_module, cuda_ordering_rules  # noqa: E402
from make_point_golden import grid_pos, subset  # noqa: E402
from oracle import point_ref as O  # noqa: E402


def order_cases():
    g = torch.Generator().manual_seed(77)
    yield "peano_27x27_grid", "peano", 27, 27, grid_pos(27, 27).float().unsqueeze(0)
    yield "peano_81x81_sub", "peano", 81, 81, subset(81, 81, 1500, 2, g)
    yield "peano_9x18_wide", "peano", 9, 18, grid_pos(9, 18).float().unsqueeze(0)
    yield "peano_27x45_wide_sub", "peano", 27, 45, subset(27, 45, 700, 2, g)
    yield "peano_20x20_nonpow", "peano", 20, 20, subset(20, 20, 300, 2, g)
    yield "hilbert_32x32_grid", "hilbert", 32, 32, grid_pos(32, 32).float().unsqueeze(0)
    yield "hilbert_64x64_sub", "hilbert", 64, 64, subset(64, 64, 1200, 2, g)
    yield "hilbert_56x56_nonpow", "hilbert", 56, 56, subset(56, 56, 627, 2, g)
    yield "hilbert_16x16_frac", "hilbert", 16, 16, torch.rand(2, 400, 2, generator=g) * 15.99


def cluster_cases():
    g = torch.Generator().manual_seed(78)
    for sf in ("peano", "hilbert"):
        for ua in (True, False):
            for nr in (False, True):
                tag = "%s_ua%d_nr%d" % (sf, ua, nr)
                yield "grid56_m8_" + tag, grid_pos(56, 56).float().unsqueeze(0).expand(2, -1, -1).contiguous(), 8, 56, 56, sf, nr, ua
                yield "sub56_627_m8_" + tag, subset(56, 56, 627, 3, g), 8, 56, 56, sf, nr, ua
        yield "sub64_2000_m20_%s" % sf, subset(64, 64, 2000, 2, g), 20, 64, 64, sf, False, True
    yield "peano_rect27x54_900_m8", subset(27, 54, 900, 2, g), 8, 27, 54, "peano", False, True


def main():
    pu = load_reference_module("point_utils")
    out = {"orders": {}, "clusters": {}}
    for name, sf, h, w, pos in order_cases():
        fn = pu.calculate_peano_order if sf == "peano" else pu.calculate_hilbert_order
        with cuda_ordering_rules():
            rank, order = fn(h, w, pos)
        r2, o2 = O.curve_order_ref(h, w, pos, sf)
        assert torch.equal(rank, r2) and torch.equal(order, o2), name
        out["orders"][name] = dict(curve=sf, h=h, w=w, pos=pos, rank=rank.to(torch.int32), order=order.to(torch.int32))
        print("%-28s oracle == reference" % name)
    for name, pos, m, h, w, sf, nr, ua in cluster_cases():
        with cuda_ordering_rules():
            ref = pu.space_filling_cluster(pos, m, h, w, no_reorder=nr, sf_type=sf, use_anchor=ua)
        mine = O.sfc_curve_ref(pos, m, h, w, sf, no_reorder=nr, use_anchor=ua)
        assert len(ref) == len(mine)
        for a, b in zip(ref, mine):
            if a is None or b is None:
                assert a is None and b is None, name
            else:
                assert a.shape == b.shape and torch.equal(a.to(b.dtype), b), name
        out["clusters"][name] = {"pos": pos.to(torch.int16), "m": m, "h": h, "w": w, "sf_type": sf, "no_reorder": nr,
                                 "use_anchor": ua,
                                 "out": [None if t is None else (t.to(torch.int32) if t.dtype == torch.int64 else t)
                                         for t in ref]}
        print("%-36s oracle == reference" % name)
    path = os.path.join(HERE, "point_sfc_curves.pt")
    torch.save(out, path)
    print("wrote point_sfc_curves.pt (%d bytes)" % os.path.getsize(path))


if __name__ == "__main__":
    main()
