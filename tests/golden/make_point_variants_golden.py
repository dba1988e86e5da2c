"""Generates tests/golden/point_sfc_variants.pt: the other call forms of the REFERENCE's space_filling_cluster
(/root/reference/models/point_utils.py:161-201, imported unmodified) with the default curve — `use_anchor=False`
and/or `no_reorder=True` (SURVEY §8 a4') — and checks the oracle restatement against them.
Run:  python tests/golden/make_point_variants_golden.py"""
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from ref_stubs import load_reference_module, cuda_ordering_rules  # noqa: E402
from make_point_golden import grid_pos, subset  # noqa: E402
from oracle import point_ref as O  # noqa: E402


def cases():
    g = torch.Generator().manual_seed(4321)
    for nr in (False, True):
        for ua in (False, True):
            if not nr and ua:
                continue   # the models' form: tests/golden/point_sfc.pt
            tag = "nr%d_ua%d" % (nr, ua)
            yield "grid56_m8_" + tag, grid_pos(56, 56).float().unsqueeze(0).expand(2, -1, -1).contiguous(), 8, 56, 56, nr, ua
            yield "sub56_627_m8_" + tag, subset(56, 56, 627, 3, g), 8, 56, 56, nr, ua
            yield "rect64x128_2000_dup_m8_" + tag, subset(64, 128, 2000, 2, g, dup=300), 8, 64, 128, nr, ua


def main():
    pu = load_reference_module("point_utils")
    out = {}
    for name, pos, m, h, w, nr, ua in cases():
        with cuda_ordering_rules():
            ref = pu.space_filling_cluster(pos, m, h, w, no_reorder=nr, use_anchor=ua)
        mine = O.sfc_variant_ref(pos, m, h, w, no_reorder=nr, use_anchor=ua)
        assert len(ref) == len(mine)
        for a, b in zip(ref, mine):
            if a is None or b is None:
                assert a is None and b is None, name
            else:
                assert a.shape == b.shape and torch.equal(a.to(b.dtype), b), name
        out[name] = {"pos": pos.to(torch.int16), "m": m, "h": h, "w": w, "no_reorder": nr, "use_anchor": ua,
                     "out": [None if t is None else (t.to(torch.int32) if t.dtype == torch.int64 else t) for t in ref]}
        print("%-36s oracle == reference" % name)
    torch.save(out, os.path.join(HERE, "point_sfc_variants.pt"))
    print("wrote point_sfc_variants.pt (%d bytes)" % os.path.getsize(os.path.join(HERE, "point_sfc_variants.pt")))


if __name__ == "__main__":
    main()
