"""Generates tests/golden/point_*.pt by running the REFERENCE's own space_filling_cluster
(/root/reference/models/point_utils.py:61-208, imported unmodified) in the build container, and checks the oracle
restatement (oracle/point_ref.py) against it while doing so.  Run:  python tests/golden/make_point_golden.py

Cases (SURVEY.md §8c fixtures): the stage-1 grids of configs 1-4, the reference's test_cluster.py input shape
(h=w=100, n=2499, m=20), adaptive-downsampled subsets whose length is not a multiple of the cluster size (padded last
cluster + mask), a rectangular canvas with duplicate positions (config 5), and a tie-free float input that is run
through the unpatched reference (no ordering rule involved).
"""
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from ref_stubs import load_reference_module, cuda_ordering_rules  # noqa: E402
from oracle import point_ref as O  # noqa: E402


def grid_pos(h, w):
    ys, xs = torch.meshgrid(torch.arange(h), torch.arange(w), indexing="ij")
    return torch.stack([xs, ys], 2).reshape(-1, 2)


def subset(h, w, n, B, g, dup=0):
    out = []
    for _ in range(B):
        p = grid_pos(h, w)[torch.randperm(h * w, generator=g)[:n - dup]]
        if dup:
            p = torch.cat([p, p[torch.randint(p.shape[0], (dup,), generator=g)]])
            p = p[torch.randperm(n, generator=g)]
        out.append(p)
    return torch.stack(out).float()


def cases():
    g = torch.Generator().manual_seed(1234)
    yield "grid56_m8", grid_pos(56, 56).float().unsqueeze(0).expand(2, -1, -1).contiguous(), 8, 56, 56, True
    yield "grid96_m24", grid_pos(96, 96).float().unsqueeze(0), 24, 96, 96, True
    yield "testcluster_100_2499_m20", subset(100, 100, 2499, 2, g), 20, 100, 100, True
    yield "sub56_627_m8", subset(56, 56, 627, 3, g), 8, 56, 56, True
    yield "sub56_784_m8", subset(56, 56, 784, 2, g), 8, 56, 56, True
    yield "sub56_125_m8", subset(56, 56, 125, 4, g), 8, 56, 56, True
    yield "sub56_196_m8", subset(56, 56, 196, 2, g), 8, 56, 56, True
    yield "sub96_2304_m24", subset(96, 96, 2304, 1, g), 24, 96, 96, True
    yield "rect64x128_2000_dup_m8", subset(64, 128, 2000, 2, g, dup=300), 8, 64, 128, True
    yield "float_tiefree_500_m8", torch.rand(2, 500, 2, generator=g) * torch.tensor([39.0, 29.0]), 8, 30, 40, False


def main():
    pu = load_reference_module("point_utils")
    out = {}
    for name, pos, m, h, w, impose in cases():
        if impose:
            with cuda_ordering_rules():
                ref = pu.space_filling_cluster(pos, m, h, w, no_reorder=False)
        else:
            ref = pu.space_filling_cluster(pos, m, h, w, no_reorder=False)
            key = O.sfc_key_ref(pos, m, h, w)[0]
            assert all(key[b].unique().numel() == key.shape[1] for b in range(key.shape[0])), "not tie-free"
        mine = O.sfc_ref(pos, m, h, w)
        for a, b, what in zip(ref, mine, ("pos", "mean", "member_idx", "mask", "ranking")):
            if a is None or b is None:
                assert a is None and b is None, (name, what)
            else:
                assert a.shape == b.shape and a.dtype == b.dtype and torch.equal(a, b), (name, what)
        p_s, mean, member, mask, ranking = ref
        out[name] = {"pos": pos.to(torch.float32 if not impose else torch.int16), "m": m, "h": h, "w": w,
                     "ordering_rule_imposed": impose,
                     "ranking": ranking.squeeze(2).to(torch.int32), "mean": mean, "has_mask": mask is not None}
        print("%-28s B=%d n=%d k=%d  oracle == reference" % (name, pos.shape[0], pos.shape[1], mean.shape[1]))
    torch.save(out, os.path.join(HERE, "point_sfc.pt"))
    print("wrote point_sfc.pt (%d bytes)" % os.path.getsize(os.path.join(HERE, "point_sfc.pt")))


if __name__ == "__main__":
    main()
