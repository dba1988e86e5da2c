"""Probe (run on the GPU box) of the library-defined ordering rules the reference inherits from torch's CUDA kernels
(SURVEY.md §8c): the oracle freezes the rule each probe confirms.  Writes gpurun_out/torch_rules.json.

  (i)   x.sort(dim=1) on CUDA == x.sort(dim=1, stable=True) on keys with many ties (point_utils.py:159)
  (ii)  x.topk(k, sorted=False)[1] on CUDA == [indices with x > kth, ascending] ++ [indices with x == kth, ascending]
        (aff_transformer.py:320), single-block and multi-block paths
  (iii) x.mean(dim) on CUDA vs x.sum(dim) / m  for m = 8, 20, 24 (cluster centres, point_utils.py:185)
  (iv)  torch.sort(scores, descending=False) for the MixRes split selection (mixres_neighbour.py:735)
"""
import json
import os
import sys

import torch


def topk_rule(x, k):
    """Rule A, evaluated with stable sorts only."""
    B, n = x.shape
    kth = x.sort(dim=1, descending=True, stable=True)[0][:, k - 1:k]
    gt = x > kth
    eq = x == kth
    ar = torch.arange(n, device=x.device).expand(B, n)
    key = torch.where(gt, ar, torch.where(eq, ar + n, ar + 2 * n))
    return key.sort(dim=1, stable=True)[1][:, :k]


def main():
    dev = torch.device("cuda")
    g = torch.Generator(device="cpu").manual_seed(0)
    out = {"torch": torch.__version__, "gpu": torch.cuda.get_device_name(0)}
    # (i) sort stability: keys quantised to few distinct values
    res = []
    for (B, n) in [(256, 3136), (256, 627), (256, 125), (16, 9216), (1, 30720), (4, 784), (2, 2499), (1, 2048), (3, 8192)]:
        for levels in (7, 300, 100000):
            x = (torch.randint(levels, (B, n), generator=g).float() * 1.25).to(dev)
            a = x.sort(dim=1)[1]
            b = x.sort(dim=1, stable=True)[1]
            res.append({"B": B, "n": n, "levels": levels, "equal_to_stable": bool(torch.equal(a, b))})
            x3 = x.unsqueeze(2)
            a3 = x3.sort(dim=1)[1]
            res[-1]["dim1_of_3d_equal_to_stable"] = bool(torch.equal(a3.squeeze(2), b))
    out["sort"] = res
    # (ii) topk order
    res = []
    for (B, n, k) in [(256, 3136, 431), (256, 3136, 588), (256, 627, 76), (256, 784, 147), (256, 125, 9), (256, 196, 33),
                      (2, 3136, 431), (16, 9216, 1728), (16, 2304, 432), (1, 3136, 431), (64, 3136, 431), (512, 3136, 431),
                      (300, 5000, 100), (30, 12000, 1000), (4, 30000, 5000)]:
        for levels in (5, 50, 0):
            x = torch.rand(B, n, generator=g) if levels == 0 else torch.randint(levels, (B, n), generator=g).float() / levels
            x = x.to(dev)
            got = x.topk(k, dim=1, sorted=False)[1]
            want = topk_rule(x, k)
            same_set = bool(torch.equal(got.sort(dim=1)[0], want.sort(dim=1)[0]))
            res.append({"B": B, "n": n, "k": k, "levels": levels, "rule_A": bool(torch.equal(got, want)),
                        "same_set": same_set})
    out["topk"] = res
    # (iii) mean vs sum/m
    res = []
    for m in (8, 20, 24):
        x = torch.randint(0, 512, (64, 500, m, 2), generator=g).float().to(dev)
        a = x.mean(2)
        b = x.sum(2) / m
        c = x.sum(2) * (1.0 / m)
        res.append({"m": m, "mean_eq_div": bool(torch.equal(a, b)), "mean_eq_mulrecip": bool(torch.equal(a, c)),
                    "n_diff_div": int((a != b).sum().item())})
    out["mean"] = res
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open("gpurun_out/torch_rules.json", "w"), indent=1)
    bad = [r for r in out["sort"] if not (r["equal_to_stable"] and r["dim1_of_3d_equal_to_stable"])]
    badk = [r for r in out["topk"] if not r["rule_A"]]
    print("sort: %d cases, %d not equal to stable" % (len(out["sort"]), len(bad)), bad[:5])
    print("topk: %d cases, %d not rule A" % (len(out["topk"]), len(badk)), badk[:8])
    print("mean:", out["mean"])


if __name__ == "__main__":
    sys.exit(main())
