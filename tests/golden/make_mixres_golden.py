"""Generates tests/golden/mixres_stage.pt by running the REFERENCE's own MixResNeighbour stage
(/root/reference/models/mixres_neighbour.py `BasicLayer`, imported unmodified through tests/golden/ref_stubs.py) and its
token split (`MixResNeighbour.divide_tokens_to_split_and_keep`, :731-755, called unbound) on CPU in the build
container.  Cases: a ragged off-grid token set with a padded last cluster (local attention), an on-grid set in
training mode (cached clustering), and a set small enough for global attention.  Stores the index tensors seen by the
blocks, the output positions/features and gradients.  Also checks that the host glue in
ml-autofocusformermod_b200/affb200/mixres_neighbour.py with the oracle's operator set reproduces the reference.
Run:  python tests/golden/make_mixres_golden.py"""
import os
import sys
import types

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(ROOT, "ml-autofocusformermod_b200"))
from ref_stubs import load_reference_module, cuda_ordering_rules  # noqa: E402
from oracle.aff_ops_ref import OracleOps, fill_parameters_by_name  # noqa: E402

LAYER = dict(dim=64, cluster_size=8, nbhd_size=48, depth=2, num_heads=2, mlp_ratio=2.)
GRAD_KEYS = ["blocks.0.attn.blank_k", "blocks.0.attn.pos_embed.weight", "blocks.1.attn.kv.weight",
             "blocks.1.mlp.fc1.weight", "blocks.0.norm1.weight"]


def make_case(name, seed):
    g = torch.Generator().manual_seed(seed)
    if name == "ragged":          # 2 x 901 tokens off a 40 x 64 canvas (x < 64, y < 40), mixed scales
        h, w, n, b = 40, 64, 901, 2
        pos = torch.stack([torch.randperm(h * w, generator=g)[:n] for _ in range(b)])
        on_grid = False
    elif name == "grid":          # every cell of a 24 x 32 canvas, training-mode cache path
        h, w, b = 24, 32, 2
        n = h * w
        pos = torch.arange(n).unsqueeze(0).expand(b, -1)
        on_grid = True
    else:                         # "global": 40 tokens <= nbhd 48
        h, w, n, b = 16, 16, 40, 2
        pos = torch.stack([torch.randperm(h * w, generator=g)[:n] for _ in range(b)])
        on_grid = False
    xy = torch.stack([pos % w, pos // w], dim=2).float()
    scale = torch.randint(0, 3, (b, n, 1), generator=g).float()
    feat = torch.randn(b, n, LAYER["dim"], generator=g)
    return dict(pos=torch.cat([scale, xy], 2), feat=feat, h=h, w=w, on_grid=on_grid)


def run(layer, case, capture):
    def pre(mod, args, kwargs):
        if not capture:
            a = dict(kwargs)
            for nm, v in zip(["feat", "member_idx", "cluster_mask", "pe_idx", "global_attn"], args):
                a[nm] = v
            for k in ("member_idx", "cluster_mask", "pe_idx"):
                capture[k] = None if a.get(k) is None else a[k].detach().clone()
            capture["global_attn"] = bool(a["global_attn"])
    hk = layer.blocks[0].register_forward_pre_hook(pre, with_kwargs=True)
    feat = case["feat"].clone().requires_grad_(True)
    pos, out = layer(case["pos"].clone(), feat, case["h"], case["w"], case["on_grid"])
    hk.remove()
    wgt = torch.linspace(-1, 1, out.numel()).reshape(out.shape)
    (out * wgt).sum().backward()
    grads = {k: p.grad.detach().clone() for k, p in layer.named_parameters() if k in GRAD_KEYS}
    for p in layer.parameters():
        p.grad = None
    return pos.detach(), out.detach(), feat.grad.detach(), grads


def main():
    ref_mod = load_reference_module("mixres_neighbour")
    from affb200 import mixres_neighbour as mine_mod
    err = lambda p, q: ((p - q).norm() / q.norm()).item()
    out = {"layer": LAYER, "seed": 5, "cases": {}}
    for ci, name in enumerate(["ragged", "grid", "global"]):
        case = make_case(name, 100 + ci)
        torch.manual_seed(0)
        ref = ref_mod.BasicLayer(**LAYER)
        fill_parameters_by_name(ref, seed=5)
        ref.train()
        cap_ref = {}
        with cuda_ordering_rules():
            pos_r, out_r, dx_r, grads_r = run(ref, case, cap_ref)

        mine = mine_mod.BasicLayer(ops=OracleOps(), **LAYER)
        assert [n for n, _ in mine.named_parameters()] == [n for n, _ in ref.named_parameters()], "parameter names"
        fill_parameters_by_name(mine, seed=5)
        mine.train()
        mine.compact = False     # the reference's own table indexing, for a bit-exact index comparison
        cap = {}
        pos_m, out_m, dx_m, grads_m = run(mine, case, cap)
        for k in ("member_idx", "cluster_mask", "pe_idx"):
            assert (cap_ref[k] is None) == (cap[k] is None), (name, k)
            assert cap_ref[k] is None or torch.equal(cap_ref[k], cap[k]), (name, k)
        assert cap_ref["global_attn"] == cap["global_attn"]
        assert torch.equal(pos_r, pos_m), name
        e = (err(out_m, out_r), err(dx_m, dx_r), max(err(grads_m[k], grads_r[k]) for k in grads_r))
        print("%-7s glue+oracle ops vs reference: out %.2e  dx %.2e  grads %.2e   idx %s" %
              (name, *e, None if cap_ref["member_idx"] is None else tuple(cap_ref["member_idx"].shape)))
        assert e[0] < 1e-5 and e[1] < 1e-4 and e[2] < 1e-4
        mine.compact = True      # compacted table: same numbers
        pos_c, out_c, dx_c, _ = run(mine, case, {})
        assert torch.equal(pos_c, pos_r) and err(out_c, out_r) < 1e-5 and err(dx_c, dx_r) < 1e-4
        out["cases"][name] = dict(
            case=dict(pos=case["pos"], feat=case["feat"], h=case["h"], w=case["w"], on_grid=case["on_grid"]),
            index={k: (None if cap_ref[k] is None else cap_ref[k].to(torch.int32)) for k in ("member_idx", "cluster_mask", "pe_idx")},
            global_attn=cap_ref["global_attn"], pos=pos_r, out=out_r, dx=dx_r, grads=grads_r)

    # token split (reference :731-755) on seeded scores incl. ties
    g = torch.Generator().manual_seed(9)
    B, N, C = 3, 257, 16
    scores = torch.randn(B, N, generator=g)
    scores[:, ::5] = scores[:, 1::5][:, : scores[:, ::5].shape[1]]       # ties
    feat = torch.randn(B, N, C, generator=g)
    pos = torch.randn(B, N, 3, generator=g)
    holder = types.SimpleNamespace(upscale_ratio=0.3)
    with cuda_ordering_rules():
        res = ref_mod.MixResNeighbour.divide_tokens_to_split_and_keep(holder, feat, pos, scores)
    mine_res = mine_mod.divide_tokens_to_split_and_keep(
        feat, pos, scores, 0.3, argsort=lambda s: s.sort(dim=1, stable=True)[1])
    for a, b_ in zip(res, mine_res):
        assert torch.equal(a, b_)
    out["split"] = dict(scores=scores, feat=feat, pos=pos, upscale_ratio=0.3, result=[r.clone() for r in res])
    path = os.path.join(HERE, "mixres_stage.pt")
    torch.save(out, path)
    print("wrote mixres_stage.pt (%d bytes)" % os.path.getsize(path))


if __name__ == "__main__":
    main()
