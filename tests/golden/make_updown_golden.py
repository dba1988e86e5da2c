"""Generates tests/golden/updown_small.pt by running the REFERENCE's own seven-backbone `UpDownBackbone`
(/root/reference/models/maskfiner_up_down.py with MixResViT / MixResNeighbour from models/mixres_vit.py and
models/mixres_neighbour.py, imported unmodified through tests/golden/ref_stubs.py and wired as models/build.py:86-160
wires them) on CPU in the build container, at a reduced canvas (2 x 3 x 128 x 256) and reduced widths/depths with the
small config's structure (scales 0 1 2 3 2 1 0, patch sizes 32 16 8 4 8 16 32, upscale ratio .5, keep_old_scale,
layer scale, head dim 32).  Two things are imposed on the CPU run: the `.to('cuda')` of the reference's three
hard-coded device strings is mapped to CPU, and the importance scores `torch.randn(B, N)` it draws
(maskfiner_up_down.py:147-149) are recorded so that the test can replay them.
Stores: image (uint8), the recorded scores, logits, the last stage's res2 tokens, per-stage token counts, and the
gradient of a weighted logit sum for a few parameters.  Also checks that ml-autofocusformermod_b200/affb200/
maskfiner_up_down.py with the oracle's operator set reproduces the reference.
Run:  python tests/golden/make_updown_golden.py"""
import contextlib
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(ROOT, "ml-autofocusformermod_b200"))
from ref_stubs import load_reference_module, cuda_ordering_rules  # noqa: E402
from oracle.aff_ops_ref import OracleOps, fill_parameters_by_name  # noqa: E402

CFG = dict(
    names=["MixResViT", "MixResNeighbour", "MixResNeighbour", "MixResNeighbour", "MixResNeighbour", "MixResNeighbour", "MixResViT"],
    embed_dim=[128, 64, 64, 64, 64, 64, 128], depths=[1, 1, 1, 2, 1, 2, 1], num_heads=[4, 2, 2, 2, 2, 2, 4],
    patch_sizes=[32, 16, 8, 4, 8, 16, 32], drop_path_rate=0.0, upscale_ratio=[0.0, 0.5, 0.5, 0.5, 0.0, 0.0, 0.0],
    split_ratio=4, mlp_ratio=2.0, cluster_size=8, nbhd_size=48, out_features=["res2", "res3", "res4", "res5"],
    keep_old_scale=True, n_scales=4, layer_scale=1e-5)
NUM_CLASSES, SEED, B, H, W = 10, 11, 2, 128, 256
GRAD_KEYS = ["backbones.0.layers.blocks.0.attn.qkv.weight", "backbones.1.rel_pos_emb", "backbones.1.image_patch_projection.weight",
             "backbones.2.old_feat_importance", "backbones.3.layers.blocks.1.attn.pos_embed.weight",
             "backbones.3.layers.blocks.0.attn.blank_k", "backbones.5.token_projection.weight",
             "backbones.6.layers.blocks.0.mlp.dwconv.dwconv.weight", "head.weight"]


def build_reference(c):
    """models/build.py:86-160 on the reference's classes, from the dict form of the yaml."""
    vit, nb, ud = (load_reference_module(n) for n in ("mixres_vit", "mixres_neighbour", "maskfiner_up_down"))
    n_scales, n_layers = c["n_scales"], len(c["names"])
    min_ps = c["patch_sizes"][n_scales - 1]
    bb_in_feats = [[None], ["res5"], ["res5", "res4"], ["res5", "res4", "res3"], ["res5", "res4", "res3"],
                   ["res5", "res4"], ["res5"], [None]]
    dpr = [x.item() for x in torch.linspace(0, c["drop_path_rate"], sum(c["depths"]))]
    bbs = []
    for li, name in enumerate(c["names"]):
        first = li == 0
        in_ch = 3 if first else c["embed_dim"][li - 1]
        if li >= n_scales:
            scale, patch_sizes = n_layers - li - 1, c["patch_sizes"][li:]
            out_features = c["out_features"][-(n_layers - li):]
            in_ch = c["embed_dim"][li - 1] + c["embed_dim"][n_layers - li - 1]
        else:
            scale, patch_sizes, out_features = li, c["patch_sizes"][:li + 1], c["out_features"][-(li + 1):]
        dp = dpr[sum(c["depths"][:li]):sum(c["depths"][:li + 1])]
        kw = dict(patch_sizes=patch_sizes, n_layers=c["depths"][li], d_model=c["embed_dim"][li], n_heads=c["num_heads"][li],
                  mlp_ratio=c["mlp_ratio"], dropout=0.0, drop_path_rate=dp, split_ratio=c["split_ratio"], channels=in_ch,
                  n_scales=n_scales, min_patch_size=min_ps, upscale_ratio=c["upscale_ratio"][li], out_features=out_features,
                  first_layer=first, layer_scale=c["layer_scale"])
        if name == "MixResViT":
            bbs.append(vit.MixResViT(num_register_tokens=0, **kw))
        else:
            bbs.append(nb.MixResNeighbour(attn_drop_rate=0.0, cluster_size=c["cluster_size"], nbhd_size=c["nbhd_size"],
                                          keep_old_scale=c["keep_old_scale"], scale=scale, add_image_data_to_all=False, **kw))
    return ud.UpDownBackbone(backbones=bbs, backbone_dims=c["embed_dim"], out_dim=256, all_out_features=c["out_features"],
                             n_scales=n_scales, num_classes=NUM_CLASSES, bb_in_feats=bb_in_feats, aux_loss=False)


@contextlib.contextmanager
def cpu_device_strings(recorded):
    """`.to('cuda')` -> no-op; torch.randn(B, N) draws of the forward are recorded."""
    o_to, o_randn = torch.Tensor.to, torch.randn

    def to(self, *a, **kw):
        if a and a[0] == "cuda":
            return self
        return o_to(self, *a, **kw)

    def randn(*a, **kw):
        r = o_randn(*a, **kw)
        if len(a) == 2 and all(isinstance(v, int) for v in a):
            recorded.append(r.clone())
        return r

    torch.Tensor.to, torch.randn = to, randn
    try:
        yield
    finally:
        torch.Tensor.to, torch.randn = o_to, o_randn


def run(model, im, keys=GRAD_KEYS):
    cap = {}
    hk = model.backbones[3].register_forward_hook(lambda m, a, out: cap.update(res2=out["res2"].detach().clone(),
                                                                              res2_pos=out["res2_pos"].detach().clone()))
    counts = []
    hks = [bb.layers.register_forward_pre_hook(lambda m, a, kw: counts.append(int(a[-1].shape[1])),
                                               with_kwargs=True) for bb in model.backbones]
    logits = model(im)
    hk.remove()
    [h.remove() for h in hks]
    wgt = torch.linspace(-1, 1, logits.numel()).reshape(logits.shape)
    (logits * wgt).sum().backward()
    grads = {k: p.grad.detach().clone() for k, p in model.named_parameters() if k in keys}
    for p in model.parameters():
        p.grad = None
    return logits.detach(), cap, counts, grads


def main():
    from affb200 import maskfiner_up_down as mine_mod
    err = lambda p, q: ((p - q).norm() / q.norm().clamp_min(1e-30)).item()
    g = torch.Generator().manual_seed(SEED)
    im_u8 = torch.randint(0, 256, (B, 3, H, W), generator=g, dtype=torch.uint8)
    im = im_u8.float() / 255. - 0.5

    torch.manual_seed(0)
    ref = build_reference(CFG)
    fill_parameters_by_name(ref, seed=SEED)
    ref.train()
    scores = []
    torch.manual_seed(SEED)
    with cuda_ordering_rules(), cpu_device_strings(scores):
        logits_r, cap_r, counts_r, grads_r = run(ref, im)
    assert len(scores) == 3 and set(grads_r) == set(GRAD_KEYS), (len(scores), sorted(grads_r))

    it = iter(scores)
    cpu_order = lambda a, b_: torch.cdist(a.float(), b_.float(), p=1).argmin(dim=2)
    mine = mine_mod.build_maskfiner_up_down(CFG, num_classes=NUM_CLASSES, ops=OracleOps(),
                                            argsort=lambda s: s.sort(dim=1, stable=True)[1], find_order=cpu_order,
                                            mask_fn=lambda b, n, dev: next(it).to(dev))
    names_r = [n for n, _ in ref.named_parameters()]
    names_m = [n for n, _ in mine.named_parameters()]
    assert names_m == names_r, sorted(set(names_m) ^ set(names_r))
    assert [tuple(p.shape) for p in mine.parameters()] == [tuple(p.shape) for p in ref.parameters()]
    missing, unexpected = mine.load_state_dict(ref.state_dict(), strict=False)
    assert not missing, missing      # (unexpected: none of the reference's buffers matter here)
    mine.train()
    logits_m, cap_m, counts_m, grads_m = run(mine, im)
    assert counts_m == counts_r, (counts_m, counts_r)
    assert torch.equal(cap_m["res2_pos"], cap_r["res2_pos"])
    e = (err(logits_m, logits_r), err(cap_m["res2"], cap_r["res2"]), max(err(grads_m[k], grads_r[k]) for k in grads_r))
    print("glue + oracle ops vs reference UpDownBackbone: logits %.2e  res2 %.2e  grads %.2e   tokens/stage %s" % (*e, counts_r))
    assert e[0] < 1e-5 and e[1] < 1e-5 and e[2] < 1e-4, e
    torch.save(dict(cfg=CFG, num_classes=NUM_CLASSES, seed=SEED, image_u8=im_u8, scores=scores, logits=logits_r,
                    res2=cap_r["res2"], res2_pos=cap_r["res2_pos"], counts=counts_r, grads=grads_r),
               os.path.join(HERE, "updown_small.pt"))
    print("wrote", os.path.join(HERE, "updown_small.pt"), os.path.getsize(os.path.join(HERE, "updown_small.pt")) // 1024, "KiB")


if __name__ == "__main__":
    main()
