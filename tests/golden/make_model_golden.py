"""Generates tests/golden/model_aff_small.pt by running the REFERENCE's own AutoFocusFormer
(/root/reference/models/aff_transformer.py, imported unmodified through tests/golden/ref_stubs.py) on CPU in the build
container: logits, selected gradients, and the index tensors of every stage (nbhd_idx, cluster_mask, pe_idx as seen by
the first block of each stage; token positions after every merge).  Also checks that the host glue in
ml-autofocusformermod_b200/affb200/aff_transformer.py with the oracle's operator set reproduces the reference.
Run:  python tests/golden/make_model_golden.py"""
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(ROOT, "ml-autofocusformermod_b200"))
from ref_stubs import load_reference_module, cuda_ordering_rules  # noqa: E402
from oracle.aff_ops_ref import OracleOps, fill_parameters_by_name  # noqa: E402

CFG = dict(in_chans=3, num_classes=10, embed_dim=[32, 64, 128, 160], cluster_size=8, nbhd_size=[48, 48, 48, 49],
           alpha=4.0, ds_rate=0.21, depths=[2, 2, 2, 1], num_heads=[2, 4, 8, 10], mlp_ratio=2., drop_path_rate=0.0,
           img_size=160)
GRAD_KEYS = ["patch_embed.proj1.weight", "layers.0.blocks.0.attn.blank_k", "layers.0.blocks.1.attn.pos_embed.weight",
             "layers.0.downsample.weight_net.0.weight", "layers.0.prob_net.weight", "layers.1.blocks.0.attn.kv.weight",
             "layers.1.downsample.linear.bias", "layers.2.blocks.1.mlp.fc1.weight", "layers.3.blocks.0.attn.blank_v",
             "head.weight"]


def run(model, x, capture):
    hooks = []
    for li, layer in enumerate(model.layers):
        def pre(mod, args, kwargs, li=li):
            if ("stage%d" % li) not in capture:
                a = dict(kwargs)
                names = ["feat", "member_idx", "cluster_mask", "pe_idx", "global_attn"]
                for nm, v in zip(names, args):
                    a[nm] = v
                capture["stage%d" % li] = {k: (None if a.get(k) is None else a[k].detach().clone())
                                           for k in ("member_idx", "cluster_mask", "pe_idx")}
        hooks.append(layer.blocks[0].register_forward_pre_hook(pre, with_kwargs=True))
        if layer.downsample is not None:
            def post(mod, args, kwargs, out, li=li):
                capture["merge%d_pos" % li] = out[0].detach().clone()
            hooks.append(layer.downsample.register_forward_hook(post, with_kwargs=True))
    x = x.clone().requires_grad_(True)
    logits = model(x)
    loss = (logits * torch.linspace(-1, 1, logits.numel()).reshape(logits.shape)).sum()
    loss.backward()
    for h in hooks:
        h.remove()
    grads = {k: p.grad.detach().clone() for k, p in model.named_parameters() if k in GRAD_KEYS}
    return logits.detach(), x.grad.detach(), grads


def main():
    ref_mod = load_reference_module("aff_transformer")
    torch.manual_seed(0)
    ref = ref_mod.AutoFocusFormer(**CFG)
    fill_parameters_by_name(ref, seed=7)
    ref.train()
    g = torch.Generator().manual_seed(11)
    x = torch.randn(2, 3, 160, 160, generator=g)
    cap_ref = {}
    with cuda_ordering_rules():
        logits, gx, grads = run(ref, x, cap_ref)

    from affb200.aff_transformer import AutoFocusFormer
    mine = AutoFocusFormer(ops=OracleOps(), **CFG)
    assert [n for n, _ in mine.named_parameters()] == [n for n, _ in ref.named_parameters()], "parameter names differ"
    fill_parameters_by_name(mine, seed=7)
    mine.train()
    cap = {}
    l2, gx2, grads2 = run(mine, x, cap)
    for k in cap_ref:
        a, b = cap_ref[k], cap[k]
        if isinstance(a, dict):
            for kk in a:
                assert (a[kk] is None) == (b[kk] is None), (k, kk)
                assert a[kk] is None or torch.equal(a[kk], b[kk]), (k, kk)
        else:
            assert torch.equal(a, b), k
    err = lambda p, q: ((p - q).norm() / q.norm()).item()
    print("glue+oracle ops vs reference: logits %.2e  dx %.2e  grads %.2e" %
          (err(l2, logits), err(gx2, gx), max(err(grads2[k], grads[k]) for k in grads)))
    assert err(l2, logits) < 1e-5 and err(gx2, gx) < 1e-4 and all(err(grads2[k], grads[k]) < 1e-4 for k in grads)

    out = {"cfg": CFG, "seed": 7, "x_seed": 11, "x_shape": tuple(x.shape), "x_sum": x.double().sum().item(), "logits": logits, "dx": gx, "grads": grads,
           "index": {k: ({kk: (None if vv is None else vv.to(torch.int16)) for kk, vv in v.items()}
                         if isinstance(v, dict) else v) for k, v in cap_ref.items()}}
    torch.save(out, os.path.join(HERE, "model_aff_small.pt"))
    for k, v in cap_ref.items():
        print(k, {kk: (None if vv is None else tuple(vv.shape)) for kk, vv in v.items()} if isinstance(v, dict) else tuple(v.shape))
    print("wrote model_aff_small.pt (%d bytes)" % os.path.getsize(os.path.join(HERE, "model_aff_small.pt")))


if __name__ == "__main__":
    main()
