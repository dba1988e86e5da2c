"""The seven-backbone U-shaped classifier (affb200/maskfiner_up_down.py, BASELINE config 5) against the golden produced by
the REFERENCE's own `UpDownBackbone` (tests/golden/updown_small.pt, generator make_updown_golden.py; reduced canvas and
widths, the small config's structure):
  * CPU: the assembly + the oracle's operator set reproduces logits, the finest stage's tokens, token counts and
    gradients (fp32, 1e-5 / 1e-4);
  * GPU: the assembly + the CUDA operator set does (positions bit-exact; fp32 1e-4 on logits and tokens, 1e-3 on
    gradients), and under bf16 autocast with the fused attention operator (bf16 tolerance, stated below);
  * GPU: the full small config at 1024 x 2048 builds with the reference's parameter names and runs one forward."""
import os

import pytest
import torch

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "updown_small.pt")


def rel(a, b):
    return ((a.float().cpu() - b).norm() / b.norm().clamp_min(1e-30)).item()


def _build(g, dev, ops, cpu_hooks):
    from affb200.maskfiner_up_down import build_maskfiner_up_down
    from oracle.aff_ops_ref import fill_parameters_by_name
    it = iter(g["scores"])
    kw = dict(mask_fn=lambda b, n, d: next(it).to(d))
    if cpu_hooks:
        kw.update(argsort=lambda s: s.sort(dim=1, stable=True)[1],
                  find_order=lambda a, b: torch.cdist(a.float(), b.float(), p=1).argmin(dim=2))
    model = build_maskfiner_up_down(g["cfg"], num_classes=g["num_classes"], ops=ops, **kw)
    fill_parameters_by_name(model, seed=g["seed"])
    return model.to(dev).train()


def _run(model, g, dev, autocast=False):
    cap, counts = {}, []
    hk = model.backbones[3].register_forward_hook(lambda m, a, out: cap.update(res2=out["res2"].detach(),
                                                                              res2_pos=out["res2_pos"].detach()))
    hks = [bb.layers.register_forward_pre_hook(lambda m, a: counts.append(int(a[-1].shape[1]))) for bb in model.backbones]
    im = (g["image_u8"].float() / 255. - 0.5).to(dev)
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
        logits = model(im)
    hk.remove()
    [h.remove() for h in hks]
    wgt = torch.linspace(-1, 1, logits.numel(), device=dev).reshape(logits.shape)
    (logits.float() * wgt).sum().backward()
    grads = {k: p.grad.detach() for k, p in model.named_parameters() if k in g["grads"]}
    return logits.detach(), cap, counts, grads


def _check(g, res, tol, tol_grad, pos_exact=True):
    logits, cap, counts, grads = res
    assert counts == g["counts"], (counts, g["counts"])
    if pos_exact:
        assert torch.equal(cap["res2_pos"].cpu(), g["res2_pos"])
        assert rel(cap["res2"], g["res2"]) < tol, rel(cap["res2"], g["res2"])
    assert rel(logits, g["logits"]) < tol, rel(logits, g["logits"])
    for k, v in g["grads"].items():
        assert rel(grads[k], v) < tol_grad, (k, rel(grads[k], v))


def test_assembly_with_oracle_ops_reproduces_reference_updown():
    from oracle.aff_ops_ref import OracleOps
    g = torch.load(GOLD)
    _check(g, _run(_build(g, "cpu", OracleOps(), True), g, "cpu"), 1e-5, 1e-4)


def test_parameter_names_of_the_small_config():
    """names / shapes a reference checkpoint of configs/maskfiner_up_down_small.yaml carries (spot checks)"""
    from affb200.maskfiner_up_down import build_maskfiner_up_down
    from oracle.aff_ops_ref import OracleOps
    with torch.device("meta"):
        m = build_maskfiner_up_down(num_classes=1000, ops=OracleOps())
    sd = dict(m.named_parameters())
    assert sd["backbones.0.patch_embed.conv_layers.4.conv.weight"].shape == (512, 256, 3, 3)
    assert sd["backbones.1.rel_pos_emb"].shape == (1, 4, 512)
    assert sd["backbones.1.image_patch_projection.weight"].shape == (512, 3 * 16 * 16)
    assert sd["backbones.1.token_projection.weight"].shape == (256, 512)
    assert sd["backbones.3.high_res_mlp.layers.2.linear.weight"].shape == (128, 128)
    assert sd["backbones.4.token_norm.weight"].shape == (64 + 128,)
    assert sd["backbones.5.layers.blocks.23.attn.pos_embed.weight"].shape == (8, 5)
    assert sd["backbones.6.token_projection.weight"].shape == (512, 256 + 512)
    assert sd["backbones.6.layers.blocks.2.gamma2"].shape == (512,)
    assert sd["head.weight"].shape == (1000, 512)
    assert [len(b.layers.blocks) for b in m.backbones] == [2, 2, 2, 4, 6, 24, 3]
    assert m.bb_scales == [0, 1, 2, 3, 2, 1, 0]


@pytest.mark.gpu
def test_cuda_updown_matches_reference_golden_fp32():
    from affb200.aff_transformer import CudaOps
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    g = torch.load(GOLD)
    ops = CudaOps()
    ops.fused_attention = None
    _check(g, _run(_build(g, "cuda", ops, False), g, "cuda"), 1e-4, 1e-3)


@pytest.mark.gpu
def test_cuda_updown_bf16_fused_close_to_golden():
    """bf16 autocast + the fused attention operator through all five off-grid stages.  Tolerance: 3e-2 relative on the
    logits (13 blocks at ~1e-2 per bf16 operator, layer scale keeps the residual stream close); gradients 1.5e-1."""
    from affb200.aff_transformer import CudaOps
    from affb200 import fused as fa
    g = torch.load(GOLD)
    hits = []
    ops = CudaOps()
    real = ops.fused_attention or fa.fused_cluster_attention

    def counted(*a, **kw):
        out = real(*a, **kw)
        hits.append(out is not None)
        return out
    ops.fused_attention = counted
    res = _run(_build(g, "cuda", ops, False), g, "cuda", autocast=True)
    assert hits and all(hits), "fused attention declined a stage: %s" % hits
    _check(g, res, 3e-2, 1.5e-1, pos_exact=False)
    assert torch.equal(res[1]["res2_pos"].cpu(), g["res2_pos"])


@pytest.mark.gpu
def test_cuda_updown_small_config_full_canvas_forward():
    """configs/maskfiner_up_down_small.yaml at 1024 x 2048, batch 1 (BASELINE config 5): token counts per stage are
    2048 / 6144 / 14336 / 30720 / 14336 / 6144 / 2048 and the logits are finite."""
    from affb200.aff_transformer import CudaOps
    from affb200.maskfiner_up_down import build_maskfiner_up_down
    torch.manual_seed(0)
    model = build_maskfiner_up_down(num_classes=1000, ops=CudaOps()).cuda().eval()
    counts = []
    [bb.layers.register_forward_pre_hook(lambda m, a: counts.append(int(a[-1].shape[1]))) for bb in model.backbones]
    im = torch.randn(1, 3, 1024, 2048, device="cuda")
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        out = model(im)
    assert counts == [2048, 6144, 14336, 30720, 14336, 6144, 2048], counts
    assert out.shape == (1, 1000) and torch.isfinite(out.float()).all()
