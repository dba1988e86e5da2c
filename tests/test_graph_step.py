"""GPU: the CUDA-graph training step (affb200/graph_step.py) computes what the eager step computes, and a failed
index-plan speculation is reported by check()."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu


def _tiny(ops):
    from affb200.aff_transformer import AutoFocusFormer
    # the shape of the model golden (tests/golden/make_model_golden.py): four stages, masked last clusters
    return AutoFocusFormer(in_chans=3, num_classes=10, embed_dim=[32, 64, 128, 160], cluster_size=8,
                           nbhd_size=[48, 48, 48, 49], alpha=4.0, ds_rate=0.21, depths=[2, 2, 2, 1],
                           num_heads=[2, 4, 8, 10], mlp_ratio=2., drop_path_rate=0.0, img_size=160, ops=ops)


def test_graphed_step_matches_eager():
    from affb200.aff_transformer import CudaOps
    from affb200 import fused
    from affb200.graph_step import GraphedTrainStep
    from oracle.aff_ops_ref import fill_parameters_by_name
    ops = CudaOps()
    ops.fused_attention = fused.fused_cluster_attention
    torch.manual_seed(0)
    m1 = _tiny(ops)
    fill_parameters_by_name(m1, seed=1)
    m1 = m1.cuda().train()
    m2 = copy.deepcopy(m1)
    g = torch.Generator().manual_seed(5)
    xs = [torch.randn(8, 3, 160, 160, generator=g).cuda() for _ in range(6)]
    ys = [torch.randint(0, 10, (8,), generator=g).cuda() for _ in range(6)]
    lossf = torch.nn.functional.cross_entropy

    o1 = torch.optim.AdamW(m1.parameters(), lr=1e-3, fused=True, capturable=True)
    step = GraphedTrainStep(m1, o1, lossf, xs[0], ys[0], warmup=2)   # two real steps on batch 0
    o2 = torch.optim.AdamW(m2.parameters(), lr=1e-3, fused=True)

    def eager(x, y):
        o2.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = m2(x)
        loss = lossf(out.float(), y)
        loss.backward()
        o2.step()
        return loss.detach()

    for _ in range(2):
        eager(xs[0], ys[0])
    assert step.kernels_per_step > 10
    for x, y in zip(xs[1:], ys[1:]):
        l1 = step(x, y).clone()
        l2 = eager(x, y)
        assert torch.isfinite(l1)
        assert abs(float(l1) - float(l2)) < 2e-2 * max(1.0, abs(float(l2))), (float(l1), float(l2))
    step.check()
    # parameters stay together over the 5 steps (bf16 arithmetic, atomics order in the table gradient)
    worst = max(float((a - b).abs().max() / (b.abs().max() + 1e-6)) for a, b in zip(m1.parameters(), m2.parameters()))
    assert worst < 5e-2, worst


def test_guard_reports_failed_speculation():
    from affb200 import plan
    dev = torch.device("cuda", 0)
    plan.guard_tensor(dev).zero_()
    plan.guard_add(torch.tensor([1], device=dev))
    with pytest.raises(RuntimeError, match="speculated index plan"):
        plan.check_guard(dev)
    plan.check_guard(dev)  # cleared


def test_static_mode_needs_an_eager_pass_first():
    from affb200 import plan
    idx = torch.randint(0, 77, (3, 77, 5), device="cuda")
    with plan.static_plans():
        with pytest.raises(RuntimeError, match="eager pass"):
            plan.get_plan(idx).qplan(idx, 77)
