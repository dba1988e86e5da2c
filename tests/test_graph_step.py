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


def _models():
    from affb200.aff_transformer import CudaOps
    from affb200 import fused
    from oracle.aff_ops_ref import fill_parameters_by_name
    ops = CudaOps()
    ops.fused_attention = fused.fused_cluster_attention
    torch.manual_seed(0)
    m1 = _tiny(ops)
    fill_parameters_by_name(m1, seed=1)
    m1 = m1.cuda().train()
    return m1, copy.deepcopy(m1)


def _batches(n):
    g = torch.Generator().manual_seed(5)
    xs = [torch.randn(8, 3, 160, 160, generator=g).cuda() for _ in range(n)]
    ys = [torch.randint(0, 10, (8,), generator=g).cuda() for _ in range(n)]
    return xs, ys


def _eager_step(model, opt, x, y):
    opt.zero_grad(set_to_none=True)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        out = model(x)
    loss = torch.nn.functional.cross_entropy(out.float(), y)
    loss.backward()
    opt.step()
    return loss.detach()


def test_graphed_step_matches_eager_loss_and_gradients():
    """With a zero learning rate both models keep identical weights, so every replay must reproduce the eager
    step's loss and gradients on a batch the capture never saw (new token selections, new index plans)."""
    from affb200.graph_step import GraphedTrainStep
    m1, m2 = _models()
    for m in (m1, m2):   # BatchNorm statistics: no momentum so that the two models see the same statistics path
        for mod in m.modules():
            if isinstance(mod, torch.nn.BatchNorm2d):
                mod.momentum = 0.0
    xs, ys = _batches(5)
    o1 = torch.optim.AdamW(m1.parameters(), lr=0.0, weight_decay=0.0, fused=True, capturable=True)
    o2 = torch.optim.AdamW(m2.parameters(), lr=0.0, weight_decay=0.0, fused=True)
    step = GraphedTrainStep(m1, o1, torch.nn.functional.cross_entropy, xs[0], ys[0], warmup=2)
    assert step.kernels_per_step > 10
    for x, y in zip(xs[1:], ys[1:]):
        l1 = float(step(x, y))
        l2 = float(_eager_step(m2, o2, x, y))
        assert abs(l1 - l2) < 2e-3 * max(1.0, abs(l2)), (l1, l2)
        g1 = torch.cat([p.grad.reshape(-1) for p in m1.parameters() if p.requires_grad])   # views of step.flat_grad
        g2 = torch.cat([p.grad.reshape(-1) for p in m2.parameters() if p.requires_grad])
        rel = float((g1 - g2).norm() / g2.norm())
        assert rel < 2e-2, rel
    step.check()
    for a, b in zip(m1.parameters(), m2.parameters()):
        assert torch.equal(a, b)
        assert a.grad.data_ptr() >= step.flat_grad.data_ptr() and a.data_ptr() >= step.flat_param.data_ptr()
        assert torch.equal(a._affb200_lp, a.detach().bfloat16())     # the shadow the Linears read


def test_linear_gradient_sink_accumulates_in_place():
    """`_affb200_sink`: the weight-gradient GEMM and the bias column sums add into the existing fp32 .grad (fp32 output
    from bf16 operands) and autograd gets None; the result matches the stock Linear backward."""
    from affb200.linear_ops import linear
    torch.manual_seed(5)
    lin = torch.nn.Linear(64, 128).cuda()
    ref = torch.nn.Linear(64, 128).cuda()
    ref.load_state_dict(lin.state_dict())
    x = torch.randn(300, 64, device="cuda")
    go = torch.randn(300, 128, device="cuda").bfloat16()
    for p in lin.parameters():
        p.grad = torch.full_like(p, 0.25)
        p._affb200_sink = True
        p._affb200_lp = p.detach().bfloat16()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        linear(x, lin.weight, lin.bias).backward(go)
        ref(x).backward(go)
    for a, b in zip(lin.parameters(), ref.parameters()):
        assert a.grad.dtype == torch.float32
        err = float((a.grad - 0.25 - b.grad).abs().max() / b.grad.abs().max())
        assert err < 1e-2, err


def test_graphed_step_trains():
    """A few real optimisation steps through the graphs reduce the loss on a fixed batch."""
    from affb200.graph_step import GraphedTrainStep
    m1, _ = _models()
    xs, ys = _batches(1)
    opt = torch.optim.AdamW(m1.parameters(), lr=2e-3, fused=True, capturable=True)
    step = GraphedTrainStep(m1, opt, torch.nn.functional.cross_entropy, xs[0], ys[0], warmup=1)
    first = float(step(xs[0], ys[0]))
    for _ in range(10):
        last = float(step(xs[0], ys[0]))
    step.check()
    assert last < first, (first, last)


def test_failed_speculation_is_redone_eagerly(monkeypatch):
    """Capture with a row bound far below what the batches need (every stage-2 tile breaks it): each replay trips the
    device guard, the step is redone eagerly before the optimizer runs, and losses / gradients still match eager."""
    from affb200 import plan
    from affb200.graph_step import GraphedTrainStep
    m1, m2 = _models()
    for m in (m1, m2):
        for mod in m.modules():
            if isinstance(mod, torch.nn.BatchNorm2d):
                mod.momentum = 0.0
    xs, ys = _batches(4)
    o1 = torch.optim.AdamW(m1.parameters(), lr=0.0, weight_decay=0.0, fused=True, capturable=True)
    o2 = torch.optim.AdamW(m2.parameters(), lr=0.0, weight_decay=0.0, fused=True)
    real_cap = plan.observed_umax_cap
    monkeypatch.setattr(plan, "observed_umax_cap", lambda kind, key: min(real_cap(kind, key) or 31, 31))
    # the block-structured kernels size their launches from the recorded largest union of 8-row blocks: cap it at 8
    from affb200 import fused_cl
    real_obs = fused_cl.observed
    monkeypatch.setattr(fused_cl, "observed",
                        lambda skey: (lambda o: o if not (o and o.get("ok")) else dict(o, ucap=min(o["ucap"], 8)))(real_obs(skey)))
    step = GraphedTrainStep(m1, o1, torch.nn.functional.cross_entropy, xs[0], ys[0], warmup=1)
    step.max_recaptures = 1
    for i, (x, y) in enumerate(zip(xs[1:], ys[1:])):
        l1 = float(step(x, y))
        l2 = float(_eager_step(m2, o2, x, y))
        assert step.fallbacks == i + 1
        assert abs(l1 - l2) < 2e-3 * max(1.0, abs(l2)), (l1, l2)
        g1 = torch.cat([p.grad.reshape(-1) for p in m1.parameters() if p.requires_grad])
        g2 = torch.cat([p.grad.reshape(-1) for p in m2.parameters() if p.requires_grad])
        assert float((g1 - g2).norm() / g2.norm()) < 2e-2
        assert all(bool(torch.isfinite(p).all()) for p in m1.parameters())
    assert step.recaptures == 1
    monkeypatch.setattr(plan, "observed_umax_cap", real_cap)
    monkeypatch.setattr(fused_cl, "observed", real_obs)
    step.max_recaptures = 2
    step(xs[0], ys[0])                    # falls back once more, re-captures with the real bounds ...
    n = step.fallbacks
    step(xs[1], ys[1])                    # ... and replays cleanly from then on
    assert step.fallbacks == n


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


def test_graphed_forward_matches_eager_inference():
    """GraphedForward (the reference's throughput() loop as one CUDA graph): replays on batches the capture never saw
    reproduce the eager eval-mode forward; token selections differ per batch, so the speculated plans are exercised."""
    from affb200.graph_step import GraphedForward
    m1, _ = _models()
    m1.eval()
    xs, _ = _batches(4)
    fwd = GraphedForward(m1, xs[0])
    for x in xs[1:]:
        got = fwd(x).float().clone()
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
            want = m1(x).float()
        assert torch.isfinite(got).all()
        # bit-identical kernels on identical inputs; a fallback (eager redo) would also be exact
        assert torch.equal(got, want), (got - want).abs().max()
    assert fwd.kernels_per_step > 0


def test_flat_optimizer_performs_the_per_tensor_update():
    """AdamW over one flat tensor per param group (the default of GraphedTrainStep) and AdamW over the individual
    parameter tensors: the same element-wise update.  On a deterministic toy network (no atomics anywhere) the
    parameters stay bit-identical over several real steps with two param groups (weights decayed, biases / norms not,
    as the reference's build_optimizer does); on the AFF model the two trajectories agree to gradient noise."""
    from affb200.graph_step import GraphedTrainStep
    xs, ys = _batches(5)

    def opt_for(m, lr):
        decay = [p for p in m.parameters() if p.dim() > 1]
        rest = [p for p in m.parameters() if p.dim() <= 1]
        return torch.optim.AdamW([{"params": decay, "weight_decay": 0.05}, {"params": rest, "weight_decay": 0.0}],
                                 lr=lr, fused=True, capturable=True)

    def toy():
        torch.manual_seed(3)
        return torch.nn.Sequential(torch.nn.Flatten(), torch.nn.Linear(3 * 160 * 160, 64), torch.nn.LayerNorm(64),
                                   torch.nn.GELU(), torch.nn.Linear(64, 10)).cuda()
    t1, t2 = toy(), toy()
    s1 = GraphedTrainStep(t1, opt_for(t1, 1e-3), torch.nn.functional.cross_entropy, xs[0], ys[0], warmup=1)
    s2 = GraphedTrainStep(t2, opt_for(t2, 1e-3), torch.nn.functional.cross_entropy, xs[0], ys[0], warmup=1,
                          flat_optimizer=False)
    assert [len(g["params"]) for g in s1.opt.param_groups] == [1, 1]
    assert [len(g["params"]) for g in s2.opt.param_groups] == [2, 4]
    for x, y in zip(xs[1:], ys[1:]):
        assert float(s1(x, y)) == float(s2(x, y))
    fresh = toy()
    for (k, a), b, c in zip(t1.named_parameters(), t2.parameters(), fresh.parameters()):
        assert torch.equal(a, b), k
        assert torch.equal(a._affb200_lp, b._affb200_lp), k
        assert not torch.equal(a, c), k          # the steps were real

    m1, m2 = _models()
    s1 = GraphedTrainStep(m1, opt_for(m1, 1e-5), torch.nn.functional.cross_entropy, xs[0], ys[0], warmup=1)
    s2 = GraphedTrainStep(m2, opt_for(m2, 1e-5), torch.nn.functional.cross_entropy, xs[0], ys[0], warmup=1,
                          flat_optimizer=False)
    assert [len(g["params"]) for g in s1.opt.param_groups] == [1, 1]
    for x, y in zip(xs[1:], ys[1:]):
        l1, l2 = float(s1(x, y)), float(s2(x, y))
        assert abs(l1 - l2) < 2e-2 * abs(l2), (l1, l2)


def test_caller_optimizer_hyperparameters_clip_and_state_dict():
    """The caller's optimizer stays authoritative: a learning rate changed on it (what an LR scheduler does every
    iteration, reference main.py:262-267) reaches the captured optimizer step; clip_grad_norm bounds the flat gradient
    between the two graphs (utils.py:186); the stepping optimizer's state can be saved and restored (utils.py:55-77)."""
    from affb200.graph_step import GraphedTrainStep
    m1, m2 = _models()
    xs, ys = _batches(4)
    o1 = torch.optim.AdamW(m1.parameters(), lr=0.0, weight_decay=0.0, fused=True, capturable=True)
    o2 = torch.optim.AdamW(m2.parameters(), lr=0.0, weight_decay=0.0, fused=True)
    step = GraphedTrainStep(m1, o1, torch.nn.functional.cross_entropy, xs[0], ys[0], warmup=1, clip_grad_norm=0.05)
    before = step.flat_param.clone()
    step(xs[1], ys[1])
    assert torch.equal(step.flat_param, before)                       # lr = 0: nothing moves
    assert float(step.flat_grad.norm()) <= 0.05 * 1.001               # clipped in place before the step
    for o in (o1, o2):
        for g in o.param_groups:
            g["lr"] = 1e-3                                            # "scheduler": the caller's optimizer only
    sd = step.optimizer_state_dict()
    n_steps = float(list(sd["state"].values())[0]["step"])
    step(xs[2], ys[2])
    _eager_step(m2, o2, xs[2], ys[2])                                  # (eager: unclipped, Adam normalises the scale)
    moved = float((step.flat_param - before).abs().max())
    assert 5e-4 < moved < 2e-3, moved                                 # one Adam step of size lr
    assert float(list(step.optimizer_state_dict()["state"].values())[0]["step"]) == n_steps + 1
    step.load_optimizer_state_dict(sd)                                # back to the saved moment estimates
    assert float(list(step.optimizer_state_dict()["state"].values())[0]["step"]) == n_steps
    step(xs[3], ys[3])
    assert bool(torch.isfinite(step.flat_param).all())


def test_redone_step_does_not_update_batchnorm_statistics_twice(monkeypatch):
    """A replay whose speculation failed has already moved the BatchNorm running statistics; the eager redo must start
    from the values before the replay (one momentum update per step, as an eager step makes)."""
    from affb200 import fused_cl
    from affb200.graph_step import GraphedTrainStep
    m1, m2 = _models()
    xs, ys = _batches(3)
    o1 = torch.optim.AdamW(m1.parameters(), lr=0.0, weight_decay=0.0, fused=True, capturable=True)
    o2 = torch.optim.AdamW(m2.parameters(), lr=0.0, weight_decay=0.0, fused=True)
    real_obs = fused_cl.observed
    monkeypatch.setattr(fused_cl, "observed",
                        lambda skey: (lambda o: o if not (o and o.get("ok")) else dict(o, ucap=min(o["ucap"], 8)))(real_obs(skey)))
    step = GraphedTrainStep(m1, o1, torch.nn.functional.cross_entropy, xs[0], ys[0], warmup=1)
    step.max_recaptures = 0
    _eager_step(m2, o2, xs[0], ys[0])          # the warm-up pass of the graphed model
    for x, y in zip(xs[1:], ys[1:]):
        step(x, y)
        _eager_step(m2, o2, x, y)
    assert step.fallbacks == 2
    b1 = {k: v for k, v in m1.named_buffers()}
    for k, v in m2.named_buffers():
        assert torch.allclose(b1[k].float(), v.float(), rtol=2e-3, atol=1e-5), k
