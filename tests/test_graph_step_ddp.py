"""The product's own multi-GPU step (affb200/graph_step.py: flat gradient buffer, ONE all-reduce between the
forward+backward graph and the optimizer graph, rank-0 broadcast at construction, eager redo when a rank's speculated
index plans fail) at world size 2: both ranks end with identical parameters, equal to a single-process step on the
mean of the two shards' gradients; a guard failure forced on rank 1 only still issues the same number of collectives.
Two processes share cuda:0 over gloo (host-mediated reduction: nothing on the device waits for another rank); with two
GPUs (`gpurun --gpus 2`) each rank takes its own device over NCCL."""
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, out_dir, force_fail_rank):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "ml-autofocusformermod_b200"))
    import torch.distributed as dist
    from tests.test_graph_step import _tiny
    from affb200.aff_transformer import CudaOps
    from affb200 import fused_cl
    from affb200.graph_step import GraphedTrainStep
    from oracle.aff_ops_ref import fill_parameters_by_name
    ndev = torch.cuda.device_count()
    dev = torch.device("cuda", rank if ndev >= world else 0)
    torch.cuda.set_device(dev)
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("nccl" if ndev >= world else "gloo", rank=rank, world_size=world)
    torch.manual_seed(0)
    model = _tiny(CudaOps())
    fill_parameters_by_name(model, seed=1 + rank)        # ranks start differently: the constructor must broadcast rank 0's
    model = model.to(dev).train()
    g = torch.Generator().manual_seed(5)
    xs = [torch.randn(8, 3, 160, 160, generator=g) for _ in range(3)]
    ys = [torch.randint(0, 10, (8,), generator=g) for _ in range(3)]
    shard = lambda t: t[rank * 4:(rank + 1) * 4].to(dev)
    opt = torch.optim.SGD(model.parameters(), lr=0.0)
    if rank == force_fail_rank:
        real = fused_cl.observed
        fused_cl.observed = lambda skey: (lambda o: o if not (o and o.get("ok")) else dict(o, ucap=min(o["ucap"], 8)))(real(skey))
    step = GraphedTrainStep(model, opt, torch.nn.functional.cross_entropy, shard(xs[0]), shard(ys[0]), world=world, warmup=1)
    step.max_recaptures = 0
    for g_ in opt.param_groups:
        g_["lr"] = 0.05
    losses = [float(step(shard(xs[i]), shard(ys[i]))) for i in (1, 2)]
    torch.save({"params": step.flat_param.cpu(), "fallbacks": step.fallbacks, "losses": losses,
                "names": [n for n, _ in model.named_parameters()]}, os.path.join(out_dir, "rank%d.pt" % rank))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("force_fail_rank", [-1, 1])
def test_graphed_step_world_2(tmp_path, force_fail_rank):
    import torch.multiprocessing as mp
    port = 29500 + (os.getpid() % 2000) + (7 if force_fail_rank >= 0 else 0)
    mp.spawn(_worker, args=(2, port, str(tmp_path), force_fail_rank), nprocs=2, join=True)
    r0, r1 = torch.load(tmp_path / "rank0.pt"), torch.load(tmp_path / "rank1.pt")
    assert torch.equal(r0["params"], r1["params"]), "ranks diverged"
    if force_fail_rank == 1:      # (a rank may also fall back on its own: its shard's neighbourhoods are data-dependent)
        assert r1["fallbacks"] == 2
    # single process: the same two steps on the full batches with rank 0's initial weights; SGD on the mean gradient
    sys.path.insert(0, ROOT)
    from tests.test_graph_step import _tiny
    from affb200.aff_transformer import CudaOps
    from oracle.aff_ops_ref import fill_parameters_by_name
    torch.manual_seed(0)
    model = _tiny(CudaOps())
    fill_parameters_by_name(model, seed=1)
    model = model.cuda().train()
    g = torch.Generator().manual_seed(5)
    xs = [torch.randn(8, 3, 160, 160, generator=g) for _ in range(3)]
    ys = [torch.randint(0, 10, (8,), generator=g) for _ in range(3)]
    opt = torch.optim.SGD(model.parameters(), lr=0.0)

    def one(x, y, lr):
        for g_ in opt.param_groups:
            g_["lr"] = lr
        opt.zero_grad(set_to_none=True)
        grads = None
        for r in range(2):   # per-shard forward (BatchNorm statistics and clustering are per rank, main.py:102)
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out = model(x[r * 4:(r + 1) * 4].cuda())
            torch.nn.functional.cross_entropy(out.float(), y[r * 4:(r + 1) * 4].cuda()).backward()
        for p in model.parameters():
            if p.grad is not None:
                p.grad.mul_(0.5)
        opt.step()
    bn = {k: v.clone() for k, v in model.named_buffers()}
    one(xs[0], ys[0], 0.0)     # the constructor's warm-up step (lr 0)
    one(xs[1], ys[1], 0.05)
    one(xs[2], ys[2], 0.05)
    ref = torch.cat([torch.nn.functional.pad(p.detach().reshape(-1), (0, (-p.numel()) % 8)) for p in model.parameters()]).cpu()
    rel = float((r0["params"] - ref).norm() / ref.norm())
    assert rel < 2e-3, rel
