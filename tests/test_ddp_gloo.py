"""CPU, world_size 2, gloo: the batch-sharded data-parallel path (reference main.py:102:
DistributedDataParallel(model, broadcast_buffers=False), one process per shard).  The host glue runs with the
oracle's operator set on CPU; after backward every rank must hold the same gradients, equal to the mean of the
per-shard gradients computed without DDP (no data-path collective exists; the gradient all-reduce is the only one)."""
import os
import socket
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CFG = dict(in_chans=3, num_classes=10, embed_dim=[32, 64, 128, 160], cluster_size=8, nbhd_size=[48, 48, 48, 49],
           ds_rate=0.21, depths=[1, 1, 1, 1], num_heads=[2, 4, 8, 10], mlp_ratio=2., drop_path_rate=0.0, img_size=160)


def _model():
    for p in (ROOT, os.path.join(ROOT, "ml-autofocusformermod_b200")):
        if p not in sys.path:
            sys.path.insert(0, p)
    from affb200.aff_transformer import AutoFocusFormer
    from oracle.aff_ops_ref import OracleOps, fill_parameters_by_name
    m = AutoFocusFormer(ops=OracleOps(), **CFG)
    fill_parameters_by_name(m, seed=3)
    return m.train()


def _shard(rank):
    g = torch.Generator().manual_seed(100 + rank)  # per-rank seed, as the reference (main.py:428)
    return torch.randn(2, 3, 160, 160, generator=g), torch.randint(0, 10, (2,), generator=g)


def _grads(model, x, y):
    model.zero_grad(set_to_none=True)
    torch.nn.functional.cross_entropy(model(x), y).backward()
    return {k: p.grad.detach().clone() for k, p in model.named_parameters() if p.grad is not None}


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(2)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from torch.nn.parallel import DistributedDataParallel as DDP
    model = DDP(_model(), broadcast_buffers=False)
    x, y = _shard(rank)
    g = _grads(model, x, y)
    torch.save({k.replace("module.", "", 1): v for k, v in g.items()}, os.path.join(out_dir, "g%d.pt" % rank))
    dist.barrier()
    dist.destroy_process_group()


def test_ddp_gradients_are_the_mean_of_the_shard_gradients(tmp_path):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    g0, g1 = torch.load(tmp_path / "g0.pt"), torch.load(tmp_path / "g1.pt")
    for k in g0:
        assert torch.equal(g0[k], g1[k]), k
    torch.set_num_threads(4)
    model = _model()
    singles = [_grads(model, *_shard(r)) for r in range(2)]
    for k in g0:
        want = (singles[0][k] + singles[1][k]) / 2
        # (the conv bias in front of BatchNorm has a zero true gradient: absolute floor)
        err = (g0[k] - want).abs().max().item()
        assert err <= 1e-5 * want.abs().max().item() + 2e-6, (k, err)
