"""Dev probe: NCCL all-reduce captured inside a CUDA graph — on the capture stream, on a forked side stream, and issued
from an autograd backward hook on a side stream.  torchrun --nproc-per-node 2 tools/dev_nccl_graph.py"""
import os
import torch
import torch.distributed as dist


def main():
    rank = int(os.environ["RANK"])
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl")
    buf = torch.ones(1 << 20, device=dev) * (rank + 1)
    dist.all_reduce(buf)           # communicator up before capture
    torch.cuda.synchronize()
    side = torch.cuda.Stream(device=dev)

    def variant(name, fn):
        x = torch.ones(1 << 20, device=dev) * (rank + 1)
        work = torch.zeros(1 << 20, device=dev)
        s = torch.cuda.Stream(device=dev)
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            fn(x, work)            # eager warm-up
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        x.fill_(rank + 1)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn(x, work)
        x.fill_(rank + 1)
        g.replay()
        torch.cuda.synchronize()
        print("rank %d %-28s x[0] = %.3f (expect 1.5)  work[0] = %.1f" % (rank, name, x[0].item(), work[0].item()), flush=True)

    def on_main(x, work):
        work.add_(1)
        dist.all_reduce(x, op=dist.ReduceOp.AVG)

    def on_side(x, work):
        cur = torch.cuda.current_stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            dist.all_reduce(x, op=dist.ReduceOp.AVG)
        work.add_(1)
        cur.wait_stream(side)

    lin1, lin2 = torch.nn.Linear(64, 64).to(dev), torch.nn.Linear(64, 64).to(dev)
    state = {}

    def hook(mod, gi, go):
        cur = torch.cuda.current_stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            dist.all_reduce(state["x"], op=dist.ReduceOp.AVG)
    lin2.register_full_backward_hook(hook)

    def in_hook(x, work):
        state["x"] = x
        inp = torch.ones(8, 64, device=dev, requires_grad=True)
        out = lin2(lin1(inp)).sum()
        out.backward()
        work.add_(1)
        torch.cuda.current_stream().wait_stream(side)

    variant("capture stream", on_main)
    variant("forked side stream", on_side)
    variant("backward hook, side stream", in_hook)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
