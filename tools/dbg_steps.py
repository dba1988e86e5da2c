import os, sys, torch
sys.path.insert(0, "."); sys.path.insert(0, "ml-autofocusformermod_b200")
import bench_workloads as W
dev = torch.device("cuda", 0)
wl = W.get("")
run = wl.build(wl.default_batch, dev, 1)
batch = run.to_device(run.make_host_batch(0))
for i in range(8):
    l = run.step(batch)
    torch.cuda.synchronize()
    bad = sum(int(not torch.isfinite(p).all()) for p in run.model.parameters())
    print(i, float(l), "nonfinite params", bad, flush=True)
    if run.gstep is not None:
        run.gstep.check()
