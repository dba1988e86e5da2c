"""Dev harness: the config-5 model (maskfiner_up_down_small, 1 x 3 x 1024 x 2048) on one GPU — eager forward time per
backbone, which attention kernels each stage ran, the graph-replayed forward, and the stand-alone cost of the index
build pieces at the three token counts.  Run on the GPU box:  python tools/dev_updown.py"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ml-autofocusformermod_b200"))


def ev_time(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def main():
    from affb200.aff_transformer import CudaOps
    from affb200.maskfiner_up_down import build_maskfiner_up_down
    from affb200 import fused, point_utils as PU
    torch.backends.cudnn.benchmark = True
    torch.manual_seed(0)
    ops = CudaOps()
    model = build_maskfiner_up_down(num_classes=1000, ops=ops).cuda().eval()
    im = torch.randn(1, 3, 1024, 2048, device="cuda")

    def fwd():
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
            return model(im)

    if "--profile-one" in sys.argv:      # under ncu --profile-from-start off: one eager forward inside the range
        for _ in range(3):
            fwd()
        torch.cuda.synchronize()
        torch.cuda.cudart().cudaProfilerStart()
        fwd()
        torch.cuda.synchronize()
        torch.cuda.cudart().cudaProfilerStop()
        return
    fused.trace = []
    fwd()
    calls, fused.trace = fused.trace, None
    stages = []
    for n, H, fam in calls:
        if not stages or stages[-1][0] != n:
            stages.append([n, H, 0, fam or "unfused"])
        stages[-1][2] += 1
    print("attention per stage (tokens, heads, blocks, kernels):", stages)

    # how often a stage's tile unions exceed the 64-block limit (importance scores are redrawn every forward)
    from affb200 import plan as _plan
    fam = {}
    for _ in range(20):
        fused.trace = []
        fwd()
        calls, fused.trace = fused.trace, None
        for n, H, f in calls:
            fam.setdefault(n, []).append(f or "unfused")
    print("kernel family over 20 forwards:", {n: {f: v.count(f) for f in set(v)} for n, v in fam.items()})
    print("observed:", {k: v for k, v in _plan._observed.items() if k[0] == "cl"})

    # per-backbone eager time
    marks = []
    hooks = []
    for j, bb in enumerate(model.backbones):
        def pre(m, a, j=j):
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            marks.append(("pre", j, e))

        def post(m, a, o, j=j):
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            marks.append(("post", j, e))
        hooks += [bb.register_forward_pre_hook(pre), bb.register_forward_hook(post)]
    for _ in range(3):
        marks.clear()
        t0 = time.perf_counter()
        fwd()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
    pres = {j: e for k, j, e in marks if k == "pre"}
    posts = {j: e for k, j, e in marks if k == "post"}
    per = [pres[j].elapsed_time(posts[j]) for j in range(len(model.backbones))]
    print("eager forward wall %.1f ms; per backbone ms: %s  sum %.1f" % (wall * 1e3, ["%.2f" % v for v in per], sum(per)))
    for h in hooks:
        h.remove()
    print("eager forward (events, 5 iters): %.2f ms" % ev_time(fwd))

    # graph replay
    try:
        from affb200.graph_step import GraphedForward
        from affb200 import plan as _plan
        from affb200 import fused_cl
        fused_cl.debug_static = []
        g = GraphedForward(model, im)
        print("observed after warm-up:", {k: v for k, v in _plan._observed.items() if k[0] == "cl"})
        ts = []
        for _ in range(24):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fb = g.fallbacks
            g(im)
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
            if g.fallbacks != fb:
                print("  fallback: static plan stats", [(k, u, r, st.tolist(), int(bd)) for k, u, r, st, bd in fused_cl.debug_static[-5:]])
        print("graphed forward per call ms:", ["%.1f" % t for t in ts])
        print("observed at the end:", {k: v for k, v in _plan._observed.items() if k[0] == "cl"})
        print("graphed forward: median %.2f ms  kernels/step %d  fallbacks %d recaptures %d"
              % (sorted(ts)[len(ts) // 2], g.kernels_per_step, g.fallbacks, g.recaptures))
    except Exception as e:  # noqa: BLE001
        import traceback
        traceback.print_exc()
        print("graphed forward failed:", repr(e)[:300])

    # index build pieces
    gen = torch.Generator().manual_seed(1)
    for n in (6144, 14336, 30720):
        cell = torch.randperm(256 * 512, generator=gen)[:n]
        pos = torch.stack([(cell % 512).float(), (cell // 512).float()], 1).unsqueeze(0).cuda()
        t_sfc = ev_time(lambda: PU.space_filling_cluster(pos, 8, 256, 512))
        ps, mean, _, _, _ = PU.space_filling_cluster(pos, 8, 256, 512)
        t_nb = ev_time(lambda: PU.build_nbhd(ps, mean, 8, 48, 511, clamp=True))
        sc = torch.randn(1, n, device="cuda")
        t_sort = ev_time(lambda: PU.argsort_stable(sc))
        t_tsort = ev_time(lambda: sc.sort(dim=1, stable=True))
        t_l1 = ev_time(lambda: PU.find_pos_org_order(pos, ps))
        print("n=%5d  space_filling_cluster %.3f ms  build_nbhd %.3f ms  argsort_stable %.3f ms (torch.sort %.3f)  nearest_l1 %.3f ms"
              % (n, t_sfc, t_nb, t_sort, t_tsort, t_l1))


if __name__ == "__main__":
    main()
