"""Workloads of bench.py.  Each workload offers
    build(B, device, world) -> run   with  run.make_host_batch(i) / run.to_device(host) / run.step(dev) -> loss
                                           run.roofline(peak, src) -> dict
    cpu_reference(sample_b) -> callable   (oracle port on the host cores; the only oracle use besides tests/smoke)
    config(B, world), metric, unit, dtype_name, default_batch
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))

# AFF-Tiny-1/5 (reference configs/aff_tiny_1_5th.yaml:6-13): local-attention stages (N, C, heads, depth) and the
# three merges N -> N' = int(N*0.2) (aff_transformer.py:292).  Stage 4 (25 tokens) is global/dense attention.
TINY15_STAGES = [(3136, 64, 2, 3), (627, 128, 4, 4), (125, 256, 8, 18)]
TINY15_MERGES = [(627, 3136, 64), (125, 627, 128), (25, 125, 256)]
NBHD, CLUSTER, INNER = 48, 8, 4


def block_structured_idx(B, Nq, Nk, M, m, device, seed):
    """Neighbourhoods of the model's form (aff_transformer.py:480): nnc = M/m nearest clusters x m consecutive rows;
    cluster choice jittered per token; rows >= Nk (padded last cluster) -> 0 (point_utils.py:204)."""
    g = torch.Generator().manual_seed(seed)
    nnc = M // m
    k = (Nk + m - 1) // m
    own = (torch.arange(Nq) * Nk // max(Nq, 1)) // m
    win = nnc + 2
    perm = torch.rand(B, Nq, win, generator=g).argsort(-1)[..., :nnc]
    cl = (own[None, :, None] + perm - win // 2).clamp_(0, k - 1)
    idx = (cl[..., None] * m + torch.arange(m)).reshape(B, Nq, M)
    idx = torch.where(idx >= Nk, torch.zeros_like(idx), idx)
    return idx.contiguous().to(device)


class ClustenChain:
    """v0 workload: every CLUSTEN call of one AFF-Tiny-1/5 forward+backward (25 blocks of QK -> softmax -> AV, and
    the three WF merges) at the model's shapes, driven through the autograd.Function surface.  The linears, MLPs and
    index builders of the full model are not part of this workload."""
    metric = "AFF-Tiny-1/5 224x224 off-grid attention path (CLUSTEN QK/AV/WF fwd+bwd) img/s"
    unit = "img/s"
    dtype_name = "bf16"
    default_batch = 256

    def config(self, B, world):
        return {"workload": "clusten_chain_aff_tiny_1_5th_224", "batch_per_gpu": B, "global_batch": B * world,
                "stages": TINY15_STAGES, "merges": TINY15_MERGES, "nbhd": NBHD, "cluster_size": CLUSTER,
                "parallelism": "dp%d" % world, "l2": "inputs larger than L2 (stage-1 working set 0.67 GB)",
                "idx": "block-structured synthetic neighbourhoods (int64)"}

    def build(self, B, device, world):
        return _ChainRun(B, device, torch.bfloat16)

    def cpu_reference(self, sample_b):
        from oracle import clusten_ref as R  # the checker, timed here as the CPU baseline only
        run = _ChainRun(sample_b, torch.device("cpu"), torch.float32, fns=(R.qk_ref, R.av_ref, R.wf_ref))
        host = run.make_host_batch(0)

        def go():
            run.step(host)
        return go


class _ChainRun:
    def __init__(self, B, device, dtype, fns=None):
        self.B, self.device, self.dtype = B, device, dtype
        if fns is None:
            from clusten import CLUSTENQKFunction, CLUSTENAVFunction, CLUSTENWFFunction
            fns = (CLUSTENQKFunction.apply, CLUSTENAVFunction.apply, CLUSTENWFFunction.apply)
        self.qk, self.av, self.wf = fns
        self.idx = [block_structured_idx(B, n, n, NBHD, CLUSTER, device, 10 + s)
                    for s, (n, _, _, _) in enumerate(TINY15_STAGES)]
        self.idx_m = [block_structured_idx(B, nq, nk, NBHD, CLUSTER, device, 20 + s)
                      for s, (nq, nk, _) in enumerate(TINY15_MERGES)]
        g = torch.Generator().manual_seed(5)
        self.w = [torch.randn(B, nq, NBHD, INNER, generator=g).to(dtype).to(device).requires_grad_(True)
                  for (nq, _, _) in TINY15_MERGES]

    def make_host_batch(self, i):
        g = torch.Generator().manual_seed(100 + i)
        ts = [torch.randn(self.B, n, c, generator=g).to(self.dtype) for (n, c, _, _) in TINY15_STAGES]
        return [t.pin_memory() if self.device.type == "cuda" else t for t in ts]

    def to_device(self, host):
        return [t.to(self.device, non_blocking=True) for t in host]

    def step(self, feats):
        loss = None
        for s, (n, c, h, depth) in enumerate(TINY15_STAGES):
            x = feats[s].detach().requires_grad_(True)
            kv = x.view(self.B, n, h, c // h).permute(0, 2, 1, 3)
            q = kv * (c // h) ** -0.5
            for _ in range(depth):
                attn = self.qk(q, kv, self.idx[s]).float().softmax(-1).to(self.dtype)
                q = q + self.av(attn, kv, self.idx[s])
            merged = self.wf(self.w[s], x, self.idx_m[s])
            l = q.float().square().mean() + merged.float().square().mean()
            loss = l if loss is None else loss + l
        loss.backward()
        return loss

    def roofline(self, peak, peak_src):
        """Per-op CUDA-event times at every stage shape (through the C ABI, on the current stream); the headline
        entry is the op with the largest share of the step (time x calls per step).  Algorithmic bytes are
        SURVEY.md §8(d)'s: every API-visible input read once + output written once, nbhd_idx at 8 B/element."""
        return clusten_roofline(self.B, self.device, self.dtype, self.idx, peak, peak_src)


def _time_op(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) / iters


def clusten_roofline(B, device, dtype, idx_list, peak, peak_src):
    from affb200 import _native as N
    from affb200.plan import get_plan
    s = torch.empty((), dtype=dtype).element_size()
    rows = []
    for st, (n, c_full, h, depth) in enumerate(TINY15_STAGES):
        idx = idx_list[st]
        c, M, n = c_full // h, idx.shape[2], idx.shape[1]
        q = torch.randn(B, h, n, c, device=device).to(dtype)
        k = torch.randn(B, h, n, c, device=device).to(dtype)
        attn = torch.softmax(torch.randn(B, h, n, M, device=device), -1).to(dtype)
        o_attn, o_feat, d_q, d_k = torch.empty_like(attn), torch.empty_like(q), torch.empty_like(q), torch.empty_like(k)
        pl = get_plan(idx)
        dense = dtype in (torch.float16, torch.bfloat16) and c in (16, 32)
        pf, pb = pl.struct(idx, n, dense, False), pl.struct(idx, n, dense, True)
        A = (B, h, n, n, M, c, N.dtype_enum(q), N.stream_ptr(device))
        P = N.p
        ops = {
            "qk_fwd": (lambda: N.LIB.affb200_qk_fwd(P(q), P(k), P(idx), pf, P(o_attn), *A),
                       s * (2 * B * h * n * c + B * h * n * M) + 8 * B * n * M),
            "qk_bwd": (lambda: N.LIB.affb200_qk_bwd(P(attn), P(q), P(k), P(idx), pb, P(d_q), P(d_k), *A),
                       s * (B * h * n * M + 4 * B * h * n * c) + 8 * B * n * M),
            "av_fwd": (lambda: N.LIB.affb200_av_fwd(P(attn), P(k), P(idx), pf, P(o_feat), *A),
                       s * (B * h * n * M + 2 * B * h * n * c) + 8 * B * n * M),
            "av_bwd": (lambda: N.LIB.affb200_av_bwd(P(q), P(attn), P(k), P(idx), pb, P(o_attn), P(d_k), *A),
                       s * (2 * B * h * n * M + 4 * B * h * n * c) + 8 * B * n * M),
        }
        for name, (fn, nbytes) in ops.items():
            ms = _time_op(fn)
            rows.append({"kernel": "%s@stage%d" % (name, st + 1), "ms": ms, "calls_per_step": depth,
                         "bytes": nbytes, "GBps": nbytes / ms / 1e6})
        del q, k, attn, o_attn, o_feat, d_q, d_k
    top = max(rows, key=lambda r: r["ms"] * r["calls_per_step"])
    return {"bound": "hbm", "kernel": top["kernel"], "achieved": top["GBps"], "peak": peak, "unit": "GB/s",
            "frac": top["GBps"] / peak, "traffic": None, "peak_source": peak_src,
            "algorithmic_bytes_per_launch": top["bytes"], "avg_launch_ms": top["ms"],
            "all": [{"kernel": r["kernel"], "ms": round(r["ms"], 4), "GBps": round(r["GBps"], 1),
                     "frac": round(r["GBps"] / peak, 3), "calls_per_step": r["calls_per_step"]} for r in rows]}


WORKLOADS = {"clusten_chain": ClustenChain}
DEFAULT = "clusten_chain"


def get(name):
    return WORKLOADS[name or DEFAULT]()


def _load_host_module(name):
    """A module of the host-side tree (affb200/<name>.py: written against an operator table, no top-level dependency
    on the CUDA library) loaded under a stand-in package, so that the CPU reference arm never maps libaffb200.so —
    importing the real `affb200` package would dlopen it."""
    import importlib
    import types
    pkg = "affb200_host"
    if pkg not in sys.modules:
        m = types.ModuleType(pkg)
        m.__path__ = [os.path.join(ROOT, "ml-autofocusformermod_b200", "affb200")]
        sys.modules[pkg] = m
    return importlib.import_module(pkg + "." + name)


def _host_glue():
    return _load_host_module("aff_transformer")


def merge_rows(merges, infer):
    """CUDA-event times of the merge operators at the model's three merges with the model's own neighbourhoods: the
    fused merge (affb200_merge_fwd/_bwd, what the step runs) and the drop-in WF operator on materialised weights
    (affb200_wf_fwd/_bwd).  Algorithmic bytes are SURVEY.md 8(d)'s WF figures (weights at the activation type, idx at
    8 B); the fused merge reads the same feature rows and writes the same output without the weights tensor, so the
    same byte count is used for both (it over-credits the fused form by the weights it never touches)."""
    from affb200 import _native as N
    from affb200.plan import get_plan
    rows = []
    for st, (table, pe_idx, lp, mask, feat, member) in enumerate(merges):
        B, Nq, M = member.shape
        Nk, C = feat.shape[1], feat.shape[2]
        dev = feat.device
        s = feat.element_size()
        featc = feat.detach().contiguous()
        lpc = lp.detach().reshape(B, Nk).contiguous()
        tab = table.detach().float().contiguous()
        pe_c, nb_c = pe_idx.contiguous(), member.contiguous()
        mk_c = None if mask is None else mask.contiguous()
        out = torch.empty((B, Nq, 4 * C), dtype=torch.float32, device=dev)
        d_out = torch.randn_like(out)
        off, ent = get_plan(member).csr(nb_c, Nk)
        d_wt = torch.empty((B, Nq, M, 4), dtype=torch.float32, device=dev)
        d_lp = torch.empty((B, Nk), dtype=torch.float32, device=dev)
        d_feat = torch.empty_like(featc)
        sp, de, le = N.stream_ptr(dev), N.dtype_enum(featc), N.DTYPE_ENUM[lpc.dtype]
        P = N.p
        fwd_b = s * (B * Nq * M * 4 + B * Nk * C + B * Nq * 4 * C) + 8 * B * Nq * M
        bwd_b = s * (B * Nq * 4 * C + 2 * B * Nq * M * 4 + 2 * B * Nk * C) + 8 * B * Nq * M
        w = (tab[pe_c] * lpc.float().gather(1, nb_c.reshape(B, -1)).reshape(B, Nq, M, 1)).to(featc.dtype).contiguous()
        o_wf, dw, go = torch.empty((B, Nq, 4, C), dtype=featc.dtype, device=dev), torch.empty_like(w), None
        go = torch.randn_like(o_wf)
        pl = get_plan(member)
        pb = pl.struct(member, Nk, False, True)
        A = (B, Nq, Nk, M, 4, C, de, sp)
        ops = [("merge_fwd", lambda: N.LIB.affb200_merge_fwd(P(tab), P(pe_c), P(nb_c), P(mk_c), P(lpc), P(featc), P(out),
                                                             B, Nq, Nk, M, C, le, de, sp), fwd_b),
               ("wf_fwd", lambda: N.LIB.affb200_wf_fwd(P(w), P(featc), P(nb_c), None, P(o_wf), *A), fwd_b)]
        if not infer:
            ops += [("merge_bwd", lambda: N.LIB.affb200_merge_bwd(P(d_out), P(tab), P(pe_c), P(nb_c), P(mk_c), P(lpc),
                                                                  P(featc), P(off), P(ent), P(d_wt), P(d_lp), P(d_feat),
                                                                  B, Nq, Nk, M, C, le, de, sp), bwd_b),
                    ("wf_bwd", lambda: N.LIB.affb200_wf_bwd(P(go), P(w), P(featc), P(nb_c), pb, P(dw), P(d_feat), *A), bwd_b)]
        for name, fn, nbytes in ops:
            ms = _time_op(fn)
            rows.append({"kernel": "%s@merge%d" % (name, st + 1), "ms": ms, "calls_per_step": 1 if name.startswith("merge") else 0,
                         "bytes": nbytes, "GBps": nbytes / ms / 1e6})
    return rows


class AffTrain:
    """The workload BASELINE.json's metric is quoted on: AFF-Tiny-1/5 224x224 forward + backward (+ AdamW step) in
    bf16 autocast, random-init weights, synthetic images, per-GPU batch 256 (configs/aff_tiny_1_5th.yaml:17),
    batch-sharded DDP for N > 1 (reference main.py:102)."""
    model_name = "aff_tiny_1_5th"
    metric = "AFF-Tiny-1/5 224x224 fwd+bwd img/s"
    unit = "img/s"
    dtype_name = "bf16"
    default_batch = 256
    img = 224

    def config(self, B, world):
        return {"workload": "%s_%d_train_step" % (self.model_name, self.img), "batch_per_gpu": B,
                "global_batch": B * world, "parallelism": "dp%d" % world,
                "step": "forward + cross-entropy + backward + AdamW step, bf16 autocast, drop_path as configured",
                "l2": "activations per step (GBs) far exceed the 126 MB L2; inputs alternate between two batches",
                "attention": os.environ.get("AFFB200_ATTN", "fused"),
                "launch": "eager" if os.environ.get("AFFB200_GRAPH", "1") == "0" else
                          "CUDA graphs (fwd+loss+bwd, optimizer); one flat NCCL all-reduce between them for dp > 1"}

    def build(self, B, device, world):
        return _AffRun(self, B, device, world)

    def cpu_reference(self, sample_b):
        build_aff = _host_glue().build_aff
        from oracle.aff_ops_ref import OracleOps  # the checker, timed here as the CPU baseline only
        torch.manual_seed(0)
        model = build_aff(self.model_name, ops=OracleOps()).train()
        g = torch.Generator().manual_seed(1)
        x = torch.randn(sample_b, 3, self.img, self.img, generator=g)
        y = torch.randint(0, 1000, (sample_b,), generator=g)

        def go():
            model.zero_grad(set_to_none=True)
            torch.nn.functional.cross_entropy(model(x), y).backward()
        return go


class _AffRun:
    def __init__(self, wl, B, device, world):
        from affb200.aff_transformer import build_aff, CudaOps
        self.wl, self.B, self.device, self.world = wl, B, device, world
        torch.backends.cudnn.benchmark = True     # as the reference's entry point does (main.py:434): the stem convs
        ops = CudaOps()
        if os.environ.get("AFFB200_ATTN", "fused") == "fused":
            try:
                from affb200 import fused
                ops.fused_attention = fused.fused_cluster_attention
            except ImportError:
                pass
        self.ops = ops
        self.infer = bool(getattr(wl, "infer", False))
        model = (wl.build_model(ops) if hasattr(wl, "build_model") else build_aff(wl.model_name, ops=ops)).to(device)
        model = model.eval() if self.infer else model.train()
        self.hw = getattr(wl, "img_hw", (wl.img, wl.img))
        self.stage_layers = wl.stage_layers(model) if hasattr(wl, "stage_layers") else list(model.layers[:3])
        self.model = model
        # default: the step replayed as CUDA graphs (affb200/graph_step.py; gradients averaged with one flat NCCL
        # all-reduce).  AFFB200_GRAPH=0: eager launches, DistributedDataParallel for world > 1 as the reference does.
        self.graphed = os.environ.get("AFFB200_GRAPH", "1") != "0"
        self.gstep = None
        if world > 1 and not self.graphed and not self.infer:
            from torch.nn.parallel import DistributedDataParallel as DDP
            model = DDP(model, device_ids=[device.index], broadcast_buffers=False, gradient_as_bucket_view=True)
        self.net = model
        self.opt = None if self.infer else torch.optim.AdamW(model.parameters(), lr=1e-4, weight_decay=0.05,
                                                             fused=True, capturable=self.graphed)

    def make_host_batch(self, i):
        g = torch.Generator().manual_seed(1000 + i)
        x = torch.randn(self.B, 3, *self.hw, generator=g).pin_memory()
        y = torch.randint(0, 1000, (self.B,), generator=g).pin_memory()
        return [x, y]

    def to_device(self, host):
        return [t.to(self.device, non_blocking=True) for t in host]

    def step(self, batch):
        x, y = batch
        if self.infer:       # replicas only: no collective on the inference path
            if self.graphed:
                if self.gstep is None:
                    from affb200.graph_step import GraphedForward
                    self.gstep = GraphedForward(self.model, x)
                return self.gstep(x)[0, 0]
            with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
                return self.model(x)[0, 0]
        if self.graphed:
            if self.gstep is None:
                from affb200.graph_step import GraphedTrainStep
                self.gstep = GraphedTrainStep(self.model, self.opt, torch.nn.functional.cross_entropy, x, y,
                                              world=self.world)
            return self.gstep(x, y)
        self.opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            logits = self.net(x)
        loss = torch.nn.functional.cross_entropy(logits.float(), y)
        loss.backward()
        self.opt.step()
        return loss

    def roofline(self, peak, peak_src):
        """The attention kernels at this model's stage shapes with the model's own neighbourhoods (captured from a
        forward): the fused neighbourhood-attention forward kernel and backward kernel group (what the step runs),
        followed by the unfused CLUSTEN operators for reference.  Algorithmic bytes of the fused op = its API-visible
        operands once: fwd q + kv + out at s bytes, nbhd_idx + pe_idx (+ mask) at 8 B; bwd q, kv, out, d_out read and
        d_q, d_kv written, plus the same index tensors."""
        cap, merges = [], []
        hooks = [layer.blocks[0].register_forward_pre_hook(lambda m, a: cap.append((a[1], a[2], a[3], a[5])))
                 for layer in self.stage_layers]
        real_merge = getattr(self.ops, "merge_wf", None)
        if real_merge is not None:     # operands of the three merges (ClusterMerging: aff_transformer.py:343-361)
            def spy(table, pe_idx, lp, mask, feat, member):
                merges.append((table, pe_idx, lp, mask, feat, member))
                return real_merge(table, pe_idx, lp, mask, feat, member)
            self.ops.merge_wf = spy
        from affb200 import fused as _fused
        _fused.trace = []
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
            self.model(torch.randn(self.B, 3, *self.hw, device=self.device))
        calls, _fused.trace = _fused.trace, None
        if real_merge is not None:
            self.ops.merge_wf = real_merge
        for h in hooks:
            h.remove()
        # which kernels each stage's blocks ran (stages in forward order; one entry per distinct token count)
        stages = []
        for n, H, fam in calls:
            if not stages or stages[-1]["tokens"] != n:
                stages.append({"tokens": n, "heads": H, "blocks": 0, "kernels": fam or "unfused operators"})
            stages[-1]["blocks"] += 1
        rows = []
        if self.ops.fused_attention is not None:
            from affb200 import fused
            for st, (nbhd, mask, pe_idx, stage_table) in enumerate(cap):   # stage_table: the (compacted) rows pe_idx indexes
                layer = self.stage_layers[st]
                attn = layer.blocks[0].attn
                B, n, M = nbhd.shape
                H, hc = attn.num_heads, attn.dim
                q = torch.randn(B, n, hc, device=self.device).bfloat16()
                kv = torch.randn(B, n, 2 * hc, device=self.device).bfloat16()
                table = torch.randn(stage_table.shape[0], H, device=self.device).bfloat16()
                with torch.set_grad_enabled(not self.infer):     # inference builds forward-only plans
                    calls = fused.raw_calls(q, kv, H, attn.scale, nbhd, pe_idx, mask, table, attn.blank_k, attn.blank_v,
                                            pre_table=stage_table)
                if calls is None:
                    continue
                nidx = 8 * B * n * M * (3 if mask is not None else 2)
                for nm, fn, nbytes in (("fused_attn_fwd", calls[0], 2 * 4 * B * n * hc + nidx),
                                       ("fused_attn_bwd", calls[1], 2 * 8 * B * n * hc + nidx)):
                    if fn is None or (self.infer and nm.endswith("bwd")):
                        continue
                    ms = _time_op(fn)
                    rows.append({"kernel": "%s@stage%d" % (nm, st + 1), "ms": ms, "calls_per_step": layer.depth,
                                 "bytes": nbytes, "GBps": nbytes / ms / 1e6})
                del q, kv, calls
        rows += merge_rows(merges, self.infer)
        top = max(rows, key=lambda r: r["ms"] * r["calls_per_step"]) if rows else None
        unfused = clusten_roofline(self.B, self.device, torch.bfloat16, [c[0] for c in cap], peak, peak_src)
        if top is None:
            return unfused
        fmt = lambda r: {"kernel": r["kernel"], "ms": round(r["ms"], 4), "GBps": round(r["GBps"], 1),
                         "frac": round(r["GBps"] / peak, 3), "calls_per_step": r["calls_per_step"]}
        # measured DRAM bytes (ncu dram__bytes_read + write) of the same launch group: from the capture of THIS build's
        # kernels committed under profiles/ (the file names the git commit it was taken at); null when that file does
        # not list the group — never a number from another build
        traffic, traffic_src = None, None
        try:
            import json
            tj = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r02_ncu_traffic.json")))
            if tj.get("model") == self.wl.model_name and tj.get("batch_per_gpu") == self.B:    # the capture's own shapes only
                traffic = tj["per_launch_group"].get(top["kernel"], {}).get("dram_bytes")
                traffic_src = "profiles/r02_ncu_traffic.json @ %s" % tj.get("commit")
        except Exception:
            traffic = None
        return {"bound": "hbm", "kernel": top["kernel"], "achieved": top["GBps"], "peak": peak, "unit": "GB/s",
                "frac": top["GBps"] / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": top["bytes"], "avg_launch_ms": top["ms"],
                "all": [fmt(r) for r in rows], "unfused_clusten_ops": unfused["all"], "fused_stages": stages}


class AffTinyTrain(AffTrain):
    model_name = "aff_tiny"
    metric = "AFF-Tiny 224x224 fwd+bwd img/s"


class AffInfer(AffTrain):
    """Inference throughput as the reference's own `throughput()` measures it (main.py:387-414): model.eval(), autocast
    forward on a fixed batch.  BASELINE config 2: AFF-Mini 224x224, synthetic batch 256, one B200."""
    model_name = "aff_mini"
    metric = "AFF-Mini 224x224 inference img/s"
    default_batch = 256
    infer = True

    def config(self, B, world):
        c = super().config(B, world)
        c.update(workload="%s_%d_inference" % (self.model_name, self.img),
                 step="model.eval() forward, bf16 autocast (reference main.py:387-414)",
                 launch="eager" if os.environ.get("AFFB200_GRAPH", "1") == "0" else "one CUDA graph per forward")
        return c

    def cpu_reference(self, sample_b):
        build_aff = _host_glue().build_aff
        from oracle.aff_ops_ref import OracleOps  # the checker, timed here as the CPU baseline only
        torch.manual_seed(0)
        model = build_aff(self.model_name, ops=OracleOps()).eval()
        x = torch.randn(sample_b, 3, self.img, self.img, generator=torch.Generator().manual_seed(1))

        def go():
            with torch.no_grad():
                model(x)
        return go


class AffBase384Infer(AffInfer):
    """BASELINE config 4: AFF-Base 384x384 (22kto1k) inference, yaml batch 16."""
    model_name = "aff_base_22kto1k_384"
    metric = "AFF-Base 384x384 inference img/s"
    default_batch = 16
    img = 384


class AffTinyInfer(AffInfer):
    model_name = "aff_tiny_1_5th"
    metric = "AFF-Tiny-1/5 224x224 inference img/s"


class MaskfinerUDSmallInfer(AffInfer):
    """BASELINE config 5: the fork's seven-backbone U-shaped classifier (configs/maskfiner_up_down_small.yaml) at
    1024 x 2048, batch 1, forward only (model.eval(), bf16 autocast).  The reference publishes no number for it."""
    model_name = "maskfiner_up_down_small"
    metric = "maskfiner_up_down_small 1024x2048 inference img/s"
    default_batch = 1
    ref_batch = 1
    img_hw = (1024, 2048)

    def config(self, B, world):
        c = super().config(B, world)
        c.update(workload="maskfiner_ud_small_1024x2048_infer",
                 tokens_per_stage=[2048, 6144, 14336, 30720, 14336, 6144, 2048],
                 l2="the patch-embedding activations (512 x 1024 x 32 fp32 = 64 MB, read and written by three passes) and "
                    "~80 MB of bf16 weights exceed the 126 MB L2; inputs alternate between two images",
                 step="model.eval() forward of the 7-backbone chain, bf16 autocast; importance scores ~ N(0,1) per step "
                      "as the reference draws them (maskfiner_up_down.py:147-149)")
        return c

    def build_model(self, ops):
        from affb200.maskfiner_up_down import build_maskfiner_up_down
        return build_maskfiner_up_down(num_classes=1000, ops=ops)

    def stage_layers(self, model):
        return [bb.layers for bb in model.backbones if hasattr(bb.layers, "nbhd_size")]

    def cpu_reference(self, sample_b):
        from oracle.aff_ops_ref import OracleOps  # the checker, timed here as the CPU baseline only
        torch.manual_seed(0)
        ud = _load_host_module("maskfiner_up_down")
        stable = lambda s: s.sort(dim=1, stable=True)[1]
        l1 = lambda a, b: torch.cat([torch.cdist(a[:, i:i + 1024].float(), b.float(), p=1).argmin(dim=2)
                                     for i in range(0, a.shape[1], 1024)], dim=1)
        model = ud.build_maskfiner_up_down(num_classes=1000, ops=OracleOps(), argsort=stable, find_order=l1).eval()
        x = torch.randn(sample_b, 3, *self.img_hw, generator=torch.Generator().manual_seed(1))

        def go():
            with torch.no_grad():
                model(x)
        return go


WORKLOADS.update({"maskfiner_ud_small_1024x2048_infer": MaskfinerUDSmallInfer, "aff_tiny_1_5th": AffTrain, "aff_tiny": AffTinyTrain, "aff_mini_infer": AffInfer,
                  "aff_base_384_infer": AffBase384Infer, "aff_tiny_1_5th_infer": AffTinyInfer})
DEFAULT = "aff_tiny_1_5th"
