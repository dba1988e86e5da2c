"""A training step of the off-grid model as CUDA graphs (the stages are launch-latency-bound once the
kernels are fast; the reference's step is ~3000 eager launches plus host syncs such as `nonzero`,
aff_transformer.py:323).

    step = GraphedTrainStep(model, optimizer, loss_fn, example_x, example_y, world=1)
    loss = step(x, y)           # device tensor
    step.fallbacks              # how many steps had to be redone eagerly (a speculated index plan did not hold)

Shapes are static (batch, image size, and therefore every stage's token count `int(n * ds_rate)`), so the whole
forward + loss + backward is captured once and replayed; the optimizer step is a second graph.  Between the two, for
world_size > 1, the gradients — all views of ONE flat fp32 buffer — are averaged with a single NCCL all-reduce over
NVLink (the batch-sharded data parallelism of the reference's DDP, main.py:102, with one bucket).

What normally needs a device->host read while the index structures are built (tile-plan statistics, the range check
of the packed bias indices) runs in `plan.static_plans()` mode: the sizes observed in the eager warm-up passes are
assumed, and violations are accumulated in a device-side guard.  The guard (4 bytes) is read once per step between
the two graphs; a step whose speculation failed is redone eagerly with exact plans before the optimizer runs, so a
failed speculation costs time, never correctness.

The warm-up passes are real optimisation steps on the example batch (the optimizer state must exist before capture).

Parameters and gradients are views of flat fp32 buffers ordered by optimizer param group; for Adam / AdamW / SGD the
step itself then runs on ONE tensor per group (`flat_optimizer=True`, see `_flat_optimizer`): same update, one launch.
"""
import torch
import torch.distributed as dist

from . import plan as _plan


class GraphedTrainStep:
    """See the module docstring.  The caller's optimizer stays the place where hyper-parameters live: its param groups'
    `lr` (and any other hyper-parameter) are read before every step, so an LR scheduler attached to it works as in the
    reference's loop (main.py:262-267: `lr_scheduler.step_update`); `clip_grad_norm` reproduces `clip_grad` between
    backward and step (utils.py:186); `optimizer_state_dict()` / `load_optimizer_state_dict()` checkpoint the state of
    the optimizer that actually steps (utils.py:55-77 saves `optimizer.state_dict()`)."""

    def __init__(self, model, optimizer, loss_fn, example_x, example_y, world=1, autocast_dtype=torch.bfloat16,
                 warmup=3, flat_optimizer=True, clip_grad_norm=None):
        self.model, self.opt, self.loss_fn, self.world = model, optimizer, loss_fn, world
        self.user_opt = optimizer
        self.clip_grad_norm = clip_grad_norm
        self.autocast_dtype = autocast_dtype
        dev = example_x.device
        self.device = dev
        self.params = [p for p in model.parameters() if p.requires_grad]
        if any(p.dtype != torch.float32 for p in self.params):
            raise RuntimeError("GraphedTrainStep expects fp32 master parameters (autocast training)")
        if world > 1:
            for p in self.params:            # same start on every rank, as DDP's constructor does
                dist.broadcast(p.data, 0)
            for b in model.buffers():
                dist.broadcast(b.data, 0)
        # parameters, gradients and the low-precision shadow each live in ONE flat buffer (32-byte aligned slices):
        # one all-reduce, one cast kernel per step instead of one per use of every weight, and gradient GEMMs that
        # accumulate in place (linear_ops: `_affb200_lp`, `_affb200_sink`)
        # (the optimizer's param groups first, in order, so that each group is one contiguous range: see _flat_optimizer)
        ids = {id(p) for p in self.params}
        ordered = [p for g in optimizer.param_groups for p in g["params"] if id(p) in ids and p.requires_grad]
        seen = {id(p) for p in ordered}
        group_counts = [sum(1 for p in g["params"] if id(p) in ids and p.requires_grad) for g in optimizer.param_groups]
        self.params = ordered + [p for p in self.params if id(p) not in seen]
        offs, n = [], 0
        for p in self.params:
            offs.append(n)
            n += (p.numel() + 7) // 8 * 8
        self.flat_param = torch.zeros(n, dtype=torch.float32, device=dev)
        self.flat_grad = torch.zeros(n, dtype=torch.float32, device=dev)
        self.flat_lp = torch.zeros(n, dtype=autocast_dtype, device=dev)
        for p, off in zip(self.params, offs):
            sl = slice(off, off + p.numel())
            self.flat_param[sl].copy_(p.data.reshape(-1))
            p.data = self.flat_param[sl].view_as(p)
            p.grad = self.flat_grad[sl].view_as(p)
            p._affb200_lp = self.flat_lp[sl].view_as(p)
            p._affb200_sink = True
        if flat_optimizer:
            self.opt = self._flat_optimizer(optimizer, group_counts, offs, n) or optimizer
        # learning rates as device tensors: the captured optimizer step reads them, __call__ refreshes them from the
        # caller's optimizer (a Python float would be baked into the graph)
        self._lr_seen = []
        for g in self.opt.param_groups:
            self._lr_seen.append(float(g["lr"]))
            if type(self.opt) in (torch.optim.Adam, torch.optim.AdamW) and g.get("capturable", False):
                g["lr"] = torch.tensor(float(g["lr"]), dtype=torch.float32, device=dev)
        self._hp_seen = self._hyper()
        # buffers the training forward updates in place (BatchNorm running statistics): saved at the start of every
        # replay so that a step redone eagerly does not apply the update twice
        self._bufs = [b for b in model.buffers()]
        self._buf_snap = [torch.empty_like(b) for b in self._bufs]
        self.sync_params()
        self.x = example_x.clone()
        self.y = example_y.clone()
        self.loss = None
        _plan.guard_tensor(dev)              # allocated outside the graphs' memory pool

        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                self._fwd_bwd()
                self._reduce()
                self._clip()
                self.opt.step()
                self.sync_params()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)

        self.fallbacks = 0        # steps whose speculated index plans did not hold and were redone eagerly
        self.recaptures = 0
        self.max_recaptures = 8
        self._capture()

    def _flat_optimizer(self, opt, group_counts, offs, n):
        """Adam / AdamW / SGD update every element independently with its group's hyper-parameters, so an optimizer
        over one flat tensor per param group (a contiguous range of the flat buffers, padding included: its gradient
        is always zero) performs the same update as one over the ~400 parameter tensors — in one kernel launch per
        group at streaming bandwidth instead of dozens of multi-tensor chunks.  Returns that optimizer (same class, the
        groups' hyper-parameters), or None when the given optimizer is of another class, already has state, or owns
        parameters outside the model.  The optimizer passed in then only supplies class and hyper-parameters; the
        state lives in `self.opt` (one tensor per group)."""
        import inspect
        if type(opt) not in (torch.optim.Adam, torch.optim.AdamW, torch.optim.SGD) or len(opt.state) != 0:
            return None
        if sum(group_counts) != len(self.params) or any(len(g["params"]) != c for g, c in zip(opt.param_groups, group_counts)):
            return None
        groups, k = [], 0
        for g, c in zip(opt.param_groups, group_counts):
            if c == 0:
                continue
            lo = offs[k]
            hi = offs[k + c] if k + c < len(offs) else n
            k += c
            fp = torch.nn.Parameter(self.flat_param[lo:hi])
            fp.grad = self.flat_grad[lo:hi]
            groups.append({**{kk: v for kk, v in g.items() if kk != "params"}, "params": [fp]})
        accepted = set(inspect.signature(type(opt).__init__).parameters)
        return type(opt)(groups, **{kk: v for kk, v in opt.defaults.items() if kk in accepted})

    def _capture(self):
        from . import _native as N
        l0 = N.LIB.affb200_launch_count()
        self.g_fb = torch.cuda.CUDAGraph()
        with _plan.static_plans(), torch.cuda.graph(self.g_fb):
            self.loss = self._fwd_bwd(snapshot=True)
        # libaffb200 kernels recorded in the graph = launched by every replay
        self.kernels_per_step = int(N.LIB.affb200_launch_count() - l0)
        self.g_opt = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.g_opt, pool=self.g_fb.pool()):
            self._clip()
            self.opt.step()
            self.sync_params()

    def sync_params(self):
        """Refresh the low-precision shadow from the fp32 parameters (call after changing parameters by hand, e.g.
        load_state_dict; the optimizer graph does it after every step)."""
        self.flat_lp.copy_(self.flat_param)
        from .linear_ops import mark_shadow_fresh
        mark_shadow_fresh(self.params)

    def _groups(self):
        """(caller's group, stepping group) pairs; the flat optimizer keeps the caller's non-empty groups in order"""
        ug = [g for g in self.user_opt.param_groups if any(p.requires_grad for p in g["params"])] \
            if self.opt is not self.user_opt else list(self.user_opt.param_groups)
        return list(zip(ug, self.opt.param_groups)) if len(ug) == len(self.opt.param_groups) else []

    def _hyper(self):
        return [tuple(sorted((k, v) for k, v in g.items() if k not in ("params", "lr") and isinstance(v, (int, float, bool, tuple))))
                for g, _ in self._groups()]

    def _sync_hyper(self):
        """Bring the stepping optimizer's hyper-parameters up to date with the caller's before a step.  Learning rates
        are device tensors (one scalar fill when a value changed); anything else that changed (weight decay, betas ...)
        is a constant of the captured step, so the optimizer graph is captured again."""
        pairs = self._groups()
        for i, (ug, sg) in enumerate(pairs):
            lr = float(ug["lr"])
            if lr != self._lr_seen[i]:
                self._lr_seen[i] = lr
                if torch.is_tensor(sg["lr"]):
                    sg["lr"].fill_(lr)
                else:
                    sg["lr"] = lr
                    self._hp_seen = None
        hp = self._hyper()
        if hp != self._hp_seen:
            for ug, sg in pairs:
                if sg is not ug:
                    for k, v in ug.items():
                        if k not in ("params", "lr"):
                            sg[k] = v
            self._hp_seen = hp
            torch.cuda.synchronize(self.device)
            self.g_opt = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.g_opt, pool=self.g_fb.pool()):
                self._clip()
                self.opt.step()
                self.sync_params()

    def optimizer_state_dict(self):
        """state_dict (a copy) of the optimizer that steps: one flat tensor per param group when flat_optimizer took
        effect, laid out as the model's parameters in optimizer-group order, each padded to 8 elements."""
        import copy
        return copy.deepcopy(self.opt.state_dict())

    def load_optimizer_state_dict(self, sd):
        """Restore a state saved by optimizer_state_dict(): the values are copied INTO the existing state tensors (the
        captured optimizer step holds their addresses).  Hyper-parameters keep coming from the caller's optimizer."""
        params = [p for g in self.opt.param_groups for p in g["params"]]
        for idx, st in sd["state"].items():
            cur = self.opt.state[params[int(idx)]]
            for k, v in st.items():
                if torch.is_tensor(v) and torch.is_tensor(cur.get(k)):
                    cur[k].copy_(v)
                else:
                    cur[k] = v

    def _clip(self):
        if self.clip_grad_norm is not None:      # clip_grad_norm_ on the flat gradient: no host read
            norm = torch.linalg.vector_norm(self.flat_grad)
            self.flat_grad.mul_(torch.clamp(self.clip_grad_norm / (norm + 1e-6), max=1.0))

    def _fwd_bwd(self, snapshot=False):
        if snapshot:
            for b, sn in zip(self._bufs, self._buf_snap):
                sn.copy_(b)
        self.flat_grad.zero_()
        with torch.autocast("cuda", dtype=self.autocast_dtype):
            out = self.model(self.x)
        loss = self.loss_fn(out.float(), self.y)
        loss.backward()
        return loss.detach()

    def _reduce(self):
        if self.world > 1:
            if dist.get_backend() == "nccl":
                dist.all_reduce(self.flat_grad, op=dist.ReduceOp.AVG)
            else:   # gloo (tests on one device): no AVG
                dist.all_reduce(self.flat_grad, op=dist.ReduceOp.SUM)
                self.flat_grad.div_(self.world)

    def __call__(self, x, y):
        self.x.copy_(x, non_blocking=True)
        self.y.copy_(y, non_blocking=True)
        self.g_fb.replay()
        # One 4-byte read between the two graphs: did every speculated index plan of this replay hold?  If not, the
        # replay skipped tiles (zeros, see csrc): the gradients are discarded and the step is redone eagerly with exact
        # plans before the optimizer sees anything; the graphs are then re-captured with the larger sizes now on record.
        guard = _plan.guard_tensor(self.device)
        self._sync_hyper()
        if int(guard.item()) != 0:
            guard.zero_()
            self.fallbacks += 1
            for b, sn in zip(self._bufs, self._buf_snap):     # undo the failed replay's in-place buffer updates
                b.copy_(sn)
            loss = self._fwd_bwd()
            self._reduce()
            self._clip()
            self.opt.step()
            self.sync_params()
            if self.recaptures < self.max_recaptures:
                self.recaptures += 1
                torch.cuda.synchronize(self.device)
                self._capture()              # allocates a new self.loss
            self.loss.copy_(loss)
            return self.loss
        self._reduce()
        self.g_opt.replay()
        return self.loss

    def check(self):
        """Kept for callers of the earlier interface: every replay is now verified (and redone if needed) inside
        __call__, so this only reports a guard raised by kernels run outside the step."""
        _plan.check_guard(self.device)


class GraphedForward:
    """Inference forward of the off-grid model as one CUDA graph (the reference's `throughput()` loop, main.py:387-414,
    is `model.eval()` + autocast forward on a fixed batch shape).

        fwd = GraphedForward(model, example_x)      # eval mode, autocast
        logits = fwd(x)                             # static output tensor, overwritten by the next call

    Index plans are speculated from the eager warm-up passes exactly as in GraphedTrainStep; the guard is read after the
    replay and a failed speculation is redone eagerly (and the graph re-captured) before the result is returned."""

    def __init__(self, model, example_x, autocast_dtype=torch.bfloat16, warmup=2):
        self.model, self.autocast_dtype = model, autocast_dtype
        self.device = example_x.device
        self.x = example_x.clone()
        self.out = None
        self.fallbacks = 0
        self.recaptures = 0
        self.max_recaptures = 8
        _plan.guard_tensor(self.device)
        self.refresh()
        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                self._fwd()
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)
        self._capture()

    def refresh(self):
        """Low-precision shadows of the (frozen) parameters, so that the forward does not cast every weight at every
        use (linear_ops.low_precision).  Call again after loading new weights."""
        for p in self.model.parameters():
            if p.dtype == torch.float32:
                sh = getattr(p, "_affb200_lp", None)
                if sh is not None and sh.dtype == self.autocast_dtype and sh.shape == p.shape:
                    sh.copy_(p.detach())
                else:
                    p._affb200_lp = p.detach().to(self.autocast_dtype)
        from .linear_ops import mark_shadow_fresh
        mark_shadow_fresh(self.model.parameters())

    def _fwd(self):
        with torch.no_grad(), torch.autocast("cuda", dtype=self.autocast_dtype):
            return self.model(self.x)

    def _capture(self):
        from . import _native as N
        l0 = N.LIB.affb200_launch_count()
        self._snap = _plan.observed_snapshot()
        self.g = torch.cuda.CUDAGraph()
        with _plan.static_plans(), torch.cuda.graph(self.g):
            self.out = self._fwd()
        self.kernels_per_step = int(N.LIB.affb200_launch_count() - l0)

    def __call__(self, x):
        self.x.copy_(x, non_blocking=True)
        self.g.replay()
        guard = _plan.guard_tensor(self.device)
        if int(guard.item()) != 0:
            guard.zero_()
            self.fallbacks += 1
            out = self._fwd()
            # re-capture only when the eager pass put larger sizes on record: a draw that no captured size could hold
            # (a tile whose union exceeds the kernels' limit) is simply served by this eager pass
            if self.recaptures < self.max_recaptures and _plan.observed_snapshot() != self._snap:
                self.recaptures += 1
                torch.cuda.synchronize(self.device)
                self._capture()
            self.out.copy_(out)
        return self.out
