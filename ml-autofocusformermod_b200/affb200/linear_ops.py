"""nn.Linear with the bias gradient as one column-sum pass (csrc/col_sum.cu).

  linear(x, weight, bias) -> x @ weight.T + bias      (reference: the nn.Linear layers of aff_transformer.py:85-100,285)
  linear_gelu(x, weight, bias) -> GELU(linear(...))    (fc1 + act of the Mlp, aff_transformer.py:35-51)
  gelu(x)                                              (exact GELU; table lookups for f16/bf16, csrc/gelu.cu)

The two GEMMs of the backward and the forward addmm are the stock cuBLAS calls nn.Linear makes (same dtypes under
autocast); only `grad_out.sum(0)`, which autograd runs as a generic strided reduction, is replaced — by fp32 column
sums accumulated in a fixed order.  `col_sum(x2d)` is exported for tests.

Two optional per-parameter attributes, set by `graph_step.GraphedTrainStep`, remove the per-call copies around the
GEMMs (under autocast the reference casts every weight to bf16 on every use and rounds every weight gradient to bf16
before adding it to the fp32 .grad):
  `_affb200_lp`    a low-precision shadow of the parameter, refreshed after each optimizer step, used instead of
                   `param.to(autocast dtype)`;
  `_affb200_sink`  True: the gradient GEMM accumulates straight into the existing fp32 `.grad` (cuBLAS beta = 1, fp32
                   output from bf16 operands: no bf16 rounding of dW, no separate add), the bias column sums likewise;
                   autograd then receives None for that parameter.
"""
import os

import torch
from torch.autograd import Function

from . import _native as N

_OK = (torch.float32, torch.float16, torch.bfloat16)


def col_sum(x2d, into=None):
    """fp32 column sums of a contiguous CUDA matrix [R, C] (C a multiple of 8); `into` (fp32 [C], contiguous): the sums
    are added to it in place and it is returned."""
    if not x2d.is_cuda:
        raise RuntimeError("x must be a CUDA tensor")
    R, C = x2d.shape
    if x2d.dtype not in _OK or C % 8 != 0:
        s = x2d.sum(0, dtype=torch.float32)   # stock CUDA reduction for widths the kernel does not cover
        return s if into is None else into.add_(s)
    x2d = x2d.contiguous()
    dev = x2d.device
    with torch.cuda.device(dev):
        out = torch.empty(C, dtype=torch.float32, device=dev) if into is None else into
        work = torch.empty(max(1, N.LIB.affb200_col_sum_work_floats(R, C)), dtype=torch.float32, device=dev)
        N.check(N.LIB.affb200_col_sum(N.p(x2d), N.p(out), N.p(work), R, C, N.dtype_enum(x2d),
                                      0 if into is None else 1, N.stream_ptr(dev)), "col_sum")
    return out


def low_precision(p, dt):
    """The parameter in the compute dtype: its registered shadow when there is one and the parameter has not been written
    since the shadow was refreshed (`_affb200_lp_ver` = the parameter's version counter at that time: load_state_dict, an
    EMA copy or another optimizer bump it), else a cast."""
    s = getattr(p, "_affb200_lp", None)
    if s is not None and s.dtype == dt and getattr(p, "_affb200_lp_ver", -1) == p._version:
        return s
    return p.to(dt)


def mark_shadow_fresh(params):
    """record that the `_affb200_lp` shadows of these parameters hold their current values"""
    for p in params:
        if getattr(p, "_affb200_lp", None) is not None:
            p._affb200_lp_ver = p._version


def _sink(p):
    """The fp32 .grad to accumulate into directly, or None."""
    if getattr(p, "_affb200_sink", False):
        g = p.grad
        if g is not None and g.dtype == torch.float32 and g.is_contiguous():
            return g
    return None


_gelu_tabs = {}   # (device index, dtype) -> (fwd_tab, bwd_tab)


def gelu_tables(dev, dtype):
    """The two 128 KB lookup tables of csrc/gelu.cu for this device and 2-byte dtype (built on first use; under CUDA
    graph capture they must already exist: the eager warm-up passes of GraphedTrainStep build them)."""
    key = (dev.index, dtype)
    t = _gelu_tabs.get(key)
    if t is None:
        with torch.cuda.device(dev):
            fwd = torch.empty(65536, dtype=torch.int16, device=dev)
            bwd = torch.empty(32768, dtype=torch.float32, device=dev)
            N.check(N.LIB.affb200_gelu_tables(N.DTYPE_ENUM[dtype], N.p(fwd), N.p(bwd), N.stream_ptr(dev)), "gelu_tables")
        t = _gelu_tabs[key] = (fwd, bwd)
    return t


def _gelu_ok(x):
    return x.is_cuda and x.dtype in (torch.float16, torch.bfloat16) and x.shape[-1] % 8 == 0 and x.shape[-1] <= 8192


def _gelu_fwd(h):
    """h contiguous f16/bf16 -> GELU(h)"""
    dev = h.device
    fwd, _ = gelu_tables(dev, h.dtype)
    with torch.cuda.device(dev):
        y = torch.empty_like(h)
        N.check(N.LIB.affb200_gelu_fwd(N.p(h), N.p(y), h.numel(), N.dtype_enum(h), N.p(fwd), N.stream_ptr(dev)), "gelu_fwd")
    return y


def _gelu_bwd(dy2, h2, col_into=None, want_cols=False):
    """dy2, h2 contiguous [R, C] -> (dh, column sums of dh or None); col_into: fp32 [C] the sums are added to."""
    dev = h2.device
    R, C = h2.shape
    _, bwd = gelu_tables(dev, h2.dtype)
    with torch.cuda.device(dev):
        dh = torch.empty_like(h2)
        cols = col_into
        if cols is None and want_cols:
            cols = torch.empty(C, dtype=torch.float32, device=dev)
        work = None
        if cols is not None:
            work = torch.empty(max(1, N.LIB.affb200_gelu_bwd_work_floats(C)), dtype=torch.float32, device=dev)
        N.check(N.LIB.affb200_gelu_bwd(N.p(dy2), N.p(h2), N.p(dh), R, C, N.dtype_enum(h2), N.p(bwd), N.p(cols),
                                       1 if col_into is not None else 0, N.p(work), N.stream_ptr(dev)), "gelu_bwd")
    return dh, cols


class GeluFunction(Function):
    @staticmethod
    def forward(ctx, x):
        xc = x.contiguous()
        ctx.save_for_backward(xc)
        return _gelu_fwd(xc)

    @staticmethod
    def backward(ctx, dy):
        (xc,) = ctx.saved_tensors
        C = xc.shape[-1]
        dy2 = dy.to(xc.dtype).contiguous().reshape(-1, C)
        dh, _ = _gelu_bwd(dy2, xc.reshape(-1, C))
        return dh.reshape(xc.shape)


def gelu(x):
    """Exact GELU.  f16/bf16 CUDA tensors whose last dim is a multiple of 8 take the lookup kernels; everything else
    (fp32 activations) the stock CUDA operator."""
    if not x.is_cuda:
        raise RuntimeError("x must be a CUDA tensor")
    if not _gelu_ok(x) or x.numel() == 0:
        return torch.nn.functional.gelu(x)
    return GeluFunction.apply(x)


class LinearFunction(Function):
    @staticmethod
    def forward(ctx, x, weight, bias):
        if not x.is_cuda:
            raise RuntimeError("x must be a CUDA tensor")
        dt = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled("cuda") else x.dtype
        x2 = x.reshape(-1, x.shape[-1]).to(dt)
        wb = low_precision(weight, dt)
        with torch.autocast("cuda", enabled=False):
            y = torch.addmm(low_precision(bias, dt), x2, wb.t())
        ctx.save_for_backward(x2, wb)
        ctx.meta = (x.shape, x.dtype, weight.dtype, bias.dtype)
        ctx.params = (weight, bias)
        return y.reshape(*x.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, wb = ctx.saved_tensors
        xshape, xdt, wdt, bdt = ctx.meta
        dy2 = dy.reshape(-1, dy.shape[-1]).to(wb.dtype).contiguous()
        dx = dw = db = None
        with torch.cuda.device(dy2.device):   # autograd's worker thread may not have a current device yet
            if ctx.needs_input_grad[0]:
                dx = (dy2 @ wb).reshape(xshape).to(xdt)
            weight, bias = ctx.params
            if ctx.needs_input_grad[1]:
                g = _sink(weight) if dy2.dtype != torch.float32 else None
                if g is not None:
                    torch.ops.aten.addmm.dtype_out(g, dy2.t(), x2, torch.float32, out=g)   # g += dy^T x, fp32 out
                else:
                    dw = (dy2.t() @ x2).to(wdt)
            if ctx.needs_input_grad[2]:
                g = _sink(bias)
                if g is not None and dy2.dtype in _OK and dy2.shape[1] % 8 == 0 and g.data_ptr() % 16 == 0:
                    col_sum(dy2, into=g)
                else:
                    db = col_sum(dy2).to(bdt)
        return dx, dw, db


def linear(x, weight, bias=None):
    if bias is None:
        return torch.nn.functional.linear(x, weight)
    return LinearFunction.apply(x, weight, bias)


def _param_grads(dy2, x2, weight, bias, wdt, bdt, need_w, need_b):
    """(dw, db) of y = x W^T + b for the caller to return — None where the gradient went straight into the
    parameter's fp32 .grad (weight: GEMM accumulating into it; bias: column sums added to it)."""
    dw = db = None
    if need_w:
        g = _sink(weight) if dy2.dtype != torch.float32 else None
        if g is not None:
            torch.ops.aten.addmm.dtype_out(g, dy2.t(), x2, torch.float32, out=g)   # g += dy^T x, fp32 out
        else:
            dw = (dy2.t() @ x2).to(wdt)
    if need_b:
        g = _sink(bias)
        if g is not None and dy2.dtype in _OK and dy2.shape[1] % 8 == 0 and g.data_ptr() % 16 == 0:
            col_sum(dy2, into=g)
        else:
            db = col_sum(dy2).to(bdt)
    return dw, db


class LinearPairFunction(Function):
    """(x W1^T + b1, x W2^T + b2) for two Linears on one input (q and kv of ClusterAttention, reference
    aff_transformer.py:85-86,96-100).  The backward forms dx = dy1 W1 + dy2 W2 by letting the second GEMM accumulate
    onto the first one's output, where autograd would write two dx tensors and add them in another pass."""

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2):
        if not x.is_cuda:
            raise RuntimeError("x must be a CUDA tensor")
        dt = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled("cuda") else x.dtype
        x2 = x.reshape(-1, x.shape[-1]).to(dt)
        wb1, wb2 = low_precision(w1, dt), low_precision(w2, dt)
        with torch.autocast("cuda", enabled=False):
            y1 = torch.addmm(low_precision(b1, dt), x2, wb1.t())
            y2 = torch.addmm(low_precision(b2, dt), x2, wb2.t())
        ctx.save_for_backward(x2, wb1, wb2)
        ctx.meta = (x.shape, x.dtype, w1.dtype, b1.dtype, w2.dtype, b2.dtype)
        ctx.params = (w1, b1, w2, b2)
        return y1.reshape(*x.shape[:-1], w1.shape[0]), y2.reshape(*x.shape[:-1], w2.shape[0])

    @staticmethod
    def backward(ctx, dy1, dy2):
        x2, wb1, wb2 = ctx.saved_tensors
        xshape, xdt, w1dt, b1dt, w2dt, b2dt = ctx.meta
        w1, b1, w2, b2 = ctx.params
        g1 = dy1.reshape(-1, dy1.shape[-1]).to(wb1.dtype).contiguous()
        g2 = dy2.reshape(-1, dy2.shape[-1]).to(wb2.dtype).contiguous()
        dx = None
        with torch.cuda.device(g1.device):   # autograd's worker thread may not have a current device yet
            if ctx.needs_input_grad[0]:
                dx = (g1 @ wb1).addmm_(g2, wb2).reshape(xshape).to(xdt)
            dw1, db1 = _param_grads(g1, x2, w1, b1, w1dt, b1dt, ctx.needs_input_grad[1], ctx.needs_input_grad[2])
            dw2, db2 = _param_grads(g2, x2, w2, b2, w2dt, b2dt, ctx.needs_input_grad[3], ctx.needs_input_grad[4])
        return dx, dw1, db1, dw2, db2


def linear_pair(x, w1, b1, w2, b2):
    """(linear(x, w1, b1), linear(x, w2, b2)) with one dx in the backward."""
    if b1 is None or b2 is None:
        return linear(x, w1, b1), linear(x, w2, b2)
    return LinearPairFunction.apply(x, w1, b1, w2, b2)


class LinearGeluFunction(Function):
    """GELU(x @ W^T + b): the addmm, then the lookup GELU; the backward produces d(pre-activation) and its column sums
    (the bias gradient) in one pass."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        dt = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled("cuda") else x.dtype
        x2 = x.reshape(-1, x.shape[-1]).to(dt)
        wb = low_precision(weight, dt)
        with torch.autocast("cuda", enabled=False):
            h = torch.addmm(low_precision(bias, dt), x2, wb.t())
        y = _gelu_fwd(h)
        ctx.save_for_backward(x2, wb, h)
        ctx.meta = (x.shape, x.dtype, weight.dtype, bias.dtype)
        ctx.params = (weight, bias)
        return y.reshape(*x.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, wb, h = ctx.saved_tensors
        xshape, xdt, wdt, bdt = ctx.meta
        weight, bias = ctx.params
        dy2 = dy.reshape(-1, dy.shape[-1]).to(h.dtype).contiguous()
        dx = dw = db = None
        with torch.cuda.device(dy2.device):
            g = _sink(bias) if ctx.needs_input_grad[2] else None
            if g is not None and g.data_ptr() % 16 != 0:
                g = None
            dh, cols = _gelu_bwd(dy2, h, col_into=g, want_cols=ctx.needs_input_grad[2])
            if ctx.needs_input_grad[2] and g is None:
                db = cols.to(bdt)
            if ctx.needs_input_grad[0]:
                dx = (dh @ wb).reshape(xshape).to(xdt)
            if ctx.needs_input_grad[1]:
                g = _sink(weight)
                if g is not None:
                    torch.ops.aten.addmm.dtype_out(g, dh.t(), x2, torch.float32, out=g)
                else:
                    dw = (dh.t() @ x2).to(wdt)
        return dx, dw, db


def linear_gelu(x, weight, bias):
    """GELU(nn.Linear(x)) as the Mlp computes it (fc1 -> act)."""
    dt = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled("cuda") else x.dtype
    if bias is None or dt not in (torch.float16, torch.bfloat16) or weight.shape[0] % 8 != 0 or weight.shape[0] > 8192 \
            or x.numel() == 0 or os.environ.get("AFFB200_GELU_LUT", "1") == "0":
        return torch.nn.functional.gelu(linear(x, weight, bias))
    return LinearGeluFunction.apply(x, weight, bias)
