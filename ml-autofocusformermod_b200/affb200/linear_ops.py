"""nn.Linear with the bias gradient as one column-sum pass (csrc/col_sum.cu).

  linear(x, weight, bias) -> x @ weight.T + bias      (reference: the nn.Linear layers of aff_transformer.py:85-100,285)

The two GEMMs of the backward and the forward addmm are the stock cuBLAS calls nn.Linear makes (same dtypes under
autocast); only `grad_out.sum(0)`, which autograd runs as a generic strided reduction, is replaced — by fp32 column
sums accumulated in a fixed order.  `col_sum(x2d)` is exported for tests.
"""
import torch
from torch.autograd import Function

from . import _native as N

_OK = (torch.float32, torch.float16, torch.bfloat16)


def col_sum(x2d):
    """fp32 column sums of a contiguous CUDA matrix [R, C] (C a multiple of 8)."""
    if not x2d.is_cuda:
        raise RuntimeError("x must be a CUDA tensor")
    R, C = x2d.shape
    if x2d.dtype not in _OK or C % 8 != 0:
        return x2d.sum(0, dtype=torch.float32)   # stock CUDA reduction for widths the kernel does not cover
    x2d = x2d.contiguous()
    dev = x2d.device
    with torch.cuda.device(dev):
        out = torch.empty(C, dtype=torch.float32, device=dev)
        work = torch.empty(max(1, N.LIB.affb200_col_sum_work_floats(R, C)), dtype=torch.float32, device=dev)
        N.check(N.LIB.affb200_col_sum(N.p(x2d), N.p(out), N.p(work), R, C, N.dtype_enum(x2d), N.stream_ptr(dev)), "col_sum")
    return out


class LinearFunction(Function):
    @staticmethod
    def forward(ctx, x, weight, bias):
        if not x.is_cuda:
            raise RuntimeError("x must be a CUDA tensor")
        dt = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled("cuda") else x.dtype
        x2 = x.reshape(-1, x.shape[-1]).to(dt)
        wb = weight.to(dt)
        with torch.autocast("cuda", enabled=False):
            y = torch.addmm(bias.to(dt), x2, wb.t())
        ctx.save_for_backward(x2, wb)
        ctx.meta = (x.shape, x.dtype, weight.dtype, bias.dtype)
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
            if ctx.needs_input_grad[1]:
                dw = (dy2.t() @ x2).to(wdt)
            if ctx.needs_input_grad[2]:
                db = col_sum(dy2).to(bdt)
        return dx, dw, db


def linear(x, weight, bias=None):
    if bias is None:
        return torch.nn.functional.linear(x, weight)
    return LinearFunction.apply(x, weight, bias)
