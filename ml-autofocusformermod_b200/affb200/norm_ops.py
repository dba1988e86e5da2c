"""LayerNorm over the channel dim of token rows (csrc/layer_norm.cu).

  layer_norm(x, weight, bias, eps, out_dtype=None) -> LayerNorm(x) in `out_dtype`

The reference uses nn.LayerNorm (models/aff_transformer.py:199-200,284), which autocast evaluates in fp32: the bf16
residual stream is copied to fp32, normalised, and copied back to bf16 by the next Linear.  This operator reads x in
its storage dtype, keeps the statistics and the arithmetic in fp32 and writes the consumer's dtype directly — the value
the next Linear sees is the reference's fp32 result rounded once, as in the reference.  `out_dtype=None` follows
autocast: the autocast dtype when autocast is on (what the next Linear would cast to), else x's dtype.
"""
import torch
from torch.autograd import Function

from . import _native as N

_OK = (torch.float32, torch.float16, torch.bfloat16)


def supported(x, weight, bias, out_dtype):
    C = x.shape[-1]
    return (x.is_cuda and x.dtype in _OK and out_dtype in _OK and weight is not None and bias is not None
            and C % 8 == 0 and C <= 1024 and x.numel() > 0)


class LayerNormFunction(Function):
    @staticmethod
    def forward(ctx, x, weight, bias, eps, out_dtype):
        if not x.is_cuda:
            raise RuntimeError("x must be a CUDA tensor")
        C = x.shape[-1]
        xc = x.contiguous()
        R = xc.numel() // C
        dev = x.device
        g = weight.detach().float().contiguous()
        b = bias.detach().float().contiguous()
        with torch.cuda.device(dev):
            y = torch.empty(x.shape, dtype=out_dtype, device=dev)
            mean = torch.empty(R, dtype=torch.float32, device=dev)
            rstd = torch.empty(R, dtype=torch.float32, device=dev)
            N.check(N.LIB.affb200_layer_norm_fwd(N.p(xc), N.p(g), N.p(b), N.p(y), N.p(mean), N.p(rstd), R, C, float(eps),
                                                 N.dtype_enum(xc), N.DTYPE_ENUM[out_dtype], N.stream_ptr(dev)),
                    "layer_norm_fwd")
        ctx.save_for_backward(xc, g, mean, rstd)
        ctx.meta = (weight.dtype, bias.dtype)
        return y

    @staticmethod
    def backward(ctx, dy):
        xc, g, mean, rstd = ctx.saved_tensors
        wdt, bdt = ctx.meta
        C = xc.shape[-1]
        R = xc.numel() // C
        dev = xc.device
        dy = dy.contiguous()
        if dy.dtype not in _OK:
            dy = dy.float()
        with torch.cuda.device(dev):
            dx = torch.empty_like(xc)
            dg = torch.empty(C, dtype=torch.float32, device=dev)
            db = torch.empty(C, dtype=torch.float32, device=dev)
            work = torch.empty(max(1, N.LIB.affb200_layer_norm_bwd_work_floats(R, C)), dtype=torch.float32, device=dev)
            N.check(N.LIB.affb200_layer_norm_bwd(N.p(dy), N.p(xc), N.p(g), N.p(mean), N.p(rstd), N.p(dx), N.p(dg), N.p(db),
                                                 N.p(work), R, C, N.dtype_enum(xc), N.dtype_enum(dy), N.stream_ptr(dev)),
                    "layer_norm_bwd")
        return dx, dg.to(wdt), db.to(bdt), None, None


def layer_norm(x, weight, bias, eps=1e-5, out_dtype=None):
    if not x.is_cuda:
        raise RuntimeError("x must be a CUDA tensor")
    if out_dtype is None:
        out_dtype = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled("cuda") else x.dtype
    if not supported(x, weight, bias, out_dtype):
        # shapes the kernel does not cover (C not a multiple of 8, C > 1024, no affine): stock CUDA LayerNorm
        return torch.nn.functional.layer_norm(x, (x.shape[-1],), weight, bias, eps)
    return LayerNormFunction.apply(x, weight, bias, eps, out_dtype)
