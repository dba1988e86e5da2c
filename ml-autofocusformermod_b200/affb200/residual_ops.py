"""Residual connection with stochastic depth (csrc/residual.cu).

  residual_add(feat, x, p, training) -> feat + drop_path(x, p)     (reference aff_transformer.py:224-229, timm DropPath)

The per-sample keep draw is timm's (`x.new_empty(B,1,..).bernoulli_(keep)`: B values from the device generator) and
the scale is timm's `.div_(keep)` value, 1/keep rounded to x's dtype; the draw stays 0/1 in fp32 (one kernel instead of
draw + divide + cast) and the kernel multiplies by the scale; multiply and add run as one vectorised pass, and the
backward of the x branch as another.
"""
import torch
from torch.autograd import Function

from . import _native as N

_OK = (torch.float32, torch.float16, torch.bfloat16)


def _scale_add(a, x, s, scale):
    B = x.shape[0]
    per = x.numel() // max(B, 1)
    dev = x.device
    with torch.cuda.device(dev):
        out = torch.empty_like(x)
        N.check(N.LIB.affb200_scale_add(N.p(a), N.p(x), N.p(s), scale, N.p(out), B, per, N.dtype_enum(x), N.stream_ptr(dev)),
                "scale_add")
    return out


class ScaleAddFunction(Function):
    @staticmethod
    def forward(ctx, feat, x, s, scale):
        ctx.save_for_backward(s)
        ctx.scale = scale
        return _scale_add(feat.contiguous(), x.contiguous(), s, scale)

    @staticmethod
    def backward(ctx, g):
        (s,) = ctx.saved_tensors
        dx = _scale_add(None, g.contiguous(), s, ctx.scale) if ctx.needs_input_grad[1] else None
        return (g if ctx.needs_input_grad[0] else None), dx, None, None


def residual_add(feat, x, p, training):
    if not x.is_cuda:
        raise RuntimeError("x must be a CUDA tensor")
    if p == 0. or not training:
        return feat + x
    keep = 1 - p
    B = x.shape[0]
    per = x.numel() // max(B, 1)
    if feat.dtype != x.dtype or x.dtype not in _OK or feat.shape != x.shape or (per * x.element_size()) % 16 != 0:
        s = x.new_empty(B).bernoulli_(keep).div_(keep)
        return feat + x * s.reshape((B,) + (1,) * (x.dim() - 1))   # stock CUDA ops for odd shapes
    s = torch.empty(B, dtype=torch.float32, device=x.device).bernoulli_(keep)
    return ScaleAddFunction.apply(feat, x, s, _inv_keep(keep, x.dtype))


_inv = {}


def _inv_keep(keep, dtype):
    """1/keep as `tensor.div_(keep)` leaves it in a tensor of `dtype` holding 1 (timm's scale of the kept samples)."""
    key = (keep, dtype)
    if key not in _inv:
        _inv[key] = float(torch.ones((), dtype=dtype).div_(keep))
    return _inv[key]
