"""LayerNorm over the channel dim of token rows (csrc/layer_norm.cu) and the stem's BatchNorm + GELU
(csrc/batch_norm.cu).

  layer_norm(x, weight, bias, eps, out_dtype=None) -> LayerNorm(x) in `out_dtype`
  batch_norm_act(x, bn, gelu=True)                 -> GELU(bn(x)) for a training-mode nn.BatchNorm2d `bn`
  conv_batch_norm_act(x, conv, bn, gelu=True)      -> GELU(bn(conv(x))) with the conv bias folded into the batch statistics
  conv_bias(x, conv)                               -> conv(x), bias added / its gradient reduced by streaming row kernels

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


def _ln_forward(x, weight, bias, eps, out_dtype):
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
    return xc, g, y, mean, rstd


def _sink(p):
    """fp32 .grad of a parameter registered for in-place accumulation (graph_step.GraphedTrainStep), or None."""
    if getattr(p, "_affb200_sink", False):
        g = p.grad
        if g is not None and g.dtype == torch.float32 and g.is_contiguous() and g.data_ptr() % 16 == 0:
            return g
    return None


def _ln_backward(ctx, dy, dres):
    xc, g, mean, rstd = ctx.saved_tensors
    weight, bias = ctx.params
    C = xc.shape[-1]
    R = xc.numel() // C
    dev = xc.device
    dy = dy.contiguous()
    if dy.dtype not in _OK:
        dy = dy.float()
    if dres is not None:
        dres = dres.to(xc.dtype).contiguous()
    with torch.cuda.device(dev):
        dx = torch.empty_like(xc)
        sg, sb = _sink(weight), _sink(bias)
        acc = sg is not None and sb is not None
        dg = sg if acc else torch.empty(C, dtype=torch.float32, device=dev)
        db = sb if acc else torch.empty(C, dtype=torch.float32, device=dev)
        work = torch.empty(max(1, N.LIB.affb200_layer_norm_bwd_work_floats(R, C)), dtype=torch.float32, device=dev)
        N.check(N.LIB.affb200_layer_norm_bwd(N.p(dy), N.p(xc), N.p(g), N.p(mean), N.p(rstd), N.p(dres), N.p(dx), N.p(dg),
                                             N.p(db), N.p(work), R, C, N.dtype_enum(xc), N.dtype_enum(dy),
                                             1 if acc else 0, N.stream_ptr(dev)), "layer_norm_bwd")
    if acc:
        return dx, None, None
    return dx, dg.to(weight.dtype), db.to(bias.dtype)


class LayerNormFunction(Function):
    @staticmethod
    def forward(ctx, x, weight, bias, eps, out_dtype):
        if not x.is_cuda:
            raise RuntimeError("x must be a CUDA tensor")
        xc, g, y, mean, rstd = _ln_forward(x, weight, bias, eps, out_dtype)
        ctx.save_for_backward(xc, g, mean, rstd)
        ctx.params = (weight, bias)
        return y

    @staticmethod
    def backward(ctx, dy):
        return _ln_backward(ctx, dy, None) + (None, None)


class LayerNormForkFunction(Function):
    """(x, LayerNorm(x)): the residual stream and its normalised copy leave one node, so the two gradients that come
    back — over the shortcut and through the norm — are added inside the LayerNorm backward kernel instead of by a
    separate pass of autograd's accumulation."""

    @staticmethod
    def forward(ctx, x, weight, bias, eps, out_dtype):
        if not x.is_cuda:
            raise RuntimeError("x must be a CUDA tensor")
        xc, g, y, mean, rstd = _ln_forward(x, weight, bias, eps, out_dtype)
        ctx.save_for_backward(xc, g, mean, rstd)
        ctx.params = (weight, bias)
        return xc.view_as(xc), y

    @staticmethod
    def backward(ctx, dres, dy):
        if dy is None:
            return dres, None, None, None, None
        return _ln_backward(ctx, dy, dres) + (None, None)


def layer_norm(x, weight, bias, eps=1e-5, out_dtype=None):
    if not x.is_cuda:
        raise RuntimeError("x must be a CUDA tensor")
    if out_dtype is None:
        out_dtype = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled("cuda") else x.dtype
    if not supported(x, weight, bias, out_dtype):
        # shapes the kernel does not cover (C not a multiple of 8, C > 1024, no affine): stock CUDA LayerNorm
        return torch.nn.functional.layer_norm(x, (x.shape[-1],), weight, bias, eps)
    return LayerNormFunction.apply(x, weight, bias, eps, out_dtype)


class BatchNormActFunction(Function):
    """y = act(BatchNorm(x)) with batch statistics over a channels-last [B, C, H, W] tensor; updates the running
    statistics in place (reference aff_transformer.py:524-527,541: nn.BatchNorm2d then nn.GELU)."""

    @staticmethod
    def forward(ctx, x, weight, bias, running_mean, running_var, momentum, eps, gelu):
        B, C, H, W = x.shape
        R = B * H * W
        dev = x.device
        g = weight.detach().float().contiguous()
        b = bias.detach().float().contiguous()
        with torch.cuda.device(dev):
            y = torch.empty_like(x)   # keeps the channels-last strides
            mean = torch.empty(C, dtype=torch.float32, device=dev)
            rstd = torch.empty(C, dtype=torch.float32, device=dev)
            work = torch.empty(max(1, N.LIB.affb200_batch_norm_work_floats(R, C)), dtype=torch.float32, device=dev)
            N.check(N.LIB.affb200_batch_norm_fwd(N.p(x), N.p(g), N.p(b), N.p(running_mean), N.p(running_var),
                                                 float(momentum), float(eps), 1 if gelu else 0, N.p(y), N.p(mean),
                                                 N.p(rstd), N.p(work), R, C, N.dtype_enum(x), N.stream_ptr(dev)),
                    "batch_norm_fwd")
        ctx.save_for_backward(x, g, b, mean, rstd)
        ctx.meta = (gelu, weight.dtype, bias.dtype)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, g, b, mean, rstd = ctx.saved_tensors
        gelu, wdt, bdt = ctx.meta
        B, C, H, W = x.shape
        R = B * H * W
        dev = x.device
        if dy.dtype != x.dtype:
            dy = dy.to(x.dtype)
        dy = dy.contiguous(memory_format=torch.channels_last)
        with torch.cuda.device(dev):
            dx = torch.empty_like(x)
            dg = torch.empty(C, dtype=torch.float32, device=dev)
            db = torch.empty(C, dtype=torch.float32, device=dev)
            work = torch.empty(max(1, N.LIB.affb200_batch_norm_work_floats(R, C)), dtype=torch.float32, device=dev)
            tab = None
            if gelu and x.dtype in (torch.float16, torch.bfloat16):
                from .linear_ops import gelu_tables      # GELU' as a lookup: a 2-byte BN output has 65 536 values
                tab = gelu_tables(dev, x.dtype)[1]
            N.check(N.LIB.affb200_batch_norm_bwd(N.p(dy), N.p(x), N.p(g), N.p(b), N.p(mean), N.p(rstd), 1 if gelu else 0,
                                                 N.p(dx), N.p(dg), N.p(db), N.p(work), R, C, N.dtype_enum(x),
                                                 N.p(tab), N.stream_ptr(dev)), "batch_norm_bwd")
        return dx, dg.to(wdt), db.to(bdt), None, None, None, None, None


def batch_norm_act(x, bn, gelu=True):
    """GELU(bn(x)) (or bn(x)) for an nn.BatchNorm2d module `bn` and a channels-last CUDA tensor.  Training mode with
    a fixed momentum runs the fused kernels; evaluation mode (running statistics: a per-channel affine map) and the
    layouts / channel counts the kernels do not cover use the stock CUDA operators."""
    if not x.is_cuda:
        raise RuntimeError("x must be a CUDA tensor")
    C = x.shape[1]
    ok = (bn.training and bn.affine and bn.track_running_stats and bn.momentum is not None and x.dim() == 4
          and x.dtype in _OK and C in (8, 16, 32, 64, 128) and x.numel() > 0
          and x.is_contiguous(memory_format=torch.channels_last))
    if not ok:
        y = bn(x)
        return torch.nn.functional.gelu(y) if gelu else y
    with torch.no_grad():
        bn.num_batches_tracked.add_(1)
    return BatchNormActFunction.apply(x, bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.momentum, bn.eps, gelu)


class _ZeroGradFor(torch.autograd.Function):
    """y -> y, registering `p` as an input whose gradient is exactly zero."""

    @staticmethod
    def forward(ctx, y, p):
        ctx.meta = (p.shape, p.dtype, p.device)
        return y.view_as(y)

    @staticmethod
    def backward(ctx, g):
        shape, dtype, device = ctx.meta
        return g, torch.zeros(shape, dtype=dtype, device=device)


class ConvStemFunction(Function):
    """conv 3x3 / stride 2 / padding 1 from 3 input channels, no bias, channels-last (csrc/conv_stem.cu); the weight
    gradient through the library's convolution backward (the images need no gradient)."""

    @staticmethod
    def forward(ctx, x, weight):
        B, _, H, W = x.shape
        Cout = weight.shape[0]
        dev = x.device
        w32 = weight.detach().float().contiguous()
        with torch.cuda.device(dev):
            y = torch.empty((B, Cout, H // 2, W // 2), dtype=x.dtype, device=dev, memory_format=torch.channels_last)
            N.check(N.LIB.affb200_conv_stem_fwd(N.p(x), N.p(w32), N.p(y), B, H, W, Cout, N.dtype_enum(x),
                                                N.stream_ptr(dev)), "conv_stem_fwd")
        ctx.save_for_backward(x)
        ctx.weight = weight
        return y

    @staticmethod
    def backward(ctx, g):
        from .linear_ops import low_precision, _sink
        (x,) = ctx.saved_tensors
        weight = ctx.weight
        if not ctx.needs_input_grad[1]:
            return None, None
        with torch.cuda.device(g.device):
            wl = low_precision(weight, x.dtype)
            g = g.to(x.dtype).contiguous(memory_format=torch.channels_last)
            _, dw, _ = torch.ops.aten.convolution_backward(g, x, wl, None, [2, 2], [1, 1], [1, 1], False, [0, 0], 1,
                                                           [False, True, False])
            sink = _sink(weight)
            if sink is not None:
                sink.add_(dw)
                return None, None
        return None, dw.to(weight.dtype)


def _stem_conv_ok(x, conv):
    return (x.is_cuda and x.dim() == 4 and x.dtype in (torch.float16, torch.bfloat16) and not x.requires_grad
            and conv.in_channels == 3 and conv.kernel_size == (3, 3) and conv.stride == (2, 2)
            and conv.padding == (1, 1) and conv.dilation == (1, 1) and conv.groups == 1
            and conv.padding_mode == "zeros" and conv.out_channels % 16 == 0
            and x.shape[2] % 2 == 0 and x.shape[3] % 2 == 0 and (x.shape[3] * 6) % 16 == 0
            and (x.shape[3] // 2) * (conv.out_channels // 16) <= 256 and 0 < x.shape[0] <= 65535
            and x.is_contiguous(memory_format=torch.channels_last) and conv.weight.dtype == torch.float32)


def _conv_bn_eval(x, conv, bn, gelu):
    """Inference: evaluation-mode BatchNorm is a per-channel affine map, so conv bias, running statistics, affine and
    GELU run as ONE pass over the bias-free convolution output (the stock chain: bias add, normalisation, GELU — three
    to five passes over the largest activation of the network).  None when the layout is not covered."""
    with torch.no_grad():
        y = ConvStemFunction.apply(x, conv.weight) if _stem_conv_ok(x, conv) else conv._conv_forward(x, conv.weight, None)
        if not (y.dtype in _OK and y.numel() > 0 and y.is_contiguous(memory_format=torch.channels_last)):
            return None
        B, C, H, W = y.shape
        dev = y.device
        mean = bn.running_mean.float()
        if conv.bias is not None:
            mean = mean - conv.bias.float()
        rstd = torch.rsqrt(bn.running_var.float() + bn.eps)
        g, b = bn.weight.float().contiguous(), bn.bias.float().contiguous()
        with torch.cuda.device(dev):
            out = torch.empty_like(y)
            N.check(N.LIB.affb200_batch_norm_apply(N.p(y), N.p(g), N.p(b), N.p(mean.contiguous()), N.p(rstd.contiguous()),
                                                   1 if gelu else 0, N.p(out), B * H * W, C, N.dtype_enum(y),
                                                   N.stream_ptr(dev)), "batch_norm_apply")
    return out


def conv_batch_norm_act(x, conv, bn, gelu=True):
    """GELU(bn(conv(x))) for the stem's conv -> BatchNorm pair (reference aff_transformer.py:541-543).  A per-channel
    constant added before a training-mode BatchNorm cancels in (x - mean): the convolution runs without its bias (no
    bias-add pass over the activation, no bias-gradient reduction over it in the backward), the batch mean that feeds
    `running_mean` gets the bias back, and the bias receives the gradient it has in exact arithmetic: zero.  Evaluation
    mode and unsupported layouts run conv (with bias) -> bn -> gelu through the stock operators."""
    if (x.is_cuda and not bn.training and bn.affine and bn.track_running_stats and conv.out_channels in (8, 16, 32, 64, 128)
            and not (torch.is_grad_enabled() and (x.requires_grad or conv.weight.requires_grad or bn.weight.requires_grad))):
        out = _conv_bn_eval(x, conv, bn, gelu)
        if out is not None:
            return out
    if conv.bias is None or not (x.is_cuda and bn.training and bn.affine and bn.track_running_stats
                                 and bn.momentum is not None and conv.out_channels in (8, 16, 32, 64, 128)):
        return batch_norm_act(conv(x), bn, gelu)
    # (the images arrive already in the compute dtype, see PatchEmbed.forward; the 3-channel stem convolution then
    # runs on this package's kernel, any other convolution on the library's)
    y = ConvStemFunction.apply(x, conv.weight) if _stem_conv_ok(x, conv) else conv._conv_forward(x, conv.weight, None)
    if not (y.dtype in _OK and y.numel() > 0 and y.is_contiguous(memory_format=torch.channels_last)):
        return batch_norm_act(y + conv.bias.to(y.dtype).view(1, -1, 1, 1), bn, gelu)
    out = batch_norm_act(y, bn, gelu)
    with torch.no_grad():
        bn.running_mean.add_(conv.bias.detach().to(bn.running_mean.dtype), alpha=bn.momentum)
    return _ZeroGradFor.apply(out, conv.bias) if conv.bias.requires_grad else out


class BiasRowsFunction(Function):
    """y = x + bias[c] on a channels-last activation; d_bias as one column-sum pass."""

    @staticmethod
    def forward(ctx, x, bias):
        B, C, H, W = x.shape
        dev = x.device
        b32 = bias.detach().float().contiguous()
        with torch.cuda.device(dev):
            y = torch.empty_like(x)      # preserves the channels-last strides
            N.check(N.LIB.affb200_bias_rows(N.p(x), N.p(b32), N.p(y), B * H * W, C, N.dtype_enum(x), N.stream_ptr(dev)),
                    "bias_rows")
        ctx.bias = bias
        return y

    @staticmethod
    def backward(ctx, g):
        from .linear_ops import col_sum, _sink
        bias = ctx.bias
        db = None
        if ctx.needs_input_grad[1]:
            B, C, H, W = g.shape
            rows = g.contiguous(memory_format=torch.channels_last).permute(0, 2, 3, 1).reshape(B * H * W, C)
            with torch.cuda.device(g.device):
                sink = _sink(bias)
                if sink is not None and sink.data_ptr() % 16 == 0:
                    col_sum(rows, into=sink)
                else:
                    db = col_sum(rows).to(bias.dtype)
        return g, db


def conv_bias(x, conv):
    """conv(x) for a channels-last CUDA activation with the bias added (and its gradient reduced) by this package's
    row kernels: the stock path adds the bias with a broadcast kernel and reduces the gradient with a generic strided
    sum, each several times slower than a streaming pass (the stem's proj2, reference aff_transformer.py:528,544)."""
    if not x.is_cuda:
        raise RuntimeError("x must be a CUDA tensor")
    if conv.bias is None:
        return conv(x)
    y = conv._conv_forward(x, conv.weight, None)
    C = y.shape[1]
    chunks = C * y.element_size() // 16
    if not (y.dim() == 4 and y.dtype in _OK and y.numel() > 0 and y.is_contiguous(memory_format=torch.channels_last)
            and (C * y.element_size()) % 16 == 0 and chunks <= 256 and 256 % chunks == 0 and C % 8 == 0):
        return y + conv.bias.to(y.dtype).view(1, -1, 1, 1)
    return BiasRowsFunction.apply(y, conv.bias)


def layer_norm_fork(x, weight, bias, eps=1e-5, out_dtype=None):
    """(x, layer_norm(x)) for a block's `shortcut = x; y = norm(x)` pair (reference aff_transformer.py:223-229); use
    the returned x as the shortcut so that its gradient is folded into the LayerNorm backward pass."""
    if not x.is_cuda:
        raise RuntimeError("x must be a CUDA tensor")
    if out_dtype is None:
        out_dtype = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled("cuda") else x.dtype
    if not supported(x, weight, bias, out_dtype) or not x.requires_grad:
        return x, layer_norm(x, weight, bias, eps, out_dtype)
    return LayerNormForkFunction.apply(x, weight, bias, eps, out_dtype)
