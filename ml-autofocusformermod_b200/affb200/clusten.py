"""The CLUSTEN operator surface: CLUSTENQKFunction / CLUSTENAVFunction / CLUSTENWFFunction.

Drop-in for the reference's clusten/clusten.py:17-92 — same `apply(a, b, nbhd_idx)` signatures, the same contiguity
and dtype coercions (the second operand is cast to the first operand's dtype), the same saved tensors and the same
`(grad_a, grad_b, None)` backward returns — but the native layer underneath is libaffb200.so (C ABI in
include/affb200.h) instead of the three pybind modules, K is consumed in its native [B,H,N,c] layout (no
transposed copy, reference clusten.py:30), bfloat16 is supported, sums are accumulated in fp32, and the scatter
halves of the backward passes walk a transposed index built once per nbhd_idx tensor (plan.py) instead of issuing
atomics, which makes gradients run-to-run deterministic.

There is no CPU path: tensors that are not on a CUDA device raise the reference's error text
(clustenqk_cuda.cpp:21 `TORCH_CHECK(x.device().is_cuda(), #x " must be a CUDA tensor")`).
"""
import torch
from torch.autograd import Function

from . import _native as N
from .plan import get_plan


def _check_cuda(**tensors):
    dev = None
    for name, t in tensors.items():
        if not t.is_cuda:
            raise RuntimeError("%s must be a CUDA tensor" % name)
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise RuntimeError("%s must be on the same device as the other arguments" % name)
    return dev


def _dense(t, c):
    """dtype / head dim combinations served by the dense-tile tensor-core kernels (csrc/clusten_dense.cu)"""
    return t.dtype in (torch.float16, torch.bfloat16) and c in (16, 32)


def _check_idx(nbhd_idx, B, Nq):
    if nbhd_idx.dtype != torch.int64:
        raise RuntimeError("nbhd_idx must be an int64 tensor")
    if nbhd_idx.dim() != 3 or nbhd_idx.shape[0] != B or nbhd_idx.shape[1] != Nq:
        raise RuntimeError("nbhd_idx must have shape [B, N, M] matching the first operand")


class CLUSTENQKFunction(Function):
    """attn[b,h,i,j] = sum_c query[b,h,i,c] * key[b,h,nbhd_idx[b,i,j],c]   (reference clusten.py:17-40)"""

    @staticmethod
    def forward(ctx, query, key, nbhd_idx):
        query = query.contiguous()
        key = key.contiguous()
        if key.dtype != query.dtype:
            key = key.to(query.dtype)
        nbhd_idx = nbhd_idx.contiguous()
        dev = _check_cuda(query=query, key=key, nbhd_idx=nbhd_idx)
        B, H, Nq, c = query.shape
        Nk = key.shape[2]
        _check_idx(nbhd_idx, B, Nq)
        M = nbhd_idx.shape[2]
        with torch.cuda.device(dev):
            attn = torch.empty((B, H, Nq, M), dtype=query.dtype, device=dev)
            plan = get_plan(nbhd_idx).struct(nbhd_idx, Nk, _dense(query, c), False)
            N.check(N.LIB.affb200_qk_fwd(N.p(query), N.p(key), N.p(nbhd_idx), plan, N.p(attn), B, H, Nq, Nk, M, c,
                                         N.dtype_enum(query), N.stream_ptr(dev)), "clusten_qk_forward")
        ctx.save_for_backward(query, key, nbhd_idx)
        return attn

    @staticmethod
    def backward(ctx, grad_attn):
        query, key, nbhd_idx = ctx.saved_tensors
        grad_attn = grad_attn.contiguous()
        if grad_attn.dtype != query.dtype:
            grad_attn = grad_attn.to(query.dtype)
        dev = _check_cuda(d_attn=grad_attn)
        B, H, Nq, c = query.shape
        Nk = key.shape[2]
        M = nbhd_idx.shape[2]
        with torch.cuda.device(dev):
            plan = get_plan(nbhd_idx).struct(nbhd_idx, Nk, _dense(query, c), True)
            d_query = torch.empty_like(query)
            d_key = torch.empty_like(key)
            N.check(N.LIB.affb200_qk_bwd(N.p(grad_attn), N.p(query), N.p(key), N.p(nbhd_idx), plan,
                                         N.p(d_query), N.p(d_key), B, H, Nq, Nk, M, c, N.dtype_enum(query),
                                         N.stream_ptr(dev)), "clusten_qk_backward")
        return d_query, d_key, None


class CLUSTENAVFunction(Function):
    """feat[b,h,i,c] = sum_j attn[b,h,i,j] * v[b,h,nbhd_idx[b,i,j],c]   (reference clusten.py:43-66)"""

    @staticmethod
    def forward(ctx, attn, v, nbhd_idx):
        attn = attn.contiguous()
        v = v.contiguous()
        nbhd_idx = nbhd_idx.contiguous()
        if attn.dtype != v.dtype:
            v = v.to(attn.dtype)
        dev = _check_cuda(attn=attn, v=v, nbhd_idx=nbhd_idx)
        B, H, Nq, M = attn.shape
        Nk, c = v.shape[2], v.shape[3]
        _check_idx(nbhd_idx, B, Nq)
        with torch.cuda.device(dev):
            feat = torch.empty((B, H, Nq, c), dtype=attn.dtype, device=dev)
            plan = get_plan(nbhd_idx).struct(nbhd_idx, Nk, _dense(attn, c), False)
            N.check(N.LIB.affb200_av_fwd(N.p(attn), N.p(v), N.p(nbhd_idx), plan, N.p(feat), B, H, Nq, Nk, M, c,
                                         N.dtype_enum(attn), N.stream_ptr(dev)), "clusten_av_forward")
        ctx.save_for_backward(attn, v, nbhd_idx)
        return feat

    @staticmethod
    def backward(ctx, grad_feat):
        attn, v, nbhd_idx = ctx.saved_tensors
        grad_feat = grad_feat.contiguous()
        if grad_feat.dtype != attn.dtype:
            grad_feat = grad_feat.to(attn.dtype)
        dev = _check_cuda(d_feat=grad_feat)
        B, H, Nq, M = attn.shape
        Nk, c = v.shape[2], v.shape[3]
        with torch.cuda.device(dev):
            plan = get_plan(nbhd_idx).struct(nbhd_idx, Nk, _dense(attn, c), True)
            d_attn = torch.empty_like(attn)
            d_v = torch.empty_like(v)
            N.check(N.LIB.affb200_av_bwd(N.p(grad_feat), N.p(attn), N.p(v), N.p(nbhd_idx), plan,
                                         N.p(d_attn), N.p(d_v), B, H, Nq, Nk, M, c, N.dtype_enum(attn),
                                         N.stream_ptr(dev)), "clusten_av_backward")
        return d_attn, d_v, None


class CLUSTENWFFunction(Function):
    """feat_new[b,i,ic,c] = sum_j weights[b,i,j,ic] * feat[b,nbhd_idx[b,i,j],c]   (reference clusten.py:69-92)"""

    @staticmethod
    def forward(ctx, weights, feat, nbhd_idx):
        weights = weights.contiguous()
        feat = feat.contiguous()
        nbhd_idx = nbhd_idx.contiguous()
        if feat.dtype != weights.dtype:
            feat = feat.to(weights.dtype)
        dev = _check_cuda(weights=weights, feat=feat, nbhd_idx=nbhd_idx)
        B, Nq, M, IC = weights.shape
        Nk, C = feat.shape[1], feat.shape[2]
        _check_idx(nbhd_idx, B, Nq)
        with torch.cuda.device(dev):
            feat_new = torch.empty((B, Nq, IC, C), dtype=weights.dtype, device=dev)
            N.check(N.LIB.affb200_wf_fwd(N.p(weights), N.p(feat), N.p(nbhd_idx), None, N.p(feat_new), B, Nq, Nk, M, IC, C,
                                         N.dtype_enum(weights), N.stream_ptr(dev)), "clusten_wf_forward")
        ctx.save_for_backward(weights, feat, nbhd_idx)
        return feat_new

    @staticmethod
    def backward(ctx, grad_feat_new):
        weights, feat, nbhd_idx = ctx.saved_tensors
        grad_feat_new = grad_feat_new.contiguous()
        if grad_feat_new.dtype != weights.dtype:
            grad_feat_new = grad_feat_new.to(weights.dtype)
        dev = _check_cuda(d_feat_new=grad_feat_new)
        B, Nq, M, IC = weights.shape
        Nk, C = feat.shape[1], feat.shape[2]
        with torch.cuda.device(dev):
            plan = get_plan(nbhd_idx).struct(nbhd_idx, Nk, False, True)
            d_weights = torch.empty_like(weights)
            d_feat = torch.empty_like(feat)
            N.check(N.LIB.affb200_wf_bwd(N.p(grad_feat_new), N.p(weights), N.p(feat), N.p(nbhd_idx), plan,
                                         N.p(d_weights), N.p(d_feat), B, Nq, Nk, M, IC, C,
                                         N.dtype_enum(weights), N.stream_ptr(dev)), "clusten_wf_backward")
        return d_weights, d_feat, None
