"""Merge half of ClusterMerging as one operator (csrc/merge.cu).

  merge_wf(weights_table, pe_idx, learned_prob, cluster_mask, feat, member_idx) -> [B, N', 4*C]

equals, in the reference's terms (models/aff_transformer.py:343-361),

  weights = weights_table[pe_idx] * (learned_prob[member_idx] * cluster_mask);  CLUSTENWFFunction(weights, feat, member_idx)

without materialising the [B, N', M, 4] weights or their gradient: the kernels form each weight from its factors.
Gradients flow to weights_table (shared-memory histogram), learned_prob and feat.  Returns None for cases the kernels
do not cover (table not fp32 x 4, channel count not in {32..512 powers of two}); the caller then composes the
unfused operators.
"""
import torch
from torch.autograd import Function

from . import _native as N
from .plan import get_plan

_OK = (torch.float32, torch.float16, torch.bfloat16)
_CH = (32, 64, 128, 256, 512)


class MergeWFFunction(Function):
    @staticmethod
    def forward(ctx, table, lp, feat, pe_idx, mask, nbhd):
        B, Nq, M = nbhd.shape
        Nk, C = feat.shape[1], feat.shape[2]
        dev = feat.device
        tab = table.detach().contiguous()
        featc = feat.contiguous()
        lpc = lp.reshape(B, Nk).contiguous() if lp is not None else None
        pe_c, nb_c = pe_idx.contiguous(), nbhd.contiguous()
        mk_c = mask.contiguous() if mask is not None else None
        with torch.cuda.device(dev):
            out = torch.empty((B, Nq, 4 * C), dtype=torch.float32, device=dev)
            N.check(N.LIB.affb200_merge_fwd(N.p(tab), N.p(pe_c), N.p(nb_c), N.p(mk_c), N.p(lpc), N.p(featc), N.p(out),
                                            B, Nq, Nk, M, C, N.DTYPE_ENUM[lpc.dtype] if lpc is not None else 0,
                                            N.dtype_enum(featc), N.stream_ptr(dev)), "merge_fwd")
        ctx.save_for_backward(tab, lpc, featc, pe_c, mk_c, nb_c)
        ctx.nbhd_ref = nbhd          # the plan (transposed index) is keyed by the caller's tensor
        ctx.meta = (table.dtype, None if lp is None else (lp.dtype, lp.shape), feat.dtype)
        return out

    @staticmethod
    def backward(ctx, d_out):
        tab, lpc, featc, pe_c, mk_c, nb_c = ctx.saved_tensors
        tdt, lpm, fdt = ctx.meta
        B, Nq, M = nb_c.shape
        Nk, C = featc.shape[1], featc.shape[2]
        dev = featc.device
        d_out = d_out.contiguous().float()
        need_f = ctx.needs_input_grad[2]
        need_lp = lpc is not None and ctx.needs_input_grad[1]
        with torch.cuda.device(dev):
            off = ent = None
            if need_f:
                off, ent = get_plan(ctx.nbhd_ref).csr(nb_c, Nk)
            d_wt = torch.empty((B, Nq, M, 4), dtype=torch.float32, device=dev)
            d_lp = torch.empty((B, Nk), dtype=torch.float32, device=dev) if need_lp else None
            d_feat = torch.empty_like(featc) if need_f else None
            N.check(N.LIB.affb200_merge_bwd(N.p(d_out), N.p(tab), N.p(pe_c), N.p(nb_c), N.p(mk_c), N.p(lpc), N.p(featc),
                                            N.p(off), N.p(ent), N.p(d_wt), N.p(d_lp), N.p(d_feat), B, Nq, Nk, M, C,
                                            N.DTYPE_ENUM[lpc.dtype] if lpc is not None else 0, N.dtype_enum(featc),
                                            N.stream_ptr(dev)), "merge_bwd")
            d_table = None
            if ctx.needs_input_grad[0]:
                P = tab.shape[0]
                d_table = torch.empty((P, 4), dtype=torch.float32, device=dev)
                N.check(N.LIB.affb200_attn_bias_bwd(N.p(d_wt), N.p(pe_c), B, 4, Nq, M, P, 1, N.p(d_table), 0,
                                                    N.stream_ptr(dev)), "attn_bias_bwd")
                d_table = d_table.to(tdt)
        if need_lp:
            d_lp = d_lp.to(lpm[0]).reshape(lpm[1])
        return d_table, d_lp, d_feat, None, None, None


def merge_wf(weights_table, pe_idx, learned_prob, cluster_mask, feat, member_idx):
    """-> [B, N', 4*C] (fp32), or None when the case must take the unfused operators."""
    if not feat.is_cuda:
        raise RuntimeError("feat must be a CUDA tensor")
    if (weights_table.dtype != torch.float32 or weights_table.dim() != 2 or weights_table.shape[1] != 4
            or feat.dtype not in _OK or feat.shape[2] not in _CH
            or (learned_prob is not None and learned_prob.dtype not in _OK) or member_idx.shape[2] > 1024):
        return None
    return MergeWFFunction.apply(weights_table, learned_prob, feat, pe_idx, cluster_mask, member_idx)


class WeightNetFunction(Function):
    """weight_net(pre_table) = GELU(LayerNorm(Linear(pre_table))) as one kernel per direction (csrc/weight_net.cu)."""

    @staticmethod
    def forward(ctx, feat, w, b, gamma, beta, eps, dt):
        dev = feat.device
        P = feat.shape[0]
        args = [t.detach().float().contiguous() for t in (feat, w, b, gamma, beta)]
        with torch.cuda.device(dev):
            out = torch.empty((P, 4), dtype=torch.float32, device=dev)
            N.check(N.LIB.affb200_weight_net_fwd(*[N.p(t) for t in args], float(eps), P, N.DTYPE_ENUM[dt], N.p(out),
                                                 N.stream_ptr(dev)), "weight_net_fwd")
        ctx.save_for_backward(*args)
        ctx.meta = (float(eps), dt, w.dtype, b.dtype, gamma.dtype, beta.dtype)
        return out

    @staticmethod
    def backward(ctx, d_out):
        args = ctx.saved_tensors
        eps, dt, wdt, bdt, gdt, bedt = ctx.meta
        dev = d_out.device
        P = args[0].shape[0]
        d_out = d_out.float().contiguous()
        with torch.cuda.device(dev):
            grads = torch.empty(32, dtype=torch.float32, device=dev)
            work = torch.empty(max(1, N.LIB.affb200_weight_net_work_floats(P)), dtype=torch.float32, device=dev)
            N.check(N.LIB.affb200_weight_net_bwd(*[N.p(t) for t in args], eps, P, N.DTYPE_ENUM[dt], N.p(d_out),
                                                 N.p(grads), N.p(work), N.stream_ptr(dev)), "weight_net_bwd")
        return (None, grads[:20].view(4, 5).to(wdt), grads[20:24].to(bdt), grads[24:28].to(gdt), grads[28:32].to(bedt),
                None, None)


def weight_net_table(weight_net, pre_table):
    """`weight_net(pre_table)` for the reference's nn.Sequential(Linear(5, 4), LayerNorm(4), GELU) (aff_transformer.py:
    262-264) -> fp32 [P, 4]; any other module structure, or a table that needs a gradient, runs the module itself."""
    if not pre_table.is_cuda:
        raise RuntimeError("pre_table must be a CUDA tensor")
    mods = list(weight_net) if isinstance(weight_net, torch.nn.Sequential) else []
    ok = (len(mods) == 3 and isinstance(mods[0], torch.nn.Linear) and isinstance(mods[1], torch.nn.LayerNorm)
          and isinstance(mods[2], torch.nn.GELU) and mods[2].approximate == "none" and mods[0].bias is not None
          and tuple(mods[0].weight.shape) == (4, 5) and tuple(mods[1].normalized_shape) == (4,)
          and mods[1].elementwise_affine and mods[1].bias is not None
          and pre_table.dim() == 2 and pre_table.shape[1] == 5 and not pre_table.requires_grad)
    dt = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled("cuda") else pre_table.dtype
    if not ok or dt not in _OK:
        return weight_net(pre_table)
    return WeightNetFunction.apply(pre_table, mods[0].weight, mods[0].bias, mods[1].weight, mods[1].bias, mods[1].eps, dt)
