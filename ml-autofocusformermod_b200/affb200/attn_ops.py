"""Relative-position bias and lookup-table operators of the attention / merging glue (csrc/attn_glue.cu).

  attn_bias_add(attn, pe_table, pe_idx, mask)  attn[b,h,i,j] += pe_table[pe_idx[b,i,j], h] (+ -100 where mask == 0),
                                               in place on the QK output   (reference aff_transformer.py:126-137)
  table_gather(table, pe_idx)                  table[pe_idx] -> [B,N,M,IC]  (reference aff_transformer.py:343-349)

Forward values equal the reference's gather + add(s); the table gradients are accumulated per CTA in shared memory
instead of through autograd's index_put / scatter-add.
"""
import torch
from torch.autograd import Function

from . import _native as N


def _table_grad(d_vals, pe_idx, H, P, channel_last, out_dtype):
    B = pe_idx.shape[0]
    n, m = pe_idx.shape[1], pe_idx.shape[2]
    dev = d_vals.device
    d_vals = d_vals.contiguous()
    with torch.cuda.device(dev):
        d_table = torch.empty((P, H), dtype=torch.float32, device=dev)
        N.check(N.LIB.affb200_attn_bias_bwd(N.p(d_vals), N.p(pe_idx), B, H, n, m, P, 1 if channel_last else 0,
                                            N.p(d_table), N.dtype_enum(d_vals), N.stream_ptr(dev)), "attn_bias_bwd")
    return d_table.to(out_dtype)


class AttnBiasAddFunction(Function):
    @staticmethod
    def forward(ctx, attn, pe_table, pe_idx, mask):
        if not attn.is_cuda:
            raise RuntimeError("attn must be a CUDA tensor")
        if not attn.is_contiguous():
            raise RuntimeError("attn must be contiguous")
        B, H, n, m = attn.shape
        pe_idx = pe_idx.contiguous()
        mask = mask.contiguous() if mask is not None else None
        table = pe_table.detach().float().contiguous()
        dev = attn.device
        with torch.cuda.device(dev):
            N.check(N.LIB.affb200_attn_bias_add(N.p(attn), N.p(table), N.p(pe_idx), N.p(mask), B, H, n, m,
                                                N.dtype_enum(attn), N.stream_ptr(dev)), "attn_bias_add")
        ctx.mark_dirty(attn)
        ctx.save_for_backward(pe_idx)
        ctx.table_meta = (pe_table.shape[0], pe_table.dtype)
        return attn

    @staticmethod
    def backward(ctx, d_attn):
        (pe_idx,) = ctx.saved_tensors
        P, dt = ctx.table_meta
        d_table = None
        if ctx.needs_input_grad[1]:
            d_table = _table_grad(d_attn, pe_idx, d_attn.shape[1], P, False, dt)
        return d_attn, d_table, None, None


class TableGatherFunction(Function):
    @staticmethod
    def forward(ctx, table, pe_idx):
        if not table.is_cuda:
            raise RuntimeError("table must be a CUDA tensor")
        ctx.save_for_backward(pe_idx)
        ctx.table_meta = (table.shape[0], table.dtype)
        tab = table.detach().contiguous()
        idx = pe_idx.contiguous()
        dev = table.device
        with torch.cuda.device(dev):
            out = torch.empty((*pe_idx.shape, table.shape[1]), dtype=table.dtype, device=dev)
            N.check(N.LIB.affb200_table_gather(N.p(tab), N.p(idx), N.p(out), idx.numel(), table.shape[1],
                                               table.element_size(), N.stream_ptr(dev)), "table_gather")
        return out

    @staticmethod
    def backward(ctx, d_out):
        (pe_idx,) = ctx.saved_tensors
        P, dt = ctx.table_meta
        return _table_grad(d_out, pe_idx.contiguous(), d_out.shape[-1], P, True, dt), None


def attn_bias_add(attn, pe_table, pe_idx, mask=None):
    return AttnBiasAddFunction.apply(attn, pe_table, pe_idx, mask)


def table_gather(table, pe_idx):
    return TableGatherFunction.apply(table, pe_idx)
