"""Fused neighbourhood attention (csrc/fused_attn.cu): the whole of ClusterAttention between the q/kv Linears and the
output projection (reference models/aff_transformer.py:103-156) as one autograd.Function.

    out = fused_cluster_attention(q_lin, kv_lin, num_heads, scale, nbhd_idx, pe_idx, cluster_mask, pe_table,
                                  blank_k, blank_v)            # -> [B, N, H*c], or None when the case is unsupported

q_lin [B,N,H*c] and kv_lin [B,N,2*H*c] are the Linear outputs themselves (kv laid out [H][2][c] per token as in
`kv.view(b, n, h, 2, c_)`, aff_transformer.py:115); no permuted copies are made in either direction.  Supported:
f16/bf16 activations, head dim 16 or 32, bias table of at most 32768 rows, neighbourhoods whose tiles reference at
most 255 distinct rows (any clustered index tensor).  Everything else returns None and the caller uses the unfused
operators.

Two kernel families sit behind the call: the per-query-tile kernels (csrc/fused_attn.cu) and, for at most 255 tokens
per image and M <= 96, the per-(image, head) kernels that keep the image's K/V in shared memory and need no scatter
pass in the backward (csrc/fused_attn_img.cu).  AFFB200_FUSED_IMG=0 forces the first family.
"""
import os
import ctypes
import weakref

import torch
from torch.autograd import Function

from .linear_ops import col_sum as _col_sum
from . import _native as N
from .plan import get_plan, is_static, guard_add


class _View(ctypes.Structure):
    _fields_ = [("p", ctypes.c_void_p), ("sB", ctypes.c_int64), ("sN", ctypes.c_int64), ("sH", ctypes.c_int64)]


def _view(t, sB, sN, sH, offset_elems=0):
    return ctypes.byref(_View(t.data_ptr() + offset_elems * t.element_size(), sB, sN, sH))


def _packed_pe(plan, pe_idx, mask):
    """uint16 bias indices (pe | masked << 15), built once per (pe_idx, mask) pair and cached on the plan."""
    key = (id(pe_idx), pe_idx._version, None if mask is None else id(mask))
    hit = plan._pe.get(key) if hasattr(plan, "_pe") else None
    if hit is not None and hit[0]() is pe_idx:
        return hit[1]
    import weakref
    dev = pe_idx.device
    pe_c = pe_idx.contiguous()
    mask_c = mask.contiguous() if mask is not None else None
    with torch.cuda.device(dev):
        out = torch.empty(pe_c.shape, dtype=torch.int16, device=dev)
        bad = torch.zeros(1, dtype=torch.int32, device=dev)
        N.check(N.LIB.affb200_fused_attn_pack_pe(N.p(pe_c), N.p(mask_c), pe_c.numel(), N.p(out), N.p(bad),
                                                 N.stream_ptr(dev)), "pack_pe")
        if is_static():
            guard_add(bad)   # no read-back under graph capture: an out-of-range bias index trips the guard instead
            res = out
        else:
            res = out if int(bad.item()) == 0 else None
    if not hasattr(plan, "_pe"):
        plan._pe = {}
    plan._pe.clear()
    plan._pe[key] = (weakref.ref(pe_idx), res)
    return res


def _img_prep(plan, nbhd_idx, q_pe):
    """(col, dupf) of affb200_fused_attn_img_prepare for this (nbhd_idx, packed bias/mask) pair, cached on the plan."""
    hit = getattr(plan, "_img", None)
    if hit is not None and hit[0] == nbhd_idx._version and hit[1] is q_pe:
        return hit[2], hit[3]
    B, n, M = nbhd_idx.shape
    dev = nbhd_idx.device
    idx = nbhd_idx.contiguous()
    with torch.cuda.device(dev):
        col = torch.empty((B, n, M), dtype=torch.uint8, device=dev)
        dupf = torch.empty((B, (n + 31) // 32), dtype=torch.uint8, device=dev)
        N.check(N.LIB.affb200_fused_attn_img_prepare(N.p(idx), N.p(q_pe), B, n, M, N.p(col), N.p(dupf),
                                                     N.stream_ptr(dev)), "fused_attn_img_prepare")
    plan._img = (nbhd_idx._version, q_pe, col, dupf)
    return col, dupf


_feat_cache = {}


def _pe_features(pre_table, dt):
    """(feature table in the compute dtype, feat8 [P, 8] fp32 = those values padded with a column of ones) for the
    Linear that generates the bias table (reference aff_transformer.py:78,125: pos_embed(pre_table)); cached."""
    key = (pre_table.data_ptr(), pre_table._version, dt)
    hit = _feat_cache.get(key)
    if hit is not None and hit[2]() is not pre_table:   # a freed (compacted) table's address reused by another one
        hit = None
    if hit is None:
        src = pre_table.detach()
        lp = src.to(dt) if dt is not None else src.float()
        P, F = lp.shape
        feat8 = torch.zeros((P, 8), dtype=torch.float32, device=src.device)
        feat8[:, :F] = lp.float()
        feat8[:, F] = 1.0
        if len(_feat_cache) > 16:
            _feat_cache.clear()
        hit = _feat_cache[key] = (lp.contiguous(), feat8, weakref.ref(pre_table))
    return hit[:2]


_sub_cache = {}


def _sub_features(pe_feat, w, tw, r):
    """rows of the feature table for offsets |dx|, |dy| <= r (the central (2r+1)^2 of the tw x tw table); cached"""
    lp = pe_feat[0]
    key = (lp.data_ptr(), lp._version, w, tw, r)
    hit = _sub_cache.get(key)
    if hit is None or hit[1]() is not lp:
        if len(_sub_cache) > 16:
            _sub_cache.clear()
        sub = lp.view(tw, tw, lp.shape[1])[w - r:w + r + 1, w - r:w + r + 1].reshape(-1, lp.shape[1]).contiguous()
        hit = _sub_cache[key] = (sub, weakref.ref(lp))
    return hit[0]


def _table_of(pe_table, pe_w, pe_b, pe_feat):
    """fp32 [P, H] bias table: given, or pos_embed evaluated the way autocast evaluates the nn.Linear."""
    if pe_w is None:
        return pe_table.detach().float().contiguous()
    from .linear_ops import low_precision
    lp, _ = pe_feat
    dt = lp.dtype
    with torch.autocast("cuda", enabled=False):
        return torch.nn.functional.linear(lp, low_precision(pe_w, dt).detach(), low_precision(pe_b, dt).detach()
                                          ).float().contiguous()


def _pe_grad(ds, q_pe, pe_feat, B, H, NM, F, w_dt, b_dt):
    """(dW [H, F], db [H]) of the table-generating Linear from dS [B, H, N*M] (affb200_pe_feat_grad)."""
    dev = ds.device
    _, feat8 = pe_feat
    out = torch.empty((H, 8), dtype=torch.float32, device=dev)
    work = torch.empty(max(1, N.LIB.affb200_pe_feat_grad_work_floats(B, H, NM)), dtype=torch.float32, device=dev)
    N.check(N.LIB.affb200_pe_feat_grad(N.p(ds), N.p(q_pe), N.p(feat8), B, H, NM, N.p(out), N.p(work), N.dtype_enum(ds),
                                       N.stream_ptr(dev)), "pe_feat_grad")
    return out[:, :F].to(w_dt), out[:, F].to(b_dt)


def _use_img(n, M):
    return n <= N.FUSED_IMG_MAX_N and M <= 96 and M % 2 == 0 and os.environ.get("AFFB200_FUSED_IMG", "1") != "0"


class FusedClusterAttentionImg(Function):
    """Small token sets: one CTA per (image, head), K/V resident, dK/dV accumulated on chip (fused_attn_img.cu)."""

    @staticmethod
    def forward(ctx, q_lin, kv_lin, pe_table, pe_w, pe_b, blank_k, blank_v, nbhd_idx, q_pe, col, dupf, pe_feat, meta):
        H, scale = meta
        B, n, hc = q_lin.shape
        c = hc // H
        M = nbhd_idx.shape[2]
        dev = q_lin.device
        q_lin = q_lin.contiguous()
        kv_lin = kv_lin.contiguous()
        table = _table_of(pe_table, pe_w, pe_b, pe_feat)
        bk = blank_k.detach().float().contiguous()
        bv = blank_v.detach().float().contiguous()
        with torch.cuda.device(dev):
            out = torch.empty_like(q_lin)
            lse = torch.empty((B, H, n), dtype=torch.float32, device=dev)
            N.check(N.LIB.affb200_fused_attn_img_fwd(
                _view(q_lin, n * hc, hc, c), _view(kv_lin, n * 2 * hc, 2 * hc, 2 * c),
                _view(kv_lin, n * 2 * hc, 2 * hc, 2 * c, c), _view(out, n * hc, hc, c), N.p(table), N.p(bk), N.p(bv),
                N.p(dupf), N.p(col), N.p(q_pe), N.p(lse), B, H, n, M, c, float(scale), N.dtype_enum(q_lin),
                N.stream_ptr(dev)), "fused_attn_img_fwd")
        ctx.save_for_backward(q_lin, kv_lin, out, lse, table, bk, bv, q_pe, col, dupf)
        ctx.M = M
        ctx.meta = (H, scale, None if pe_table is None else pe_table.dtype, blank_k.dtype, table.shape[0])
        ctx.pe_lin = None if pe_w is None else (pe_feat, pe_w.shape[1], pe_w.dtype, pe_b.dtype)
        return out

    @staticmethod
    def backward(ctx, d_out):
        q_lin, kv_lin, out, lse, table, bk, bv, q_pe, col, dupf = ctx.saved_tensors
        H, scale, table_dt, blank_dt, P = ctx.meta
        B, n, hc = q_lin.shape
        c = hc // H
        M = ctx.M
        dev = q_lin.device
        d_out = d_out.contiguous()
        if d_out.dtype != q_lin.dtype:
            d_out = d_out.to(q_lin.dtype)
        with torch.cuda.device(dev):
            dq = torch.empty_like(q_lin)
            dkv = torch.empty_like(kv_lin)
            ds = torch.empty((B, H, n, M), dtype=q_lin.dtype, device=dev)
            blank_part = torch.empty((B, H, 2, c), dtype=torch.float32, device=dev)
            d_table = torch.empty((P, H), dtype=torch.float32, device=dev) if ctx.pe_lin is None else None
            N.check(N.LIB.affb200_fused_attn_img_bwd(
                _view(q_lin, n * hc, hc, c), _view(kv_lin, n * 2 * hc, 2 * hc, 2 * c),
                _view(kv_lin, n * 2 * hc, 2 * hc, 2 * c, c), _view(out, n * hc, hc, c), _view(d_out, n * hc, hc, c),
                _view(dq, n * hc, hc, c), _view(dkv, n * 2 * hc, 2 * hc, 2 * c), _view(dkv, n * 2 * hc, 2 * hc, 2 * c, c),
                N.p(table), N.p(bk), N.p(bv), N.p(dupf), N.p(col), N.p(q_pe), N.p(lse), N.p(ds), N.p(blank_part),
                N.p(d_table), P, B, H, n, M, c, float(scale), N.dtype_enum(q_lin), N.stream_ptr(dev)),
                "fused_attn_img_bwd")
            db = _col_sum(blank_part.view(B, H * 2 * c)).view(H, 2, c)   # fixed-order fp32 column sums
            dw = dpb = None
            if ctx.pe_lin is not None:
                pe_feat, F, w_dt, b_dt = ctx.pe_lin
                dw, dpb = _pe_grad(ds, q_pe, pe_feat, B, H, n * M, F, w_dt, b_dt)
        return (dq, dkv, None if d_table is None else d_table.to(table_dt), dw, dpb,
                db[:, 0].reshape(-1).to(blank_dt), db[:, 1].reshape(-1).to(blank_dt), None, None, None, None, None, None)


class FusedClusterAttention(Function):
    @staticmethod
    def forward(ctx, q_lin, kv_lin, pe_table, pe_w, pe_b, blank_k, blank_v, nbhd_idx, q_pe, pe_feat, meta):
        H, scale, umax = meta
        B, n, hc = q_lin.shape
        c = hc // H
        M = nbhd_idx.shape[2]
        dev = q_lin.device
        q_lin = q_lin.contiguous()
        kv_lin = kv_lin.contiguous()
        table = _table_of(pe_table, pe_w, pe_b, pe_feat)
        bk = blank_k.detach().float().contiguous()
        bv = blank_v.detach().float().contiguous()
        plan = get_plan(nbhd_idx)
        with torch.cuda.device(dev):
            out = torch.empty_like(q_lin)
            lse = torch.empty((B, H, n), dtype=torch.float32, device=dev)
            st = plan.struct(nbhd_idx, n, True, False)
            N.check(N.LIB.affb200_fused_attn_fwd(
                _view(q_lin, n * hc, hc, c), _view(kv_lin, n * 2 * hc, 2 * hc, 2 * c),
                _view(kv_lin, n * 2 * hc, 2 * hc, 2 * c, c), _view(out, n * hc, hc, c), N.p(table), N.p(bk), N.p(bv), st,
                N.p(q_pe), umax, N.p(lse), B, H, n, M, c, float(scale), N.dtype_enum(q_lin), N.stream_ptr(dev)),
                "fused_attn_fwd")
        ctx.save_for_backward(q_lin, kv_lin, out, lse, table, bk, bv, nbhd_idx, q_pe)
        ctx.meta = (H, scale, umax, None if pe_table is None else pe_table.dtype, blank_k.dtype, table.shape[0])
        ctx.pe_lin = None if pe_w is None else (pe_feat, pe_w.shape[1], pe_w.dtype, pe_b.dtype)
        return out

    @staticmethod
    def backward(ctx, d_out):
        q_lin, kv_lin, out, lse, table, bk, bv, nbhd_idx, q_pe = ctx.saved_tensors
        H, scale, umax, table_dt, blank_dt, P = ctx.meta
        B, n, hc = q_lin.shape
        c = hc // H
        M = nbhd_idx.shape[2]
        dev = q_lin.device
        d_out = d_out.contiguous()
        if d_out.dtype != q_lin.dtype:
            d_out = d_out.to(q_lin.dtype)
        plan = get_plan(nbhd_idx)
        with torch.cuda.device(dev):
            st = plan.struct(nbhd_idx, n, True, True)
            T = plan.stats["q"]["T"]
            nt = (n + T - 1) // T
            dq = torch.empty_like(q_lin)
            dkv = torch.empty_like(kv_lin)
            ds = torch.empty((B, H, n, M), dtype=q_lin.dtype, device=dev)
            pw = torch.empty((B, H, n, M), dtype=q_lin.dtype, device=dev)
            blank_part = torch.empty((B, nt, H, 2, c), dtype=torch.float32, device=dev)
            d_table = torch.empty((P, H), dtype=torch.float32, device=dev) if ctx.pe_lin is None else None
            N.check(N.LIB.affb200_fused_attn_bwd(
                _view(q_lin, n * hc, hc, c), _view(kv_lin, n * 2 * hc, 2 * hc, 2 * c),
                _view(kv_lin, n * 2 * hc, 2 * hc, 2 * c, c), _view(out, n * hc, hc, c), _view(d_out, n * hc, hc, c),
                _view(dq, n * hc, hc, c), _view(dkv, n * 2 * hc, 2 * hc, 2 * c), _view(dkv, n * 2 * hc, 2 * hc, 2 * c, c),
                N.p(table), N.p(bk), N.p(bv), st, N.p(q_pe), umax, N.p(lse), N.p(ds), N.p(pw), N.p(blank_part),
                N.p(d_table), P, B, H, n, M, c, float(scale), N.dtype_enum(q_lin), N.stream_ptr(dev)), "fused_attn_bwd")
            db = _col_sum(blank_part.view(B * nt, H * 2 * c)).view(H, 2, c)   # fixed-order fp32 column sums
            dw = dpb = None
            if ctx.pe_lin is not None:
                pe_feat, F, w_dt, b_dt = ctx.pe_lin
                dw, dpb = _pe_grad(ds, q_pe, pe_feat, B, H, n * M, F, w_dt, b_dt)
        return (dq, dkv, None if d_table is None else d_table.to(table_dt), dw, dpb,
                db[:, 0].reshape(-1).to(blank_dt), db[:, 1].reshape(-1).to(blank_dt), None, None, None, None)


class FusedClusterAttentionCl(Function):
    """Block-structured neighbourhoods (fused_cl.py): union blocks staged by TMA, bias from pos_u - pos_i, dK/dV
    accumulated per key block on chip (csrc/fused_cl_fwd.cu, fused_cl_bwd.cu)."""

    @staticmethod
    def forward(ctx, q_lin, kv_lin, pe_table, pe_w, pe_b, blank_k, blank_v, nbhd_idx, pe_feat, clp, meta):
        H, scale = meta
        B, n, hc = q_lin.shape
        c = hc // H
        dev = q_lin.device
        q_lin = q_lin.contiguous()
        kv_lin = kv_lin.contiguous()
        w_tab, tw_tab = clp.w, clp.tw
        if clp.sub_r is not None and pe_w is not None and clp.q_pe is None:
            # forward only over a large table (MixRes: 1023 x 1023 rows): every offset of this index set lies within
            # clp.sub_r of the centre, so pos_embed is evaluated on those rows alone; the kernel forms its bias index
            # from position differences with the geometry it is given
            w_tab, tw_tab = clp.sub_r, 2 * clp.sub_r + 1
            table = _table_of(None, pe_w, pe_b, (_sub_features(pe_feat, clp.w, clp.tw, clp.sub_r), None))
        else:
            table = _table_of(pe_table, pe_w, pe_b, pe_feat)
        bk = blank_k.detach().float().contiguous()
        bv = blank_v.detach().float().contiguous()
        with torch.cuda.device(dev):
            out = torch.empty_like(q_lin)
            lse = torch.empty((B, H, n), dtype=torch.float32, device=dev)
            from .plan import guard_tensor
            N.check(N.LIB.affb200_fused_cl_fwd(
                _view(q_lin, n * hc, hc, c), _view(kv_lin, n * 2 * hc, 2 * hc, 2 * c),
                _view(kv_lin, n * 2 * hc, 2 * hc, 2 * c, c), _view(out, n * hc, hc, c), N.p(table), N.p(bk), N.p(bv),
                clp.byref(), N.p(lse), N.p(guard_tensor(dev)), B, H, n, c, w_tab, tw_tab, float(scale),
                N.dtype_enum(q_lin), N.stream_ptr(dev)), "fused_cl_fwd")
        if tw_tab != clp.tw:
            ctx.mark_non_differentiable(out)
            return out
        ctx.save_for_backward(q_lin, kv_lin, out, lse, table, bk, bv)
        ctx.clp = clp
        ctx.M = nbhd_idx.shape[2]
        ctx.meta = (H, scale, None if pe_table is None else pe_table.dtype, blank_k.dtype, table.shape[0])
        ctx.pe_lin = None if pe_w is None else (pe_feat, pe_w.shape[1], pe_w.dtype, pe_b.dtype)
        return out

    @staticmethod
    def backward(ctx, d_out):
        q_lin, kv_lin, out, lse, table, bk, bv = ctx.saved_tensors
        H, scale, table_dt, blank_dt, P = ctx.meta
        clp, M = ctx.clp, ctx.M
        B, n, hc = q_lin.shape
        c = hc // H
        dev = q_lin.device
        d_out = d_out.contiguous()
        if d_out.dtype != q_lin.dtype:
            d_out = d_out.to(q_lin.dtype)
        with torch.cuda.device(dev):
            from .plan import guard_tensor
            dq = torch.empty_like(q_lin)
            dkv = torch.empty_like(kv_lin)
            ds = torch.empty((B, H, n, M), dtype=q_lin.dtype, device=dev)
            delta = torch.empty((B, H, n), dtype=torch.float32, device=dev)
            # Optionally (AFFB200_BWD_L2_MB, see _bwd_chunk) the batch goes through in chunks of images small enough for the
            # 126 MB L2, so that the key-major kernel and the bias gradient find Q / K|V / dO / dS of the chunk on chip.
            cb = _bwd_chunk(B, n, hc, H, M, q_lin.element_size()) if ctx.pe_lin is not None else B
            grids = [int(N.LIB.affb200_fused_cl_bwd_grid(min(cb, B - b0), H, n)) for b0 in range(0, B, cb)]
            blank_part = torch.empty((sum(grids) * 8, H, 2, c), dtype=torch.float32, device=dev)
            dw = dpb = d_table = None
            es, row0 = q_lin.element_size(), 0
            for ci, b0 in enumerate(range(0, B, cb)):
                bn = min(cb, B - b0)
                V = lambda t, w_, off=0: _view(t, n * w_, w_, (2 * c if w_ == 2 * hc else c), b0 * n * w_ + off)
                N.check(N.LIB.affb200_fused_cl_bwd(
                    V(q_lin, hc), V(kv_lin, 2 * hc), V(kv_lin, 2 * hc, c), V(out, hc), V(d_out, hc),
                    V(dq, hc), V(dkv, 2 * hc), V(dkv, 2 * hc, c),
                    N.p(table), N.p(bk), N.p(bv), clp.byref_images(b0), N.p(lse[b0:]), N.p(delta[b0:]), N.p(ds[b0:]),
                    N.p(blank_part[row0:]), N.p(guard_tensor(dev)), bn, H, n, M, c, clp.w, clp.tw, float(scale),
                    N.dtype_enum(q_lin), N.stream_ptr(dev)), "fused_cl_bwd")
                row0 += grids[ci] * 8
                if ctx.pe_lin is not None:
                    pe_feat, F, w_dt, b_dt = ctx.pe_lin
                    w1, b1 = _pe_grad(ds[b0:b0 + bn], clp.q_pe[b0:b0 + bn], pe_feat, bn, H, n * M, F, torch.float32, torch.float32)
                    dw, dpb = (w1, b1) if dw is None else (dw + w1, dpb + b1)     # fixed order over the chunks
            db = _col_sum(blank_part.view(-1, H * 2 * c)).view(H, 2, c)   # fixed-order fp32 column sums
            if ctx.pe_lin is not None:
                dw, dpb = dw.to(ctx.pe_lin[2]), dpb.to(ctx.pe_lin[3])
            else:
                d_table = torch.empty((P, H), dtype=torch.float32, device=dev)
                N.check(N.LIB.affb200_table_grad_u16(N.p(ds), N.p(clp.q_pe), B, H, n * M, P, N.p(d_table),
                                                     N.dtype_enum(ds), N.stream_ptr(dev)), "table_grad_u16")
        return (dq, dkv, None if d_table is None else d_table.to(table_dt), dw, dpb,
                db[:, 0].reshape(-1).to(blank_dt), db[:, 1].reshape(-1).to(blank_dt), None, None, None, None)


def _table_t(table, w, tw, r, dt):
    """[H][pad8((2r+1)^2)] table of dtype dt for the resident-key dense kernels: the rows of `table` [tw*tw, H] fp32 for
    offsets |dx|, |dy| <= r, one contiguous row per head."""
    H = table.shape[1]
    n = (2 * r + 1) ** 2
    out = torch.zeros((H, (n + 7) // 8 * 8), dtype=dt, device=table.device)
    sub = table.view(tw, tw, H)[w - r:w + r + 1, w - r:w + r + 1] if r < w else table.view(tw, tw, H)
    out[:, :n] = sub.reshape(n, H).t()
    return out


class FusedClusterAttentionCld(Function):
    """Block-structured neighbourhoods of images with at most 128 tokens: dense masked attention over the image with
    keys, positions and the bias table resident in shared memory (csrc/fused_cld.cu)."""

    @staticmethod
    def forward(ctx, q_lin, kv_lin, pe_table, pe_w, pe_b, blank_k, blank_v, nbhd_idx, pe_feat, clp, meta):
        H, scale, r = meta
        B, n, hc = q_lin.shape
        c = hc // H
        M = nbhd_idx.shape[2]
        dev = q_lin.device
        q_lin = q_lin.contiguous()
        kv_lin = kv_lin.contiguous()
        table = _table_of(pe_table, pe_w, pe_b, pe_feat)
        tt = _table_t(table, clp.w, clp.tw, r, q_lin.dtype)
        bk = blank_k.detach().float().contiguous()
        bv = blank_v.detach().float().contiguous()
        blk, pos16 = clp.keep[0], clp.keep[4]
        with torch.cuda.device(dev):
            out = torch.empty_like(q_lin)
            lse = torch.empty((B, H, n), dtype=torch.float32, device=dev)
            N.check(N.LIB.affb200_fused_cld_fwd(
                _view(q_lin, n * hc, hc, c), _view(kv_lin, n * 2 * hc, 2 * hc, 2 * c),
                _view(kv_lin, n * 2 * hc, 2 * hc, 2 * c, c), _view(out, n * hc, hc, c), N.p(tt), tt.shape[1], N.p(bk),
                N.p(bv), N.p(blk), N.p(pos16), clp.npad, N.p(lse), B, H, n, M, c, r, 2 * r + 1, float(scale),
                N.dtype_enum(q_lin), N.stream_ptr(dev)), "fused_cld_fwd")
        ctx.save_for_backward(q_lin, kv_lin, out, lse, tt, bk, bv)
        ctx.clp = clp
        ctx.M = M
        ctx.meta = (H, scale, r, None if pe_table is None else pe_table.dtype, blank_k.dtype, table.shape[0])
        ctx.pe_lin = None if pe_w is None else (pe_feat, pe_w.shape[1], pe_w.dtype, pe_b.dtype)
        return out

    @staticmethod
    def backward(ctx, d_out):
        q_lin, kv_lin, out, lse, tt, bk, bv = ctx.saved_tensors
        H, scale, r, table_dt, blank_dt, P = ctx.meta
        clp, M = ctx.clp, ctx.M
        B, n, hc = q_lin.shape
        c = hc // H
        dev = q_lin.device
        d_out = d_out.contiguous()
        if d_out.dtype != q_lin.dtype:
            d_out = d_out.to(q_lin.dtype)
        blk, pos16 = clp.keep[0], clp.keep[4]
        with torch.cuda.device(dev):
            dq = torch.empty_like(q_lin)
            dkv = torch.empty_like(kv_lin)
            ds = torch.empty((B, H, n, M), dtype=q_lin.dtype, device=dev)
            blank_part = torch.empty((B, H, 2, c), dtype=torch.float32, device=dev)
            N.check(N.LIB.affb200_fused_cld_bwd(
                _view(q_lin, n * hc, hc, c), _view(kv_lin, n * 2 * hc, 2 * hc, 2 * c),
                _view(kv_lin, n * 2 * hc, 2 * hc, 2 * c, c), _view(out, n * hc, hc, c), _view(d_out, n * hc, hc, c),
                _view(dq, n * hc, hc, c), _view(dkv, n * 2 * hc, 2 * hc, 2 * c), _view(dkv, n * 2 * hc, 2 * hc, 2 * c, c),
                N.p(tt), tt.shape[1], N.p(bk), N.p(bv), N.p(blk), N.p(pos16), clp.npad, N.p(lse), N.p(ds),
                N.p(blank_part), B, H, n, M, c, r, 2 * r + 1, float(scale), N.dtype_enum(q_lin), N.stream_ptr(dev)),
                "fused_cld_bwd")
            db = _col_sum(blank_part.view(B, H * 2 * c)).view(H, 2, c)   # fixed-order fp32 column sums
            dw = dpb = d_table = None
            if ctx.pe_lin is not None:
                pe_feat, F, w_dt, b_dt = ctx.pe_lin
                dw, dpb = _pe_grad(ds, clp.q_pe, pe_feat, B, H, n * M, F, w_dt, b_dt)
            else:
                d_table = torch.empty((P, H), dtype=torch.float32, device=dev)
                N.check(N.LIB.affb200_table_grad_u16(N.p(ds), N.p(clp.q_pe), B, H, n * M, P, N.p(d_table),
                                                     N.dtype_enum(ds), N.stream_ptr(dev)), "table_grad_u16")
        return (dq, dkv, None if d_table is None else d_table.to(table_dt), dw, dpb,
                db[:, 0].reshape(-1).to(blank_dt), db[:, 1].reshape(-1).to(blank_dt), None, None, None, None)


def _bwd_chunk(B, n, hc, H, M, es):
    """images per backward chunk: the chunk's operands (q, kv, out, d_out, dq, dkv, dS) within AFFB200_BWD_L2_MB.
    Off by default (0): measured inside the graph-replayed step, chunks of 72 / 110 MB cost 36.2 / 35.7 ms per step against
    34.2 ms for the whole batch at once — the partial last wave of every chunk launch outweighs the second read of the
    operands from HBM (1.98x the float-operand bytes, profiles/r02_ncu_traffic.json)."""
    mb = float(os.environ.get("AFFB200_BWD_L2_MB", "0"))
    if mb <= 0:
        return B
    per_image = n * hc * es * 8 + n * H * M * es + n * M * 2
    return max(1, min(B, int(mb * 1e6 // per_image)))


def _cld_radius(clp, n, c):
    """table radius for the resident-key dense kernels, or None when they do not apply"""
    from . import fused_cl
    if n > fused_cl.CLD_MAX_N or c != 32 or os.environ.get("AFFB200_FUSED_CLD", "1") == "0":
        return None
    r = clp.sub_r if clp.sub_r is not None else clp.w
    return r if (2 * r + 1) ** 2 <= 16384 else None


_ctypes_sigs_done = False


trace = None   # set to a list to record (tokens, heads, kernel family | None) of every call (bench: config.fused_stages)


def _note(n, H, family):
    if trace is not None:
        trace.append((int(n), int(H), family))


def fused_cluster_attention(q_lin, kv_lin, num_heads, scale, nbhd_idx, pe_idx, cluster_mask, pe_table, blank_k,
                            blank_v, pos_embed=None):
    """Returns out [B, N, H*c], or None when this case must take the unfused operators.

    pos_embed = (pre_table [P, F<=5], weight [H, F], bias [H]) may be given instead of pe_table (pass None): the table
    is then pos_embed evaluated inside (as autocast evaluates the nn.Linear) and the backward produces d weight /
    d bias directly from dS, without the [P, H] table gradient."""
    if not q_lin.is_cuda or q_lin.dtype not in (torch.float16, torch.bfloat16) or kv_lin.dtype != q_lin.dtype:
        return None
    B, n, hc = q_lin.shape
    c = hc // num_heads
    pe_w = pe_b = pe_feat = None
    if pos_embed is not None:
        pre_table, pe_w, pe_b = pos_embed
        P = pre_table.shape[0]
        if pre_table.shape[1] > 5 or pe_b is None or (n * nbhd_idx.shape[2]) % 8 != 0 \
                or os.environ.get("AFFB200_PE_LINEAR", "1") == "0":
            pe_table = torch.nn.functional.linear(pre_table, pe_w, pe_b)   # the stock path: table through autograd
            pe_w = pe_b = None
        else:
            dt = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled("cuda") else None
            pe_feat = _pe_features(pre_table, dt)
            pe_table = None
    else:
        P = pe_table.shape[0]
    if c not in (16, 32) or nbhd_idx.shape[2] > N.PLAN_MAX_M:
        _note(n, num_heads, None)
        return None
    plan = get_plan(nbhd_idx)
    cl_first = os.environ.get("AFFB200_CL_FIRST", "0") != "0"   # block-structured kernels also for small token sets
    from . import fused_cl
    needs_grad = torch.is_grad_enabled() and any(
        t is not None and t.requires_grad for t in (q_lin, kv_lin, pe_table, pe_w, pe_b, blank_k, blank_v))
    if n <= fused_cl.CLD_MAX_N and c == 32:      # small images: dense masked attention, everything resident
        clp = fused_cl.cl_plan(nbhd_idx, cluster_mask, pe_idx, P)
        r = None if clp is None else _cld_radius(clp, n, c)
        if r is not None and (pe_w is None or (n * nbhd_idx.shape[2]) % 8 == 0) and (clp.q_pe is not None or not needs_grad):
            _note(n, num_heads, "resident-key dense")
            return FusedClusterAttentionCld.apply(q_lin, kv_lin, pe_table, pe_w, pe_b, blank_k, blank_v, nbhd_idx,
                                                  pe_feat, clp, (num_heads, scale, r))
    if (cl_first or not _use_img(n, nbhd_idx.shape[2])) and num_heads % 2 == 0 and num_heads * c <= 512:
        clp = fused_cl.cl_plan(nbhd_idx, cluster_mask, pe_idx, P)
        if clp is not None and (pe_w is None or (n * nbhd_idx.shape[2]) % 8 == 0) and (clp.q_pe is not None or not needs_grad):
            _note(n, num_heads, "block-structured")
            return FusedClusterAttentionCl.apply(q_lin, kv_lin, pe_table, pe_w, pe_b, blank_k, blank_v, nbhd_idx,
                                                 pe_feat, clp, (num_heads, scale))
    if P > 32768:     # the row-union and resident-key kernels pack the bias index into 15 bits
        _note(n, num_heads, None)
        return None
    if _use_img(n, nbhd_idx.shape[2]):
        q_pe = _packed_pe(plan, pe_idx, cluster_mask)
        if q_pe is None:
            _note(n, num_heads, None)
            return None
        col, dupf = _img_prep(plan, nbhd_idx, q_pe)
        _note(n, num_heads, "resident-key")
        return FusedClusterAttentionImg.apply(q_lin, kv_lin, pe_table, pe_w, pe_b, blank_k, blank_v, nbhd_idx, q_pe, col,
                                              dupf, pe_feat, (num_heads, scale))
    if plan.qplan(nbhd_idx, n) is None or plan.kplan(nbhd_idx, n) is None:
        _note(n, num_heads, None)
        return None
    umax = plan.stats["q"]["umax"]
    if umax > N.PLAN_UMAX - 1:
        _note(n, num_heads, None)
        return None
    q_pe = _packed_pe(plan, pe_idx, cluster_mask)
    if q_pe is None:
        _note(n, num_heads, None)
        return None
    _note(n, num_heads, "row-union")
    return FusedClusterAttention.apply(q_lin, kv_lin, pe_table, pe_w, pe_b, blank_k, blank_v, nbhd_idx, q_pe, pe_feat,
                                       (num_heads, scale, umax))


def _raw_pe_grad(pre_table, ds, q_pe, B, H, NM, keep):
    """Closure enqueueing affb200_pe_feat_grad on preallocated buffers (None when the table gradient is wanted)."""
    if pre_table is None or pre_table.shape[1] > 5 or NM % 8 != 0:
        return None
    _, feat8 = _pe_features(pre_table, ds.dtype)
    dev = ds.device
    out = torch.empty((H, 8), dtype=torch.float32, device=dev)
    work = torch.empty(max(1, N.LIB.affb200_pe_feat_grad_work_floats(B, H, NM)), dtype=torch.float32, device=dev)
    keep += (feat8, out, work)
    de, sp = N.dtype_enum(ds), N.stream_ptr(dev)
    return lambda: N.check(N.LIB.affb200_pe_feat_grad(N.p(ds), N.p(q_pe), N.p(feat8), B, H, NM, N.p(out), N.p(work), de,
                                                      sp), "pe_feat_grad")


def raw_calls(q_lin, kv_lin, num_heads, scale, nbhd_idx, pe_idx, cluster_mask, pe_table, blank_k, blank_v,
              pre_table=None):
    """(fwd, bwd) closures that enqueue the forward kernel / the backward kernel group through the C ABI on
    preallocated buffers (no autograd, no allocation) — what bench.py times for the roofline entries.  With
    pre_table (the model's call form) the backward group ends with affb200_pe_feat_grad instead of the [P, H] table
    histogram."""
    B, n, hc = q_lin.shape
    H = num_heads
    c = hc // H
    M = nbhd_idx.shape[2]
    dev = q_lin.device
    plan = get_plan(nbhd_idx)
    cl_first = os.environ.get("AFFB200_CL_FIRST", "0") != "0"
    from . import fused_cl
    if n <= fused_cl.CLD_MAX_N and c == 32:
        clp = fused_cl.cl_plan(nbhd_idx, cluster_mask, pe_idx, pe_table.shape[0])
        r = None if clp is None else _cld_radius(clp, n, c)
        if r is not None:
            return _raw_calls_cld(clp, r, q_lin, kv_lin, H, scale, nbhd_idx, pe_table, blank_k, blank_v, pre_table)
    if (cl_first or not _use_img(n, M)) and H % 2 == 0 and H * c <= 512 and c in (16, 32):
        clp = fused_cl.cl_plan(nbhd_idx, cluster_mask, pe_idx, pe_table.shape[0])
        if clp is not None:
            return _raw_calls_cl(clp, q_lin, kv_lin, H, scale, nbhd_idx, pe_table, blank_k, blank_v, pre_table)
    if _use_img(n, M):
        return _raw_calls_img(q_lin, kv_lin, H, scale, nbhd_idx, pe_idx, cluster_mask, pe_table, blank_k, blank_v,
                              pre_table)
    if plan.qplan(nbhd_idx, n) is None or plan.kplan(nbhd_idx, n) is None:
        return None
    umax = plan.stats["q"]["umax"]
    q_pe = _packed_pe(plan, pe_idx, cluster_mask)
    if q_pe is None or umax > N.PLAN_UMAX - 1:
        return None
    st_f = plan.struct(nbhd_idx, n, True, False)
    st_b = plan.struct(nbhd_idx, n, True, True)
    T = plan.stats["q"]["T"]
    table = pe_table.detach().float().contiguous()
    bk, bv = blank_k.detach().float().contiguous(), blank_v.detach().float().contiguous()
    out, dq, dkv, dout = torch.empty_like(q_lin), torch.empty_like(q_lin), torch.empty_like(kv_lin), torch.randn_like(q_lin)
    lse = torch.empty((B, H, n), dtype=torch.float32, device=dev)
    ds = torch.empty((B, H, n, M), dtype=q_lin.dtype, device=dev)
    pw = torch.empty_like(ds)
    blank_part = torch.empty((B, (n + T - 1) // T, H, 2, c), dtype=torch.float32, device=dev)
    d_table = torch.empty((table.shape[0], H), dtype=torch.float32, device=dev)
    de, sp = N.dtype_enum(q_lin), N.stream_ptr(dev)
    keep = [plan, q_pe, table, bk, bv, out, dq, dkv, dout, lse, ds, pw, blank_part, d_table]
    pe_grad = _raw_pe_grad(pre_table, ds, q_pe, B, H, n * M, keep)
    dtp = None if pe_grad is not None else N.p(d_table)
    V = lambda t, off=0, kv=False: _view(t, n * (2 * hc if kv else hc), 2 * hc if kv else hc, 2 * c if kv else c, off)

    def fwd():
        N.check(N.LIB.affb200_fused_attn_fwd(V(q_lin), V(kv_lin, 0, True), V(kv_lin, c, True), V(out), N.p(table),
                                             N.p(bk), N.p(bv), st_f, N.p(q_pe), umax, N.p(lse), B, H, n, M, c,
                                             float(scale), de, sp), "fused_attn_fwd")

    def bwd():
        N.check(N.LIB.affb200_fused_attn_bwd(V(q_lin), V(kv_lin, 0, True), V(kv_lin, c, True), V(out), V(dout), V(dq),
                                             V(dkv, 0, True), V(dkv, c, True), N.p(table), N.p(bk), N.p(bv), st_b,
                                             N.p(q_pe), umax, N.p(lse), N.p(ds), N.p(pw), N.p(blank_part), dtp,
                                             table.shape[0], B, H, n, M, c, float(scale), de, sp), "fused_attn_bwd")
        if pe_grad is not None:
            pe_grad()

    fwd._keep = keep
    return fwd, bwd


def _raw_calls_cl(clp, q_lin, kv_lin, H, scale, nbhd_idx, pe_table, blank_k, blank_v, pre_table=None):
    """(fwd, bwd) closures of the block-structured kernels on preallocated buffers; bwd = the two backward kernels and
    the bias-generator gradient from dS (what the model's backward launches, minus the column sums of the blank-token
    partials).  fwd.family names the kernels."""
    from .plan import guard_tensor
    B, n, hc = q_lin.shape
    c = hc // H
    M = nbhd_idx.shape[2]
    dev = q_lin.device
    table = pe_table.detach().float().contiguous()
    bk, bv = blank_k.detach().float().contiguous(), blank_v.detach().float().contiguous()
    out, dq, dkv, dout = torch.empty_like(q_lin), torch.empty_like(q_lin), torch.empty_like(kv_lin), torch.randn_like(q_lin)
    lse = torch.empty((B, H, n), dtype=torch.float32, device=dev)
    delta = torch.empty((B, H, n), dtype=torch.float32, device=dev)
    ds = torch.empty((B, H, n, M), dtype=q_lin.dtype, device=dev)
    cb = _bwd_chunk(B, n, hc, H, M, q_lin.element_size()) if (pre_table is not None and clp.q_pe is not None) else B
    chunks = [(b0, min(cb, B - b0)) for b0 in range(0, B, cb)]
    grids = [int(N.LIB.affb200_fused_cl_bwd_grid(bn, H, n)) for _, bn in chunks]
    blank_part = torch.empty((sum(grids) * 8, H, 2, c), dtype=torch.float32, device=dev)
    d_table = torch.empty((table.shape[0], H), dtype=torch.float32, device=dev)
    guard = guard_tensor(dev)
    de, sp = N.dtype_enum(q_lin), N.stream_ptr(dev)
    keep = [clp, table, bk, bv, out, dq, dkv, dout, lse, delta, ds, blank_part, d_table]
    pe_grads = [_raw_pe_grad(pre_table, ds[b0:b0 + bn], None if clp.q_pe is None else clp.q_pe[b0:b0 + bn], bn, H, n * M, keep)
                for b0, bn in chunks]
    pe_grad = pe_grads[0]
    V = lambda t, off=0, kv=False: _view(t, n * (2 * hc if kv else hc), 2 * hc if kv else hc, 2 * c if kv else c, off)

    def fwd():
        N.check(N.LIB.affb200_fused_cl_fwd(V(q_lin), V(kv_lin, 0, True), V(kv_lin, c, True), V(out), N.p(table), N.p(bk),
                                           N.p(bv), clp.byref(), N.p(lse), N.p(guard), B, H, n, c, clp.w, clp.tw,
                                           float(scale), de, sp), "fused_cl_fwd")

    def bwd():
        row0 = 0
        for (b0, bn), gr, pg in zip(chunks, grids, pe_grads):     # image chunks sized for the L2 (as the model's backward)
            oq, okv = b0 * n * hc, b0 * n * 2 * hc
            N.check(N.LIB.affb200_fused_cl_bwd(V(q_lin, oq), V(kv_lin, okv, True), V(kv_lin, okv + c, True), V(out, oq),
                                               V(dout, oq), V(dq, oq), V(dkv, okv, True), V(dkv, okv + c, True), N.p(table),
                                               N.p(bk), N.p(bv), clp.byref_images(b0), N.p(lse[b0:]), N.p(delta[b0:]),
                                               N.p(ds[b0:]), N.p(blank_part[row0:]), N.p(guard), bn, H, n, M, c,
                                               clp.w, clp.tw, float(scale), de, sp), "fused_cl_bwd")
            row0 += gr * 8
            if pg is not None:
                pg()
        if pe_grad is not None:
            pass
        else:
            N.check(N.LIB.affb200_table_grad_u16(N.p(ds), N.p(clp.q_pe), B, H, n * M, table.shape[0], N.p(d_table), de, sp),
                    "table_grad_u16")

    fwd._keep = keep
    fwd.family = "block-structured"
    return fwd, (bwd if clp.q_pe is not None else None)   # tables above 32 768 rows: forward only


def _raw_calls_cld(clp, r, q_lin, kv_lin, H, scale, nbhd_idx, pe_table, blank_k, blank_v, pre_table=None):
    """(fwd, bwd) closures of the resident-key dense kernels; bwd = the backward kernel and the bias-generator gradient
    from dS (what the model's backward launches, minus the column sums of the blank-token partials)."""
    B, n, hc = q_lin.shape
    c = hc // H
    M = nbhd_idx.shape[2]
    dev = q_lin.device
    table = pe_table.detach().float().contiguous()
    tt = _table_t(table, clp.w, clp.tw, r, q_lin.dtype)
    bk, bv = blank_k.detach().float().contiguous(), blank_v.detach().float().contiguous()
    out, dq, dkv, dout = torch.empty_like(q_lin), torch.empty_like(q_lin), torch.empty_like(kv_lin), torch.randn_like(q_lin)
    lse = torch.empty((B, H, n), dtype=torch.float32, device=dev)
    ds = torch.empty((B, H, n, M), dtype=q_lin.dtype, device=dev)
    blank_part = torch.empty((B, H, 2, c), dtype=torch.float32, device=dev)
    d_table = torch.empty((table.shape[0], H), dtype=torch.float32, device=dev)
    blk, pos16 = clp.keep[0], clp.keep[4]
    de, sp = N.dtype_enum(q_lin), N.stream_ptr(dev)
    keep = [clp, table, tt, bk, bv, out, dq, dkv, dout, lse, ds, blank_part, d_table]
    pe_grad = _raw_pe_grad(pre_table, ds, clp.q_pe, B, H, n * M, keep)
    V = lambda t, off=0, kv=False: _view(t, n * (2 * hc if kv else hc), 2 * hc if kv else hc, 2 * c if kv else c, off)

    def fwd():
        N.check(N.LIB.affb200_fused_cld_fwd(V(q_lin), V(kv_lin, 0, True), V(kv_lin, c, True), V(out), N.p(tt), tt.shape[1],
                                            N.p(bk), N.p(bv), N.p(blk), N.p(pos16), clp.npad, N.p(lse), B, H, n, M, c, r,
                                            2 * r + 1, float(scale), de, sp), "fused_cld_fwd")

    def bwd():
        N.check(N.LIB.affb200_fused_cld_bwd(V(q_lin), V(kv_lin, 0, True), V(kv_lin, c, True), V(out), V(dout), V(dq),
                                            V(dkv, 0, True), V(dkv, c, True), N.p(tt), tt.shape[1], N.p(bk), N.p(bv),
                                            N.p(blk), N.p(pos16), clp.npad, N.p(lse), N.p(ds), N.p(blank_part), B, H, n, M,
                                            c, r, 2 * r + 1, float(scale), de, sp), "fused_cld_bwd")
        if pe_grad is not None:
            pe_grad()
        else:
            N.check(N.LIB.affb200_table_grad_u16(N.p(ds), N.p(clp.q_pe), B, H, n * M, table.shape[0], N.p(d_table), de, sp),
                    "table_grad_u16")

    fwd._keep = keep
    fwd.family = "resident-key dense"
    return fwd, (bwd if clp.q_pe is not None else None)


def _raw_calls_img(q_lin, kv_lin, H, scale, nbhd_idx, pe_idx, cluster_mask, pe_table, blank_k, blank_v, pre_table=None):
    B, n, hc = q_lin.shape
    c = hc // H
    M = nbhd_idx.shape[2]
    dev = q_lin.device
    plan = get_plan(nbhd_idx)
    q_pe = _packed_pe(plan, pe_idx, cluster_mask)
    if q_pe is None:
        return None
    col, dupf = _img_prep(plan, nbhd_idx, q_pe)
    table = pe_table.detach().float().contiguous()
    bk, bv = blank_k.detach().float().contiguous(), blank_v.detach().float().contiguous()
    out, dq, dkv, dout = torch.empty_like(q_lin), torch.empty_like(q_lin), torch.empty_like(kv_lin), torch.randn_like(q_lin)
    lse = torch.empty((B, H, n), dtype=torch.float32, device=dev)
    ds = torch.empty((B, H, n, M), dtype=q_lin.dtype, device=dev)
    blank_part = torch.empty((B, H, 2, c), dtype=torch.float32, device=dev)
    d_table = torch.empty((table.shape[0], H), dtype=torch.float32, device=dev)
    de, sp = N.dtype_enum(q_lin), N.stream_ptr(dev)
    keep = [plan, q_pe, col, dupf, table, bk, bv, out, dq, dkv, dout, lse, ds, blank_part, d_table]
    pe_grad = _raw_pe_grad(pre_table, ds, q_pe, B, H, n * M, keep)
    dtp = None if pe_grad is not None else N.p(d_table)
    V = lambda t, off=0, kv=False: _view(t, n * (2 * hc if kv else hc), 2 * hc if kv else hc, 2 * c if kv else c, off)

    def fwd():
        N.check(N.LIB.affb200_fused_attn_img_fwd(V(q_lin), V(kv_lin, 0, True), V(kv_lin, c, True), V(out), N.p(table),
                                                 N.p(bk), N.p(bv), N.p(dupf), N.p(col), N.p(q_pe), N.p(lse), B, H, n, M,
                                                 c, float(scale), de, sp), "fused_attn_img_fwd")

    def bwd():
        N.check(N.LIB.affb200_fused_attn_img_bwd(V(q_lin), V(kv_lin, 0, True), V(kv_lin, c, True), V(out), V(dout),
                                                 V(dq), V(dkv, 0, True), V(dkv, c, True), N.p(table), N.p(bk), N.p(bv),
                                                 N.p(dupf), N.p(col), N.p(q_pe), N.p(lse), N.p(ds), N.p(blank_part),
                                                 dtp, table.shape[0], B, H, n, M, c, float(scale), de, sp),
                "fused_attn_img_bwd")
        if pe_grad is not None:
            pe_grad()

    fwd._keep = keep
    return fwd, bwd
