"""Block-structured plans and calls of the fused neighbourhood attention (csrc/fused_cl_*.cu).

The reference builds a neighbourhood from whole clusters of consecutive rows and every bias index from a position
difference (models/aff_transformer.py:476-487).  `BasicLayer` tells the plan of an index tensor where that structure
came from (`hint_geometry`): positions, rel_pos_width and table_width.  `ClPlan` then verifies the structure on the
device and holds the compact form the kernels read: union blocks per tile of 64 tokens, a 64-bit membership word per
token, packed positions, and (for the backward) per 8-row key block the list of query tokens that reference it plus
the bias indices in the slot order the backward writes dS in.  If the structure does not hold — or no hint was given —
`cl_plan()` returns None and the row-union kernels of fused.py run.
"""
import ctypes
import os

import torch

from . import _native as N
from . import plan as _plan

CL_T = 64
CL_UMAX = 64
KBCAP_STEP = 16
UCAP_SLACK = 3
CLD_MAX_N = 128        # images up to this many tokens take the resident-key dense kernels (csrc/fused_cld.cu)
UCAP_TWO_CTAS = 44     # largest union for which two CTAs of the forward kernel share an SM at head dim 32
SUB_RADII = (31, 63, 127, 255)     # forward-only plans over a large table: the bias table is evaluated on the central
                                   # (2R+1)^2 rows only, R the smallest of these that covers every (token, key) offset


class ClPlanStruct(ctypes.Structure):
    """affb200_cl_plan (include/affb200.h)"""
    _fields_ = [("ucap", ctypes.c_int32), ("q_hdr", ctypes.c_void_p), ("q_ucl", ctypes.c_void_p),
                ("q_mask", ctypes.c_void_p), ("pos16", ctypes.c_void_p), ("blk", ctypes.c_void_p),
                ("kbcap", ctypes.c_int32), ("kb_cnt", ctypes.c_void_p), ("kb_tok", ctypes.c_void_p),
                ("q_pe", ctypes.c_void_p), ("q_hdr2", ctypes.c_void_p), ("q_ucl2", ctypes.c_void_p)]


last_eager_stats = None  # stats[8] (affb200_cl_qplan_build_split) of the most recent plan built outside graph capture
debug_static = None     # set to a list: every statically built plan appends (shape, ucap, sub_r, stats, bad) tensors


def enabled():
    return os.environ.get("AFFB200_FUSED_CL", "1") != "0"


def hint_geometry(nbhd_idx, pos, rel_pos_width, table_width):
    """Called by the index build (BasicLayer._index): nbhd_idx / pe_idx were derived from `pos` [B, N, 2] with this
    table geometry (reference aff_transformer.py:483-487)."""
    p = _plan.get_plan(nbhd_idx)
    p._cl_hint = (pos, int(rel_pos_width), int(table_width))
    p._cl = None


def observed(skey):
    """What eager passes recorded for index tensors of shape skey = (B, N, M): dict(ok, ucap, kbcap) or None.  Static
    (graph-capture) builds size their launches from it; tests replace it to force a failed speculation."""
    return _plan._observed.get(("cl",) + skey)


class ClPlan:
    def __init__(self, B, n, M, w, tw):
        self.B, self.n, self.M, self.w, self.tw = B, n, M, w, tw
        self.tiles = (n + CL_T - 1) // CL_T
        self.struct = None
        self.keep = None
        self.ucap = 0
        self.kbcap = 0
        self.q_pe = None
        self.sub_r = None      # table radius the forward may use instead of w (see SUB_RADII), or None
        self.npad = self.tiles * CL_T

    def byref(self):
        return ctypes.byref(self.struct)

    def byref_images(self, b0):
        """the plan of images b0, b0+1, ... (every array is image-major: the same struct with offset pointers)"""
        if b0 == 0:
            return ctypes.byref(self.struct)
        s, t = ClPlanStruct(), self.struct
        nblk, nb = (self.n + 7) // 8, self.M // 8
        s.ucap, s.kbcap = t.ucap, t.kbcap
        s.q_hdr = t.q_hdr + 4 * b0 * self.tiles
        s.q_ucl = t.q_ucl + 2 * b0 * self.tiles * CL_UMAX
        s.q_mask = t.q_mask + 8 * b0 * self.npad
        s.pos16 = t.pos16 + 4 * b0 * self.npad
        s.blk = t.blk + 2 * b0 * self.n * nb
        s.kb_cnt = t.kb_cnt + 4 * b0 * nblk if t.kb_cnt else None
        s.kb_tok = t.kb_tok + 4 * b0 * nblk * t.kbcap if t.kb_tok else None
        s.q_pe = t.q_pe + 2 * b0 * self.n * self.M if t.q_pe else None
        s.q_hdr2 = t.q_hdr2 + 4 * b0 * self.tiles if t.q_hdr2 else None
        s.q_ucl2 = t.q_ucl2 + 2 * b0 * self.tiles * CL_UMAX if t.q_ucl2 else None
        self._chunk_structs = getattr(self, "_chunk_structs", {})
        self._chunk_structs[b0] = s          # keeps the struct alive for the (asynchronous) launch
        return ctypes.byref(s)


def _radius(maxoff, w):
    for r in SUB_RADII:
        if maxoff <= r and r < w:
            return r
    return None


def _build(nbhd_idx, mask, pe_idx, pos, w, tw, P, need_bwd=True):
    B, n, M = nbhd_idx.shape
    dev = nbhd_idx.device
    pl = ClPlan(B, n, M, w, tw)
    npad = pl.tiles * CL_T
    nblk = (n + 7) // 8
    skey = (B, n, M)
    static = _plan.is_static()
    obs = observed(skey)
    if static and obs is None:
        raise RuntimeError("affb200.fused_cl: static mode needs an eager pass over shape %s first" % (skey,))
    if static and not obs["ok"]:
        return None
    with torch.cuda.device(dev):
        sp = N.stream_ptr(dev)
        blk = torch.empty((B, n, M // 8), dtype=torch.int16, device=dev)
        hdr = torch.empty((B, pl.tiles), dtype=torch.int32, device=dev)
        ucl = torch.empty((B, pl.tiles, CL_UMAX), dtype=torch.int16, device=dev)
        qmask = torch.empty((B, npad), dtype=torch.int64, device=dev)
        pos16 = torch.empty((B, npad), dtype=torch.int32, device=dev)
        stats = torch.zeros(8, dtype=torch.int32, device=dev)
        posf = pos.detach().float().contiguous()
        nb_c, pe_c = nbhd_idx.contiguous(), pe_idx.contiguous()
        mk_c = None if mask is None else mask.contiguous()
        # tiles whose 64 tokens reference more than 64 blocks get one union per half (forward kernel only)
        hdr2 = torch.empty((B, pl.tiles), dtype=torch.int32, device=dev)
        ucl2 = torch.empty((B, pl.tiles, CL_UMAX), dtype=torch.int16, device=dev)
        N.check(N.LIB.affb200_cl_qplan_build_split(N.p(nb_c), N.p(mk_c), N.p(pe_c), N.p(posf), B, n, M, w, tw, N.p(blk),
                                                   N.p(hdr), N.p(ucl), N.p(hdr2), N.p(ucl2), N.p(qmask), N.p(pos16),
                                                   N.p(stats), sp), "cl_qplan_build")
        if not static:
            st = stats.tolist()
            if st[1]:          # the structure is absent: remembered for the shape
                _plan._observed[("cl",) + skey] = dict(ok=False)
                return None
            global last_eager_stats
            last_eager_stats = list(st)
            if st[0] or (st[3] and need_bwd):   # a tile beyond the kernels' limits (the backward takes no split tiles):
                return None                      # this call only
        kbcap, kb_cnt, kb_tok, q_pe = 0, None, None, None
        bad = torch.zeros(1, dtype=torch.int32, device=dev)
        if need_bwd:
            kbcap = obs["kbcap"] if (obs and obs.get("ok") and obs.get("kbcap")) else 96
            if static:
                kbcap += KBCAP_STEP     # room for a block that more tokens reference than in the recorded passes
            while True:
                kb_cnt = torch.empty((B, nblk), dtype=torch.int32, device=dev)
                kb_tok = torch.empty((B, nblk, kbcap), dtype=torch.int32, device=dev)
                N.check(N.LIB.affb200_cl_kplan_build(N.p(blk), B, n, M, kbcap, N.p(kb_cnt), N.p(kb_tok), N.p(stats), sp),
                        "cl_kplan_build")
                if static:
                    break
                st = stats.tolist()
                if st[2] == 0:
                    break
                stats[2] = 0
                kbcap = ((st[5] + KBCAP_STEP - 1) // KBCAP_STEP) * KBCAP_STEP
            if P <= 32768:
                q_pe = torch.empty((B, n, M), dtype=torch.int16, device=dev)
                N.check(N.LIB.affb200_cl_pack_pe_sorted(N.p(pe_c), N.p(mk_c), N.p(blk), B, n, M, N.p(q_pe), N.p(bad), sp),
                        "cl_pack_pe_sorted")
        sub_r = None
        # a table radius below w serves the forward over large tables and the resident-key kernels of small images
        want_sub = ((not need_bwd) and P > 32768) or n <= CLD_MAX_N
        if static:
            # a little room above the largest union seen: neighbourhoods of the data-dependent stages vary per batch
            ucap = min(CL_UMAX, obs["ucap"] + UCAP_SLACK)
            if ucap > UCAP_TWO_CTAS:     # one CTA per SM either way: take the kernels' limit, nothing to speculate on
                ucap = CL_UMAX
            trip = (stats[0] != 0) | (stats[1] != 0) | (stats[2] != 0) | (stats[4] > ucap) | (bad[0] != 0)
            if need_bwd:
                trip = trip | (stats[3] != 0)
            if want_sub:
                sub_r = _radius(2 * obs.get("maxoff", w), w)     # room for draws with longer offsets than recorded
                if sub_r is not None:
                    trip = trip | (stats[6] > sub_r)
            _plan.guard_add(trip)
            if debug_static is not None:
                debug_static.append((skey, ucap, sub_r, stats, bad))
        else:
            st = stats.tolist()
            if int(bad.item()) != 0:
                _plan._observed[("cl",) + skey] = dict(ok=False)
                return None
            ucap = min(CL_UMAX, st[4])
            maxoff = st[6]
            if want_sub:
                sub_r = _radius(maxoff, w)
            if obs and obs.get("ok"):
                ucap = max(ucap, obs["ucap"])
                kbcap_rec = max(kbcap, obs.get("kbcap", 0))
                maxoff = max(maxoff, obs.get("maxoff", 0))
            else:
                kbcap_rec = kbcap
            _plan._observed[("cl",) + skey] = dict(ok=True, ucap=ucap, kbcap=kbcap_rec, maxoff=maxoff)
    pl.ucap, pl.kbcap, pl.q_pe, pl.sub_r = ucap, kbcap, q_pe, sub_r
    s = ClPlanStruct()
    s.ucap = ucap
    s.q_hdr, s.q_ucl, s.q_mask, s.pos16, s.blk = (hdr.data_ptr(), ucl.data_ptr(), qmask.data_ptr(), pos16.data_ptr(),
                                                  blk.data_ptr())
    s.kbcap = kbcap
    s.kb_cnt, s.kb_tok = (0, 0) if kb_cnt is None else (kb_cnt.data_ptr(), kb_tok.data_ptr())
    s.q_pe = 0 if q_pe is None else q_pe.data_ptr()
    s.q_hdr2, s.q_ucl2 = hdr2.data_ptr(), ucl2.data_ptr()
    pl.struct = s
    pl.keep = (blk, hdr, ucl, qmask, pos16, stats, kb_cnt, kb_tok, q_pe, hdr2, ucl2)
    pl.stats = stats
    return pl


def cl_plan(nbhd_idx, mask, pe_idx, P):
    """ClPlan of this index set, or None (no geometry hint, structure absent, disabled).  P = rows of the bias table;
    above 32 768 rows the packed bias indices of the backward do not exist (plan.q_pe is None): forward only.  Built
    under torch.no_grad() the plan is forward-only as well (no key-major lists, no packed bias indices)."""
    if not enabled():
        return None
    p = _plan.get_plan(nbhd_idx)
    hint = getattr(p, "_cl_hint", None)
    if hint is None:
        return None
    hit = getattr(p, "_cl", None)
    need_bwd = torch.is_grad_enabled()     # forward-only plans (inference) carry no key-major lists
    key = (id(pe_idx), pe_idx._version, None if mask is None else id(mask), P, need_bwd)
    if hit is not None and hit[0] == key:
        return hit[1]
    pos, w, tw = hint
    B, n, M = nbhd_idx.shape
    if M % 8 != 0 or M > 8 * CL_UMAX or pos.shape[0] != B or pos.shape[1] != n or tw != 2 * w + 1 or P != tw * tw:
        res = None
    else:
        res = _build(nbhd_idx, mask, pe_idx, pos, w, tw, P, need_bwd)
    p._cl = (key, res)
    return res
