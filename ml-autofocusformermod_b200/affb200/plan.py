"""Per-index-tensor plans.

`nbhd_idx` is constant across all blocks of a stage (reference models/aff_transformer.py:489-495 passes the same
`member_idx` tensor to every block), so derived index structures are built once and shared by every forward and
backward call that receives the same tensor object:

  * tile plans (csrc/tileplan.cu): per tile of T consecutive tokens the union of referenced key rows plus 1-byte
    positions, and the transposed form per tile of key rows — what the dense-tile f16/bf16 kernels consume;
  * the transposed index (CSR: key row -> referencing (token, slot) entries) for the generic scatter halves of the
    backward passes (affb200_csr_build).

Plans are looked up by tensor identity and die with the tensor (weakref), and are invalidated by in-place writes
(`Tensor._version`).  Building a tile plan reads back one 32-byte counter block (tiles whose union exceeded 256
rows, the largest union): a plan with overflowed tiles is discarded and the generic kernels are used for that tensor.

Static mode (`static_plans()`), for CUDA-graph capture of a whole training step: nothing may be read back, so a plan
is built with the tile size and row bound *observed for the same problem shape* in earlier eager passes, and what the
read-back would have checked is accumulated on the device instead (`guard_tensor()`: non-zero = a speculation failed
in some step since the last `check_guard()`).  Kernels skip tiles that break the assumptions, so a failed speculation
produces wrong numbers but no out-of-bounds access, and `check_guard()` raises.
"""
import ctypes
import os
import threading
import weakref

import torch

from . import _native as N

_lock = threading.Lock()
_plans = {}  # id(tensor) -> (weakref, version, NbhdPlan)
_observed = {}  # (kind, B, Nq, Nk, M) -> dict(T=tile size that did not overflow, umax=largest union seen)


def observed_snapshot():
    """Hashable copy of what the eager passes have on record: a captured graph sized from it is still the best one can
    capture as long as this has not changed (a failed speculation that taught nothing new needs no re-capture)."""
    return repr(sorted((k, sorted(v.items())) for k, v in _observed.items()))
_static = False
_guards = {}  # device index -> int32[1] tensor, non-zero after a failed speculation


class static_plans:
    """Context manager: tile plans are built from recorded shapes' statistics without any device->host read."""

    def __enter__(self):
        global _static
        self.prev, _static = _static, True
        return self

    def __exit__(self, *exc):
        global _static
        _static = self.prev
        return False


def is_static():
    return _static


def guard_tensor(dev):
    dev = torch.device(dev)
    g = _guards.get(dev.index)
    if g is None:
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("affb200.plan: allocate the guard (guard_tensor(device)) before graph capture")
        g = torch.zeros(1, dtype=torch.int32, device=dev)
        _guards[dev.index] = g
    return g


def guard_add(flag_tensor):
    """Accumulate a device-side boolean/int condition into the device's guard (capturable: no host read)."""
    guard_tensor(flag_tensor.device).add_(flag_tensor.reshape(-1).ne(0).sum().to(torch.int32))


def check_guard(dev):
    """One device->host read: raises if any speculated plan of the steps since the last call was wrong."""
    g = guard_tensor(dev)
    bad = int(g.item())
    if bad:
        g.zero_()
        raise RuntimeError("affb200: %d speculated index plan(s) did not hold (a tile referenced more rows than in "
                           "the recorded passes); rerun the step eagerly" % bad)


def observed_umax_cap(kind, shape_key):
    """Largest union size the buffers sized for the observed maximum can take: the fused kernels round umax+1 up
    to a multiple of 32 rows, so any union up to that bound runs with the same launch configuration."""
    o = _observed.get((kind,) + shape_key)
    if o is None:
        return None
    return min(N.PLAN_UMAX - 1, ((o["umax"] + 1 + 31) & ~31) - 1)

NUM_SMS = 148


def pick_tile(B, n, M):
    """Tokens per tile: 64 when that still gives >= 4 CTAs per SM-wave and the neighbourhood is small enough for
    the union of a tile to stay under 256 rows; smaller tiles for small problems and wide neighbourhoods."""
    forced = os.environ.get("AFFB200_TILE")
    if forced:
        return int(forced)
    t = 64 if M <= 64 else (32 if M <= 96 else 16)
    while t > 16 and B * ((n + t - 1) // t) < 4 * NUM_SMS:
        t //= 2
    return t


class NbhdPlan:
    """Derived structures of one nbhd_idx tensor [B, Nq, M] (int64, CUDA, contiguous)."""

    def __init__(self, idx):
        self.shape = tuple(idx.shape)
        self._csr = {}    # Nk -> (off, ent)
        self._q = {}      # Nk -> (T, hdr, rows, loc) or None when not tileable
        self._k = {}      # Nk -> (T, hdr, utok, rloc) or None
        self._structs = {}
        self._lock = threading.Lock()
        self.stats = {}

    # -- transposed index -------------------------------------------------------------------------------
    def csr(self, idx, Nk, sorted_entries=True):
        """(off [B, Nk+1] int32, ent [B, Nq*M] int32) for key-row count Nk; built on first use."""
        with self._lock:
            hit = self._csr.get(Nk)
            if hit is not None:
                return hit
            B, Nq, M = self.shape
            dev = idx.device
            with torch.cuda.device(dev):
                off = torch.empty((B, Nk + 1), dtype=torch.int32, device=dev)
                ent = torch.empty((B, Nq * M), dtype=torch.int32, device=dev)
                work = torch.empty(max(1, N.LIB.affb200_csr_work_bytes(B, Nk)), dtype=torch.uint8, device=dev)
                N.check(N.LIB.affb200_csr_build(N.p(idx), B, Nq, Nk, M, N.p(off), N.p(ent), N.p(work),
                                                1 if sorted_entries else 0, N.stream_ptr(dev)), "csr_build")
            self._csr[Nk] = (off, ent)
            self._structs.clear()
            return off, ent

    # -- tile plans -------------------------------------------------------------------------------------
    def _tileable(self, Nk):
        B, Nq, M = self.shape
        return (0 < B <= 65535 and Nq > 0 and 1 <= M <= N.PLAN_MAX_M and 0 < Nk <= N.PLAN_MAX_NK
                and Nq * M < 2 ** 31)

    def qplan(self, idx, Nk):
        with self._lock:
            if Nk in self._q:
                return self._q[Nk]
            res = None
            if self._tileable(Nk):
                B, Nq, M = self.shape
                dev = idx.device
                T = pick_tile(B, Nq, M)
                skey = (B, Nq, Nk, M)
                obs = _observed.get(("q",) + skey)
                if _static and obs is None:
                    raise RuntimeError("affb200.plan: static mode needs an eager pass over shape %s first" % (skey,))
                if _static:
                    T = obs["T"]
                with torch.cuda.device(dev):
                    while T is not None:   # T None (static mode): the eager passes found this shape not tileable
                        nt = (Nq + T - 1) // T
                        hdr = torch.empty((B, nt), dtype=torch.int32, device=dev)
                        rows = torch.empty((B, nt, N.PLAN_UMAX), dtype=torch.int32, device=dev)
                        loc = torch.empty(B * Nq * M + 16, dtype=torch.uint8, device=dev)  # +16: vector tail reads
                        stats = torch.zeros(8, dtype=torch.int32, device=dev)
                        N.check(N.LIB.affb200_qplan_build(N.p(idx), B, Nq, Nk, M, T, N.p(hdr), N.p(rows), N.p(loc),
                                                          N.p(stats), N.stream_ptr(dev)), "qplan_build")
                        if _static:
                            cap = observed_umax_cap("q", skey)
                            guard_add((stats[0] != 0) | (stats[4] > cap))
                            res = (T, hdr, rows, loc)
                            self.stats["q"] = dict(T=T, overflow=0, dup_tiles=None, tiles=B * nt, umax=cap, static=True)
                            break
                        st = stats.tolist()
                        if st[0] == 0:
                            res = (T, hdr, rows, loc)
                            self.stats["q"] = dict(T=T, overflow=0, dup_tiles=st[1], tiles=B * nt, umax=st[4])
                            if obs is None or obs["T"] != T or st[4] > obs["umax"]:
                                _observed[("q",) + skey] = dict(T=T, umax=max(st[4], obs["umax"] if obs and obs["T"] == T else 0))
                            break
                        if T == 8:    # 8-token tiles (run in 16-row kernels) keep a tile inside one cluster of 24
                            self.stats["q"] = dict(T=T, overflow=st[0], tiles=B * nt)
                            _observed[("q",) + skey] = dict(T=None, umax=0)   # remembered: generic kernels for this shape
                            break
                        T //= 2
            self._q[Nk] = res
            self._structs.clear()
            return res

    def kplan(self, idx, Nk):
        with self._lock:
            if Nk in self._k:
                return self._k[Nk]
            res = None
            if self._tileable(Nk):
                B, Nq, M = self.shape
                dev = idx.device
                T = pick_tile(B, Nk, M)
                skey = (B, Nq, Nk, M)
                obs = _observed.get(("k",) + skey)
                if _static and obs is None:
                    raise RuntimeError("affb200.plan: static mode needs an eager pass over shape %s first" % (skey,))
                if _static:
                    T = obs["T"]
                with torch.cuda.device(dev):
                    while T is not None:
                        nt = (Nk + T - 1) // T
                        hdr = torch.empty((B, nt), dtype=torch.int32, device=dev)
                        utok = torch.empty((B, nt, N.PLAN_UMAX), dtype=torch.int32, device=dev)
                        rloc = torch.empty((B, nt, N.PLAN_UMAX, M), dtype=torch.uint8, device=dev)
                        cnt = torch.empty(B * nt, dtype=torch.int32, device=dev)
                        stats = torch.zeros(8, dtype=torch.int32, device=dev)
                        N.check(N.LIB.affb200_kplan_build(N.p(idx), B, Nq, Nk, M, T, N.p(hdr), N.p(utok), N.p(rloc),
                                                          N.p(cnt), N.p(stats), N.stream_ptr(dev)), "kplan_build")
                        if _static:
                            guard_add(stats[2] != 0)   # the key-tile kernels size their buffers for 256 rows
                            res = (T, hdr, utok, rloc)
                            self.stats["k"] = dict(T=T, overflow=0, dup_tiles=None, tiles=B * nt, umax=N.PLAN_UMAX,
                                                   static=True)
                            break
                        st = stats.tolist()
                        if st[2] == 0:
                            res = (T, hdr, utok, rloc)
                            self.stats["k"] = dict(T=T, overflow=0, dup_tiles=st[3], tiles=B * nt, umax=st[5])
                            _observed[("k",) + skey] = dict(T=T, umax=st[5])
                            break
                        if T == 16:
                            self.stats["k"] = dict(T=T, overflow=st[2], tiles=B * nt)
                            _observed[("k",) + skey] = dict(T=None, umax=0)
                            break
                        T //= 2
            self._k[Nk] = res
            self._structs.clear()
            return res

    # -- what the C ABI receives ------------------------------------------------------------------------
    def struct(self, idx, Nk, dense, backward):
        """ctypes affb200_nbhd_plan for a call on key-row count Nk.  `dense`: the call's dtype/head-dim can use
        the dense-tile kernels.  `backward`: the scatter half needs a key-tile plan or the transposed index."""
        q = self.qplan(idx, Nk) if dense else None
        k = self.kplan(idx, Nk) if (dense and backward) else None
        need_csr = backward and k is None
        csr = self.csr(idx, Nk) if need_csr else self._csr.get(Nk)
        key = (Nk, q is not None, k is not None, csr is not None)
        with self._lock:
            s = self._structs.get(key)
            if s is None:
                s = N.NbhdPlanStruct()
                if csr is not None:
                    s.csr_off, s.csr_ent = csr[0].data_ptr(), csr[1].data_ptr()
                if q is not None:
                    s.tq, s.q_hdr, s.q_rows, s.q_loc = q[0], q[1].data_ptr(), q[2].data_ptr(), q[3].data_ptr()
                if k is not None:
                    s.tk, s.k_hdr, s.k_utok, s.k_rloc = k[0], k[1].data_ptr(), k[2].data_ptr(), k[3].data_ptr()
                self._structs[key] = s
            return ctypes.byref(s)


def get_plan(idx):
    """Plan for this tensor object (created on first sight)."""
    key = id(idx)
    with _lock:
        hit = _plans.get(key)
        if hit is not None:
            ref, version, plan = hit
            if ref() is idx and version == idx._version and plan.shape == tuple(idx.shape):
                return plan
        plan = NbhdPlan(idx)

        def _drop(_ref, key=key):
            with _lock:
                cur = _plans.get(key)
                if cur is not None and cur[0] is _ref:
                    del _plans[key]

        _plans[key] = (weakref.ref(idx, _drop), idx._version, plan)
        return plan


def clear_plans():
    with _lock:
        _plans.clear()
