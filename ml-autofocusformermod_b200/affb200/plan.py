"""Per-index-tensor plans.

`nbhd_idx` is constant across all blocks of a stage (reference models/aff_transformer.py:489-495 passes the same
`member_idx` tensor to every block), so derived index structures are built once and shared by every forward and
backward call that receives the same tensor object:

  * the transposed index (CSR: key row -> referencing (token, slot) entries) that turns the scatter halves of the
    backward passes into gathers (affb200_csr_build in include/affb200.h).

Plans are looked up by tensor identity and die with the tensor (weakref), and are invalidated by in-place writes
(`Tensor._version`).
"""
import threading
import weakref

import torch

from . import _native as N

_lock = threading.Lock()
_plans = {}  # id(tensor) -> (weakref, version, NbhdPlan)


class NbhdPlan:
    """Derived structures of one nbhd_idx tensor [B, Nq, M] (int64, CUDA, contiguous)."""

    def __init__(self, idx):
        self.shape = tuple(idx.shape)
        self._csr = {}  # Nk -> (off, ent)
        self._lock = threading.Lock()

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
                # `work` is only used by the kernels just enqueued on the current stream; the caching allocator
                # keeps it stream-ordered, so dropping the reference here is safe.
            self._csr[Nk] = (off, ent)
            return off, ent


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
