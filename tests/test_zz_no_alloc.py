"""The allocation-free entry point affb200_col_sum_static (include/affb200.h; no counterpart in the reference):
its slab partials live in a file-scope `static __device__` array instead of a
caller workspace.  Called through the C ABI and held against the existing affb200_col_sum on the same inputs, twice in
a row (the array keeps the previous call's contents).  This file is named to be collected last:
a failure here cannot hide another test."""
import ctypes

import pytest
import torch

from tests.util import assert_close

pytestmark = pytest.mark.gpu


def _both(N, x, accumulate_into=None):
    R, C = x.shape
    dt = N.dtype_enum(x)
    st = N.stream_ptr(x.device)
    outs = []
    for static in (False, True):
        out = torch.empty(C, device=x.device, dtype=torch.float32) if accumulate_into is None \
            else accumulate_into.clone()
        acc = 0 if accumulate_into is None else 1
        if static:
            rc = N.LIB.affb200_col_sum_static(N.p(x), N.p(out), R, C, dt, acc, st)
        else:
            work = torch.empty(max(1, N.LIB.affb200_col_sum_work_floats(R, C)), device=x.device, dtype=torch.float32)
            rc = N.LIB.affb200_col_sum(N.p(x), N.p(out), N.p(work), R, C, dt, acc, st)
        N.check(rc, "col_sum static=%s" % static)
        outs.append(out)
    return outs


def test_col_sum_static_matches_col_sum_on_two_calls_in_a_row():
    from affb200 import _native as N
    g = torch.Generator().manual_seed(7)
    # first call leaves 592 x 256 partials in the shared array; the second uses a different geometry on top of them
    shapes = [((256 * 627, 256), torch.bfloat16), ((3001, 512), torch.float32), ((40, 64), torch.float16)]
    for (R, C), dt in shapes:
        x = torch.randn(R, C, generator=g).to(dt).cuda()
        want, got = _both(N, x)
        assert_close(got, want, torch.float32, "col_sum_static %dx%d" % (R, C))
        assert torch.equal(got, want)  # same kernels, same fixed summation order
        assert_close(got, x.double().sum(0), torch.float32, "col_sum_static vs fp64 sum")
    # accumulate form, and the same input again: the result does not depend on what the array held
    x = torch.randn(5000, 128, generator=g).cuda()
    base = torch.randn(128, generator=g).cuda()
    want, got = _both(N, x, accumulate_into=base)
    assert torch.equal(got, want)
    want2, got2 = _both(N, x, accumulate_into=base)
    assert torch.equal(got2, got) and torch.equal(want2, want)
    torch.cuda.synchronize()


def test_col_sum_static_allocates_nothing_and_rejects_oversize():
    from affb200 import _native as N
    x = torch.randn(4096, 256, device="cuda").bfloat16()
    out = torch.empty(256, device="cuda")
    st = N.stream_ptr(x.device)
    N.check(N.LIB.affb200_col_sum_static(N.p(x), N.p(out), 4096, 256, N.dtype_enum(x), 0, st), "warm-up")
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    stats0 = torch.cuda.memory_stats()["allocation.all.allocated"]
    for _ in range(3):
        N.check(N.LIB.affb200_col_sum_static(N.p(x), N.p(out), 4096, 256, N.dtype_enum(x), 0, st), "static")
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0  # no cudaMalloc behind the call
    assert torch.cuda.memory_stats()["allocation.all.allocated"] == stats0  # nor through torch's allocator
    # beyond the built-in array: a clear error, nothing launched
    lim = N.LIB.affb200_col_sum_static_max_floats()
    assert lim == 262144
    C = lim + 8
    assert N.LIB.affb200_col_sum_work_floats(1, C) > lim
    rc = N.LIB.affb200_col_sum_static(N.p(x), N.p(out), 1, C, 0, 0, st)
    assert rc == -3 and "built-in array" in N.last_error()
    # argument errors as in the workspace form
    assert N.LIB.affb200_col_sum_static(None, N.p(out), 8, 64, 0, 0, st) == -1
    assert N.LIB.affb200_col_sum_static(N.p(x), N.p(out), 8, 60, 0, 0, st) == -1
    assert N.LIB.affb200_col_sum_static(N.p(x), N.p(out), 8, 64, 3, 0, st) == -1  # f64
    torch.cuda.synchronize()
