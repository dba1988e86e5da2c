"""CPU checks of the drop-in boundary: libaffb200.so loads, exports every symbol include/affb200.h declares, and
the ctypes table in affb200/_native.py covers exactly that set.  No compute calls (no GPU here)."""
import ctypes
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "affb200.h")
LIB = os.path.join(ROOT, "ml-autofocusformermod_b200", "lib", "libaffb200.so")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(affb200_[a-z0-9_]+)\s*\(", src)))


def ensure_built():
    if not os.path.exists(LIB):
        import __graft_entry__ as g
        g.build()


def test_header_declares_the_operator_entry_points():
    syms = declared_symbols()
    for op in ("qk", "av", "wf"):
        assert "affb200_%s_fwd" % op in syms and "affb200_%s_bwd" % op in syms


def test_library_exports_every_declared_symbol():
    ensure_built()
    lib = ctypes.CDLL(LIB)
    for s in declared_symbols():
        assert hasattr(lib, s), "libaffb200.so does not export %s" % s
    lib.affb200_abi_version.restype = ctypes.c_int
    assert lib.affb200_abi_version() >= 1
    lib.affb200_target_arch.restype = ctypes.c_char_p
    assert lib.affb200_target_arch() == b"sm_100a"


def test_ctypes_table_matches_header():
    ensure_built()
    from affb200 import _native
    assert sorted(_native.SIGNATURES) == declared_symbols()


def test_library_is_sm100a_only():
    ensure_built()
    out = subprocess.run(["cuobjdump", "-lelf", LIB], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_argument_errors_are_reported_without_a_gpu():
    """Bad arguments are rejected before any CUDA call, with a message in affb200_last_error()."""
    ensure_built()
    from affb200 import _native as N
    rc = N.LIB.affb200_qk_fwd(None, None, None, None, None, 1, 1, -1, 1, 1, 1, 0, None)
    assert rc == -1 and "negative" in N.last_error()
    rc = N.LIB.affb200_qk_fwd(None, None, None, None, None, 1, 1, 4, 4, 4, 4, 0, None)
    assert rc == -1 and "null" in N.last_error()
    rc = N.LIB.affb200_qk_fwd(None, None, None, None, None, 0, 1, 4, 4, 4, 4, 99, None)
    assert rc == -1 and "dtype" in N.last_error()
    rc = N.LIB.affb200_qplan_build(None, 1, 8, 8, 4, 48, None, None, None, None, None)
    assert rc == -1 and "unsupported dimensions" in N.last_error()


def test_plan_struct_layout_matches_header():
    """The ctypes mirror of affb200_nbhd_plan has the header's field order and the C layout (gcc sizeof)."""
    import subprocess
    import tempfile
    from affb200 import _native as N
    fields = [f[0] for f in N.NbhdPlanStruct._fields_]
    src = open(HEADER).read()
    body = src[src.index("typedef struct {"):src.index("} affb200_nbhd_plan;")]
    order = re.findall(r"\b(\w+);", re.sub(r"/\*.*?\*/", "", body, flags=re.S))
    assert order == fields
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, "s.c")
        open(c, "w").write('#include <stdio.h>\n#include "affb200.h"\nint main(){printf("%zu %zu %zu",'
                           'sizeof(affb200_nbhd_plan),__builtin_offsetof(affb200_nbhd_plan,q_hdr),'
                           '__builtin_offsetof(affb200_nbhd_plan,k_rloc));return 0;}')
        exe = os.path.join(d, "s")
        subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), c, "-o", exe], check=True)
        size, o1, o2 = map(int, subprocess.run([exe], capture_output=True, text=True).stdout.split())
    assert size == ctypes.sizeof(N.NbhdPlanStruct)
    assert o1 == N.NbhdPlanStruct.q_hdr.offset and o2 == N.NbhdPlanStruct.k_rloc.offset
