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
    rc = N.LIB.affb200_qk_fwd(None, None, None, None, 1, 1, -1, 1, 1, 1, 0, None)
    assert rc == -1 and "negative" in N.last_error()
    rc = N.LIB.affb200_qk_fwd(None, None, None, None, 1, 1, 4, 4, 4, 4, 0, None)
    assert rc == -1 and "null" in N.last_error()
    rc = N.LIB.affb200_qk_fwd(None, None, None, None, 0, 1, 4, 4, 4, 4, 99, None)
    assert rc == -1 and "dtype" in N.last_error()
