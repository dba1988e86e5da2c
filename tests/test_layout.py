"""The product never routes through the oracle or a CPU fallback."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "ml-autofocusformermod_b200")


def test_package_does_not_import_oracle_or_reference():
    bad = []
    for dp, _, fns in os.walk(PKG):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dp, fn)).read()
                if re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M) or "/root/reference" in src:
                    bad.append(os.path.join(dp, fn))
    assert not bad, bad


def test_tools_do_not_touch_the_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline / reference leg may import oracle/."""
    bad = []
    for fn in os.listdir(os.path.join(ROOT, "tools")):
        if fn.endswith(".py"):
            src = open(os.path.join(ROOT, "tools", fn)).read()
            if re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M):
                bad.append(fn)
    assert not bad, bad


def test_oracle_says_it_is_test_infrastructure():
    for fn in os.listdir(os.path.join(ROOT, "oracle")):
        if fn.endswith((".py", ".c")):
            head = open(os.path.join(ROOT, "oracle", fn)).read(1500)
            assert "TEST INFRASTRUCTURE" in head, fn
