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


def test_low_precision_shadow_is_not_used_after_the_parameter_was_written():
    """a stale `_affb200_lp` shadow must not be served after load_state_dict / an eager update"""
    import importlib.util
    import os
    import torch
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "ml-autofocusformermod_b200", "affb200",
                        "linear_ops.py")
    src = open(path).read()
    ns = {"torch": torch}
    start = src.index("def low_precision(")
    exec(src[start:src.index("def _sink(")], ns)       # the two functions only (the module itself loads the CUDA library)
    p = torch.nn.Parameter(torch.ones(4))
    p._affb200_lp = torch.full((4,), 2.0, dtype=torch.bfloat16)
    assert ns["low_precision"](p, torch.bfloat16).float().eq(1).all()          # never marked fresh: cast of the parameter
    ns["mark_shadow_fresh"]([p])
    assert ns["low_precision"](p, torch.bfloat16) is p._affb200_lp
    with torch.no_grad():
        p.mul_(3)                                                               # an eager write bumps the version
    assert ns["low_precision"](p, torch.bfloat16).float().eq(3).all()
