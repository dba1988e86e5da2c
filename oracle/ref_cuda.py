"""TEST INFRASTRUCTURE — not product code.  Loader for the REFERENCE's own CLUSTEN CUDA extension compiled by
oracle/Makefile into oracle/_ref/ (clustenqk_cuda / clustenav_cuda / clustenwf_cuda: "the original CUDA kernels" of
BASELINE.json's north star), plus the call sequences of the reference wrapper restated so that the
compiled modules can be driven without importing /root/reference at run time:

  QK  clusten/clusten.py:21-40   forward(query, key^T contiguous, idx); backward(grad, query, key, idx)
  AV  clusten/clusten.py:47-66   forward(attn, v, idx);                 backward(grad, attn, v, idx)
  WF  clusten/clusten.py:73-92   forward(weights, feat, idx);           backward(grad, weights, feat, idx)

Only tests/ (incl. the tests/opbench.py timing script) and bench.py's comparator leg may import this module; the product package never does
(tests/test_layout.py).  `available()` is False when oracle/_ref/ holds no built modules (the build needs
/root/reference, which exists only in the build container; the built .so files travel to the GPU box)."""
import glob
import importlib.util
import os

_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")
_MODS = {}


def _path(name):
    hits = sorted(glob.glob(os.path.join(_DIR, name + "*.so")))
    return hits[0] if hits else None


def available():
    return all(_path(n) for n in ("clustenqk_cuda", "clustenav_cuda", "clustenwf_cuda"))


def module(name):
    if name not in _MODS:
        import torch  # noqa: F401  (libtorch symbols must be loaded first)
        p = _path(name)
        if p is None:
            raise RuntimeError("oracle/_ref/%s*.so not built (run `make -C oracle` where /root/reference exists)" % name)
        spec = importlib.util.spec_from_file_location(name, p)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        _MODS[name] = mod
    return _MODS[name]


def fwd_bwd(op, a, b, idx, go):
    """Forward + backward of one reference operator on CUDA tensors, in the reference wrapper's call order.
    Returns (out, grad_a, grad_b)."""
    a, b, idx, go = a.contiguous(), b.contiguous(), idx.contiguous(), go.contiguous()
    if b.dtype != a.dtype:
        b = b.to(a.dtype)
    if op == "qk":
        m = module("clustenqk_cuda")
        out = m.forward(a, b.permute(0, 1, 3, 2).contiguous(), idx)
    elif op == "av":
        m = module("clustenav_cuda")
        out = m.forward(a, b, idx)
    elif op == "wf":
        m = module("clustenwf_cuda")
        out = m.forward(a, b, idx)
    else:
        raise ValueError(op)
    ga, gb = m.backward(go, a, b, idx)
    return out, ga, gb


def forward(op, a, b, idx):
    a, b, idx = a.contiguous(), b.contiguous(), idx.contiguous()
    if op == "qk":
        return module("clustenqk_cuda").forward(a, b.permute(0, 1, 3, 2).contiguous(), idx)
    return module("clustenav_cuda" if op == "av" else "clustenwf_cuda").forward(a, b, idx)


def backward(op, go, a, b, idx):
    name = {"qk": "clustenqk_cuda", "av": "clustenav_cuda", "wf": "clustenwf_cuda"}[op]
    return module(name).backward(go.contiguous(), a.contiguous(), b.contiguous(), idx.contiguous())
