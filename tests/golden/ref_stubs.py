"""Import shims used ONLY by the golden-vector generators in the build container (they read /root/reference, which
does not exist on the GPU box; nothing in tests/, bench.py or the package imports this at run time).

The reference's modules import three things that are not installed here:
  timm.models.layers.{DropPath, trunc_normal_}  -> minimal equivalents (drop_path is identity at rate 0 / eval)
  clusten.{CLUSTENQK,AV,WF}Function              -> the pure-PyTorch gather expressions (oracle/clusten_ref.py)
  pykeops.torch.LazyTensor                       -> a dense fp32 evaluation of the one formula the reference builds
                                                    (point_utils.py:50-57) with lowest-index tie-breaking
and two torch CUDA behaviours are imposed on the CPU run, as probed on the B200 box (profiles/r01_torch_rules.json):
  Tensor.sort is stable; Tensor.topk(sorted=False) returns [> k-th ascending index] ++ [== k-th ascending index].
"""
import contextlib
import importlib.util
import os
import sys
import types

import torch
import torch.nn as nn

REF = "/root/reference"
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


class _Lazy:
    def __init__(self, t):
        self.t = t

    def __sub__(self, o):
        return _Lazy(self.t - o.t)

    def __pow__(self, p):
        if p == 2:
            return _Lazy(self.t * self.t)
        if p == 0.5:
            return _Lazy(self.t.double().sqrt().float())  # correctly rounded, as sqrtf on the GPU
        raise NotImplementedError(p)

    def sum(self, dim):
        acc = self.t[..., 0]
        for i in range(1, self.t.shape[-1]):
            acc = acc + self.t[..., i]
        return _Lazy(acc)

    # dist is b x n_db x n_query ; reduce over dim=1 (the database axis)
    def argKmin(self, k, dim):
        assert dim == 1
        return self.t.sort(dim=1, stable=True)[1][:, :k].transpose(1, 2).contiguous()

    def Kmin_argKmin(self, k, dim):
        assert dim == 1
        v, i = self.t.sort(dim=1, stable=True)
        return v[:, :k].transpose(1, 2).contiguous(), i[:, :k].transpose(1, 2).contiguous()


def _install_stub_modules():
    from oracle import clusten_ref as R
    timm = types.ModuleType("timm")
    tm = types.ModuleType("timm.models")
    tl = types.ModuleType("timm.models.layers")

    class DropPath(nn.Module):
        def __init__(self, p=0.):
            super().__init__()
            self.p = p

        def forward(self, x):
            assert self.p == 0. or not self.training
            return x

    tl.DropPath = DropPath
    tl.trunc_normal_ = nn.init.trunc_normal_
    timm.models, tm.layers = tm, tl
    sys.modules.update({"timm": timm, "timm.models": tm, "timm.models.layers": tl})

    cl = types.ModuleType("clusten")

    def _fn(f):
        return type("F", (), {"apply": staticmethod(f)})
    cl.CLUSTENQKFunction, cl.CLUSTENAVFunction, cl.CLUSTENWFFunction = _fn(R.qk_ref), _fn(R.av_ref), _fn(R.wf_ref)
    sys.modules["clusten"] = cl

    pk = types.ModuleType("pykeops")
    pkt = types.ModuleType("pykeops.torch")
    pkt.LazyTensor = _Lazy
    pk.torch = pkt
    sys.modules.update({"pykeops": pk, "pykeops.torch": pkt})


def load_reference_module(name):
    """Load /root/reference/models/<name>.py as package module ref_models.<name> (bypasses models/__init__.py)."""
    _install_stub_modules()
    if "ref_models" not in sys.modules:
        pkg = types.ModuleType("ref_models")
        pkg.__path__ = [os.path.join(REF, "models")]
        sys.modules["ref_models"] = pkg
    full = "ref_models." + name
    if full in sys.modules:
        return sys.modules[full]
    spec = importlib.util.spec_from_file_location(full, os.path.join(REF, "models", name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[full] = mod
    spec.loader.exec_module(mod)
    return mod


@contextlib.contextmanager
def cuda_ordering_rules():
    """Impose torch-CUDA's sort/topk ordering on the CPU run of the reference."""
    from oracle.point_ref import topk_unsorted_cuda_order
    o_sort, o_topk = torch.Tensor.sort, torch.Tensor.topk

    def sort(self, *a, **kw):
        kw.setdefault("stable", True)
        if a:
            kw["dim"] = a[0]
            if len(a) > 1:
                kw["descending"] = a[1]
        return o_sort(self, **kw)

    def topk(self, k, dim=-1, largest=True, sorted=True):
        if sorted or not largest or self.dim() != 2 or dim not in (1, -1):
            return o_topk(self, k, dim=dim, largest=largest, sorted=sorted)
        idx = topk_unsorted_cuda_order(self.detach().float(), k)
        return self.gather(1, idx), idx

    o_fsort = torch.sort

    def fsort(t, *a, **kw):     # torch.sort(x, dim=..., descending=...) as mixres_neighbour.py:735 calls it
        return sort(t, *a, **kw)

    torch.Tensor.sort, torch.Tensor.topk, torch.sort = sort, topk, fsort
    try:
        yield
    finally:
        torch.Tensor.sort, torch.Tensor.topk, torch.sort = o_sort, o_topk, o_fsort
