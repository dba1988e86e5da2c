"""Pure-PyTorch gather oracle for the three CLUSTEN operators.  TEST INFRASTRUCTURE (see oracle/__init__.py).

Each function restates the "pytorch equivalent" expression of the reference's own kernel test:
  qk_ref : clusten/test_qk_kernel.py:39-40   key.gather(...) ; (query.unsqueeze(3) * key_gather).sum(-1)
  av_ref : clusten/test_av_kernel.py:39-40   val.gather(...) ; (attn.unsqueeze(4) * val_gather).sum(3)
  wf_ref : clusten/test_wf_kernel.py:40-41   feat.gather(...) ; weights.transpose(-1,-2) @ feat_gather
They are differentiable through autograd, run on CPU (and on CUDA), and accept any float dtype; the parity tests
evaluate them in fp32/fp64 on inputs already rounded to the kernel's storage dtype.

Parity pin: these ARE the reference's oracle expressions (no separate golden vectors exist in the reference:
its tests use unseeded randn and print norms).  They are held against the reference's own compiled CUDA kernels
(oracle/_ref, tests/test_ref_cuda_parity.py) on the GPU box, and the parity tests evaluate them on seeded inputs
generated in the test itself (tests/util.py) — there are no stored CLUSTEN fixtures.
"""
import torch


def qk_ref(query, key, nbhd_idx):
    """attn[b,h,i,j] = sum_c query[b,h,i,c] * key[b,h,nbhd_idx[b,i,j],c]"""
    b, h, n, c = query.shape
    m = nbhd_idx.shape[-1]
    gidx = nbhd_idx.reshape(b, 1, n * m, 1).expand(-1, h, -1, c)
    key_gather = key.gather(index=gidx, dim=2).reshape(b, h, n, m, c)
    return (query.unsqueeze(3) * key_gather).sum(-1)


def av_ref(attn, v, nbhd_idx):
    """feat[b,h,i,c] = sum_j attn[b,h,i,j] * v[b,h,nbhd_idx[b,i,j],c]"""
    b, h, n, m = attn.shape
    c = v.shape[-1]
    gidx = nbhd_idx.reshape(b, 1, n * m, 1).expand(-1, h, -1, c)
    val_gather = v.gather(index=gidx, dim=2).reshape(b, h, n, m, c)
    return (attn.unsqueeze(4) * val_gather).sum(3)


def wf_ref(weights, feat, nbhd_idx):
    """feat_new[b,i,ic,c] = sum_j weights[b,i,j,ic] * feat[b,nbhd_idx[b,i,j],c]"""
    b, n_, m, ic = weights.shape
    c = feat.shape[-1]
    feat_gather = feat.gather(index=nbhd_idx.reshape(b, n_ * m, 1).expand(-1, -1, c), dim=1).reshape(b, n_, m, c)
    return weights.transpose(-1, -2) @ feat_gather


REFS = {"qk": qk_ref, "av": av_ref, "wf": wf_ref}


def fwd_bwd(op, a, b, nbhd_idx, grad_out=None):
    """Forward + backward through the oracle.  Returns (out, grad_a, grad_b).  With grad_out=None the loss is
    out.mean(), exactly what the reference tests differentiate (test_qk_kernel.py:30)."""
    a = a.detach().clone().requires_grad_(True)
    b = b.detach().clone().requires_grad_(True)
    out = REFS[op](a, b, nbhd_idx)
    if grad_out is None:
        out.mean().backward()
    else:
        out.backward(grad_out)
    return out.detach(), a.grad, b.grad
