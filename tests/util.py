import torch

# north_star tolerances (BASELINE.json): 1e-5 relative in fp32, 1e-2 in bf16 (fp16 is held to the bf16 bound's
# tighter sibling 2e-3; fp64 to 1e-12).  Relative error = ||x - y||_2 / ||y||_2 per tensor.
TOL = {torch.float32: 1e-5, torch.bfloat16: 1e-2, torch.float16: 2e-3, torch.float64: 1e-12}


def rel_err(x, y):
    x = x.detach().double().cpu()
    y = y.detach().double().cpu()
    den = torch.linalg.norm(y).item()
    num = torch.linalg.norm(x - y).item()
    return num / den if den > 0 else num


def assert_close(x, y, dtype, what=""):
    assert x.shape == y.shape, "%s: shape %s vs %s" % (what, tuple(x.shape), tuple(y.shape))
    e = rel_err(x, y)
    assert e <= TOL[dtype], "%s: relative error %.3e > %.1e" % (what, e, TOL[dtype])


def make_inputs(op, dims, dtype, seed, device="cpu", idx_mode="uniform"):
    """Seeded operands in the reference tests' distributions (randn operands, randint indices:
    clusten/test_qk_kernel.py:21-25), rounded to `dtype`.  dims: qk/av (B,H,Nq,Nk,M,c); wf (B,Nq,Nk,M,IC,C)."""
    g = torch.Generator().manual_seed(seed)
    if op in ("qk", "av"):
        B, H, Nq, Nk, M, c = dims
        if op == "qk":
            a = torch.randn(B, H, Nq, c, generator=g)
        else:
            a = torch.softmax(torch.randn(B, H, Nq, M, generator=g), dim=-1)
        b = torch.randn(B, H, Nk, c, generator=g)
        out_shape = (B, H, Nq, M) if op == "qk" else (B, H, Nq, c)
    else:
        B, Nq, Nk, M, IC, C = dims
        a = torch.randn(B, Nq, M, IC, generator=g)
        b = torch.randn(B, Nk, C, generator=g)
        out_shape = (B, Nq, IC, C)
    if Nk > 0:
        idx = torch.randint(Nk, (B, Nq, M), generator=g)
        if idx_mode == "dup" and M > 1:
            idx[..., 1::2] = idx[..., 0::2][..., : idx[..., 1::2].shape[-1]]  # every neighbour listed twice
        elif idx_mode == "hot":
            idx = idx % max(1, min(Nk, 3))  # three rows referenced by everybody (long CSR rows)
    else:
        idx = torch.zeros((B, Nq, M), dtype=torch.int64)
    go = torch.randn(out_shape, generator=g)
    return a.to(dtype).to(device), b.to(dtype).to(device), idx.to(device), go.to(dtype).to(device)


def clustered_idx(B, Nq, Nk, M, m, seed, pad_to_zero=True):
    """Block-structured neighbourhoods like the model builds (reference aff_transformer.py:476-480):
    nbhd_idx[b,i,q*m+r] = cluster[b,i,q]*m + r for M//m clusters near the token's own position, with a per-batch
    random choice among nearby clusters; rows >= Nk (padded last cluster) become 0 as in point_utils.py:204."""
    g = torch.Generator().manual_seed(seed)
    nnc = max(1, M // m)
    k = (Nk + m - 1) // m
    own = (torch.arange(Nq) * Nk // max(Nq, 1)) // m
    jitter = torch.randint(-1, 2, (B, Nq, 1), generator=g)
    base = own[None, :, None] + jitter
    # nnc distinct clusters around base: a random subset of a window of nnc+1 clusters
    win = nnc + 1
    perm = torch.rand(B, Nq, win, generator=g).argsort(-1)[..., :nnc]
    cl = (base + perm - win // 2).clamp_(0, k - 1)
    idx = (cl[..., None] * m + torch.arange(m)).reshape(B, Nq, nnc * m)
    if idx.shape[-1] < M:
        idx = torch.cat([idx, idx[..., : M - idx.shape[-1]]], -1)
    idx = torch.where(idx >= Nk, torch.zeros_like(idx) if pad_to_zero else idx.clamp(max=Nk - 1), idx)
    return idx.contiguous()
