"""GPU parity of the fused neighbourhood attention (csrc/fused_attn.cu) against the reference's expression chain
(aff_transformer.py:103-156) evaluated in fp64 from the oracle's operators on the same bf16/f16-rounded inputs, with
neighbourhoods produced by the real index pipeline (grid / subsampled positions -> clustering -> kNN -> nbhd build)."""
import pytest
import torch

from tests.util import rel_err

pytestmark = pytest.mark.gpu


def _positions(h, w, n, B, seed):
    g = torch.Generator().manual_seed(seed)
    ys, xs = torch.meshgrid(torch.arange(h), torch.arange(w), indexing="ij")
    allp = torch.stack([xs, ys], 2).reshape(-1, 2).float()
    if n == h * w:
        return allp.unsqueeze(0).expand(B, -1, -1).contiguous()
    return torch.stack([allp[torch.randperm(h * w, generator=g)[:n]] for _ in range(B)])


def _reference(q_lin, kv_lin, H, scale, nbhd, pe_idx, mask, table, bk, bv, go):
    from oracle.aff_ops_ref import OracleOps as O
    t = [x.detach().double().cpu().requires_grad_(True) for x in (q_lin, kv_lin, table, bk, bv)]
    q_lin, kv_lin, table, bk, bv = t
    b, n, hc = q_lin.shape
    c = hc // H
    q = (q_lin * scale).reshape(b, n, H, c).permute(0, 2, 1, 3)
    kv = kv_lin.view(b, n, H, 2, c).permute(3, 0, 2, 1, 4)
    key, v = kv[0], kv[1]
    nb, pe, mk = nbhd.cpu(), pe_idx.cpu(), (None if mask is None else mask.cpu())
    attn = O.qk(q, key, nb)
    attn = O.attn_bias_add(attn, table, pe, mk)
    blank = (q * bk.reshape(1, H, 1, c)).sum(-1, keepdim=True)
    attn = torch.cat([attn, blank], -1).softmax(-1)
    out = O.av(attn[..., :-1], v, nb) + attn[..., -1:] * bv.reshape(1, H, 1, c)
    out = out.permute(0, 2, 1, 3).reshape(b, n, hc)
    out.backward(go.double().cpu())
    return [out.detach()] + [x.grad for x in t]


CASES = [  # (h, w, n, B, H, c, m, nbhd, img)
    (56, 56, 3136, 2, 2, 32, 8, 48, 224),
    (56, 56, 627, 3, 4, 32, 8, 48, 224),      # padded last cluster -> mask + duplicate row 0
    (56, 56, 125, 5, 8, 32, 8, 48, 224),
    (56, 56, 784, 2, 4, 16, 8, 48, 224),      # head dim 16
    (40, 40, 70, 2, 8, 16, 8, 48, 160),
    (96, 96, 2304, 1, 8, 32, 24, 144, 384),   # AFF-Base geometry (nbhd 144); table 191^2 > 32768 -> must decline
    # small token sets: the per-(image, head) kernels (fused_attn_img.cu); each also run through the tile kernels
    (56, 56, 196, 3, 8, 32, 8, 48, 224),      # 13 sixteen-row slices of dK/dV: two per warp
    (56, 56, 255, 2, 4, 16, 8, 48, 224),      # the largest image the resident kernels take
    (56, 56, 250, 2, 2, 32, 16, 96, 224),     # M = 96: 12 slots per thread
    (56, 56, 33, 4, 8, 32, 8, 48, 224),       # M = 40 > tokens of a tile's last rows; two tiles, second nearly empty
    (28, 28, 32, 3, 4, 32, 8, 48, 224),       # exactly one full tile, no padding
]


def _families(case):
    return ["tile", "img"] if case[2] <= 255 else ["tile"]


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("case,family", [(c, f) for c in CASES for f in _families(c)])
def test_fused_attention_matches_reference_chain(case, family, dtype, monkeypatch):
    from affb200.point_utils import space_filling_cluster, build_nbhd
    from affb200 import fused
    from affb200.fused import fused_cluster_attention
    monkeypatch.setenv("AFFB200_FUSED_IMG", "1" if family == "img" else "0")
    h, w, n, B, H, c, m, nbhd_size, img = case
    assert fused._use_img(n, m * min(round(nbhd_size / m), -(-n // m))) == (family == "img")
    rpw = img // 4 - 1
    P = (2 * rpw + 1) ** 2
    pos = _positions(h, w, n, B, seed=n).cuda()
    p_s, mean, _, _, _ = space_filling_cluster(pos, m, h, w)
    nbhd, mask, pe_idx, _ = build_nbhd(p_s, mean, m, nbhd_size, rpw)
    g = torch.Generator().manual_seed(7 + n)
    hc = H * c
    q_lin = torch.randn(B, n, hc, generator=g).to(dtype).cuda().requires_grad_(True)
    kv_lin = torch.randn(B, n, 2 * hc, generator=g).to(dtype).cuda().requires_grad_(True)
    table = (0.5 * torch.randn(P, H, generator=g)).to(dtype).cuda().requires_grad_(True)
    bk = torch.randn(hc, generator=g).cuda().requires_grad_(True)
    bv = torch.randn(hc, generator=g).cuda().requires_grad_(True)
    go = torch.randn(B, n, hc, generator=g).to(dtype).cuda()
    scale = c ** -0.5
    out = fused_cluster_attention(q_lin, kv_lin, H, scale, nbhd, pe_idx, mask, table, bk, bv)
    if P > 32768:
        assert out is None
        return
    assert out is not None and out.shape == (B, n, hc) and out.dtype == dtype
    out.backward(go)
    # the kernel rounds blank_k / blank_v to the activation dtype for the tensor-core contraction
    ref = _reference(q_lin, kv_lin, H, scale, nbhd, pe_idx, mask, table, bk.to(dtype), bv.to(dtype), go)
    tol = 1e-2 if dtype == torch.bfloat16 else 2e-3
    names = ["out", "d_q", "d_kv", "d_table", "d_blank_k", "d_blank_v"]
    got = [out, q_lin.grad, kv_lin.grad, table.grad, bk.grad, bv.grad]
    for nm, a, r in zip(names, got, ref):
        e = rel_err(a, r)
        assert e < tol, (nm, e)
    if n % m:
        assert mask is not None


def test_fused_attention_is_deterministic():
    from affb200.point_utils import space_filling_cluster, build_nbhd
    from affb200.fused import fused_cluster_attention
    pos = _positions(56, 56, 627, 3, seed=1).cuda()
    p_s, mean, _, _, _ = space_filling_cluster(pos, 8, 56, 56)
    nbhd, mask, pe_idx, _ = build_nbhd(p_s, mean, 8, 48, 55)
    g = torch.Generator().manual_seed(3)
    args = [torch.randn(3, 627, 128, generator=g).bfloat16().cuda(), torch.randn(3, 627, 256, generator=g).bfloat16().cuda()]
    table = torch.randn(111 * 111, 4, generator=g).bfloat16().cuda()
    bk, bv = torch.randn(128, generator=g).cuda(), torch.randn(128, generator=g).cuda()
    outs = []
    for _ in range(2):
        a = [x.clone().requires_grad_(True) for x in args]
        o = fused_cluster_attention(a[0], a[1], 4, 32 ** -0.5, nbhd, pe_idx, mask, table, bk, bv)
        o.square().sum().backward()
        outs.append((o.detach(), a[0].grad, a[1].grad))
    for x, y in zip(*outs):
        assert torch.equal(x, y)


@pytest.mark.parametrize("family", ["tile", "img"])
def test_fused_attention_with_repeated_unmasked_neighbours(family, monkeypatch):
    """An index tensor drawn with replacement: tokens list the same key row several times with different bias rows,
    some of the repeats masked.  Both kernel families must accumulate the repeats (sequential operand build)."""
    from affb200.fused import fused_cluster_attention
    monkeypatch.setenv("AFFB200_FUSED_IMG", "1" if family == "img" else "0")
    B, n, H, c, M, P = 3, 100, 4, 32, 48, 6241
    g = torch.Generator().manual_seed(41)
    nbhd = torch.randint(0, n, (B, n, M), generator=g).cuda()
    nbhd[0, :40] = torch.sort(torch.randperm(n, generator=g)[:M])[0].cuda()     # rows without repeats: one clean tile
    pe_idx = torch.randint(0, P, (B, n, M), generator=g).cuda()
    mask = (torch.rand(B, n, M, generator=g) > 0.2).long().cuda()
    hc = H * c
    q_lin = torch.randn(B, n, hc, generator=g).bfloat16().cuda().requires_grad_(True)
    kv_lin = torch.randn(B, n, 2 * hc, generator=g).bfloat16().cuda().requires_grad_(True)
    table = (0.5 * torch.randn(P, H, generator=g)).bfloat16().cuda().requires_grad_(True)
    bk = torch.randn(hc, generator=g).cuda().requires_grad_(True)
    bv = torch.randn(hc, generator=g).cuda().requires_grad_(True)
    go = torch.randn(B, n, hc, generator=g).bfloat16().cuda()
    out = fused_cluster_attention(q_lin, kv_lin, H, c ** -0.5, nbhd, pe_idx, mask, table, bk, bv)
    assert out is not None
    out.backward(go)
    ref = _reference(q_lin, kv_lin, H, c ** -0.5, nbhd, pe_idx, mask, table, bk.bfloat16(), bv.bfloat16(), go)
    for nm, a, r in zip(["out", "d_q", "d_kv", "d_table", "d_blank_k", "d_blank_v"],
                        [out, q_lin.grad, kv_lin.grad, table.grad, bk.grad, bv.grad], ref):
        assert rel_err(a, r) < 1e-2, (nm, rel_err(a, r))


@pytest.mark.parametrize("autocast", [False, True])
@pytest.mark.parametrize("case,family", [(CASES[1], "tile"), (CASES[2], "img"), (CASES[2], "tile"), (CASES[0], "tile")])
def test_fused_attention_pos_embed_form(case, family, autocast, monkeypatch):
    """pos_embed=(pre_table, W, b) instead of a table (what ClusterAttention passes, reference aff_transformer.py:78,125):
    same output and d_q / d_kv as with the table handed in, and dW / db equal to the chain rule through the [P, H] table
    gradient of the table form (d_table^T · [feat | 1]), which the kernel never materialises."""
    from affb200.point_utils import space_filling_cluster, build_nbhd
    from affb200.fused import fused_cluster_attention
    from affb200.aff_transformer import build_pe_lookup
    monkeypatch.setenv("AFFB200_FUSED_IMG", "1" if family == "img" else "0")
    h, w, n, B, H, c, m, nbhd_size, img = case
    rpw = img // 4 - 1
    pos = _positions(h, w, n, B, seed=n).cuda()
    p_s, mean, _, _, _ = space_filling_cluster(pos, m, h, w)
    nbhd, mask, pe_idx, _ = build_nbhd(p_s, mean, m, nbhd_size, rpw)
    pre_table = build_pe_lookup(img)[2].cuda()
    g = torch.Generator().manual_seed(11 + n)
    hc = H * c
    dt = torch.bfloat16
    q0 = torch.randn(B, n, hc, generator=g).to(dt).cuda()
    kv0 = torch.randn(B, n, 2 * hc, generator=g).to(dt).cuda()
    W = (0.2 * torch.randn(H, 5, generator=g)).cuda().requires_grad_(True)
    b = (0.2 * torch.randn(H, generator=g)).cuda().requires_grad_(True)
    bk, bv = torch.randn(hc, generator=g).cuda(), torch.randn(hc, generator=g).cuda()
    go = torch.randn(B, n, hc, generator=g).to(dt).cuda()
    scale = c ** -0.5

    def run(form):
        q, kv = q0.clone().requires_grad_(True), kv0.clone().requires_grad_(True)
        with torch.autocast("cuda", dtype=dt, enabled=autocast):
            if form == "linear":
                out = fused_cluster_attention(q, kv, H, scale, nbhd, pe_idx, mask, None, bk, bv,
                                              pos_embed=(pre_table, W, b))
                tab = None
            else:
                # the stock module's table (bf16 under autocast), as an fp32 leaf so that its gradient is not
                # rounded to bf16 before the chain rule below
                tab = torch.nn.functional.linear(pre_table, W, b).detach().float().requires_grad_(True)
                out = fused_cluster_attention(q, kv, H, scale, nbhd, pe_idx, mask, tab, bk, bv)
        out.backward(go)
        if tab is not None:
            return out.detach(), q.grad, kv.grad, None, None, tab.grad
        res = (out.detach(), q.grad, kv.grad, W.grad.clone(), b.grad.clone(), None)
        W.grad = b.grad = None
        return res

    lin, tab = run("linear"), run("table")
    assert torch.equal(lin[0], tab[0]) and torch.equal(lin[1], tab[1]) and torch.equal(lin[2], tab[2])
    # chain rule in fp64 from the table form's own table gradient
    feat = (pre_table.to(dt) if autocast else pre_table).double()
    dT = tab[5].double()
    tol = 1e-4
    assert rel_err(lin[3], dT.t() @ feat) < tol, rel_err(lin[3], dT.t() @ feat)
    assert rel_err(lin[4], dT.sum(0)) < tol, rel_err(lin[4], dT.sum(0))


@pytest.mark.parametrize("n,heads,dim", [(25, 16, 512), (32, 4, 64), (7, 2, 64)])
def test_global_attention_runs_through_the_fused_operator(n, heads, dim):
    """A stage whose token count is below the neighbourhood size attends globally (reference aff_transformer.py:
    112-113,141-142: q @ k^T, softmax with the blank token, attn @ v).  ClusterAttention hands that to the fused
    operator as a neighbourhood of all tokens (padded to a multiple of 8 with masked slots); output and gradients
    match the stock chain in bf16 (1e-2 per operator; chain of three)."""
    from affb200.aff_transformer import ClusterAttention, CudaOps, build_pe_lookup
    from affb200 import fused as fa
    B = 6
    g = torch.Generator().manual_seed(n)
    w, tw, table = build_pe_lookup(64)
    table = table.cuda()
    cells = torch.stack([torch.randperm(15 * 15, generator=g)[:n] for _ in range(B)])
    pos = torch.stack([cells % 15, cells // 15], 2).float().cuda()
    rel = (pos[:, None, :, :] + w) - pos[:, :, None, :]
    pe_idx = (rel[..., 1] * tw + rel[..., 0]).long()
    feat = torch.randn(B, n, dim, generator=g).cuda()
    go = torch.randn(B, n, dim, generator=g).cuda()
    torch.manual_seed(1)
    attn = ClusterAttention(dim, heads).cuda()
    res, hits = [], []

    def counting(*a, **kw):
        out = fa.fused_cluster_attention(*a, **kw)
        hits.append(out is not None)
        return out
    for fused in (False, True):
        ops = CudaOps()
        ops.fused_attention = counting if fused else None
        for p in attn.parameters():
            p.grad = None
        x = feat.clone().requires_grad_(True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            y = attn(x, None, None, pe_idx, True, table, ops)
        y.backward(go.to(y.dtype))
        res.append((y.float(), x.grad, {k: p.grad.clone() for k, p in attn.named_parameters()}))
    assert hits == [True]
    (y0, dx0, g0), (y1, dx1, g1) = res
    assert rel_err(y1, y0) < 2e-2, rel_err(y1, y0)
    assert rel_err(dx1, dx0) < 3e-2, rel_err(dx1, dx0)
    for k in g0:
        assert rel_err(g1[k], g0[k]) < 3e-2, (k, rel_err(g1[k], g0[k]))
