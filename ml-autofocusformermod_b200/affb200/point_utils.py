"""Index builders of the off-grid attention path, host side (mirrors reference models/point_utils.py and the index
code inlined in models/aff_transformer.py / models/mixres_neighbour.py).  Every function enqueues hand-written
sm_100a kernels of libaffb200.so on the current CUDA stream (C ABI: include/affb200.h); there is no CPU path.

  space_filling_cluster(pos, m, h, w, no_reorder, sf_type, use_anchor)   reference point_utils.py:61-208
  calculate_peano_order / calculate_hilbert_order(h, w, pos)               reference point_utils.py:211-349
  knn_keops(query, database, k, return_dist=False)           reference point_utils.py:27-58
  argsort_stable(x, descending=False)                        torch.sort(...)[1] on CUDA (mixres_neighbour.py:735)
  build_nbhd(pos, cluster_mean_pos, m, nbhd_size, rel_pos_width, clamp=False)
                                                             reference aff_transformer.py:446-487
  select_merge_centers(pos, learned_prob, stride, alpha, ds_rate, reserve_num, reserve_on=True)
                                                             reference aff_transformer.py:289-326
  gather_rows(src, idx)                                      reference aff_transformer.py:335-338
  find_pos_org_order(pos_org, pos_shuffled)                  reference maskfiner_up_down.py:150-154
  reorder_tokens(feat, pos_ranking)                          reference aff_transformer.py:466,472
"""
import math

import torch

from . import _native as N


def _f32_pos(pos, name="pos"):
    if not pos.is_cuda:
        raise RuntimeError("%s must be a CUDA tensor" % name)
    if pos.dim() != 3 or pos.shape[2] != 2:
        raise RuntimeError("%s must have shape [B, N, 2]" % name)
    pos = pos.detach()
    if pos.dtype != torch.float32:
        pos = pos.to(torch.float32)  # point_utils.py:87-88 / :45-48
    return pos.contiguous()


def _work(nbytes, dev):
    return torch.empty(max(16, nbytes), dtype=torch.uint8, device=dev)


def space_filling_cluster(pos, m, h, w, no_reorder=False, sf_type='', use_anchor=True):
    """Balanced clustering along a space-filling scan (reference models/point_utils.py:61-208).  Returns, like the
    reference, (pos_sorted b×n×2 fp32, cluster_mean_pos b×k×2 fp32, member_idx b×k×m int64,
    cluster_mask b×k×m int64 | None, pos_ranking b×n×1 int64), or with no_reorder=True
    (cluster_mean_pos, member_idx = the ranking cut into clusters, cluster_mask).

    use_anchor=True (what every model calls): anchor grid + distance-ratio key, one fused kernel chain
    (`affb200_sfc_cluster`).  use_anchor=False: the tokens' own row-scan order with alternating direction
    (point_utils.py:166-179), ranked by the same stable rank sort (`affb200_argsort_f32`).  sf_type 'peano' /
    'hilbert' (point_utils.py:98-122,162-165): curve keys from `affb200_curve_key`; with anchors the anchor grid is
    ordered along the curve and the token kernels take the order as tables (`affb200_sfc_cluster_tables`).  Any other
    sf_type means the default scan, as in the reference."""
    curve = {'peano': 1, 'hilbert': 2}.get(sf_type, 0)
    pos = _f32_pos(pos)
    b, n, d = pos.shape
    dev = pos.device
    k = int(math.ceil(n / m))
    with torch.cuda.device(dev):
        if use_anchor:
            ranking = torch.empty((b, n), dtype=torch.int64, device=dev)
            pos_sorted = torch.empty_like(pos)
            mean = torch.empty((b, k, 2), dtype=torch.float32, device=dev)
            work = _work(N.LIB.affb200_sfc_work_bytes(b, n), dev)
            if curve:
                npw, nump, plw, plh, grid_rank, prev, nxt = _curve_anchor_tables(curve, k, h, w, dev)
                N.check(N.LIB.affb200_sfc_cluster_tables(N.p(pos), b, n, m, npw, nump, plw, plh, N.p(grid_rank),
                                                         N.p(prev), N.p(nxt), N.p(ranking), N.p(pos_sorted),
                                                         N.p(mean), N.p(work), N.stream_ptr(dev)),
                        "space_filling_cluster")
            else:
                N.check(N.LIB.affb200_sfc_cluster(N.p(pos), b, n, m, h, w, N.p(ranking), N.p(pos_sorted), N.p(mean),
                                                  N.p(work), N.stream_ptr(dev)), "space_filling_cluster")
        else:
            if curve:
                key = curve_key(pos, h, w, curve)
            else:
                x, y = pos[..., 0].long(), pos[..., 1].long()
                key = torch.where(y % 2 == 1, (w - 1 - x) + y * w, x + y * w)
                if h * w >= 2 ** 24:
                    raise RuntimeError("space_filling_cluster(use_anchor=False): canvas too large for an exact fp32 key")
            ranking = argsort_stable(key.to(torch.float32))
            pos_sorted = pos.gather(1, ranking.unsqueeze(2).expand(-1, -1, d))
            if k * m == n:
                mean = pos_sorted.reshape(b, k, m, d).mean(2)
            else:
                pad = torch.zeros(b, k * m, d, dtype=torch.float32, device=dev)
                pad[:, :n] = pos_sorted
                cnt = (torch.arange(k * m, device=dev) < n).reshape(1, k, m).sum(2, keepdim=True)
                mean = pad.reshape(b, k, m, d).sum(2) / cnt
        if k * m == n:
            cluster_mask = None
        else:
            cluster_mask = (torch.arange(k * m, device=dev) < n).long().reshape(1, k, m).expand(b, -1, -1)
        if no_reorder:
            member_idx = torch.zeros(b, k * m, dtype=torch.int64, device=dev)
            member_idx[:, :n] = ranking
            if cluster_mask is not None:
                cluster_mask = cluster_mask.contiguous()
            return mean, member_idx.reshape(b, k, m), cluster_mask
        member_idx = torch.arange(k * m, device=dev)
        if k * m != n:
            member_idx[n:] = 0
        member_idx = member_idx.reshape(1, k, m).expand(b, -1, -1)
    return pos_sorted, mean, member_idx, cluster_mask, ranking.unsqueeze(2)


def curve_key(pos, h, w, curve):
    """fp32 `final_order` of calculate_peano_order (curve=1) / calculate_hilbert_order (curve=2) per token
    (reference point_utils.py:211-282, :285-349)."""
    pos = _f32_pos(pos)
    b, n, _ = pos.shape
    dev = pos.device
    with torch.cuda.device(dev):
        key = torch.empty((b, n), dtype=torch.float32, device=dev)
        N.check(N.LIB.affb200_curve_key(N.p(pos), b, n, h, w, curve, N.p(key), N.stream_ptr(dev)), "curve_key")
    return key


def calculate_peano_order(h, w, pos):
    """(rank of every token b x n, tokens in curve order b x n), reference point_utils.py:211-282."""
    return _curve_order(h, w, pos, 1)


def calculate_hilbert_order(h, w, pos):
    """reference point_utils.py:285-349."""
    return _curve_order(h, w, pos, 2)


def _curve_order(h, w, pos, curve):
    order_index = argsort_stable(curve_key(pos, h, w, curve))
    src = torch.arange(order_index.shape[1], device=order_index.device).unsqueeze(0).expand_as(order_index)
    return torch.zeros_like(order_index).scatter_(1, order_index, src), order_index


def _curve_anchor_tables(curve, k, h, w, dev):
    """Anchor grid of a Peano / Hilbert clustering (reference point_utils.py:93-146): grid size snapped to a power of
    the curve's base, anchors ordered along the curve, centres of the previous / next anchor per rank.  The tables are a
    few hundred entries; they are formed with the reference's own fp32 tensor expressions on the device."""
    base = 3 if curve == 1 else 2
    patch_len = (h * w / k) ** 0.5
    nph = int(round(h / patch_len))
    nph = max(base, int(base ** round(math.log(nph, base))))
    plh = h / nph
    npw = int(round(w / h * base) * (nph / base))
    plw = w / npw
    ys, xs = torch.meshgrid(torch.arange(nph, device=dev), torch.arange(npw, device=dev), indexing="ij")
    grid_pos = torch.stack([xs, ys], dim=2).reshape(-1, 2)
    grid_rank, order_idx = _curve_order(nph, npw, grid_pos.unsqueeze(0).float(), curve)
    ordered_grid = grid_pos[order_idx[0]]
    patch_len_hw = torch.tensor([plw, plh], dtype=torch.float32, device=dev)
    init = ordered_grid * patch_len_hw + patch_len_hw / 2 - 0.5
    nump = init.shape[0]
    if nump < 3:
        raise RuntimeError("space_filling_cluster: anchor grid too small")
    prev = torch.zeros_like(init)
    prev[1:] = init[:nump - 1]
    prev[0] = prev[1] - (prev[2] - prev[1])
    nxt = torch.zeros_like(init)
    nxt[:nump - 1] = init[1:]
    nxt[-1] = nxt[-2] + (nxt[-2] - nxt[-3])
    return npw, nump, float(patch_len_hw[0].item()), float(patch_len_hw[1].item()), \
        grid_rank[0].to(torch.int32).contiguous(), prev.contiguous(), nxt.contiguous()


def knn_keops(query, database, k, return_dist=False):
    """k nearest database points per query (2-D, fp32 Euclidean distance, ascending, ties -> lowest index)."""
    query = _f32_pos(query, "query")
    database = _f32_pos(database, "database")
    b, nq, _ = query.shape
    nd = database.shape[1]
    dev = query.device
    with torch.cuda.device(dev):
        idx = torch.empty((b, nq, k), dtype=torch.int64, device=dev)
        dist = torch.empty((b, nq, k), dtype=torch.float32, device=dev) if return_dist else None
        N.check(N.LIB.affb200_knn2d(N.p(query), N.p(database), b, nq, nd, k, N.p(idx), N.p(dist), N.stream_ptr(dev)),
                "knn")
    if return_dist:
        return idx, dist
    return idx


def find_pos_org_order(pos_org, pos_shuffled):
    """Index in pos_shuffled of the L1-nearest point of every pos_org point, first index on ties: the reference's
    `argmin(cdist(pos_org, pos_shuffled, p=1), dim=2)` (maskfiner_up_down.py:150-154) without the distance matrix."""
    q = _f32_pos(pos_org, "pos_org")
    d = _f32_pos(pos_shuffled, "pos_shuffled")
    b, nq, _ = q.shape
    dev = q.device
    with torch.cuda.device(dev):
        idx = torch.empty((b, nq), dtype=torch.int64, device=dev)
        N.check(N.LIB.affb200_nearest_l1(N.p(q), N.p(d), b, nq, d.shape[1], N.p(idx), N.stream_ptr(dev)), "nearest_l1")
    return idx


def argsort_stable(x, descending=False):
    """Stable argsort along dim 1 of a [B, N] fp32 tensor."""
    if not x.is_cuda:
        raise RuntimeError("x must be a CUDA tensor")
    x = x.detach().to(torch.float32).contiguous()
    b, n = x.shape
    dev = x.device
    with torch.cuda.device(dev):
        order = torch.empty((b, n), dtype=torch.int64, device=dev)
        work = _work(N.LIB.affb200_argsort_work_bytes(b, n), dev)
        N.check(N.LIB.affb200_argsort_f32(N.p(x), b, n, 1 if descending else 0, N.p(order), N.p(work),
                                          N.stream_ptr(dev)), "argsort")
    return order


def build_nbhd(pos, cluster_mean_pos, m, nbhd_size, rel_pos_width, clamp=False):
    """Neighbourhood of every token = the members of its nnc nearest clusters.  pos must be in cluster order.
    Returns (nbhd_idx b×n×M int64, cluster_mask b×n×M int64 | None, pe_idx b×n×M int64, nearest_cluster b×n×nnc)."""
    pos = _f32_pos(pos)
    b, n, _ = pos.shape
    dev = pos.device
    k = int(math.ceil(n / float(m)))
    nnc = min(int(round(nbhd_size / float(m))), k)
    M = m * nnc
    nearest = knn_keops(pos, cluster_mean_pos, nnc)
    with torch.cuda.device(dev):
        nbhd = torch.empty((b, n, M), dtype=torch.int64, device=dev)
        pe_idx = torch.empty((b, n, M), dtype=torch.int64, device=dev)
        mask = torch.empty((b, n, M), dtype=torch.int64, device=dev) if k * m != n else None
        N.check(N.LIB.affb200_nbhd_build(N.p(pos), N.p(nearest), b, n, m, nnc, rel_pos_width, 2 * rel_pos_width + 1,
                                         1 if clamp else 0, N.p(nbhd), N.p(mask), N.p(pe_idx), N.stream_ptr(dev)),
                "nbhd_build")
    return nbhd, mask, pe_idx, nearest


def select_merge_centers(pos, learned_prob, stride, alpha, ds_rate, reserve_num, reserve_on=True, check=False):
    """Indices of the tokens kept by adaptive downsampling, in the reference's order: the topk-sampled centres
    (torch CUDA topk(sorted=False) order) followed by the reserved coarse-grid tokens (ascending).
    Returns (idx b×keep int64, final_prob b×n fp32).  `check=True` synchronises and raises when the number of
    reserved grid tokens differs from reserve_num (the reference's reshape would raise: aff_transformer.py:323)."""
    pos = _f32_pos(pos)
    b, n, _ = pos.shape
    dev = pos.device
    keep = int(n * ds_rate)
    sample = keep - reserve_num if reserve_on else keep
    lp = None
    if learned_prob is not None:
        lp = learned_prob.detach().reshape(b, n).contiguous()
        if not lp.is_cuda:
            raise RuntimeError("learned_prob must be a CUDA tensor")
    with torch.cuda.device(dev):
        idx = torch.empty((b, keep), dtype=torch.int64, device=dev)
        final = torch.empty((b, n), dtype=torch.float32, device=dev)
        status = torch.empty((b,), dtype=torch.int32, device=dev)
        work = _work(N.LIB.affb200_merge_select_work_bytes(b, n), dev) if stride != 2 else None
        N.check(N.LIB.affb200_merge_select(N.p(pos), N.p(lp), N.dtype_enum(lp) if lp is not None else 0, b, n, stride,
                                           float(alpha), sample, reserve_num, 1 if reserve_on else 0, N.p(final),
                                           N.p(idx), N.p(status), N.p(work), N.stream_ptr(dev)), "merge_select")
    if check and int(status.max().item()) != 0:
        raise RuntimeError("merge_select: reserved-token count differs from reserve_num, or duplicate positions")
    return idx, final


def gather_rows(src, idx):
    """src [B, N, M] int64, idx [B, K] int64 -> [B, K, M]  (member_idx / pe_idx / mask at the merge centres)."""
    b, n, M = src.shape
    K = idx.shape[1]
    if M % 2:
        return src.gather(1, idx.unsqueeze(2).expand(-1, -1, M))
    src = src.contiguous()
    idx = idx.contiguous()
    dev = src.device
    with torch.cuda.device(dev):
        out = torch.empty((b, K, M), dtype=torch.int64, device=dev)
        N.check(N.LIB.affb200_gather_rows_i64(N.p(src), N.p(idx), b, n, K, M, N.p(out), N.stream_ptr(dev)),
                "gather_rows")
    return out


class PermuteRowsFunction(torch.autograd.Function):
    """feat[b, perm[b]] with the adjoint scatter as backward (perm is a permutation per image)."""

    @staticmethod
    def forward(ctx, feat, perm):
        ctx.save_for_backward(perm)
        return _permute_rows(feat, perm, 0)

    @staticmethod
    def backward(ctx, d_out):
        (perm,) = ctx.saved_tensors
        return _permute_rows(d_out.contiguous(), perm, 1), None


def _permute_rows(x, perm, scatter):
    b, n, c = x.shape
    dev = x.device
    with torch.cuda.device(dev):
        out = torch.empty_like(x)
        N.check(N.LIB.affb200_permute_rows(N.p(x), N.p(perm), N.p(out), b, n, c * x.element_size(), scatter,
                                           N.stream_ptr(dev)), "permute_rows")
    return out


def reorder_tokens(feat, pos_ranking):
    """feat [B, N, C] reordered by pos_ranking [B, N, 1] (or [B, N]) as `feat[b, reorder]` does in the reference
    (aff_transformer.py:466,472).  pos_ranking must be a permutation per image (space_filling_cluster's is)."""
    if not feat.is_cuda:
        raise RuntimeError("feat must be a CUDA tensor")
    b, n, c = feat.shape
    perm = pos_ranking.reshape(b, n)
    if (c * feat.element_size()) % 16 != 0 or feat.numel() == 0:
        return feat.gather(1, perm.unsqueeze(2).expand(-1, -1, c))
    return PermuteRowsFunction.apply(feat.contiguous(), perm.contiguous())
