"""CPU restatement of the reference's index builders.  TEST INFRASTRUCTURE (see oracle/__init__.py).

  sfc_ref        space_filling_cluster, default variant (sf_type='', use_anchor=True, no_reorder=False), the only one
                 any model calls                              reference models/point_utils.py:61-208
  knn_ref        knn_keops                                    reference models/point_utils.py:27-58
  nbhd_ref       neighbourhood / mask / pe_idx build          reference models/aff_transformer.py:446-487
  select_ref     ClusterMerging selection half                reference models/aff_transformer.py:289-326
  split_ref      MixRes split/keep ordering                   reference models/mixres_neighbour.py:731-741

All arithmetic is fp32 in the reference's operation order (no fused multiply-add), on CPU tensors.

Parity pin
  * sfc_ref is checked against the reference function itself, imported from /root/reference in the build container
    (tests/golden/make_point_golden.py -> tests/golden/point_*.pt), with `Tensor.sort` made stable: the reference
    calls `key.sort(dim=1)` (point_utils.py:159) on CUDA, where slices of these lengths are sorted stably (probed on
    the B200 box: tools/probe_torch_rules.py -> profiles/r01_torch_rules.json); torch's CPU sort is not stable, and
    ties are common (2 922 distinct keys for 3 136 tokens on the stage-1 grid).  On tie-free inputs the unpatched
    reference is used.
  * knn_ref: parity UNPINNED.  The arithmetic lives in pykeops 2.1.1 (create_env.sh:10), which is absent from
    /root/reference and from this image; no reference test exercises it.  The published algorithm is restated:
    fp32 d = sqrt(sum_dim (q - x)^2), the k smallest per query in ascending distance, ties -> lowest database index.
  * select_ref: `topk(sorted=False)` order is torch's CUDA radix-select order: indices with value > k-th value
    ascending, then indices equal to the k-th value ascending (probed on the box, same file).
"""
import math

import torch

F32 = torch.float32


def sfc_grid(n, m, h, w):
    """Anchor-grid geometry (point_utils.py:89-97): k clusters, patch grid nph x npw, fp32 patch lengths (x, y)."""
    k = int(math.ceil(n / m))
    patch_len = (h * w / k) ** 0.5
    nph = int(round(h / patch_len))
    npw = int(round(w / patch_len))
    plh, plw = h / nph, w / npw
    return k, nph, npw, torch.tensor([plw, plh], dtype=F32)


def sfc_anchor_tables(nph, npw, w, pl):
    """Boustrophedon scan of the patch grid (point_utils.py:124-146).
    Returns rank_of_patch [nph*npw] int64 (patch y*npw+x -> scan rank), prev/next centre per rank [nump,2] fp32."""
    ys = torch.arange(nph).view(-1, 1).expand(nph, npw)
    xs = torch.arange(npw).view(1, -1).expand(nph, npw)
    odd = (ys % 2) == 1
    scan_key = torch.where(odd, -xs + ys * w + (w - 1), xs + ys * w).reshape(-1)  # distinct values
    order = scan_key.argsort(stable=True)            # rank -> patch
    rank_of_patch = torch.empty_like(order)
    rank_of_patch[order] = torch.arange(order.numel())
    grid = torch.stack([xs, ys], 2).reshape(-1, 2)[order]     # int64, in scan order
    centre = grid * pl + pl / 2 - 0.5                         # fp32: ((g*pl) + pl/2) - 0.5
    nump = centre.shape[0]
    prev = torch.zeros_like(centre)
    prev[1:] = centre[:nump - 1]
    prev[0] = prev[1] - (prev[2] - prev[1])
    nxt = torch.zeros_like(centre)
    nxt[:nump - 1] = centre[1:]
    nxt[-1] = nxt[-2] + (nxt[-2] - nxt[-3])
    return rank_of_patch, prev, nxt


def sfc_key_ref(pos, m, h, w):
    """fp32 sort key of every token (point_utils.py:148-158) and its parts."""
    pos = pos.detach().to(F32)
    b, n, d = pos.shape
    k, nph, npw, pl = sfc_grid(n, m, h, w)
    rank_of_patch, prev, nxt = sfc_anchor_tables(nph, npw, w, pl)
    cell = (pos / pl).floor()
    patch = (cell[..., 0] + cell[..., 1] * npw).long()
    rank = rank_of_patch[patch]                               # b x n
    dp = (pos - prev[rank]).pow(2).sum(-1)
    dn = (pos - nxt[rank]).pow(2).sum(-1)
    ratio = dp / (dn + 1e-5)
    key = rank.unsqueeze(2) * (ratio.max() + 1) + ratio.unsqueeze(2)   # b x n x 1, fp32
    return key.squeeze(2), rank, ratio


def sfc_ref(pos, m, h, w):
    """Returns (pos_sorted b×n×2 fp32, cluster_mean_pos b×k×2 fp32, member_idx b×k×m int64,
    cluster_mask b×k×m int64 | None, pos_ranking b×n×1 int64)."""
    pos = pos.detach().to(F32)
    b, n, d = pos.shape
    k = int(math.ceil(n / m))
    key, _, _ = sfc_key_ref(pos, m, h, w)
    ranking = key.sort(dim=1, stable=True)[1]                 # ascending (key, original index)
    pos_s = pos.gather(1, ranking.unsqueeze(2).expand(-1, -1, d))
    if k * m == n:
        mask = None
        mean = pos_s.reshape(b, k, m, d).mean(2)
    else:
        pad = torch.zeros(b, k * m, d, dtype=F32)
        pad[:, :n] = pos_s
        mask = torch.zeros(b, k * m, dtype=torch.int64)
        mask[:, :n] = 1
        mask = mask.reshape(b, k, m)
        mean = pad.reshape(b, k, m, d).sum(2) / mask.sum(2, keepdim=True)
    member = torch.arange(k * m)
    member[n:] = 0
    member = member.unsqueeze(0).expand(b, -1).reshape(b, k, m)
    return pos_s, mean, member, mask, ranking.unsqueeze(2)


def scan_key_ref(pos, w):
    """Row-scan order with alternating direction of the token positions themselves (`use_anchor=False`, default curve;
    reference point_utils.py:166-179): key = x + y*w on even rows, (w-1-x) + y*w on odd rows (integer positions)."""
    p = pos.detach().to(F32)
    x, y = p[..., 0].long(), p[..., 1].long()
    return torch.where(y % 2 == 1, (w - 1 - x) + y * w, x + y * w)


def sfc_variant_ref(pos, m, h, w, no_reorder=False, use_anchor=True):
    """The other call forms of space_filling_cluster with the default curve (reference point_utils.py:161-201):
    use_anchor=False ranks the tokens by `scan_key_ref`; no_reorder=True returns (cluster_mean_pos, member_idx,
    cluster_mask) with member_idx = the ranking cut into clusters (padding -> 0) instead of reordered positions."""
    pos = pos.detach().to(F32)
    b, n, d = pos.shape
    k = int(math.ceil(n / m))
    key = sfc_key_ref(pos, m, h, w)[0] if use_anchor else scan_key_ref(pos, w)
    ranking = key.sort(dim=1, stable=True)[1]
    pos_s = pos.gather(1, ranking.unsqueeze(2).expand(-1, -1, d))
    if k * m == n:
        mask = None
        mean = pos_s.reshape(b, k, m, d).mean(2)
    else:
        pad = torch.zeros(b, k * m, d, dtype=F32)
        pad[:, :n] = pos_s
        mask = torch.zeros(b, k * m, dtype=torch.int64)
        mask[:, :n] = 1
        mask = mask.reshape(b, k, m)
        mean = pad.reshape(b, k, m, d).sum(2) / mask.sum(2, keepdim=True)
    if no_reorder:
        member = torch.zeros(b, k * m, dtype=torch.int64)
        member[:, :n] = ranking
        return mean, member.reshape(b, k, m), mask
    member = torch.arange(k * m)
    member[n:] = 0
    member = member.unsqueeze(0).expand(b, -1).reshape(b, k, m)
    return pos_s, mean, member, mask, ranking.unsqueeze(2)


# ---- Peano / Hilbert curve orders (reference models/point_utils.py:211-349) --------------------------------------
_PEANO = torch.tensor([2, 3, 8, 1, 4, 7, 0, 5, 6])                    # digit of cell x + 3y (:224)
_HIL_ROT = torch.tensor([[-1, 1, 0, 0], [0, -1, 0, 1], [1, 0, -1, 0]])  # rotation step by state%3, cell x + 2y (:303)
_HIL_RES = torch.tensor([[0, 3, 1, 2], [2, 3, 1, 0], [2, 1, 3, 0], [0, 1, 3, 2]])  # digit by state%4, cell (:308)


def curve_key_ref(pos, h, w, curve):
    """fp32 `final_order` of calculate_peano_order (curve='peano', :211-277) / calculate_hilbert_order ('hilbert',
    :285-344), restated level by level with the orientation carried as running state instead of the reference's
    re-multiplied products: per level cell = floor(rem / cur_h), rem %= cur_h (fp32 tensor / Python scalar, :237-242);
    Peano: a level whose y is 1 mirrors x of all finer levels, whose x is 1 (x odd on a widened first level) mirrors
    y (:225,233,247-254); Hilbert: digit = RES[state%4][cell], state += ROT[state%3][cell] (:318-338)."""
    base = 3 if curve == 'peano' else 2
    p = pos.detach().to(F32)
    levels = math.ceil(math.log(h, base))
    assert levels >= 1, "h too short"
    first_w = None
    if h != w:
        first_w = round(base * (w / h))
        if first_w == base:
            first_w = None
    if curve == 'hilbert' and first_w is not None:
        raise IndexError("hilbert with round(2*w/h) != 2 indexes a 4-entry table out of range in the reference")
    rem, last = p, h
    cells = []
    for _ in range(levels):
        cur = last / base
        cells.append((rem / cur).floor().long())
        rem = rem % cur
        last = cur
    orders = []
    if curve == 'peano':
        sx = torch.ones(p.shape[:2], dtype=torch.long)
        sy = torch.ones(p.shape[:2], dtype=torch.long)
        for i, c in enumerate(cells):
            x = torch.where(sx > 0, c[..., 0], 2 - c[..., 0])
            y = torch.where(sy > 0, c[..., 1], 2 - c[..., 1])
            if i == 0 and first_w is not None:
                c6 = x % 6
                d = torch.where(c6 < 3, _PEANO[(y * 3 + c6.clamp(max=2))], _PEANO[((2 - y) * 3 + (c6 - 3).clamp(min=0))])
                orders.append((d + (x // 3) * 9).to(F32))
                sy = sy * torch.where(c[..., 0] % 2 == 1, -1, 1)
            else:
                orders.append(_PEANO[y * 3 + x].to(F32))
                sy = sy * torch.where(c[..., 0] == 1, -1, 1)
            sx = sx * torch.where(c[..., 1] == 1, -1, 1)
    else:
        state = torch.zeros(p.shape[:2], dtype=torch.long)
        for c in cells:
            cell = c[..., 0] + 2 * c[..., 1]
            orders.append(_HIL_RES[state % 4, cell].to(F32))
            state = state + _HIL_ROT[state % 3, cell]
    key = orders[-1]
    for i in range(levels - 1):
        key = key + orders[i] * float(base * base) ** (levels - i - 1)
    return key


def curve_order_ref(h, w, pos, curve):
    """(rank of each token, tokens in curve order) = calculate_*_order's two returns (:278-282)."""
    idx = curve_key_ref(pos, h, w, curve).sort(dim=1, stable=True)[1]
    src = torch.arange(idx.shape[1]).unsqueeze(0).expand_as(idx)
    return torch.zeros_like(idx).scatter_(1, idx, src), idx


def sfc_curve_ref(pos, m, h, w, sf_type, no_reorder=False, use_anchor=True):
    """space_filling_cluster with sf_type 'peano' / 'hilbert' (reference point_utils.py:98-122,135-166)."""
    pos = pos.detach().to(F32)
    b, n, d = pos.shape
    k = int(math.ceil(n / m))
    base = 3 if sf_type == 'peano' else 2
    if use_anchor:
        patch_len = (h * w / k) ** 0.5
        nph = int(round(h / patch_len))
        nph = max(base, int(base ** round(math.log(nph, base))))
        plh = h / nph
        npw = int(round(w / h * base) * (nph / base))
        plw = w / npw
        ys, xs = torch.meshgrid(torch.arange(nph), torch.arange(npw), indexing="ij")
        grid = torch.stack([xs, ys], 2).reshape(-1, 2)
        rank_of_cell, order = curve_order_ref(nph, npw, grid.unsqueeze(0), sf_type)
        pl = torch.tensor([plw, plh], dtype=F32)
        centre = grid[order[0]] * pl + pl / 2 - 0.5
        prev = torch.cat([(centre[0] - (centre[1] - centre[0]))[None], centre[:-1]])
        nxt = torch.cat([centre[1:], (centre[-1] + (centre[-1] - centre[-2]))[None]])
        cell = (pos / pl).floor()
        rank = rank_of_cell[0][(cell[..., 0] + cell[..., 1] * npw).long()]
        ratio = (pos - prev[rank]).pow(2).sum(-1) / ((pos - nxt[rank]).pow(2).sum(-1) + 1e-5)
        key = (rank.unsqueeze(2) * (ratio.max() + 1) + ratio.unsqueeze(2)).squeeze(2)
    else:
        key = curve_key_ref(pos, h, w, sf_type)
    ranking = key.sort(dim=1, stable=True)[1]
    pos_s = pos.gather(1, ranking.unsqueeze(2).expand(-1, -1, d))
    if k * m == n:
        mask, mean = None, pos_s.reshape(b, k, m, d).mean(2)
    else:
        pad = torch.zeros(b, k * m, d, dtype=F32)
        pad[:, :n] = pos_s
        mask = torch.zeros(b, k * m, dtype=torch.int64)
        mask[:, :n] = 1
        mask = mask.reshape(b, k, m)
        mean = pad.reshape(b, k, m, d).sum(2) / mask.sum(2, keepdim=True)
    if no_reorder:
        member = torch.zeros(b, k * m, dtype=torch.int64)
        member[:, :n] = ranking
        return mean, member.reshape(b, k, m), mask
    member = torch.arange(k * m)
    member[n:] = 0
    return pos_s, mean, member.unsqueeze(0).expand(b, -1).reshape(b, k, m), mask, ranking.unsqueeze(2)


def find_pos_org_order_ref(pos_org, pos_shuffled):
    """reference maskfiner_up_down.py:150-154 (torch.argmin returns the first minimal index)"""
    dists = torch.cdist(pos_org.float(), pos_shuffled.float(), p=1)
    return torch.argmin(dists, dim=2)


def knn_ref(query, database, k, return_dist=False):
    """idx b×nq×k int64 (and dist fp32): k smallest fp32 sqrt(dx²+dy²+…), ascending, ties -> lowest index."""
    q = query.detach().to(F32)
    x = database.detach().to(F32)
    diff = q[:, :, None, :] - x[:, None, :, :]
    d2 = diff[..., 0] * diff[..., 0]
    for t in range(1, diff.shape[-1]):
        d2 = d2 + diff[..., t] * diff[..., t]
    dist = d2.double().sqrt().float()   # IEEE correctly rounded fp32 sqrt (torch's CPU fp32 sqrt is not: 1 ulp off)
    sd, si = dist.sort(dim=2, stable=True)
    if return_dist:
        return si[..., :k].contiguous(), sd[..., :k].contiguous()
    return si[..., :k].contiguous()


def nbhd_ref(pos, mean, m, nbhd_size, rel_pos_width, clamp=False):
    """Neighbourhood build of BasicLayer.forward (aff_transformer.py:446-487; mixres_neighbour.py:518-528).
    pos b×n×2 (cluster-sorted), mean b×k×2.  Returns (nbhd_idx b×n×M int64, mask b×n×M int64 | None,
    pe_idx b×n×M int64, nearest_cluster b×n×nnc int64).  clamp=True is the MixRes variant (rel_pos clamped to the
    table before the index is formed)."""
    b, n, d = pos.shape
    k = int(math.ceil(n / float(m)))
    nnc = min(int(round(nbhd_size / float(m))), k)
    M = m * nnc
    table_width = 2 * rel_pos_width + 1
    member = torch.arange(k * m)
    member[n:] = 0
    member = member.reshape(1, k, m).expand(b, -1, -1)
    cmask = None
    if k * m != n:
        cmask = torch.zeros(k * m, dtype=torch.int64)
        cmask[:n] = 1
        cmask = cmask.reshape(1, k, m).expand(b, -1, -1)
    near = knn_ref(pos, mean, nnc)
    gi = near.reshape(b, -1, 1).expand(-1, -1, m)
    nbhd = member.gather(1, gi).reshape(b, n, M)
    mask = cmask.gather(1, gi).reshape(b, n, M) if cmask is not None else None
    pos_n = pos.gather(1, nbhd.reshape(b, -1, 1).expand(-1, -1, d)).reshape(b, n, M, d)
    rel = pos_n - (pos.unsqueeze(2) - rel_pos_width)
    if clamp:
        rel = rel.clamp(0, table_width - 1)
    pe_idx = (rel[..., 1] * table_width + rel[..., 0]).long()
    return nbhd, mask, pe_idx, near


def topk_unsorted_cuda_order(x, k):
    """Index order of torch.topk(k, dim=1, sorted=False) on CUDA: values above the k-th value in ascending index
    order, then values equal to it in ascending index order (radix select + ordered compaction)."""
    B, n = x.shape
    if k == 0:
        return torch.zeros(B, 0, dtype=torch.int64)
    kth = x.sort(dim=1, descending=True, stable=True)[0][:, k - 1:k]
    ar = torch.arange(n).expand(B, n)
    cls = torch.where(x > kth, ar, torch.where(x == kth, ar + n, ar + 2 * n))
    return cls.sort(dim=1, stable=True)[1][:, :k].contiguous()


def select_ref(pos, learned_prob, stride, alpha, ds_rate, reserve_num, reserve_on=True):
    """Merge-centre selection (aff_transformer.py:289-326).  pos b×n×2, learned_prob b×n×1 (any float dtype) or None.
    Returns (idx b×keep int64 — sampled centres in topk order followed by the reserved grid tokens ascending,
    final_prob b×n fp32)."""
    b, n, d = pos.shape
    keep = int(n * ds_rate)
    if stride == 2:
        grid = ((pos % stride).sum(-1) == 0).float()
    else:
        _, md = knn_ref(pos, pos, 2, return_dist=True)
        md = md[:, :, 1]
        ada = 2 ** (md.log2().ceil() + 1)
        grid = ((pos.long() % ada.unsqueeze(2).long()).sum(-1) == 0).float()
    final = grid
    if learned_prob is not None:
        lp = learned_prob.detach().view(b, n)
        lp = lp * alpha                      # stays in learned_prob's dtype (bf16 under autocast)
        final = final + lp
    if reserve_on:
        rmask = ((pos % (stride * 2)).sum(-1) == 0).float()
        final = final + (rmask * (-100))
        sample = keep - reserve_num
    else:
        sample = keep
    final = final.float()
    sidx = topk_unsorted_cuda_order(final, sample)
    if reserve_on:
        ridx = rmask.nonzero(as_tuple=True)[1].reshape(b, reserve_num)
        return torch.cat([sidx, ridx], -1), final
    return sidx, final


def split_ref(scores, k_split):
    """MixRes divide_tokens_to_split_and_keep ordering (mixres_neighbour.py:735-741): ascending stable sort;
    the first n-k_split indices are kept, the last k_split are split.  Returns (keep_idx, split_idx)."""
    order = scores.sort(dim=1, descending=False, stable=True)[1]
    n = scores.shape[1]
    return order[:, : n - k_split].contiguous(), order[:, n - k_split:].contiguous()
