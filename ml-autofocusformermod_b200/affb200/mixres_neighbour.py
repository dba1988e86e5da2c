"""MixResNeighbour stage — host-side glue of the off-grid attention path as the fork's segmentation backbone uses it
(mirror of the hot-path parts of reference models/mixres_neighbour.py: `BasicLayer` :408-543, `ClusterAttention`
:173-283, `ClusterTransformerBlock` :286-361, and the importance-score token split :731-755), written against this
package's CUDA operators.  Module and parameter names follow the reference, so a stage's `state_dict` loads unchanged.

Differences from the AFF stage (aff_transformer.py): token positions carry a leading scale channel (b x n x 3:
scale, x, y) that is reordered with the tokens but takes no part in clustering; the relative-position table is built
for a 2048-pixel canvas (rel_pos_width 511, table 1023 x 1023, mixres_neighbour.py:13-27) and relative offsets are
clamped into it (:527); there is no merging.  Differences from the reference that do not change results:
  * the reference evaluates `pos_embed` (Linear 5->heads) on all 1 046 529 table rows in every block (:249); here the
    stage keeps only the table rows its pe_idx actually references (`compact_table`), which also brings the table
    under the fused attention kernel's 32 768-row limit, so the fused operator runs;
  * the on-grid clustering cache also keeps the neighbourhood tensors (as in aff_transformer.BasicLayer).
The dense image-patch MLPs, `split_pos` / `split_features` arithmetic and `MixResViT` are stock PyTorch in the reference
and are not part of the path."""
import math

import torch
import torch.nn as nn

from .aff_transformer import ClusterTransformerBlock, _gather_tokens, build_pe_lookup, compact_table, default_ops

# assumes largest input resolution is 2048 x 2048 (reference mixres_neighbour.py:13-15)
REL_POS_WIDTH = 2048 // 4 - 1
TABLE_WIDTH = 2 * REL_POS_WIDTH + 1
_tables = {}


def pre_table(device):
    """(dx, dy, dist, sin, cos) lookup rows [TABLE_WIDTH^2, 5] on `device` (one copy per device, like the reference's
    module global :17-27)."""
    key = str(device)
    if key not in _tables:
        w, tw, table = build_pe_lookup(2048)
        assert w == REL_POS_WIDTH and tw == TABLE_WIDTH
        _tables[key] = table.to(device)
    return _tables[key]


class BasicLayer(nn.Module):
    """One MixResNeighbour stage: cluster the tokens along the space-filling scan, build neighbourhoods of the nnc
    nearest clusters, run the cluster-attention blocks (reference mixres_neighbour.py:408-543)."""

    def __init__(self, dim, cluster_size, nbhd_size, depth, num_heads, mlp_ratio, drop=0., attn_drop=0., drop_path=0.,
                 layer_scale=0.0, ops=None):
        super().__init__()
        self.dim, self.nbhd_size, self.cluster_size, self.depth = dim, nbhd_size, cluster_size, depth
        self.blocks = nn.ModuleList([
            ClusterTransformerBlock(dim, num_heads, mlp_ratio, drop, attn_drop,
                                    drop_path[i] if isinstance(drop_path, (list, tuple)) else drop_path, layer_scale)
            for i in range(depth)])
        self.ops = ops
        self.no_reorder = False   # never flipped by the reference either (:457)
        self.compact = True
        self._grid_cache = {}

    def _index(self, pos, h, w, ops):
        """-> (pos in cluster order, reorder | None, nbhd_idx, mask, pe_idx) for local attention (:486-528)."""
        b, n, _ = pos.shape
        m = self.cluster_size
        k = int(math.ceil(n / float(m)))
        if k == n:
            mean, reorder = pos, None                       # clusters of one token (:489-492)
        else:
            pos, mean, _, _, reorder = ops.space_filling_cluster(pos, m, h, w, no_reorder=False)
        nbhd, mask, pe_idx, _ = ops.build_nbhd(pos, mean, m, self.nbhd_size, REL_POS_WIDTH, clamp=True)
        if getattr(ops, 'hint_geometry', None) is not None:   # where nbhd / pe_idx came from: block-structured kernels
            ops.hint_geometry(nbhd, pos, REL_POS_WIDTH, TABLE_WIDTH)
        return pos, reorder, nbhd, mask, pe_idx

    def forward(self, pos, feat, h, w, on_grid):
        """pos b x n x 3 (scale, x, y), feat b x n x c -> (pos, feat) in cluster order."""
        ops = self.ops if self.ops is not None else default_ops()
        pos_scale, pos = pos[:, :, :1], pos[:, :, 1:]
        b, n, d = pos.shape
        c = feat.shape[2]
        assert self.cluster_size > 0, 'self.cluster_size must be positive'
        table = pre_table(feat.device)
        global_attn = self.nbhd_size >= n
        if global_attn:
            member_idx, cluster_mask = None, None
            rel_pos = (pos[:, None, :, :] + REL_POS_WIDTH) - pos[:, :, None, :]              # b x n x n x d
            rel_pos = rel_pos.clamp(0, TABLE_WIDTH - 1)
            pe_idx = (rel_pos[..., 1] * TABLE_WIDTH + rel_pos[..., 0]).long()
        else:
            k = int(math.ceil(n / float(self.cluster_size)))
            nnc = min(int(round(self.nbhd_size / float(self.cluster_size))), k)
            self.nbhd_size = self.cluster_size * nnc      # rounded to a multiple of the cluster size (:482)
            if on_grid and self.training:
                key = (b, n, h, w, str(pos.device), id(ops), torch.is_grad_enabled())
                if key not in self._grid_cache:
                    self._grid_cache.clear()
                    self._grid_cache[key] = self._index(pos, h, w, ops)
                pos, reorder, member_idx, cluster_mask, pe_idx = self._grid_cache[key]
            else:
                pos, reorder, member_idx, cluster_mask, pe_idx = self._index(pos, h, w, ops)
            if reorder is not None:
                feat = getattr(ops, 'reorder_tokens', _gather_tokens)(feat, reorder)
                pos_scale = pos_scale.gather(1, reorder)
        # inference with the block-structured forward keeps the full table: that kernel takes the bias from position
        # differences and has no row limit, and the unique() of the compaction is a host sync (no graph capture)
        keep_full = (not torch.is_grad_enabled() and not global_attn and feat.is_cuda
                     and getattr(ops, 'hint_geometry', None) is not None and getattr(ops, 'fused_attention', None) is not None)
        if self.compact and not keep_full:
            pe_idx, table = compact_table(pe_idx, table)
        for blk in self.blocks:
            feat = blk(feat, member_idx, cluster_mask, pe_idx, global_attn, table, ops)
        return torch.cat([pos_scale.to(pos.dtype), pos], dim=2), feat

    def extra_repr(self):
        return "dim=%d, depth=%d" % (self.dim, self.depth)


def divide_tokens_to_split_and_keep(feat_at_curr_scale, pos_at_curr_scale, importance_scores, upscale_ratio,
                                    argsort=None):
    """Importance-score token split (reference mixres_neighbour.py:731-755): ascending sort of the scores, the top
    k_split = int(N * upscale_ratio) tokens are split, the rest kept, each group in sort order; hard masks with the
    straight-through softmax term.  The ordering comes from `affb200_argsort_f32` (stable, = torch's CUDA sort)."""
    if argsort is None:
        from .point_utils import argsort_stable as argsort
    B, N, C = feat_at_curr_scale.shape
    k_split = int(N * upscale_ratio)
    sorted_indices = argsort(importance_scores)
    bottomk_idx = sorted_indices[:, :-k_split]
    topk_idx = sorted_indices[:, -k_split:]
    mask_split_hard = torch.zeros_like(importance_scores).scatter(1, topk_idx, 1.0)
    mask_keep_hard = torch.zeros_like(importance_scores).scatter(1, bottomk_idx, 1.0)
    soft = torch.softmax(importance_scores, dim=1)
    mask_split = mask_split_hard + (soft - soft.detach())
    mask_keep = mask_keep_hard + ((1.0 - soft) - (1.0 - soft).detach())
    tokens_split = feat_at_curr_scale * mask_split.unsqueeze(-1)
    tokens_keep = feat_at_curr_scale * mask_keep.unsqueeze(-1)
    gi = lambda t, i: t.gather(1, i.unsqueeze(2).expand(-1, -1, t.shape[2]))
    return gi(tokens_split, topk_idx), gi(pos_at_curr_scale, topk_idx), gi(tokens_keep, bottomk_idx), \
        gi(pos_at_curr_scale, bottomk_idx)
