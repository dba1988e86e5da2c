"""AutoFocusFormer backbone — host-side glue of the off-grid attention path (mirror of reference
models/aff_transformer.py: same module tree and parameter names, so a reference state_dict loads unchanged), written
against this package's CUDA operators:

  index building   affb200.point_utils.space_filling_cluster / build_nbhd / select_merge_centers / gather_rows
  attention        CLUSTENQKFunction / CLUSTENAVFunction, or the fused neighbourhood-attention Function
  merging          CLUSTENWFFunction

The linears, MLPs, LayerNorms and the conv stem are stock PyTorch (cuBLAS/cuDNN).  Differences from the reference
that do not change results: the relative-position lookup table is a per-model buffer instead of a module global
(aff_transformer.py:14-32); the stage-1 (on-grid) clustering AND neighbourhood build are cached per (batch, h, w)
(the reference caches only the clustering, :462-464, and redoes kNN + gathers every forward); timm is not needed.

`ops` lets a caller substitute the operator set (tests and the CPU baseline inject the oracle's pure-PyTorch
functions from outside the package); the default is the CUDA set and there is no CPU fallback inside the package.
"""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F


class CudaOps:
    """The operator set of the product path (all hand-written sm_100a kernels behind the C ABI)."""

    def __init__(self):
        from .clusten import CLUSTENQKFunction, CLUSTENAVFunction, CLUSTENWFFunction
        from . import point_utils as P
        from . import attn_ops as A
        from . import norm_ops as L
        from . import linear_ops as LI
        from . import residual_ops as RO
        self.layer_norm = L.layer_norm
        self.layer_norm_fork = L.layer_norm_fork
        self.batch_norm_act = L.batch_norm_act
        self.conv_batch_norm_act = L.conv_batch_norm_act
        self.conv_bias = L.conv_bias
        self.linear = LI.linear
        self.linear_gelu = LI.linear_gelu
        self.linear_pair = LI.linear_pair
        self.residual_add = RO.residual_add
        from . import merge_ops as MO
        self.merge_wf = MO.merge_wf
        self.weight_net_table = MO.weight_net_table
        self.attn_bias_add = A.attn_bias_add
        self.table_gather = A.table_gather
        self.qk = CLUSTENQKFunction.apply
        self.av = CLUSTENAVFunction.apply
        self.wf = CLUSTENWFFunction.apply
        self.space_filling_cluster = P.space_filling_cluster
        self.build_nbhd = P.build_nbhd
        self.select_merge_centers = P.select_merge_centers
        self.gather_rows = P.gather_rows
        self.reorder_tokens = P.reorder_tokens
        from . import fused as FU
        from . import fused_cl as FC
        # the whole of ClusterAttention between the Linears as one operator; declines (None) what it does not cover
        self.fused_attention = FU.fused_cluster_attention
        self.hint_geometry = FC.hint_geometry


_default_ops = None


def default_ops():
    global _default_ops
    if _default_ops is None:
        _default_ops = CudaOps()
    return _default_ops


def build_pe_lookup(img_size):
    """(dx, dy, distance, sin, cos) for every relative offset in [-(s/4-1), s/4-1]^2; 0/0 entries set to 0
    (reference aff_transformer.py:14-32).  Returns (rel_pos_width, table_width, table [(2w+1)^2, 5])."""
    rel_pos_width = img_size // 4 - 1
    table_width = 2 * rel_pos_width + 1
    off = torch.arange(table_width).float() - rel_pos_width
    ys, xs = torch.meshgrid(off, off, indexing="ij")
    dis = (ys ** 2 + xs ** 2) ** 0.5
    table = torch.stack([xs, ys, dis, ys / dis, xs / dis], dim=2)
    table[~torch.isfinite(table)] = 0
    return rel_pos_width, table_width, table.reshape(-1, 5)


def compact_table(pe_idx, table):
    """Rows of `table` that pe_idx references, and pe_idx re-expressed in them: table[pe_idx] == table_c[pe_idx_c].
    Neighbourhoods are local, so a stage touches a few thousand of the (2w+1)^2 relative offsets; evaluating
    pos_embed / weight_net on those rows only is SURVEY §8 f2, and it brings the 191 x 191 table of the 384-pixel
    configs (and the 1023 x 1023 one of MixRes) under the fused attention kernel's 32 768-row limit."""
    uniq, inv = torch.unique(pe_idx, return_inverse=True)
    return inv.reshape(pe_idx.shape), table.index_select(0, uniq)


COMPACT_ROWS_ABOVE = 32768   # tables larger than this are compacted per stage (0 = always)


def _gather_tokens(feat, reorder):
    """feat[b, reorder] (reference aff_transformer.py:466,472) with stock indexing: operator sets without a
    `reorder_tokens` entry (the CPU oracle set) use this."""
    return feat.gather(1, reorder.expand(-1, -1, feat.shape[2]))


def _ln(ops, mod, x):
    """nn.LayerNorm `mod` applied through the operator set (one fused pass on the product path)."""
    return ops.layer_norm(x, mod.weight, mod.bias, mod.eps)


def _lin(ops, mod, x):
    """nn.Linear `mod` applied through the operator set (bias gradient as one column-sum pass on the product path)."""
    return ops.linear(x, mod.weight, mod.bias)


def drop_path(x, p, training):
    """Stochastic depth per sample (timm DropPath semantics: scale kept paths by 1/(1-p))."""
    if p == 0. or not training:
        return x
    keep = 1 - p
    mask = x.new_empty((x.shape[0],) + (1,) * (x.dim() - 1)).bernoulli_(keep)
    return x * mask.div_(keep)


class Mlp(nn.Module):
    def __init__(self, in_features, hidden_features, drop=0.):
        super().__init__()
        self.fc1 = nn.Linear(in_features, hidden_features)
        self.act = nn.GELU()
        self.fc2 = nn.Linear(hidden_features, in_features)
        self.drop = nn.Dropout(drop)

    def forward(self, x, ops):
        h = ops.linear_gelu(x, self.fc1.weight, self.fc1.bias)    # act(fc1(x))
        return self.drop(_lin(ops, self.fc2, self.drop(h)))


_global_nbhd = {}


def global_as_neighbourhood(pe_idx):
    """Global attention (every token attends to all n tokens; reference aff_transformer.py:112-113,141-142) written as
    neighbourhood attention so that the fused operator can run it: -> (member_idx, mask, pe_idx), each b x n x M with
    M = n rounded up to a multiple of 8, member_idx[., ., j] = j, the padding slots masked (they get the -100 bias the
    reference gives padded cluster slots, i.e. weight exp(-100)), or None when n is too large for the operator.
    Cached on the pe_idx tensor (one stage's blocks share it); the index / mask tensors are static per (b, n)."""
    hit = getattr(pe_idx, "_affb200_gnb", None)
    if hit is not None and hit[0] == pe_idx._version:
        return hit[1] if hit[1] is None or hit[1][2] is not None else (hit[1][0], hit[1][1], pe_idx)
    b, n, n2 = pe_idx.shape
    M = (n + 7) // 8 * 8
    res = None
    if n == n2 and M <= 96 and pe_idx.is_cuda:
        key = (b, n, str(pe_idx.device))
        if key not in _global_nbhd:
            ar = torch.arange(M, device=pe_idx.device)
            member = torch.where(ar < n, ar, torch.zeros_like(ar)).expand(b, n, M).contiguous()
            mask = (ar < n).long().expand(b, n, M).contiguous() if M != n else None
            if len(_global_nbhd) > 8:
                _global_nbhd.clear()
            _global_nbhd[key] = (member, mask)
        member, mask = _global_nbhd[key]
        res = (member, mask, torch.nn.functional.pad(pe_idx, (0, M - n)) if M != n else None)
    pe_idx._affb200_gnb = (pe_idx._version, res)      # (no reference to pe_idx itself: no cycle)
    return res if res is None or res[2] is not None else (member, mask, pe_idx)


class ClusterAttention(nn.Module):
    """Attention of every token over the members of its nearest clusters (+ a learned blank token), with a
    relative-position bias looked up by pe_idx (reference aff_transformer.py:54-160)."""

    def __init__(self, dim, num_heads, attn_drop=0., proj_drop=0.):
        super().__init__()
        self.dim, self.num_heads = dim, num_heads
        self.scale = (dim // num_heads) ** -0.5
        self.q = nn.Linear(dim, dim)
        self.kv = nn.Linear(dim, 2 * dim)
        self.blank_k = nn.Parameter(torch.randn(dim))
        self.blank_v = nn.Parameter(torch.randn(dim))
        self.pos_embed = nn.Linear(5, num_heads)
        self.attn_drop = nn.Dropout(attn_drop)
        self.proj = nn.Linear(dim, dim)
        self.proj_drop = nn.Dropout(proj_drop)

    def forward(self, feat, member_idx, cluster_mask, pe_idx, global_attn, pre_table, ops):
        b, n, c = feat.shape
        h = self.num_heads
        c_ = c // h
        if hasattr(ops, 'linear_pair'):     # one dx in the backward instead of two and their sum
            q_lin, kv = ops.linear_pair(feat, self.q.weight, self.q.bias, self.kv.weight, self.kv.bias)
        else:
            q_lin = _lin(ops, self.q, feat)
            kv = _lin(ops, self.kv, feat)
        if ops.fused_attention is not None and self.attn_drop.p == 0.:
            # global attention over a small token set is neighbourhood attention with member_idx = all tokens
            if not global_attn:
                nb = (member_idx, cluster_mask, pe_idx)
            else:
                nb = global_as_neighbourhood(pe_idx) if q_lin.dtype in (torch.float16, torch.bfloat16) else None
            if nb is not None:
                # the bias table pos_embed(pre_table) is evaluated inside; its gradient goes straight to weight / bias
                out = ops.fused_attention(q_lin, kv, h, self.scale, nb[0], nb[2], nb[1], None,
                                          self.blank_k, self.blank_v,
                                          pos_embed=(pre_table, self.pos_embed.weight, self.pos_embed.bias))
                if out is not None:                                     # b n c; None when unsupported
                    return self.proj_drop(_lin(ops, self.proj, out))
        pe_table = self.pos_embed(pre_table)                       # T*T x h
        q = (q_lin * self.scale).reshape(b, n, h, c_).permute(0, 2, 1, 3)   # b h n c_
        kv = kv.view(b, n, h, 2, c_).permute(3, 0, 2, 1, 4)        # 2 b h n c_
        key, v = kv[0], kv[1]
        if global_attn:
            attn = q @ key.transpose(-1, -2)
        else:
            attn = ops.qk(q, key, member_idx)                      # b h n m
        # attn += pe_table[pe_idx] (+ -100 at padded slots): one in-place pass (reference :126-137)
        attn = ops.attn_bias_add(attn, pe_table, pe_idx, None if global_attn else cluster_mask)
        blank_attn = (q * self.blank_k.reshape(1, h, 1, c_)).sum(-1, keepdim=True)
        attn = torch.cat([attn, blank_attn], dim=-1).softmax(dim=-1)
        attn = self.attn_drop(attn)
        blank_attn, attn = attn[..., -1:], attn[..., :-1]
        blank_v = blank_attn * self.blank_v.reshape(1, h, 1, c_)
        if global_attn:
            out = attn.to(v.dtype) @ v
        else:
            # under autocast the softmax output is fp32; the reference then runs AV in fp32 (clusten.py:52-53).
            # Here the probabilities are rounded to v's dtype so that AV stays on the bf16 tensor-core path.
            out = ops.av(attn.to(v.dtype), v, member_idx)
        feat = (out + blank_v.to(out.dtype)).permute(0, 2, 1, 3).reshape(b, n, c)
        return self.proj_drop(_lin(ops, self.proj, feat))


class ClusterTransformerBlock(nn.Module):
    def __init__(self, dim, num_heads, mlp_ratio=2., drop=0., attn_drop=0., drop_path=0., layer_scale=0.0):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim)
        self.attn = ClusterAttention(dim, num_heads, attn_drop=attn_drop, proj_drop=drop)
        self.drop_path_p = float(drop_path)
        self.norm2 = nn.LayerNorm(dim)
        self.mlp = Mlp(dim, int(dim * mlp_ratio), drop=drop)
        self.layer_scale = isinstance(layer_scale, (int, float)) and layer_scale > 0
        if self.layer_scale:
            self.gamma1 = nn.Parameter(layer_scale * torch.ones(dim))
            self.gamma2 = nn.Parameter(layer_scale * torch.ones(dim))

    def forward(self, feat, member_idx, cluster_mask, pe_idx, global_attn, pre_table, ops):
        # (shortcut, norm(shortcut)) leave one autograd node: the two gradients meet inside the LayerNorm backward
        feat, y = ops.layer_norm_fork(feat, self.norm1.weight, self.norm1.bias, self.norm1.eps)
        x = self.attn(y, member_idx, cluster_mask, pe_idx, global_attn, pre_table, ops)
        if self.layer_scale:
            x = self.gamma1 * x
        feat = ops.residual_add(feat, x, self.drop_path_p, self.training)
        feat, y = ops.layer_norm_fork(feat, self.norm2.weight, self.norm2.bias, self.norm2.eps)
        x = self.mlp(y, ops)
        if self.layer_scale:
            x = self.gamma2 * x
        return ops.residual_add(feat, x, self.drop_path_p, self.training)


class ClusterMerging(nn.Module):
    """Adaptive downsampling: keep int(n*ds_rate) tokens (importance top-k + reserved coarse grid) and merge each
    kept token's neighbourhood with a PointConv-style weighted sum (reference aff_transformer.py:247-365)."""

    def __init__(self, dim, out_dim, alpha=4.0, ds_rate=0.25, reserve_on=True):
        super().__init__()
        self.dim, self.alpha, self.ds_rate, self.reserve_on = dim, alpha, ds_rate, reserve_on
        inner_ch = 4
        self.weight_net = nn.Sequential(nn.Linear(5, inner_ch, bias=True), nn.LayerNorm(inner_ch), nn.GELU())
        self.norm = nn.LayerNorm(inner_ch * dim)
        self.linear = nn.Linear(dim * inner_ch, out_dim)

    def forward(self, pos, feat, member_idx, cluster_mask, learned_prob, stride, pe_idx, reserve_num, pre_table, ops):
        b, n, c = feat.shape
        idx, _ = ops.select_merge_centers(pos, learned_prob, stride, self.alpha, self.ds_rate, reserve_num,
                                          self.reserve_on)                               # b x n'
        n_new = idx.shape[1]
        assert n_new == int(n * self.ds_rate), "n not equal to keep num!"
        pos = pos.gather(1, idx.unsqueeze(2).expand(-1, -1, pos.shape[2]))
        member_idx = ops.gather_rows(member_idx, idx)                                    # b x n' x m
        pe_idx = ops.gather_rows(pe_idx, idx)
        if cluster_mask is not None:
            cluster_mask = ops.gather_rows(cluster_mask, idx)
        m = member_idx.shape[-1]
        wn = getattr(ops, "weight_net_table", None)                                      # T*T x ic
        weights_table = self.weight_net(pre_table) if wn is None else wn(self.weight_net, pre_table)
        fused = getattr(ops, "merge_wf", None)
        if fused is not None:    # weights formed inside the kernels from table[pe_idx], lp[member_idx] and the mask
            out = fused(weights_table, pe_idx, learned_prob, cluster_mask, feat, member_idx)
            if out is not None:
                return pos, _lin(ops, self.linear, _ln(ops, self.norm, out))
        weights = ops.table_gather(weights_table, pe_idx)                                # b x n' x m x ic
        if learned_prob is not None:
            lp = learned_prob.gather(1, member_idx.reshape(b, -1, 1)).reshape(b, n_new, m, 1)
            if cluster_mask is not None:
                lp = lp * cluster_mask.unsqueeze(3)
            weights = weights * lp
        elif cluster_mask is not None:
            weights = weights * cluster_mask.unsqueeze(3)
        feat = ops.wf(weights, feat, member_idx).reshape(b, n_new, -1)                   # b x n' x ic*c
        return pos, _lin(ops, self.linear, _ln(ops, self.norm, feat))


class BasicLayer(nn.Module):
    """One stage: (re)cluster the tokens, build neighbourhoods, run the blocks, optionally downsample
    (reference aff_transformer.py:368-508)."""

    def __init__(self, dim, out_dim, cluster_size, nbhd_size, depth, num_heads, mlp_ratio, alpha=4.0, ds_rate=0.25,
                 reserve_on=True, drop=0., attn_drop=0., drop_path=0., layer_scale=0.0, downsample=True):
        super().__init__()
        self.dim, self.nbhd_size, self.cluster_size, self.depth = dim, nbhd_size, cluster_size, depth
        self.blocks = nn.ModuleList([
            ClusterTransformerBlock(dim, num_heads, mlp_ratio, drop, attn_drop,
                                    drop_path[i] if isinstance(drop_path, (list, tuple)) else drop_path, layer_scale)
            for i in range(depth)])
        self.downsample = ClusterMerging(dim, out_dim, alpha, ds_rate, reserve_on) if downsample else None
        if downsample:
            self.prob_net = nn.Linear(dim, 1)
        self._grid_cache = {}

    def _index(self, pos, h, w, rel_pos_width, ops, pre_table):
        """-> (pos_sorted, reorder | None, nbhd_idx, mask, pe_idx, table) for local attention; `table` is pre_table
        or, for large tables, its rows referenced by this stage (pe_idx then indexes those)."""
        b, n, _ = pos.shape
        m = self.cluster_size
        k = int(math.ceil(n / float(m)))
        if k == n:
            mean, reorder = pos, None
        else:
            pos, mean, _, _, reorder = ops.space_filling_cluster(pos, m, h, w, no_reorder=False)
        nbhd, mask, pe_idx, _ = ops.build_nbhd(pos, mean, m, self.nbhd_size, rel_pos_width)
        if getattr(ops, 'hint_geometry', None) is not None:   # where nbhd / pe_idx came from: block-structured kernels
            ops.hint_geometry(nbhd, pos, rel_pos_width, 2 * rel_pos_width + 1)
        if pre_table.shape[0] > COMPACT_ROWS_ABOVE and not (pe_idx.is_cuda and torch.cuda.is_current_stream_capturing()) \
                and not (not torch.is_grad_enabled() and getattr(ops, 'hint_geometry', None) is not None
                         and pre_table.shape[0] == (2 * rel_pos_width + 1) ** 2 and pre_table.shape[0] <= 65536):
            # (inference keeps the full table of the 384-pixel configs: the block-structured forward takes the bias
            # from position differences and has no row limit)
            pe_idx, pre_table = compact_table(pe_idx, pre_table)
        return pos, reorder, nbhd, mask, pe_idx, pre_table

    def forward(self, pos, feat, h, w, on_grid, stride, rel_pos_width, table_width, pre_table, ops):
        b, n, d = pos.shape
        c = feat.shape[2]
        global_attn = self.nbhd_size >= n
        if global_attn:
            member_idx, cluster_mask = None, None
            rel_pos = (pos[:, None, :, :] + rel_pos_width) - pos[:, :, None, :]          # b x n x n x d
            pe_idx = (rel_pos[..., 1] * table_width + rel_pos[..., 0]).long()
            if getattr(ops, 'hint_geometry', None) is not None and pe_idx.is_cuda:
                # all tokens as one neighbourhood of whole 8-row blocks: the block-structured kernels take it
                nb = global_as_neighbourhood(pe_idx)
                if nb is not None:
                    ops.hint_geometry(nb[0], pos, rel_pos_width, table_width)
        else:
            if on_grid:
                key = (b, n, h, w, str(pos.device), id(ops), COMPACT_ROWS_ABOVE, id(pre_table), pre_table.dtype,
                       torch.is_grad_enabled())   # (the cached tuple holds a table derived from pre_table; its version
                # counter is left out on purpose: captured graphs hold the addresses of the cached index tensors)
                if key not in self._grid_cache:
                    self._grid_cache.clear()
                    self._grid_cache[key] = self._index(pos, h, w, rel_pos_width, ops, pre_table)
                pos, reorder, member_idx, cluster_mask, pe_idx, pre_table = self._grid_cache[key]
            else:
                pos, reorder, member_idx, cluster_mask, pe_idx, pre_table = self._index(pos, h, w, rel_pos_width, ops,
                                                                                        pre_table)
            if reorder is not None:  # store tokens of a cluster together
                feat = getattr(ops, 'reorder_tokens', _gather_tokens)(feat, reorder)
        for blk in self.blocks:
            feat = blk(feat, member_idx, cluster_mask, pe_idx, global_attn, pre_table, ops)
        if self.downsample is not None:
            learned_prob = self.prob_net(feat).sigmoid()                                   # b x n x 1
            reserve_num = math.ceil(h / (stride * 2)) * math.ceil(w / (stride * 2))
            if global_attn:
                raise NotImplementedError("downsampling after a global-attention stage (no config does this)")
            pos, feat = self.downsample(pos, feat, member_idx, cluster_mask, learned_prob, stride, pe_idx,
                                        reserve_num, pre_table, ops)
        return pos, feat


class PatchEmbed(nn.Module):
    """conv3x3/s2 -> BN -> GELU -> conv3x3/s2 (-> LN); token positions are the (x, y) grid coordinates
    (reference aff_transformer.py:511-553)."""

    def __init__(self, in_chans=3, embed_dim=32, patch_norm=True):
        super().__init__()
        self.proj1 = nn.Conv2d(in_chans, embed_dim // 2, kernel_size=3, stride=2, padding=1)
        self.bn = nn.BatchNorm2d(embed_dim // 2)
        self.act1 = nn.GELU()
        self.proj2 = nn.Conv2d(embed_dim // 2, embed_dim, kernel_size=3, stride=2, padding=1)
        self.norm = nn.LayerNorm(embed_dim) if patch_norm else None

    def forward(self, x, ops=None):
        if x.is_cuda:
            # NHWC through the stem: cuDNN's native layout (no NCHW<->NHWC conversion kernels around the convs), the
            # per-channel BatchNorm reductions run over contiguous channel vectors, and the token matrix [b, h*w, c]
            # below is then a view instead of a transpose copy
            # (under autocast the cast to the compute dtype rides on the same copy: one pass over the images, not two)
            dt = torch.get_autocast_dtype("cuda") if torch.is_autocast_enabled("cuda") else x.dtype
            x = x.to(dtype=dt, memory_format=torch.channels_last)
        if ops is not None and hasattr(ops, 'conv_batch_norm_act'):
            x = ops.conv_batch_norm_act(x, self.proj1, self.bn)
        else:
            x = self.proj1(x)
            x = self.act1(self.bn(x)) if ops is None else ops.batch_norm_act(x, self.bn)
        x = ops.conv_bias(x, self.proj2) if ops is not None and hasattr(ops, 'conv_bias') and x.is_cuda else self.proj2(x)
        b, c, h, w = x.shape
        x = x.flatten(2).transpose(1, 2)
        if self.norm is not None:
            x = self.norm(x) if ops is None else _ln(ops, self.norm, x)
        ys, xs = torch.meshgrid(torch.arange(h, device=x.device), torch.arange(w, device=x.device), indexing="ij")
        pos = torch.stack([xs, ys], dim=2).reshape(1, -1, 2).expand(b, -1, -1).to(x.dtype)
        return pos, x, h, w


class AutoFocusFormer(nn.Module):
    """reference aff_transformer.py:556-670 (same constructor arguments; `downsample`/`norm_layer` are fixed to the
    only values the reference uses)."""

    def __init__(self, in_chans=3, num_classes=1000, embed_dim=(32, 128, 256, 512), cluster_size=8,
                 nbhd_size=(48, 48, 48, 49), alpha=4.0, ds_rate=0.25, reserve_on=True, depths=(2, 2, 6, 2),
                 num_heads=(3, 6, 12, 24), mlp_ratio=2., drop_rate=0., attn_drop_rate=0., drop_path_rate=0.1,
                 patch_norm=True, layer_scale=0.0, img_size=224, ops=None, **kwargs):
        super().__init__()
        self.num_classes, self.num_layers = num_classes, len(depths)
        self.num_features = embed_dim[-1]
        self.ops = ops
        self.patch_embed = PatchEmbed(in_chans, embed_dim[0], patch_norm)
        self.pos_drop = nn.Dropout(p=drop_rate)
        self.rel_pos_width, self.table_width, table = build_pe_lookup(img_size)
        self.register_buffer("pre_table", table, persistent=False)
        dpr = [x.item() for x in torch.linspace(0, drop_path_rate, sum(depths))]
        self.layers = nn.ModuleList()
        for i in range(self.num_layers):
            last = i == self.num_layers - 1
            self.layers.append(BasicLayer(
                dim=int(embed_dim[i]), out_dim=None if last else int(embed_dim[i + 1]), cluster_size=cluster_size,
                nbhd_size=nbhd_size[i], depth=depths[i], num_heads=num_heads[i], mlp_ratio=mlp_ratio, alpha=alpha,
                ds_rate=ds_rate, reserve_on=reserve_on, drop=drop_rate, attn_drop=attn_drop_rate,
                drop_path=dpr[sum(depths[:i]):sum(depths[:i + 1])], layer_scale=layer_scale, downsample=not last))
        self.norm = nn.LayerNorm(self.num_features)
        self.head = nn.Linear(self.num_features, num_classes) if num_classes > 0 else nn.Identity()
        self.apply(self._init_weights)

    @staticmethod
    def _init_weights(m):
        if isinstance(m, nn.Linear):
            nn.init.trunc_normal_(m.weight, std=.02)
            if m.bias is not None:
                nn.init.constant_(m.bias, 0)
        elif isinstance(m, nn.LayerNorm):
            nn.init.constant_(m.bias, 0)
            nn.init.constant_(m.weight, 1.0)

    def forward_features(self, x):
        ops = self.ops if self.ops is not None else default_ops()
        pos, x, h, w = self.patch_embed(x, ops)
        x = self.pos_drop(x)
        for i, layer in enumerate(self.layers):
            pos, x = layer(pos, x, h, w, i == 0, 2 ** (i + 1), self.rel_pos_width, self.table_width,
                           self.pre_table, ops)
        return self.norm(x).mean(1)

    def forward(self, x):
        return self.head(self.forward_features(x))


# hyper-parameters of the reference's configs (configs/aff_*.yaml)
CONFIGS = {
    "aff_mini": dict(embed_dim=[32, 128, 256, 384], depths=[2, 2, 6, 2], num_heads=[2, 4, 8, 16], mlp_ratio=2.,
                     nbhd_size=[48, 48, 48, 49], cluster_size=8, ds_rate=0.25, drop_path_rate=0.0, img_size=224),
    "aff_mini_1_5th": dict(embed_dim=[32, 128, 256, 384], depths=[2, 2, 6, 2], num_heads=[2, 4, 8, 16], mlp_ratio=2.,
                           nbhd_size=[48, 48, 48, 49], cluster_size=8, ds_rate=0.2, drop_path_rate=0.0, img_size=224),
    "aff_tiny": dict(embed_dim=[64, 128, 256, 512], depths=[3, 4, 18, 5], num_heads=[2, 4, 8, 16], mlp_ratio=3.,
                     nbhd_size=[48, 48, 48, 49], cluster_size=8, ds_rate=0.25, drop_path_rate=0.2, img_size=224),
    "aff_tiny_1_5th": dict(embed_dim=[64, 128, 256, 512], depths=[3, 4, 18, 5], num_heads=[2, 4, 8, 16], mlp_ratio=3.,
                           nbhd_size=[48, 48, 48, 49], cluster_size=8, ds_rate=0.2, drop_path_rate=0.2, img_size=224),
    "aff_base_22kto1k_384": dict(embed_dim=[128, 256, 512, 1024], depths=[3, 4, 18, 2], num_heads=[4, 8, 16, 32],
                                 mlp_ratio=3., nbhd_size=[144, 144, 144, 144], cluster_size=24, ds_rate=0.25,
                                 drop_path_rate=0.2, img_size=384, layer_scale=1e-5),
}


def build_aff(name, **overrides):
    cfg = dict(CONFIGS[name])
    cfg.update(overrides)
    return AutoFocusFormer(**cfg)
