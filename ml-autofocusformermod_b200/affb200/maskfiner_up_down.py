"""maskfiner_up_down — the fork's U-shaped multi-resolution classifier (BASELINE config 5) assembled around this package's
MixResNeighbour stage (mixres_neighbour.BasicLayer: clustering, kNN neighbourhoods and cluster attention on the CUDA
operators).  Mirrors what the reference wires together in models/maskfiner_up_down.py:28-128 (UpDownBackbone),
models/mixres_neighbour.py:580-973 (MixResNeighbour) and models/mixres_vit.py:325-421 (MixResViT), with the reference's
module and parameter names so that its state_dict loads unchanged.

Data flow (7 backbones for the "small" config, scales 0 1 2 3 2 1 0; positions are (scale, x, y) in units of the finest
patch):
    j = 0      MixResViT on the coarsest patch grid (dense attention over H/32 x W/32 tokens) -> res5
    j = 1..3   MixResNeighbour: the half of the tokens of the previous scale with the largest importance score is split
               into four children each (positions offset by half the parent's patch, features = parent + learned
               child / scale embeddings + an MLP of the child's image patch); parents are kept (keep_old_scale); all
               tokens of all scales then go through one off-grid attention stage together  -> res5 .. res(5 - j)
    j = 4..6   no splitting: the stage input is the previous output concatenated channel-wise with the same tokens'
               features from the mirror stage on the way up (U-Net skip), re-matched by position
The importance score is a standard normal draw per token (reference maskfiner_up_down.py:147-149); pass `mask_fn` to
make it reproducible.  The token re-matching between stages (`find_pos_org_order`, :150-154: L1 cdist + argmin) is
affb200_nearest_l1 — no N x N matrix.

Everything outside BasicLayer is stock PyTorch here as in the reference: conv patch embeddings,
LayerNorms, the image-patch MLPs and MixResViT's dense attention (torch SDPA).
"""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from .mixres_neighbour import BasicLayer, divide_tokens_to_split_and_keep

SMALL = dict(  # reference configs/maskfiner_up_down_small.yaml:5-24
    names=["MixResViT", "MixResNeighbour", "MixResNeighbour", "MixResNeighbour", "MixResNeighbour", "MixResNeighbour", "MixResViT"],
    embed_dim=[512, 256, 128, 64, 128, 256, 512], depths=[2, 2, 2, 4, 6, 24, 3], num_heads=[16, 8, 4, 2, 4, 8, 16],
    patch_sizes=[32, 16, 8, 4, 8, 16, 32], drop_path_rate=0.3, upscale_ratio=[0.0, 0.5, 0.5, 0.5, 0.0, 0.0, 0.0],
    split_ratio=4, mlp_ratio=4.0, cluster_size=8, nbhd_size=48, out_features=["res2", "res3", "res4", "res5"],
    keep_old_scale=True, n_scales=4, layer_scale=1e-5)


_grid_pos = {}


def grid_positions(height, width, patch_size, min_patch_size, scale, device="cpu"):
    """(scale, x, y) of every patch of size `patch_size`, row-major over y then x, in units of the finest patch
    (reference mixres_vit.py:66-77).  One copy per geometry and device (no host copy inside a captured forward)."""
    key = (height, width, patch_size, min_patch_size, scale, str(device))
    if key not in _grid_pos:
        step = patch_size // min_patch_size
        xs = torch.arange(0, width // min_patch_size, step)
        ys = torch.arange(0, height // min_patch_size, step)
        yy, xx = torch.meshgrid(ys, xs, indexing="ij")
        xy = torch.stack([xx, yy], dim=2).reshape(-1, 2)
        _grid_pos[key] = torch.cat([torch.full((xy.shape[0], 1), scale, dtype=xy.dtype), xy], dim=1).to(device)
    return _grid_pos[key]


def group_norm_1(x, weight, bias, eps):
    """nn.GroupNorm(1, C) of the reference's patch-embedding convs (mixres_vit.py:219-232): statistics over all of
    (C, H, W) per image.  Stock F.group_norm gives one thread block per (image, group) — at batch 1 on a 512 x 1024
    feature map that is a single block reducing 16 M values (measured 17 ms of the 19 ms first backbone); this form
    uses torch's multi-block reductions.  fp32 like the autocast policy of group_norm."""
    xf = x.float()
    var, mean = torch.var_mean(xf, dim=(1, 2, 3), unbiased=False, keepdim=True)
    scale = torch.rsqrt(var + eps) * weight.float().view(1, -1, 1, 1)
    return torch.addcmul(bias.float().view(1, -1, 1, 1) - mean * scale, xf, scale)


class PositionEmbeddingSine(nn.Module):
    """sin/cos embedding of (x, y), coordinates normalised by their maximum over the whole batch (mixres_vit.py:11-63)."""

    def __init__(self, num_pos_feats, temperature=10000):
        super().__init__()
        self.num_pos_feats, self.temperature = num_pos_feats, temperature

    def forward(self, pos):
        x, y = pos[:, :, 0], pos[:, :, 1]
        x = torch.clamp(x / (x.max() + 1e-6), 0, 1) * (2 * math.pi)
        y = torch.clamp(y / (y.max() + 1e-6), 0, 1) * (2 * math.pi)
        k = torch.arange(self.num_pos_feats, dtype=torch.float32, device=pos.device)
        dim_t = self.temperature ** (2 * k.div(2, rounding_mode="floor") / self.num_pos_feats)
        px, py = x[:, :, None] / dim_t, y[:, :, None] / dim_t
        px = torch.cat((px[:, :, 0::2].sin(), px[:, :, 1::2].cos()), dim=2)
        py = torch.cat((py[:, :, 0::2].sin(), py[:, :, 1::2].cos()), dim=2)
        return torch.cat((px, py), dim=2)


class _DownConv(nn.Module):
    """3x3 stride-2 conv -> LeakyReLU -> norm; the ViT stage normalises with GroupNorm(1) (`g_norm`, mixres_vit.py:
    219-232), the neighbour stage with BatchNorm (`b_norm`, mixres_neighbour.py:546-559)."""

    def __init__(self, cin, cout, batch_norm):
        super().__init__()
        self.conv = nn.Conv2d(cin, cout, kernel_size=3, stride=2, padding=1)
        if batch_norm:
            self.b_norm = nn.BatchNorm2d(cout)
        else:
            self.g_norm = nn.GroupNorm(1, cout)
        self.relu = nn.LeakyReLU()

    def forward(self, x):
        x = self.relu(self.conv(x))
        if hasattr(self, "b_norm"):
            return self.b_norm(x)
        return group_norm_1(x, self.g_norm.weight, self.g_norm.bias, self.g_norm.eps)


class OverlapPatchEmbedding(nn.Module):
    """log2(patch) stride-2 convs, channels doubling up to embed_dim, then LayerNorm over tokens."""

    def __init__(self, patch_size, embed_dim, channels, batch_norm):
        super().__init__()
        n = int(round(math.log2(patch_size)))
        dims = [channels] + [int(embed_dim // 2 ** (n - 1 - i)) for i in range(n)]
        self.conv_layers = nn.Sequential(*[_DownConv(dims[i], dims[i + 1], batch_norm) for i in range(n)])
        self.out_norm = nn.LayerNorm(embed_dim)

    def forward(self, im):
        return self.out_norm(self.conv_layers(im).flatten(2).transpose(1, 2))


def _init(m):
    """reference mixres_vit.py:234-246"""
    if isinstance(m, nn.Linear):
        nn.init.trunc_normal_(m.weight, std=0.02)
        if m.bias is not None:
            nn.init.constant_(m.bias, 0)
    elif isinstance(m, nn.LayerNorm):
        nn.init.constant_(m.bias, 0)
        nn.init.constant_(m.weight, 1.0)
    elif isinstance(m, nn.Conv2d):
        nn.init.kaiming_normal_(m.weight, a=0.2, mode="fan_out", nonlinearity="leaky_relu")
        if m.bias is not None:
            nn.init.zeros_(m.bias)


def _drop_path(x, p, training):
    if p == 0. or not training:
        return x
    keep = x.new_empty((x.shape[0],) + (1,) * (x.ndim - 1)).bernoulli_(1 - p)
    return x * keep.div_(1 - p)


# ---- dense stage (first and last backbone) ------------------------------------------------------------------------
class _DWConv(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.dwconv = nn.Conv2d(dim, dim, 3, 1, 1, bias=True, groups=dim)

    def forward(self, x, h, w):
        b, n, c = x.shape
        return self.dwconv(x.transpose(1, 2).reshape(b, c, h, w)).flatten(2).transpose(1, 2)


class _FeedForward(nn.Module):
    def __init__(self, dim, hidden):
        super().__init__()
        self.fc1 = nn.Linear(dim, hidden)
        self.dwconv = _DWConv(hidden)
        self.fc2 = nn.Linear(hidden, dim)

    def forward(self, x, h, w):
        return self.fc2(F.gelu(self.dwconv(self.fc1(x), h, w)))


class _Attention(nn.Module):
    def __init__(self, dim, heads):
        super().__init__()
        self.heads = heads
        self.qkv = nn.Linear(dim, dim * 3)
        self.proj = nn.Linear(dim, dim)

    def forward(self, x):
        b, n, c = x.shape
        q, k, v = self.qkv(x).reshape(b, n, 3, self.heads, c // self.heads).permute(2, 0, 3, 1, 4)
        return self.proj(F.scaled_dot_product_attention(q, k, v).transpose(1, 2).reshape(b, n, c))


class _Block(nn.Module):
    def __init__(self, dim, heads, hidden, drop_path, layer_scale):
        super().__init__()
        self.norm1, self.norm2 = nn.LayerNorm(dim), nn.LayerNorm(dim)
        self.attn = _Attention(dim, heads)
        self.mlp = _FeedForward(dim, hidden)
        self.p = drop_path
        self.layer_scale = bool(layer_scale and layer_scale > 0)
        if self.layer_scale:
            self.gamma1 = nn.Parameter(layer_scale * torch.ones(dim))
            self.gamma2 = nn.Parameter(layer_scale * torch.ones(dim))

    def forward(self, x, h, w):
        y = self.attn(self.norm1(x))
        x = x + _drop_path(self.gamma1 * y if self.layer_scale else y, self.p, self.training)
        y = self.mlp(self.norm2(x), h, w)
        return x + _drop_path(self.gamma2 * y if self.layer_scale else y, self.p, self.training)


class _TransformerLayer(nn.Module):
    def __init__(self, n, dim, heads, hidden, drop_path, layer_scale):
        super().__init__()
        self.blocks = nn.ModuleList([_Block(dim, heads, hidden, drop_path[i], layer_scale) for i in range(n)])

    def forward(self, x, h, w):
        for blk in self.blocks:
            x = blk(x, h, w)
        return x


class MixResViT(nn.Module):
    """Dense-attention stage on one regular patch grid (reference mixres_vit.py:325-421)."""

    def __init__(self, patch_sizes, n_layers, d_model, n_heads, mlp_ratio=4.0, drop_path_rate=(0.0,), channels=3,
                 n_scales=2, min_patch_size=4, upscale_ratio=0.0, first_layer=True, layer_scale=0.0, out_features=("res5",)):
        super().__init__()
        self.patch_size, self.patch_sizes = patch_sizes[-1], patch_sizes
        self.min_patch_size, self.upscale_ratio, self.first_layer = min_patch_size, upscale_ratio, first_layer
        self._out_features = list(out_features)
        if first_layer:
            self.pe_layer = PositionEmbeddingSine(d_model // 2)
            self.patch_embed = OverlapPatchEmbedding(self.patch_size, d_model, channels, batch_norm=False)
        else:
            self.token_norm = nn.LayerNorm(channels)
            self.token_projection = nn.Linear(channels, d_model) if channels != d_model else nn.Identity()
        self.layers = _TransformerLayer(n_layers, d_model, n_heads, int(d_model * mlp_ratio), list(drop_path_rate), layer_scale)
        self.norm_out = nn.LayerNorm(d_model)
        self.apply(_init)

    def forward(self, im, scale, features, features_pos, upsampling_mask):
        B, _, H, W = im.shape
        grid = (H // self.patch_size, W // self.patch_size)
        if self.first_layer:
            pos = grid_positions(H, W, self.patch_size, self.min_patch_size, scale, im.device).repeat(B, 1, 1)
            x = self.patch_embed(im)
            x = x + self.pe_layer(pos[:, :, 1:])
        else:
            x, pos = self.token_projection(self.token_norm(features)), features_pos
        x = self.layers(x, h=grid[0], w=grid[1])
        name = self._out_features[0]
        return {name: self.norm_out(x.float()).to(x.dtype), name + "_pos": pos[:, :, 1:], name + "_scale": pos[:, :, 0],
                name + "_spatial_shape": grid, "min_spatial_shape": (H // self.min_patch_size, W // self.min_patch_size)}


# ---- off-grid stage -------------------------------------------------------------------------------------------------
class _MLPBlock(nn.Module):
    def __init__(self, cin, cout):
        super().__init__()
        self.linear = nn.Linear(cin, cout)
        self.norm = nn.LayerNorm(cout)

    def forward(self, x):
        return self.norm(F.gelu(self.linear(x)))


class _MLPDeepNorm(nn.Module):
    def __init__(self, cin, hidden, cout, num_layers=3):
        super().__init__()
        dims = [cin] + [hidden] * (num_layers - 1) + [cout]
        self.layers = nn.ModuleList([_MLPBlock(dims[i], dims[i + 1]) for i in range(num_layers)])

    def forward(self, x):
        for layer in self.layers:
            x = layer(x)
        return x


_group_sizes = {}


def _split_rows(tensors, key_col, n_groups, cache_key):
    """Rows of each tensor in `tensors` ([B, N, ...]) grouped by the small non-negative integer key_col [B, N] (a
    scale or a 0/1 flag), every group in the original row order (the reference's torch.where scan order,
    mixres_neighbour.py:756-776,955-962): -> list over key values of lists of tensors.  Every image has the same group
    sizes (the token budget per scale is a function of the shapes alone), so they are read once per `cache_key` and
    reused: later passes make no host round trip and can be captured in a CUDA graph."""
    k8 = key_col.to(torch.uint8)
    sizes = _group_sizes.get(cache_key)
    if sizes is None:
        counts = [torch.bincount(r.long(), minlength=n_groups) for r in k8]
        assert all(torch.equal(c, counts[0]) for c in counts[1:]), "every image must carry the same number of tokens per scale"
        sizes = counts[0].tolist()
        assert len(sizes) == n_groups, sizes
        _group_sizes[cache_key] = sizes
    if k8.is_cuda:        # affb200_argsort_f32 (tile sort + rank merge): a fraction of torch's radix-sort launch chain
        from .point_utils import argsort_stable
        order = argsort_stable(k8.float())
    else:
        order = k8.argsort(dim=1, stable=True)
    out, at = [], 0
    for n in sizes:
        idx = order[:, at:at + n]
        out.append([t.gather(1, idx.reshape(*idx.shape, *([1] * (t.dim() - 2))).expand(-1, -1, *t.shape[2:])) for t in tensors])
        at += n
    return out


class MixResNeighbour(nn.Module):
    """Off-grid attention stage over the tokens of all scales so far, optionally preceded by the token split
    (reference mixres_neighbour.py:580-973).  `ops`: the operator set of the stage (default: the CUDA operators)."""

    def __init__(self, patch_sizes, n_layers, d_model, n_heads, drop_path_rate=(0.0,), channels=1, mlp_ratio=4.0,
                 split_ratio=4, n_scales=4, cluster_size=8, nbhd_size=48, layer_scale=0.0, min_patch_size=4,
                 upscale_ratio=0.25, keep_old_scale=False, scale=1, first_layer=False, out_features=("res5",), ops=None,
                 argsort=None):
        super().__init__()
        self.argsort = argsort      # stable ascending argsort of the scores; None = affb200_argsort_f32
        self.patch_size, self.patch_sizes = patch_sizes[-1], patch_sizes
        self.split_ratio, self.n_scales, self.min_patch_size = split_ratio, n_scales, min_patch_size
        self.upscale_ratio, self.keep_old_scale, self.scale, self.first_layer = upscale_ratio, keep_old_scale, scale, first_layer
        self.do_upsample = not (upscale_ratio == 0 or first_layer)
        self._out_features = list(out_features)
        self.layers = BasicLayer(dim=int(d_model), cluster_size=cluster_size, nbhd_size=nbhd_size, depth=n_layers,
                                 num_heads=n_heads, mlp_ratio=mlp_ratio, drop_path=list(drop_path_rate),
                                 layer_scale=layer_scale, ops=ops)
        if first_layer:
            self.pos_embed = PositionEmbeddingSine(d_model // 2)
            self.patch_embed = OverlapPatchEmbedding(self.patch_size, d_model, channels, batch_norm=True)
        else:
            if self.do_upsample:
                self.rel_pos_emb = nn.Parameter(torch.randn(1, split_ratio, channels))
                self.scale_emb = nn.Parameter(torch.randn(1, 1, channels))
                self.image_patch_projection = nn.Linear(3 * self.patch_size ** 2, channels)
                self.high_res_norm1 = nn.LayerNorm(channels)
                self.high_res_mlp = _MLPDeepNorm(channels, channels, channels)
                self.high_res_norm2 = nn.LayerNorm(channels)
                self.image_feat_importance = nn.Parameter(torch.ones(1))
                self.old_feat_importance = nn.Parameter(torch.ones(1))
            self.token_norm = nn.LayerNorm(channels)
            self.token_projection = nn.Linear(channels, d_model) if channels != d_model else nn.Identity()
        self.norm_out = nn.LayerNorm(d_model)
        self.apply(_init)

    # -- token split (:885-933) ------------------------------------------------------------------------------------
    def _children(self, feat, pos, scale):
        """four children per split token: features (:780-786) and positions (:788-804: the parent's corner, +half a
        parent patch in x, in y, in both; scale column = the new scale)."""
        b, n, c = feat.shape
        f = (feat.unsqueeze(2) + self.rel_pos_emb + self.scale_emb).reshape(b, n * self.split_ratio, c)
        half = 2 ** (self.n_scales - scale - 1)
        x, y = pos[:, :, 1], pos[:, :, 2]
        xy = torch.stack([torch.stack([x, y], 2), torch.stack([x + half, y], 2), torch.stack([x, y + half], 2),
                          torch.stack([x + half, y + half], 2)], dim=2).reshape(b, n * self.split_ratio, 2)
        sc = torch.full((b, xy.shape[1], 1), scale, dtype=torch.int32, device=pos.device)
        return f, torch.cat([sc, xy], dim=2)

    def _image_patches(self, tokens, xy, im):
        """adds an MLP of each child's own image patch (:807-829)"""
        b, n, _ = xy.shape
        ps = self.patch_size
        org = xy * self.min_patch_size
        off = torch.arange(ps, device=xy.device)
        oy, ox = torch.meshgrid(off, off, indexing="ij")                       # patch pixels row-major: (dy, dx)
        px = (org[:, :, None, 0] + ox.reshape(1, 1, -1)).long()                # b x n x ps^2
        py = (org[:, :, None, 1] + oy.reshape(1, 1, -1)).long()
        bi = torch.arange(b, device=xy.device).view(b, 1, 1).expand_as(px)
        patch = im[bi, :, py, px].reshape(b, n, ps * ps * 3)                    # (pixel, channel) fastest-last
        h = self.high_res_norm2(self.high_res_mlp(self.high_res_norm1(F.gelu(self.image_patch_projection(patch)))))
        return self.old_feat_importance * tokens + self.image_feat_importance * h

    def upsample_features(self, im, scale, features, features_pos, upsampling_mask):
        B, N, _ = features.shape
        at_old = features_pos[:, :, 0] == scale - 1
        per_token = [features, features_pos] + ([upsampling_mask] if upsampling_mask.shape[1] == N else [])
        (curr, old) = _split_rows(per_token, ~at_old, 2, (id(self), "split", B, N))
        feat_curr, pos_curr, feat_old, pos_old = curr[0], curr[1], old[0], old[1]
        mask = curr[2] if len(curr) > 2 else upsampling_mask
        f_split, p_split, f_keep, p_keep = divide_tokens_to_split_and_keep(
            feat_curr, pos_curr, mask, self.upscale_ratio, argsort=self.argsort)
        f_child, p_child = self._children(f_split.clone(), p_split.clone(), scale)
        f_child = self._image_patches(f_child, p_child[:, :, 1:], im)
        feats, poss = [feat_old, f_keep], [pos_old, p_keep]
        if self.keep_old_scale:
            feats.append(f_split)
            poss.append(p_split)
        feats.append(f_child)
        poss.append(p_child)
        all_feat = torch.cat(feats, dim=1)
        all_pos = torch.cat(poss, dim=1)
        return self.token_projection(self.token_norm(all_feat)), all_pos

    def forward(self, im, scale, features, features_pos, upsampling_mask):
        B, _, H, W = im.shape
        grid_min = (H // self.min_patch_size, W // self.min_patch_size)
        if self.first_layer:
            pos = grid_positions(H, W, self.patch_size, self.min_patch_size, scale, im.device).repeat(B, 1, 1)
            x = self.patch_embed(im)
            x = x + self.pos_embed(pos[:, :, 1:])
        elif self.do_upsample:
            x, pos = self.upsample_features(im, scale, features, features_pos, upsampling_mask)
        else:
            x, pos = self.token_projection(self.token_norm(features)), features_pos
        pos, x = self.layers(pos, x, h=grid_min[0], w=grid_min[1], on_grid=False)
        outs = {}
        groups = _split_rows([pos, x], pos[:, :, 0], scale + 1, (id(self), "out", B, x.shape[1]))
        for s, (p_s, x_s) in enumerate(groups):
            name = "res%d" % (self.n_scales - s + 1)
            outs[name] = self.norm_out(x_s.float()).to(x_s.dtype)
            outs[name + "_pos"], outs[name + "_scale"] = p_s[:, :, 1:], p_s[:, :, 0]
            outs[name + "_spatial_shape"] = (H // self.patch_sizes[s], W // self.patch_sizes[s])
        outs["min_spatial_shape"] = grid_min
        return outs


# ---- the chain ------------------------------------------------------------------------------------------------------
class UpDownBackbone(nn.Module):
    """reference maskfiner_up_down.py:28-154.  `find_order(pos_org, pos_shuffled) -> indices` re-matches a stage's
    tokens of one scale to the order that scale had when it first appeared (default: affb200_nearest_l1);
    `mask_fn(B, N, device) -> [B, N]` draws the importance scores (default: standard normal, as the reference)."""

    def __init__(self, backbones, backbone_dims, all_out_features, n_scales, num_classes, bb_in_feats, aux_loss=False,
                 find_order=None, mask_fn=None):
        super().__init__()
        self.backbones = nn.ModuleList(backbones)
        self.final_upsampling_ratios = [b.upscale_ratio for b in backbones]
        self.all_out_features, self.n_scales, self.bb_in_feats = list(all_out_features), n_scales, bb_in_feats
        up = list(range(n_scales))
        self.bb_scales = up + up[-2::-1]
        self.aux_loss = aux_loss
        if aux_loss:
            self.heads = nn.ModuleList([nn.Linear(backbone_dims[i - n_scales], num_classes) for i in range(n_scales)])
        else:
            self.head = nn.Linear(backbone_dims[-1], num_classes)
        self.find_order, self.mask_fn = find_order, mask_fn
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

    def _order(self, pos_org, pos_new):
        if self.find_order is not None:
            return self.find_order(pos_org, pos_new)
        from .point_utils import find_pos_org_order
        return find_pos_org_order(pos_org, pos_new)

    def forward(self, im):
        outs, features, features_pos, mask, output = {}, None, None, None, None
        n_bb = len(self.backbones)
        for j, bb in enumerate(self.backbones):
            output = bb(im, self.bb_scales[j], features, features_pos, mask)
            feats, poss, scales = [], [], []
            for f in bb._out_features:
                feat, fpos, fscale = output[f], output[f + "_pos"], output[f + "_scale"]
                B, N, _ = feat.shape
                if f + "_pos" in outs:      # seen before: back to the order of its first appearance
                    order = self._order(outs[f + "_pos"], fpos)
                    gi = lambda t: t.gather(1, order.reshape(B, N, *([1] * (t.dim() - 2))).expand(-1, -1, *t.shape[2:]))
                    feat, fpos, fscale = gi(feat), gi(fpos), gi(fscale)
                    outs[f].append(feat)
                else:
                    outs[f], outs[f + "_pos"], outs[f + "_scale"] = [feat], fpos, fscale
                    outs[f + "_spatial_shape"] = output[f + "_spatial_shape"]
                if j + 1 < len(self.bb_in_feats) and f in self.bb_in_feats[j + 1]:
                    if j >= self.n_scales - 1:      # way down: concatenate the mirror stage's features of this scale
                        feat = torch.cat([feat, outs[f][-((j - self.n_scales + 1) * 2 + 2)]], dim=2)
                    feats.append(feat)
                    poss.append(fpos)
                    scales.append(fscale)
            if j < self.n_scales - 1:               # still going up: scores for the next split
                B, N, _ = feats[0].shape
                mask = self.mask_fn(B, N, im.device) if self.mask_fn is not None else torch.randn(B, N, device=im.device)
            if j < n_bb - 1:
                p, s = torch.cat(poss, dim=1), torch.cat(scales, dim=1)
                features_pos = torch.cat([s.unsqueeze(2), p], dim=2)
                features = torch.cat(feats, dim=1)
        if self.aux_loss:
            return [self.heads[i](outs[f][-1].mean(1)) for i, f in enumerate(self.all_out_features)]
        return self.head(output[self.all_out_features[-1]].mean(1))


def build_maskfiner_up_down(cfg=None, num_classes=1000, ops=None, argsort=None, **kw):
    """The seven-backbone model of reference models/build.py:86-160 for a config dict like SMALL."""
    c = dict(SMALL if cfg is None else cfg)
    n_scales, n_layers = c["n_scales"], len(c["names"])
    min_ps = c["patch_sizes"][n_scales - 1]
    bb_in_feats = [[None], ["res5"], ["res5", "res4"], ["res5", "res4", "res3"], ["res5", "res4", "res3"],
                   ["res5", "res4"], ["res5"], [None]]
    dpr = [x.item() for x in torch.linspace(0, c["drop_path_rate"], sum(c["depths"]), device="cpu")]
    bbs = []
    for li, name in enumerate(c["names"]):
        first = li == 0
        if li >= n_scales:
            scale = n_layers - li - 1
            patch_sizes = c["patch_sizes"][li:]
            out_features = c["out_features"][-(n_layers - li):]
            in_ch = c["embed_dim"][li - 1] + c["embed_dim"][n_layers - li - 1]
        else:
            scale = li
            patch_sizes = c["patch_sizes"][:li + 1]
            out_features = c["out_features"][-(li + 1):]
            in_ch = 3 if first else c["embed_dim"][li - 1]
        dp = dpr[sum(c["depths"][:li]):sum(c["depths"][:li + 1])]
        common = dict(patch_sizes=patch_sizes, n_layers=c["depths"][li], d_model=c["embed_dim"][li], n_heads=c["num_heads"][li],
                      mlp_ratio=c["mlp_ratio"], drop_path_rate=dp, channels=in_ch, n_scales=n_scales, min_patch_size=min_ps,
                      upscale_ratio=c["upscale_ratio"][li], first_layer=first, layer_scale=c["layer_scale"],
                      out_features=out_features)
        if name == "MixResViT":
            bbs.append(MixResViT(**common))
        else:
            bbs.append(MixResNeighbour(split_ratio=c["split_ratio"], cluster_size=c["cluster_size"], nbhd_size=c["nbhd_size"],
                                       keep_old_scale=c["keep_old_scale"], scale=scale, ops=ops, argsort=argsort,
                                       **common))
    return UpDownBackbone(bbs, c["embed_dim"], c["out_features"], n_scales, num_classes, bb_in_feats, **kw)
