"""Operator set of the CPU oracle, shaped like affb200.aff_transformer.CudaOps.  TEST INFRASTRUCTURE (see
oracle/__init__.py): tests and bench.py's cpu_baseline / `--impl reference` legs inject it into the host glue
(`AutoFocusFormer(ops=OracleOps())`) to run the reference's pure-PyTorch path on the host cores."""
import torch

from . import clusten_ref as R
from . import point_ref as P


class OracleOps:
    fused_attention = None

    qk = staticmethod(R.qk_ref)
    av = staticmethod(R.av_ref)
    wf = staticmethod(R.wf_ref)

    @staticmethod
    def space_filling_cluster(pos, m, h, w, no_reorder=False):
        assert not no_reorder
        return P.sfc_ref(pos, m, h, w)

    @staticmethod
    def build_nbhd(pos, mean, m, nbhd_size, rel_pos_width, clamp=False):
        return P.nbhd_ref(pos, mean, m, nbhd_size, rel_pos_width, clamp)

    @staticmethod
    def select_merge_centers(pos, learned_prob, stride, alpha, ds_rate, reserve_num, reserve_on=True):
        return P.select_ref(pos, learned_prob, stride, alpha, ds_rate, reserve_num, reserve_on)

    @staticmethod
    def residual_add(feat, x, p, training):
        """`shortcut + drop_path(x)` with timm's DropPath (aff_transformer.py:224-229)"""
        if p == 0. or not training:
            return feat + x
        keep = 1 - p
        mask = x.new_empty((x.shape[0],) + (1,) * (x.dim() - 1)).bernoulli_(keep)
        return feat + x * mask.div_(keep)

    @staticmethod
    def linear(x, weight, bias=None):
        """nn.Linear (aff_transformer.py:85-100,285)"""
        return torch.nn.functional.linear(x, weight, bias)

    @staticmethod
    def batch_norm_act(x, bn):
        """nn.BatchNorm2d then nn.GELU (aff_transformer.py:524-527,541)"""
        return torch.nn.functional.gelu(bn(x))

    @staticmethod
    def linear_gelu(x, weight, bias=None):
        """fc1 then nn.GELU of the Mlp (aff_transformer.py:35-51)"""
        return torch.nn.functional.gelu(torch.nn.functional.linear(x, weight, bias))

    @staticmethod
    def layer_norm(x, weight, bias, eps=1e-5):
        """nn.LayerNorm as the reference applies it (aff_transformer.py:199-200,284)"""
        return torch.nn.functional.layer_norm(x, (x.shape[-1],), weight, bias, eps)

    @staticmethod
    def layer_norm_fork(x, weight, bias, eps=1e-5):
        """`shortcut = feat; feat = norm(feat)` (aff_transformer.py:223-229)"""
        return x, torch.nn.functional.layer_norm(x, (x.shape[-1],), weight, bias, eps)

    @staticmethod
    def attn_bias_add(attn, pe_table, pe_idx, mask=None):
        """reference aff_transformer.py:131-137"""
        h = attn.shape[1]
        pos_embed = pe_table.gather(index=pe_idx.reshape(-1, 1).expand(-1, h), dim=0).reshape(*pe_idx.shape, h)
        attn = attn + pos_embed.permute(0, 3, 1, 2)
        if mask is not None:
            attn = attn + (1 - mask.reshape(mask.shape[0], 1, mask.shape[1], -1)) * (-100)
        return attn

    @staticmethod
    def table_gather(table, pe_idx):
        """reference aff_transformer.py:347-349"""
        ic = table.shape[-1]
        return table.gather(index=pe_idx.reshape(-1, 1).expand(-1, ic), dim=0).reshape(*pe_idx.shape, ic)

    @staticmethod
    def gather_rows(src, idx):
        return src.gather(1, idx.unsqueeze(2).expand(-1, -1, src.shape[2]))


def fill_parameters_by_name(model, seed=0, scale=0.25):
    """Deterministic weights that do not depend on construction order: every parameter is drawn from a generator
    seeded by (seed, crc32(name)).  Norm weights stay near 1.  Used by the model golden and its test on both sides."""
    import zlib
    with torch.no_grad():
        for name, p in model.named_parameters():
            g = torch.Generator().manual_seed((seed * 1000003 + zlib.crc32(name.encode())) % (2 ** 31))
            r = torch.randn(p.shape, generator=g)
            if p.dim() >= 2:
                fan_in = p[0].numel()
                v = r * (scale / fan_in ** 0.5) * 4
            elif "norm" in name and name.endswith("weight") or name.endswith("bn.weight") or ".weight_net.1.weight" in name:
                v = 1 + 0.1 * r
            else:
                v = 0.1 * r
            p.copy_(v.to(p.dtype))
