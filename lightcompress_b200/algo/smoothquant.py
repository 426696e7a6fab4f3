"""SmoothQuant scales (smoothquant.py:27-59) on lcb_col_absstat."""
import torch

from .. import ops
from . import dist as lcdist


class ActMax(object):
    """Running per-channel max of |x| over calibration batches (smoothquant.py:38-51), fp32, token-shardable."""

    def __init__(self, channels, device):
        self.x_max = torch.zeros(channels, dtype=torch.float32, device=device)

    @torch.no_grad()
    def update(self, x):
        ops.col_absstat(x, self.x_max, mean=False)

    def result(self):
        return lcdist.allreduce_max(self.x_max)


@torch.no_grad()
def get_weight_scale(weights):
    """smoothquant.py:27-36: max over the layers of the column abs-max of W, clamp 1e-5 — in the weight dtype."""
    C = weights[0].shape[1]
    w_max = torch.zeros(C, dtype=torch.float32, device=weights[0].device)
    for w in weights:
        ops.col_absstat(w, w_max, mean=False)
    return w_max.to(weights[0].dtype).clamp(min=1e-5)   # a max of T-typed values is exact in T


@torch.no_grad()
def smoothquant_scale(x_max, weights, alpha=0.5):
    """smoothquant.py:54-59: scale = (x_max^alpha / w_max^(1-alpha)).clamp(1e-5), computed in the weight dtype like the
    reference (x_max is cast to w_max's dtype first)."""
    w_max = get_weight_scale(weights)
    x_max = x_max.to(dtype=w_max.dtype, device=w_max.device)
    return (x_max.pow(alpha) / w_max.pow(1 - alpha)).clamp(min=1e-5)


@torch.no_grad()
def apply_scale_ln_fcs(ln_weight, fcs, scales, ln_bias=None):
    """scale_ln_fcs (base_bq.py:750-778): ln.weight /= s (+bias), fc.weight *= s[None, :] — in place."""
    s = scales.to(ln_weight.dtype)
    ln_weight.div_(s)
    if ln_bias is not None:
        ln_bias.div_(s)
    for w in fcs:
        ops.scale_cols(w, s.to(w.dtype), out=w)
