"""AutoClip (auto_clip.py:44-282): per-(out-channel, group) search of the weight clipping range.

clip_version 'v1', w_only (the shipped AWQ recipe): the whole search is one C call, lcb_autoclip_search (csrc/autoclip.cu):
per shrink step one fused clamp + fake-quant pass over W and one pass that forms `(x * w).sum(-1)` with the reference's
roundings (products rounded to the tensor dtype, fp32 sum, output / difference / square / mean rounded to the tensor dtype)
and reduces the squared error over the sampled tokens in registers; the reference's [256, tok, C] broadcast product is
never materialised.
clip_version 'v2' and w+a configs (configs/quantization/combination/awq_comb_omni/*/step_1_awq.yml): the search taken apart
(lcb_autoclip_init / lcb_autoclip_step). The per-step weights are composed here exactly as auto_clip.py:258-274 composes
them — logit / sigmoid bounds on the [units, 1] ranges (a handful of elementwise ops in the tensor dtype), lcb_minmax,
lcb_qparams_from_range, the static fake-quant kernel — and the candidate outputs use the fake-quantized inputs
(auto_clip.py:164, 276-281)."""
import torch

from .. import _cabi, ops
from . import dist as lcdist

_SKIP = ('q_', 'k_', 'query', 'key', 'Wqkv')   # auto_clip.py:56


@torch.no_grad()
def auto_clip_layer(w, x, wquantizer, clip_sym, n_grid=20, max_shrink=0.5, n_sample_token=None):
    """Returns (best_max_val, best_min_val), each [oc, n_group, 1] in w.dtype (auto_clip.py:84-191, clip_version v1)."""
    assert w.dim() == 2
    C = w.shape[1]
    x = x.reshape(-1, C)
    if n_sample_token is None:
        n_sample_token = min(x.shape[0], 512)
    step = max(1, x.shape[0] // n_sample_token)
    x = x[0::step]
    return ops.autoclip_search(w, x.to(w.dtype), wquantizer, clip_sym, n_grid, max_shrink)


def _logit(x):
    return torch.log(x / (1 - x))   # auto_clip.py:41


def _sample_tokens(x, C, n_sample_token, padding_mask=None):
    """auto_clip.py:135-147: flatten, drop padded tokens, keep every step_size-th token."""
    x = x.reshape(-1, C)
    if padding_mask is not None and padding_mask.numel() == x.shape[0]:
        x = x[padding_mask.flatten().bool().to(x.device)]
    if n_sample_token is None:
        n_sample_token = min(x.shape[0], 512)
    step = max(1, x.shape[0] // n_sample_token)
    return x[0::step]


@torch.no_grad()
def auto_clip_layer_general(w, x, wquantizer, clip_sym, clip_version='v1', aquantizer=None, n_grid=20, max_shrink=0.5,
                            n_sample_token=None, padding_mask=None):
    """AutoClipper.auto_clip_layer (auto_clip.py:84-191) for every variant: clip_version 'v1' | 'v2', w_only
    (aquantizer None) or w+a. Returns (best_max_val, best_min_val), each [oc, n_group, 1] in w.dtype.

    v2 with per_group weights: the reference's own call raises (get_learnable_range is given the 4-D batch view and
    fake_quant_weight_static reshapes it to [oc * n_group, group]: the [oc, 1, n_group, 1] scales do not broadcast,
    quant.py:205-219, 785-817); here the ranges are per (channel, group), which is what the per_channel case — the only
    one the reference can run (awq_comb_omni) — generalises to."""
    assert w.dim() == 2 and clip_version in ('v1', 'v2')
    R, C = w.shape
    xs = _sample_tokens(x, C, n_sample_token, padding_mask).to(w.dtype)
    x2, tokens = ops.autoclip_pad_tokens(xs)
    st = ops.AutoClipState(w, wquantizer, clip_sym, x2.shape[0], tokens)
    st.original(x2)
    if aquantizer is not None:      # auto_clip.py:276-281 on the [1, tok, n_group, group] view: per_token = per (token, group)
        gsz = C // st.ng
        xq = aquantizer.fake_quant_act_dynamic(x2.reshape(x2.shape[0], st.ng, gsz)).reshape(x2.shape[0], C).contiguous()
    else:
        xq = x2
    W = st.W
    if clip_version == 'v2':
        org_min, org_max = wquantizer.get_minmax_range(W)                     # [units, 1], tensor dtype
        if clip_sym:
            org_max = torch.max(org_max.abs(), org_min.abs())                 # w.abs().amax (auto_clip.py:117)
    for i_s in range(int(max_shrink * n_grid)):
        ratio = 1 - i_s / n_grid
        if clip_version == 'v1':                                              # clamp + dynamic fake-quant, one kernel
            Wq, _, _ = wquantizer._dynamic(W, _cabi.OUT_FAKE, W.dtype, False, clip=-ratio if clip_sym else ratio)
        else:
            max_val = org_max * ratio
            min_val = -max_val if clip_sym else org_min * ratio
            rng = wquantizer.get_learnable_range(W, _logit(min_val / org_min), _logit(max_val / org_max))
            scales, zeros, qmax, qmin = wquantizer.get_qparams(rng, W.device)
            Wq = wquantizer.fake_quant_weight_static(W, {'scales': scales, 'zeros': zeros, 'qmax': qmax, 'qmin': qmin})
        st.candidate(Wq, xq, ratio)
    return st.best_max, st.best_min


@torch.no_grad()
def get_clip_factor(weight, min_val, max_val, wquantizer, clip_sym):
    """auto_clip.py:233-256 -> (up_factor, low_factor | None), shaped like the quantizer's [units, 1] ranges."""
    org_min, org_max = wquantizer.get_minmax_range(weight)
    shape = org_max.shape
    if clip_sym:
        abs_max = torch.max(org_max.abs(), org_min.abs()).clamp(min=1e-5).reshape(*max_val.shape[:2], -1)
        return _logit(max_val / abs_max).reshape(shape), None
    up = _logit(max_val / org_max.reshape(*max_val.shape[:2], -1)).reshape(shape)
    low = _logit(min_val / org_min.reshape(*min_val.shape[:2], -1)).reshape(shape)
    return up, low


@torch.no_grad()
def apply_clip_v2(layer, min_val, max_val, wquantizer, clip_sym):
    """auto_clip.py:213-229: the weight stays; the learnable bounds go to module buffers that w_qdq reads at deploy time
    (base_blockwise_quantization.py:46-65). Returns (up_factor, low_factor)."""
    up, low = get_clip_factor(layer.weight.data, min_val, max_val, wquantizer, clip_sym)
    layer.register_buffer('buf_upbound_factor', up)
    layer.register_buffer('buf_lowbound_factor', low)
    return up, low


@torch.no_grad()
def apply_clip(weight, min_val, max_val, clip_sym):
    """auto_clip.py:217-236 (v1): clamp the weight group-wise, in place on `weight.data`."""
    oc, C = weight.shape
    ng = max_val.shape[1]
    if clip_sym:
        min_val = -max_val
    w = weight.data.reshape(oc, ng, -1)
    weight.data = torch.clamp(w, min_val, max_val).reshape(oc, C)


@torch.no_grad()
def auto_clip_layer_sharded(w, x, wquantizer, clip_sym, n_sample_token=None, search=None):
    """auto_clip_layer with the OUTPUT CHANNELS dealt over the ranks (every (channel, group) is searched
    independently, auto_clip.py:108-114 already batches channels): rank r searches rows [r R / world, (r + 1) R / world) on
    identical data, one all_gather of the two [R / world, n_group, 1] results. Falls back to the full search on every rank
    when the rows do not split into multiples of 16. `search` is injectable for the CPU tests of the host logic."""
    import torch.distributed as dist
    search = search or auto_clip_layer
    rank, ws = lcdist.world()
    R = w.shape[0]
    if ws == 1 or R % (ws * 16):
        return search(w, x, wquantizer, clip_sym, n_sample_token=n_sample_token)
    rt = R // ws
    mx, mn = search(w[rank * rt:(rank + 1) * rt].contiguous(), x, wquantizer, clip_sym, n_sample_token=n_sample_token)
    both = torch.stack([mx, mn]).contiguous()
    parts = [torch.empty_like(both) for _ in range(ws)]
    dist.all_gather(parts, both)
    return torch.cat([p[0] for p in parts], 0), torch.cat([p[1] for p in parts], 0)


@torch.no_grad()
def auto_clip_block(named_linears, input_feat, wquantizer, clip_sym, n_sample_token=None, record=None, shard_rows=True,
                    clip_version='v1', aquantizer=None, padding_mask=None, factors=None):
    """AutoClipper.run (auto_clip.py:44-82): every linear except q/k-like ones. world > 1: the calibration data is the
    same on every rank (like the AWQ candidate search) and the output channels are sharded (`shard_rows`); with
    shard_rows=False every rank searches on ITS data and the ranges are averaged, as the reference does (:72-76).
    clip_version 'v2': the bounds are registered as module buffers (q/k-like linears get None buffers, :56-60) and, when
    `factors` is a dict, recorded as factors[name] = (up_factor, low_factor). aquantizer: w+a configs (None = w_only).
    padding_mask: the pipeline's list of masks; the first is applied when its size matches the tokens (:136-138)."""
    rank, ws = lcdist.world()
    general = clip_version != 'v1' or aquantizer is not None or bool(padding_mask)
    mask = padding_mask[0] if padding_mask else None

    def search(w, x, wq, sym, n_sample_token=None):
        if not general:
            return auto_clip_layer(w, x, wq, sym, n_sample_token=n_sample_token)
        return auto_clip_layer_general(w, x, wq, sym, clip_version, aquantizer, n_sample_token=n_sample_token,
                                       padding_mask=mask)

    for n, m in named_linears.items():
        if any(s in n for s in _SKIP):
            if clip_version == 'v2':
                m.register_buffer('buf_upbound_factor', None)
                m.register_buffer('buf_lowbound_factor', None)
            continue
        feats = input_feat[n]
        x = feats[0] if len(feats) == 1 else torch.cat(feats)
        if ws > 1 and shard_rows:
            max_val, min_val = auto_clip_layer_sharded(m.weight.data, x, wquantizer, clip_sym, n_sample_token, search=search)
        else:
            max_val, min_val = search(m.weight.data, x, wquantizer, clip_sym, n_sample_token=n_sample_token)
            if ws > 1:
                import torch.distributed as dist
                dist.all_reduce(max_val, op=dist.ReduceOp.SUM)
                max_val /= ws
                dist.all_reduce(min_val, op=dist.ReduceOp.SUM)
                min_val /= ws
        if record is not None:
            record[n] = (max_val, min_val)
        if clip_version == 'v1':
            apply_clip(m.weight, min_val, max_val, clip_sym)
        else:
            f = apply_clip_v2(m, min_val, max_val, wquantizer, clip_sym)
            if factors is not None:
                factors[n] = f
