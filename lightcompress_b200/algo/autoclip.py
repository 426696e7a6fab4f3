"""AutoClip v1 (auto_clip.py:44-215): per-(out-channel, group) search of the weight clipping range.

The search runs in lcb_autoclip_search (csrc/autoclip.cu): per shrink step one fused clamp + fake-quant pass over W and
one grouped tcgen05 product whose epilogue reduces the squared output error over the sampled tokens; the reference's
[256, tok, C] broadcast product is never materialised. Roundings follow the reference op by op (output, difference,
square and mean rounded to the tensor dtype); the products themselves are exact and summed in fp32 where the reference
rounds each one first, so near-tied shrink steps can resolve differently (tests/test_autoclip_gpu.py reports it)."""
import torch

from .. import ops
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
def auto_clip_block(named_linears, input_feat, wquantizer, clip_sym, n_sample_token=None, record=None, shard_rows=True):
    """AutoClipper.run (auto_clip.py:44-82): every linear except q/k-like ones. world > 1: the calibration data is the
    same on every rank (like the AWQ candidate search) and the output channels are sharded (`shard_rows`); with
    shard_rows=False every rank searches on ITS data and the ranges are averaged, as the reference does (:72-76)."""
    rank, ws = lcdist.world()
    for n, m in named_linears.items():
        if any(s in n for s in _SKIP):
            continue
        feats = input_feat[n]
        x = feats[0] if len(feats) == 1 else torch.cat(feats)
        if ws > 1 and shard_rows:
            max_val, min_val = auto_clip_layer_sharded(m.weight.data, x, wquantizer, clip_sym, n_sample_token)
        else:
            max_val, min_val = auto_clip_layer(m.weight.data, x, wquantizer, clip_sym, n_sample_token=n_sample_token)
            if ws > 1:
                import torch.distributed as dist
                dist.all_reduce(max_val, op=dist.ReduceOp.SUM)
                max_val /= ws
                dist.all_reduce(min_val, op=dist.ReduceOp.SUM)
                min_val /= ws
        if record is not None:
            record[n] = (max_val, min_val)
        apply_clip(m.weight, min_val, max_val, clip_sym)
