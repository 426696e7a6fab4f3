"""AutoClip v1 (auto_clip.py:44-215): per-(out-channel, group) search of the weight clipping range.

Interim implementation: the reference's broadcast product `(x * w).sum(-1)` — which materialises a
[256, 512, C] temporary per shrink step — becomes one batched GEMM per step over the groups ([ng, oc, gs] x [ng, gs, tok],
a plain library bmm, fp32 accumulation, output in the tensor dtype) and the fake-quant runs in the fused kernel. A
dedicated grouped tcgen05 kernel with a fused clamp + fake-quant + squared-error epilogue is the planned replacement.
Numerics differ from the reference in summation order only (its products are rounded to the tensor dtype one by one),
so near-tied (channel, group) pairs may pick a neighbouring shrink step; tests report the agreement rate."""
import torch

from . import dist as lcdist

_SKIP = ('q_', 'k_', 'query', 'key', 'Wqkv')   # auto_clip.py:56


@torch.no_grad()
def auto_clip_layer(w, x, wquantizer, clip_sym, n_grid=20, max_shrink=0.5, n_sample_token=None):
    """Returns (best_max_val, best_min_val), each [oc, n_group, 1] in w.dtype (auto_clip.py:84-191, clip_version v1)."""
    assert w.dim() == 2
    oc, C = w.shape
    gs = wquantizer.group_size if wquantizer.granularity == 'per_group' else C
    ng = C // gs
    W = w.reshape(oc, ng, gs)
    x = x.reshape(-1, C)
    if n_sample_token is None:
        n_sample_token = min(x.shape[0], 512)
    step = max(1, x.shape[0] // n_sample_token)
    x = x[0::step]
    xg = x.reshape(x.shape[0], ng, gs).permute(1, 2, 0).contiguous()          # [ng, gs, tok]
    org_out = torch.bmm(W.permute(1, 0, 2).contiguous(), xg)                  # [ng, oc, tok]
    org_max = W.abs().amax(dim=-1, keepdim=True) if clip_sym else W.amax(dim=-1, keepdim=True)
    org_min = W.amin(dim=-1, keepdim=True)
    best_max, best_min = org_max.clone(), org_min.clone()
    min_errs = torch.ones_like(org_max) * 1e9
    for i_s in range(int(max_shrink * n_grid)):
        max_val = org_max * (1 - i_s / n_grid)
        min_val = -max_val if clip_sym else org_min * (1 - i_s / n_grid)
        cur_w = torch.clamp(W, min_val, max_val)
        q_w = wquantizer.fake_quant_weight_dynamic(cur_w.reshape(oc, C)).reshape(oc, ng, gs)
        cur_out = torch.bmm(q_w.permute(1, 0, 2).contiguous(), xg)
        err = (cur_out - org_out).pow(2).mean(dim=2).t().unsqueeze(-1)        # [oc, ng, 1]
        better = err < min_errs
        min_errs[better] = err[better]
        best_max[better] = max_val[better]
        best_min[better] = min_val[better]
    return best_max, best_min


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
def auto_clip_block(named_linears, input_feat, wquantizer, clip_sym, n_sample_token=None):
    """AutoClipper.run (auto_clip.py:44-82): every linear except q/k-like ones; ranges averaged over the ranks."""
    rank, ws = lcdist.world()
    for n, m in named_linears.items():
        if any(s in n for s in _SKIP):
            continue
        feats = input_feat[n]
        x = feats[0] if len(feats) == 1 else torch.cat(feats)
        max_val, min_val = auto_clip_layer(m.weight.data, x, wquantizer, clip_sym, n_sample_token=n_sample_token)
        if ws > 1:
            import torch.distributed as dist
            dist.all_reduce(max_val, op=dist.ReduceOp.SUM)
            max_val /= ws
            dist.all_reduce(min_val, op=dist.ReduceOp.SUM)
            min_val /= ws
        apply_clip(m.weight, min_val, max_val, clip_sym)
