"""AWQ scale search on the GPU kernels: the compute of awq.py:48-278 with explicit tensors.

`awq_search_scale(x, weights, wquantizer)` covers subsets whose inspect module is the linear itself (o_proj, down_proj,
fc2, ...): 20 candidates through lcb_awq_candidates. For multi-linear inspect modules (attention, gated MLP) use
`awq_search_scale_module`, which keeps the reference's structure (scale -> fake-quant -> forward -> MSE) but with the
fused scale / fake-quant kernels and no CPU round trip of the state dict (awq.py:199, 244)."""
import torch

from .. import ops
from . import dist as lcdist

N_GRID = 20  # awq.py:198


def get_weight_scale(weights, wquantizer):
    """awq.py:48-72: mean over rows of |W| / per-group max |W|, averaged over the layers -> [C_in]."""
    total = None
    for w in weights:
        org_shape = w.shape
        r = wquantizer.reshape_tensor(w.clone())
        a = r.abs()
        ls = a.div_(a.amax(dim=1, keepdim=True)).view(org_shape)
        total = ls.mean(0) if total is None else total.add_(ls.mean(0))
    return total.div_(len(weights))


def get_act_scale(x):
    """awq.py:74-85 for one batch: mean |x| per input channel, in the activation dtype (like the reference).
    Accumulated in fp32 by lcb_col_absstat, then rounded to x.dtype (torch's mean also accumulates in fp32)."""
    C = x.shape[-1]
    acc = torch.zeros(C, dtype=torch.float32, device=x.device)
    ops.col_absstat(x, acc, mean=True)
    return (acc / float(x.numel() // C)).to(x.dtype)


def get_scales(x_mean, w_max, ratio, version='v2'):
    """awq.py:87-108."""
    if version == 'v1':
        scales = (x_mean.pow(ratio) / w_max.pow(1 - ratio)).clamp(min=1e-4).view(-1)
    else:
        scales = x_mean.pow(ratio).clamp(min=1e-4).view(-1)
    return scales / (scales.max() * scales.min()).sqrt()


@torch.no_grad()
def awq_search_scale(x, weight, wquantizer, version='v2', org_out=None, n_grid=N_GRID, x_mean=None):
    """Best per-input-channel scales for one linear whose output is the search objective.
    Returns (best_scales [C], best_index, losses float64 [n_grid]).
    Multi-GPU: the (candidate, row tile) units are dealt round-robin over the ranks on identical data — the squared error
    is a sum over output rows and the fake-quant units lie inside a row, so a row tile is an independent piece of a
    candidate's loss; ONE all-reduce(SUM) of the n_grid partial sums, then every rank takes the same argmin (first
    minimum: the strict `<` of the sequential search, awq.py:245) and computes that candidate's scales — one host sync."""
    if org_out is None:
        org_out = torch.nn.functional.linear(x, weight)
    x_mean = get_act_scale(x) if x_mean is None else x_mean
    w_max = get_weight_scale([weight], wquantizer) if version == 'v1' else None
    R = weight.shape[0]
    T = x.numel() // x.shape[-1]
    n_tiles, units = lcdist.shard_units(n_grid)
    if R % (n_tiles * 8):
        n_tiles, units = lcdist.shard_units(n_grid, 1)      # rows do not split evenly: candidate-only sharding
    sums = torch.zeros(n_grid, dtype=torch.float64, device=x.device)
    rt = R // n_tiles
    o2 = org_out.reshape(-1, R)
    mine = sorted({n for n, _ in units})          # only this rank's candidates are materialised
    cand_of = {n: get_scales(x_mean, w_max, n / n_grid, version) for n in mine}
    for tile in sorted({t for _, t in units}):
        idx = [n for n, t in units if t == tile]
        w_t = weight[tile * rt:(tile + 1) * rt]
        o_t = o2 if n_tiles == 1 else o2[:, tile * rt:(tile + 1) * rt].contiguous()
        part = ops.awq_candidates(w_t, x, o_t, torch.stack([cand_of[n] for n in idx]), wquantizer) * float(T * rt)
        sums.index_add_(0, torch.tensor(idx, device=x.device), part)                        # mean -> sum over the tile
    lcdist.reduce_sum(sums)
    losses = sums / float(T * R)
    keyed = torch.where(torch.isfinite(losses), losses, torch.full_like(losses, float('inf')))
    ar = torch.arange(n_grid, device=x.device)
    best_t = torch.where(keyed == keyed.min(), ar, torch.full_like(ar, n_grid)).min().clamp(max=n_grid - 1)   # FIRST minimum
    best = int(best_t.item())
    best_scales = cand_of[best] if best in cand_of else get_scales(x_mean, w_max, best / n_grid, version)
    if not bool(torch.isfinite(losses[best])):
        raise RuntimeError('awq_search_scale: every candidate loss is NaN / inf')
    return best_scales, best, losses.cpu()


class own_linears(object):
    """While active, every 16-bit nn.Linear inside `module` computes its F.linear on the tcgen05 contraction
    (ops.linear = lcb_linear) instead of the library GEMM: the inspect module of a multi-linear subset (the whole
    attention / MLP, awq.py:110-126) keeps its own forward — rotary embedding, softmax(QK^T)V (torch's SDPA: a library
    call, stated as such), activation — and only the projections change hands."""

    def __init__(self, module):
        self.mods = [m for m in module.modules()
                     if type(m) is torch.nn.Linear and m.weight.is_cuda and m.weight.dtype in (torch.bfloat16, torch.float16)
                     and m.in_features % 8 == 0 and m.out_features % 8 == 0]

    @staticmethod
    def _forward(m, inp):
        if inp.dtype != m.weight.dtype or not inp.is_cuda:
            return torch.nn.functional.linear(inp, m.weight, m.bias)
        return ops.linear(inp, m.weight, m.bias)

    def __enter__(self):
        import functools
        for m in self.mods:
            m.forward = functools.partial(self._forward, m)    # instance attribute: shadows nn.Linear.forward
        return self

    def __exit__(self, *exc):
        for m in self.mods:
            m.__dict__.pop('forward', None)
        return False


@torch.no_grad()
def awq_search_scale_module(x, layers, inspect_module, wquantizer, kwargs=None, version='v2', n_grid=N_GRID, own_gemm=True,
                            aquantizer=None):
    """awq.py:179-278 for any inspect module (one batch). `layers`: the nn.Linear modules that share the input x.
    Weights are restored from an on-device copy (the reference reloads a CPU state dict every candidate).
    `aquantizer` (w+a configs): the scaled input is fake-quantized before the forward (awq.py:166-177, 228-229)."""
    if own_gemm:
        with own_linears(inspect_module):
            return awq_search_scale_module(x, layers, inspect_module, wquantizer, kwargs, version, n_grid, False, aquantizer)
    kwargs = kwargs or {}

    def fwd(inp):
        out = inspect_module(inp, **kwargs)
        return out[0] if isinstance(out, tuple) else out

    org_out = fwd(x)
    x_mean = get_act_scale(x)
    w_max = get_weight_scale([m.weight.data for m in layers], wquantizer) if version == 'v1' else None
    saved = [m.weight.data.clone() for m in layers]
    mine = lcdist.shard_candidates(n_grid)
    losses = []
    for n in mine:
        scales = get_scales(x_mean, w_max, n / n_grid, version)
        for m, w0 in zip(layers, saved):
            m.weight.data = wquantizer.fake_quant_weight_dynamic(ops.scale_cols(w0, scales))
        x_tmp = ops.scale_cols(x, scales, div=True)
        if aquantizer is not None:
            x_tmp = aquantizer.fake_quant_act_dynamic(x_tmp)
        out = fwd(x_tmp)
        if out.is_cuda and out.dtype == org_out.dtype and out.dtype in (torch.bfloat16, torch.float16, torch.float32):
            losses.append(ops.sqdiff_mean(org_out, out))            # one pass, same roundings; stays on the device
        else:
            losses.append((org_out - out).float().pow(2).mean())   # stays on the device: one sync after the loop
        for m, w0 in zip(layers, saved):
            m.weight.data = w0
    losses = torch.stack(losses).tolist() if losses else []
    _, best = lcdist.argmin_across_ranks(losses, mine, x.device)
    return get_scales(x_mean, w_max, best / n_grid, version), best, losses


@torch.no_grad()
def awq_search_scale_general(inputs, layers, inspect_module, wquantizer, subset_kwargs, version='v2', aquantizer=None,
                             padding_mask=None, awq_bs=None, n_samples=None, n_grid=N_GRID, own_gemm=True):
    """Awq.search_scale_subset (awq.py:179-278) LITERALLY, for everything the single-batch paths above do not cover:
    several input batches (the running `loss_mean` is compared inside the batch loop, and `best_scales` is the running
    `scales_mean` at that moment: awq.py:240-248, reproduced as written), `awq_bs` chunking of a batch (:74-85, 110-145,
    166-177), padding masks (:233-235), w+a. Every rank searches on ITS data; the rank with the smallest error broadcasts
    its scales (:255-273). Compute: lcb_scale_cols, the fused fake-quant kernels, lcb_linear for the module's projections.
    Returns (best_scales, best_index)."""
    import torch.distributed as dist
    if own_gemm:
        with own_linears(inspect_module):
            return awq_search_scale_general(inputs, layers, inspect_module, wquantizer, subset_kwargs, version, aquantizer,
                                            padding_mask, awq_bs, n_samples, n_grid, own_gemm=False)
    bs = inputs[0].shape[0] if awq_bs is None else awq_bs

    def chunks(t):
        return [t[k * bs:(k + 1) * bs] for k in range(t.shape[0] // bs)]

    def act_scale(x):                                   # get_act_scale (:74-85)
        if x.shape[0] == bs:
            return get_act_scale(x)
        means = [get_act_scale(c) for c in chunks(x)]
        return sum(means) / len(means)

    def forward(x, kw):                                 # inspect_module_forward (:110-126)
        outs = []
        for c in ([x] if x.shape[0] == bs else chunks(x)):
            o = inspect_module(c, **kw)
            outs.append(o[0] if isinstance(o, tuple) else o)
        return outs[0] if len(outs) == 1 else torch.cat(outs, dim=0)

    def loss_of(org, out):                              # calculate_loss (:134-145)
        if out.shape[0] == bs:
            return (org - out).float().pow(2).mean().item()
        parts = [(a - b).float().pow(2).mean().item() for a, b in zip(chunks(org), chunks(out))]
        return sum(parts) / len(parts)

    w_max = get_weight_scale([m.weight.data for m in layers], wquantizer) if version == 'v1' else None
    saved = [m.weight.data.clone() for m in layers]
    best_error, best_scales, best_n = float('inf'), None, -1
    org_cache = {}
    for n in range(n_grid):
        loss_mean, scales_mean = 0, 0
        for i, x in enumerate(inputs):
            kw = subset_kwargs[i] if isinstance(subset_kwargs, list) else (subset_kwargs or {})
            if i not in org_cache:
                org_cache[i] = forward(x, kw)
            org_out = org_cache[i]
            scales = get_scales(act_scale(x), w_max, n / n_grid, version)
            for m, w0 in zip(layers, saved):
                m.weight.data = wquantizer.fake_quant_weight_dynamic(ops.scale_cols(w0, scales))
            x_tmp = ops.scale_cols(x, scales, div=True)
            if aquantizer is not None:                   # fake_quantize_input (:166-177)
                if bs == x_tmp.shape[0]:
                    x_tmp = aquantizer.fake_quant_act_dynamic(x_tmp)
                else:
                    x_tmp = torch.stack([aquantizer.fake_quant_act_dynamic(x_tmp[k]) for k in range(x_tmp.shape[0])])
            out = forward(x_tmp, kw)
            if padding_mask and org_out.shape[1] == padding_mask[i].shape[-1]:
                mask = padding_mask[i].unsqueeze(dim=-1).to(org_out.device)
                org_out, out = org_out * mask, out * mask
            loss = loss_of(org_out, out)
            ns = x.shape[0] if len(inputs) == 1 else n_samples
            loss_mean += x.shape[0] * 1.0 / ns * loss
            scales_mean += x.shape[0] * 1.0 / ns * scales
            for m, w0 in zip(layers, saved):
                m.weight.data = w0
            if loss_mean < best_error:
                best_error, best_scales, best_n = loss_mean, scales_mean, n
    rank, ws = lcdist.world()
    if ws > 1:                                           # awq.py:255-273
        dev = inputs[0].device
        err = torch.tensor([best_error], device=dev)
        dist.all_reduce(err, op=dist.ReduceOp.MIN)
        who = torch.tensor([rank if abs(best_error - err.item()) < 1e-5 else -1], device=dev)
        dist.all_reduce(who, op=dist.ReduceOp.MAX)
        src = int(who.item())
        if rank != src:
            best_scales = torch.zeros_like(best_scales)
        dist.broadcast(best_scales, src=src)
    if best_scales is None:
        raise RuntimeError('awq_search_scale_general: every candidate loss is NaN / inf')
    return best_scales, best_n
