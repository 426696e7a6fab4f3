"""CPU restatement of the reference's AWQ scale search — TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Follows /root/reference/llmc/compression/quantization/awq.py on CPU torch tensors (the reference is eager torch; bf16 /
fp16 CPU ops round like the CUDA ones, except matmul summation order and pow):
  get_weight_scale   awq.py:48-72        get_act_scale  awq.py:74-85       get_scales  awq.py:87-108
  search loop        awq.py:202-248 for a subset whose inspect module is the linear itself, one batch (bs = -1)
Parity pinning: tests/golden/make_awq_golden.py runs the UNMODIFIED reference methods (object.__new__(Awq)) on seeded
inputs; tests/test_awq_oracle.py compares.
"""
import torch
import torch.nn.functional as F

from . import quant_oracle as qo


def get_weight_scale(weights, group_size):
    total = None
    for w in weights:
        org_shape = w.shape
        r = w.clone().reshape(-1, group_size) if group_size > 0 else w.clone()
        a = r.abs()
        mx = a.amax(dim=1, keepdim=True)
        ls = a.div_(mx).view(org_shape)
        total = ls.mean(0) if total is None else total.add_(ls.mean(0))
    return total.div_(len(weights))


def get_act_scale(x):
    return x.abs().view(-1, x.shape[-1]).mean(0)


def get_scales(x, w_max, ratio, version='v2'):
    x_mean = get_act_scale(x)
    if version == 'v1':
        scales = (x_mean.pow(ratio) / w_max.pow(1 - ratio)).clamp(min=1e-4).view(-1)
    else:
        scales = x_mean.pow(ratio).clamp(min=1e-4).view(-1)
    return scales / (scales.max() * scales.min()).sqrt()


def fake_quant_weight(w, group_size, spec):
    R, C = w.shape
    out, _, _ = qo.fakequant(w.contiguous(), R, C, 1, group_size if group_size > 0 else C, **spec)
    return out


def search_losses(W, x, cand_scales, group_size, spec):
    """loss per candidate (python floats) + index of the best (strict <, awq.py:245)."""
    org_out = F.linear(x, W)
    losses = []
    for s in cand_scales:
        wq = fake_quant_weight(W * s.view(1, -1), group_size, spec)
        x_tmp = x / s.view(1, -1)
        out = F.linear(x_tmp, wq)
        losses.append((org_out - out).float().pow(2).mean().item())
    best, best_i = float('inf'), -1
    for i, l in enumerate(losses):
        if l < best:
            best, best_i = l, i
    return losses, best_i, org_out
