"""CPU restatement of the reference's AutoClip v1 search — TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Follows /root/reference/llmc/compression/quantization/auto_clip.py on CPU torch tensors:
  auto_clip_layer        auto_clip.py:84-191   (one input tensor, w_only, clip_version 'v1', oc batching kept)
  fake_quantize_weight   auto_clip.py:258-262  (clamp, then IntegerQuantizer.fake_quant_weight_dynamic via oracle/lc_oracle.c)
Parity pinning: tests/golden/make_autoclip_golden.py runs the UNMODIFIED AutoClipper.auto_clip_layer on seeded inputs;
tests/test_autoclip_oracle.py compares (bit-exact: every op here is the reference's own torch op, in its order).
"""
import torch

from . import quant_oracle as qo


def fake_quant_clamped(w4, min_val, max_val, spec, group_size):
    """fake_quantize_weight v1 (auto_clip.py:258-262) on w4 = [oc, 1, ng, gs]."""
    cur = torch.clamp(w4, min_val, max_val)
    oc = cur.shape[0]
    flat = cur.reshape(oc, -1).contiguous()
    out, _, _ = qo.fakequant(flat, flat.shape[0], flat.shape[1], 1, group_size, **spec)
    return out.reshape(w4.shape)


def auto_clip_layer(w, x, spec, group_size, clip_sym, n_grid=20, max_shrink=0.5, n_sample_token=512):
    """Returns (best_max_val, best_min_val, errs): [oc, ng, 1] each in w.dtype, errs [steps, oc, ng] (w.dtype)."""
    gs = group_size if group_size > 0 else w.shape[1]
    w = w.reshape(w.shape[0], 1, -1, gs)
    oc_batch = 256 if w.shape[0] % 256 == 0 else 64
    assert w.shape[0] % oc_batch == 0
    x = x.view(-1, x.shape[-1])
    x = x.reshape(1, x.shape[0], -1, gs)
    if n_sample_token is None:
        n_sample_token = min(x.shape[1], 512)
    step = max(1, x.shape[1] // n_sample_token)
    x = x[:, 0::step]
    w_all, bmax_all, bmin_all, errs_all = w, [], [], []
    for i_b in range(w_all.shape[0] // oc_batch):
        w = w_all[i_b * oc_batch:(i_b + 1) * oc_batch]
        org_max = w.abs().amax(dim=-1, keepdim=True) if clip_sym else w.amax(dim=-1, keepdim=True)
        org_min = w.amin(dim=-1, keepdim=True)
        best_max, best_min = org_max.clone(), org_min.clone()
        min_errs = torch.ones_like(org_max) * 1e9
        org_out = (x * w).sum(dim=-1)
        errs = []
        for i_s in range(int(max_shrink * n_grid)):
            max_val = org_max * (1 - i_s / n_grid)
            min_val = -max_val if clip_sym else org_min * (1 - i_s / n_grid)
            q_w = fake_quant_clamped(w, min_val, max_val, spec, gs)
            cur_out = (x * q_w).sum(dim=-1)
            err = (cur_out - org_out).pow(2).mean(dim=1).view(min_errs.shape)
            err_mean = 0 + err
            err_mean /= 1
            errs.append(err_mean.clone())
            idx = err_mean < min_errs
            min_errs[idx] = err_mean[idx]
            best_max[idx] = max_val[idx]
            best_min[idx] = min_val[idx]
        bmax_all.append(best_max)
        bmin_all.append(best_min)
        errs_all.append(torch.stack(errs))
    return (torch.cat(bmax_all, 0).squeeze(1), torch.cat(bmin_all, 0).squeeze(1),
            torch.cat(errs_all, 1).squeeze(2).squeeze(-1))
