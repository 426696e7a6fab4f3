"""Generate AutoClip golden vectors by running the UNMODIFIED reference AutoClipper.auto_clip_layer on CPU.

Run in the build container only (needs /root/reference):
    PYTHONPATH=/root/reference python tests/golden/make_autoclip_golden.py
Writes tests/golden/autoclip_golden.pt.gz: best_max / best_min of auto_clip.py:84-191 (clip_version v1, w_only) and, to
pin the clamped fake-quant on its own, fake_quantize_weight (auto_clip.py:258-262) at two shrink ratios.
"""
import gzip
import io
import os
import sys

import torch

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'autoclip_golden.pt.gz')

CASES = [
    dict(oc=64, C=256, tok=48, dtype='bfloat16', seed=21, bit=4, sym=False, gs=128, clip_sym=False),
    dict(oc=64, C=384, tok=40, dtype='float16', seed=22, bit=4, sym=False, gs=128, clip_sym=False),
    dict(oc=64, C=256, tok=32, dtype='bfloat16', seed=23, bit=4, sym=True, gs=64, clip_sym=True),
    dict(oc=256, C=128, tok=64, dtype='bfloat16', seed=24, bit=3, sym=False, gs=128, clip_sym=False),
]


def inputs(c):
    g = torch.Generator().manual_seed(c['seed'])
    dtype = getattr(torch, c['dtype'])
    W = torch.randn(c['oc'], c['C'], generator=g) * 0.02
    flat = W.view(-1)
    idx = torch.randperm(flat.numel(), generator=g)[: flat.numel() // 200]
    flat[idx] *= 6.0                                    # outliers worth clipping
    W[1, :c['gs']] = W[1, :c['gs']].abs() + 0.01          # an all-positive group (clamp edge case)
    gain = torch.exp(0.5 * torch.randn(c['C'], generator=g))
    x = (torch.randn(1, c['tok'], c['C'], generator=g) * gain).to(dtype)
    return W.to(dtype), x


def run(c):
    from llmc.compression.quantization.auto_clip import AutoClipper
    from llmc.compression.quantization.quant import IntegerQuantizer
    W, x = inputs(c)
    wq = IntegerQuantizer(c['bit'], c['sym'], 'per_group', group_size=c['gs'])
    ac = AutoClipper(True, wq, None, 'v1', c['clip_sym'], False, None)
    bmax, bmin = ac.auto_clip_layer(0, 'fc', W.clone(), [x.clone()], n_sample_token=c['tok'])
    w4 = W.reshape(c['oc'], 1, -1, c['gs'])
    org_max = w4.abs().amax(dim=-1, keepdim=True) if c['clip_sym'] else w4.amax(dim=-1, keepdim=True)
    org_min = w4.amin(dim=-1, keepdim=True)
    qws = {}
    for i_s in (3, 9):
        mx = org_max * (1 - i_s / 20)
        mn = -mx if c['clip_sym'] else org_min * (1 - i_s / 20)
        qws[i_s] = ac.fake_quantize_weight(w4, mn, mx, org_min, org_max).reshape(c['oc'], c['C'])
    return dict(case=c, best_max=bmax, best_min=bmin, q_w=qws)


def main():
    sys.path.insert(0, '/root/reference')
    with torch.no_grad():
        recs = [run(c) for c in CASES]
    buf = io.BytesIO()
    torch.save(recs, buf)
    with gzip.open(OUT, 'wb') as f:
        f.write(buf.getvalue())
    print('wrote', OUT, os.path.getsize(OUT), 'bytes')


if __name__ == '__main__':
    main()
