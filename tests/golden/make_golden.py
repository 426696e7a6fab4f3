"""Generate golden vectors by running the UNMODIFIED reference quantizers on CPU.

Run in the build container only (needs /root/reference):
    PYTHONPATH=/root/reference python tests/golden/make_golden.py
Writes tests/golden/quant_golden.pt.gz: the case list + reference outputs (inputs are regenerated from the
seed by `make_input`, which the tests import; int32 codes are stored as int16 to keep the fixture small).

FP8: the reference delegates rounding to `qtorch.float_quantize`, which is not installable offline. For
those cases only, `float_quantize` is substituted by `x.to(torch.float8_*).float()` — the cast the
reference itself applies right after (quant.py:1183, 1211) and the rounding of its Triton path
(kernel.py:8-83). These cases are flagged `pinned: 'torch-fp8-cast'`; everything else is the reference
verbatim.
"""
import os
import sys

import torch

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'quant_golden.pt.gz')


def make_input(shape, dtype, seed, kind='weight'):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(shape, generator=g)
    if kind == 'weight':
        x = x * 0.02
        flat = x.view(-1)
        idx = torch.randperm(flat.numel(), generator=g)[: max(1, flat.numel() // 1000)]
        flat[idx] *= 20.0
    else:
        gain = torch.exp(torch.randn(shape[-1], generator=g))
        x = x * gain
    x = x.to(dtype)
    if x.dim() == 2 and x.shape[0] >= 4 and x.shape[1] >= 256:
        x[1, :128] = 0           # an all-zero group
        x[2, 128:256] = 0.03125  # a constant group
        x[3, 0] = 0.5            # values that land on .5 ties
    return x


def fp8_cast_quantize(x, e, m, rounding='nearest'):
    dt = torch.float8_e4m3fn if e == 4 else torch.float8_e5m2
    return x.to(dt).float()


def run_case(c):
    sys.path.insert(0, '/root/reference')
    from llmc.compression.quantization import quant as refq
    dtype = getattr(torch, c['dtype'])
    x = make_input(c['shape'], dtype, c['seed'], c.get('input', 'weight'))
    kw = dict(c.get('kwargs', {}))
    if c['cls'] == 'int':
        q = refq.IntegerQuantizer(c['bit'], c['sym'], c['gran'], **kw)
    else:
        refq.float_quantize = fp8_cast_quantize
        q = refq.FloatQuantizer(c['bit'], True, c['gran'], use_qtorch=True, **kw)
    rec = dict(case=c, x_sum=float(x.double().sum()))
    api = c.get('api', 'weight')
    with torch.no_grad():
        if api == 'act':
            rec['fake'] = q.fake_quant_act_dynamic(x.clone())
        else:
            rec['fake'] = q.fake_quant_weight_dynamic(x.clone())
        if x.dim() == 2:
            codes, scales, zeros = q.real_quant_weight_dynamic(x.clone())
            rec['codes_dtype'] = str(codes.dtype)
            if codes.dtype in (torch.float8_e4m3fn, torch.float8_e5m2):
                rec['codes'] = codes.view(torch.uint8)
            else:
                rec['codes'] = codes.to(torch.int16)
            rec['scales'], rec['zeros'] = scales, zeros
        _, s, z, qmax, qmin = q.get_tensor_qparams(x.clone())
        rec['qp_scales'], rec['qp_zeros'] = s, z
        rec['qmax'], rec['qmin'] = qmax, qmin
        if c.get('static'):
            # static call with fp32 qparams on the low-precision tensor (GPTQ deploy path, gptq.py:425-452)
            _, s32, z32, qmax, qmin = q.get_tensor_qparams(x.float())
            args = {'scales': s32, 'zeros': z32, 'qmax': qmax, 'qmin': qmin}
            rec['static_scales'], rec['static_zeros'] = s32, z32
            rec['static_fake'] = q.fake_quant_weight_static(x.clone(), args)
            args_t = {'scales': s, 'zeros': z, 'qmax': qmax, 'qmin': qmin}
            rec['static_fake_T'] = q.fake_quant_weight_static(x.clone(), args_t)
    return rec


def cases():
    cs = []
    seed = 100
    for dtype in ('bfloat16', 'float16', 'float32'):
        for bit in (4, 8, 3):
            for sym in (True, False):
                seed += 1
                cs.append(dict(cls='int', bit=bit, sym=sym, gran='per_group', kwargs={'group_size': 128},
                               shape=(16, 512), dtype=dtype, seed=seed, static=(bit == 4)))
        for gs in (32, 64, 256):
            seed += 1
            cs.append(dict(cls='int', bit=4, sym=False, gran='per_group', kwargs={'group_size': gs},
                           shape=(8, 512), dtype=dtype, seed=seed))
        for sym in (True, False):
            seed += 1
            cs.append(dict(cls='int', bit=8, sym=sym, gran='per_channel', shape=(24, 768), dtype=dtype, seed=seed,
                           static=True))
            seed += 1
            cs.append(dict(cls='int', bit=8, sym=sym, gran='per_token', shape=(2, 12, 1024), dtype=dtype, seed=seed,
                           api='act', input='act'))
            seed += 1
            cs.append(dict(cls='int', bit=8, sym=sym, gran='per_tensor', shape=(64, 320), dtype=dtype, seed=seed))
        seed += 1
        cs.append(dict(cls='int', bit=4, sym=False, gran='per_group', kwargs={'group_size': 128, 'round_zp': False},
                       shape=(8, 256), dtype=dtype, seed=seed))
        seed += 1
        cs.append(dict(cls='int', bit=8, sym=True, gran='per_block', kwargs={'block_size': 128}, shape=(256, 384),
                       dtype=dtype, seed=seed))
        seed += 1
        # the reference's restore_tensor (quant.py:647-651) only supports raggedness in ONE dimension
        cs.append(dict(cls='int', bit=8, sym=True, gran='per_block', kwargs={'block_size': 128}, shape=(200, 384),
                       dtype=dtype, seed=seed))
        seed += 1
        cs.append(dict(cls='int', bit=8, sym=True, gran='per_block', kwargs={'block_size': 128}, shape=(256, 264),
                       dtype=dtype, seed=seed))
        seed += 1
        cs.append(dict(cls='int', bit=4, sym=False, gran='per_channel', shape=(16, 4104), dtype=dtype, seed=seed))
        seed += 1
        cs.append(dict(cls='int', bit=4, sym=True, gran='per_group', kwargs={'group_size': 128}, shape=(5, 300),
                       dtype=dtype, seed=seed, expect='ValueError'))
        # long rows (row kernel, register cache and streaming variants)
        seed += 1
        cs.append(dict(cls='int', bit=8, sym=True, gran='per_token', shape=(1, 6, 14336), dtype=dtype, seed=seed,
                       api='act', input='act'))
        # unaligned row length (generic kernel)
        seed += 1
        cs.append(dict(cls='int', bit=8, sym=False, gran='per_channel', shape=(7, 333), dtype=dtype, seed=seed))
        for bit in ('e4m3', 'e5m2'):
            seed += 1
            cs.append(dict(cls='fp', bit=bit, gran='per_block', kwargs={'block_size': 128}, shape=(256, 384),
                           dtype=dtype, seed=seed, pinned='torch-fp8-cast'))
            seed += 1
            cs.append(dict(cls='fp', bit=bit, gran='per_block', kwargs={'block_size': 128}, shape=(200, 384),
                           dtype=dtype, seed=seed, pinned='torch-fp8-cast'))
            seed += 1
            cs.append(dict(cls='fp', bit=bit, gran='per_group', kwargs={'group_size': 128}, shape=(16, 512),
                           dtype=dtype, seed=seed, pinned='torch-fp8-cast'))
            seed += 1
            cs.append(dict(cls='fp', bit=bit, gran='per_channel', shape=(16, 512), dtype=dtype, seed=seed,
                           pinned='torch-fp8-cast'))
    return cs


def main():
    recs = []
    for c in cases():
        if c.get('expect') == 'ValueError':
            try:
                run_case(c)
            except ValueError:
                recs.append(dict(case=c, raises='ValueError'))
                continue
            raise AssertionError('reference did not raise')
        recs.append(run_case(c))
    import gzip
    import io
    buf = io.BytesIO()
    torch.save(recs, buf)
    with gzip.open(OUT, 'wb', compresslevel=9) as f:
        f.write(buf.getvalue())
    print(f'{len(recs)} cases -> {OUT} ({os.path.getsize(OUT) / 1e6:.2f} MB), torch {torch.__version__}')


if __name__ == '__main__':
    main()
