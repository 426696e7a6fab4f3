"""Generate GPTQ / AWQ golden vectors by running the UNMODIFIED reference class methods on CPU.

Run in the build container only (needs /root/reference):
    PYTHONPATH=/root/reference python tests/golden/make_algo_golden.py
Writes tests/golden/algo_golden.pt.gz. The reference objects are created with object.__new__ (no model / config needed)
and only the attributes the hot methods read are set. Two environment shims, none touching reference code:
`Tensor.cuda` is the identity (gptq.py:293 moves a scalar to CUDA for a no-op all-reduce) and a 1-rank gloo group is
initialised for `dist.all_reduce`.
"""
import gzip
import io
import os
import sys

import torch
import torch.distributed as dist
import torch.nn as nn

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'algo_golden.pt.gz')


def gptq_inputs(R, C, T, batches, seed):
    g = torch.Generator().manual_seed(seed)
    W = torch.randn(R, C, generator=g) * 0.02
    gain = torch.exp(0.5 * torch.randn(C, generator=g))
    xs = [(torch.randn(1, T, C, generator=g) * gain).to(torch.bfloat16) for _ in range(batches)]
    return W, xs


def awq_inputs(T, C, R, dtype, seed):
    g = torch.Generator().manual_seed(seed)
    gain = torch.exp(torch.randn(C, generator=g))
    x = (torch.randn(1, T, C, generator=g) * gain).to(dtype)
    W = (torch.randn(R, C, generator=g) * 0.02).to(dtype)
    return x, W


GPTQ_CASES = [
    dict(R=24, C=256, T=96, batches=2, seed=1, bit=4, sym=False, gs=128, bs=128),
    dict(R=16, C=256, T=64, batches=3, seed=2, bit=4, sym=True, gs=64, bs=128),
    dict(R=16, C=384, T=64, batches=1, seed=3, bit=3, sym=False, gs=128, bs=128),
]
AWQ_CASES = [
    dict(T=64, C=256, R=48, dtype='bfloat16', seed=11, bit=4, sym=False, gs=128, version='v2'),
    dict(T=48, C=256, R=32, dtype='float16', seed=12, bit=4, sym=False, gs=128, version='v1'),
]


def run_gptq(c):
    from llmc.compression.quantization.gptq import GPTQ
    from llmc.compression.quantization.quant import IntegerQuantizer
    W, xs = gptq_inputs(c['R'], c['C'], c['T'], c['batches'], c['seed'])
    layer = nn.Linear(c['C'], c['R'], bias=False)
    layer.weight.data = W.clone()
    g = object.__new__(GPTQ)
    g.dev = torch.device('cpu')
    g.wquantizer = IntegerQuantizer(c['bit'], c['sym'], 'per_group', group_size=c['gs'])
    g.blocksize, g.percdamp, g.chunk_num = c['bs'], 0.01, 1
    g.static_groups = g.actorder = g.owq = False
    g.layers_cache = {'l': {}}
    g.qparams, g.groups = {}, []
    g.layer_init(layer, 'l')
    for x in xs:
        g.add_batch(layer, 'l', x, None)
    H = g.layers_cache['l']['H'].clone()
    # collect_block_qparams (base_bq.py:338-365) for this layer
    _, s, z, qmax, qmin = g.wquantizer.get_tensor_qparams(layer.weight.data)
    layer.register_buffer('buf_scales', s)
    layer.register_buffer('buf_zeros', z)
    layer.register_buffer('buf_qmax', torch.tensor(qmax))
    layer.register_buffer('buf_qmin', torch.tensor(qmin))
    g.initialize_qparams_and_prepare_weights(layer, 'l')
    Wf, Hinv = g.process_hessian_and_weights(layer, 'l')
    Losses, tmp = torch.zeros_like(Wf), torch.zeros_like(Wf)
    g.weight_transform(Wf, Hinv, Losses, tmp)
    scales = torch.stack([q['scale'].reshape(-1) for q in g.groups])
    zeros = torch.stack([q['zero'].reshape(-1) for q in g.groups]) if not c['sym'] else None
    return dict(case=c, H=H, Hinv=Hinv.clone(), tmp=tmp, Losses=Losses, scales=scales, zeros=zeros, W_after=Wf)


def run_awq(c):
    from llmc.compression.quantization.awq import Awq
    from llmc.compression.quantization.quant import IntegerQuantizer
    dtype = getattr(torch, c['dtype'])
    x, W = awq_inputs(c['T'], c['C'], c['R'], dtype, c['seed'])
    a = object.__new__(Awq)
    a.wquantizer = IntegerQuantizer(c['bit'], c['sym'], 'per_group', group_size=c['gs'])
    a.trans_version = c['version']
    a._bs = x.shape[0]
    fc = nn.Linear(c['C'], c['R'], bias=False).to(dtype)
    fc.weight.data = W.clone()
    w_max = a.get_weight_scale({'fc': fc})
    org_out = a.get_original_out(x, fc, {})
    losses, all_scales = [], []
    for n in range(20):
        scales = a.get_scales(None, x, w_max, False, n / 20)
        fc.weight.data = W.clone()
        fc.weight = a.fake_quantize_weight(fc, scales, False, 'fc')
        x_tmp = a.scaling_input(x, scales, False)
        out = a.inspect_module_forward(x_tmp, fc, {})
        losses.append(a.calculate_loss(org_out, out))
        all_scales.append(scales.clone())
    best = min(range(20), key=lambda i: (losses[i], i))   # strict < in the loop: first minimum
    return dict(case=c, w_max=w_max, scales=torch.stack(all_scales), losses=torch.tensor(losses, dtype=torch.float64),
                best=best, org_out=org_out)


def main():
    sys.path.insert(0, '/root/reference')
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT='29533', RANK='0', WORLD_SIZE='1')
    dist.init_process_group('gloo', rank=0, world_size=1)
    torch.Tensor.cuda = lambda self, *a, **k: self
    with torch.no_grad():
        recs = {'gptq': [run_gptq(c) for c in GPTQ_CASES], 'awq': [run_awq(c) for c in AWQ_CASES]}
    buf = io.BytesIO()
    torch.save(recs, buf)
    with gzip.open(OUT, 'wb') as f:
        f.write(buf.getvalue())
    print('wrote', OUT, os.path.getsize(OUT), 'bytes')
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
