"""Golden vectors of the reference's Triton module (llmc/compression/quantization/kernel.py: act_quant :31-54,
weight_cast_to_fp8 :73-83, weight_cast_to_bf16 :115-138, fp8_gemm :218-242), produced by the UNMODIFIED reference on the
B200 box (Triton needs a GPU; the reference copy under oracle/_ref travels with the snapshot):

    gpurun -- 'python tests/golden/make_fp8_infer_golden.py gpurun_out/fp8_infer_golden.pt.gz'
    cp gpurun_out/fp8_infer_golden.pt.gz tests/golden/

Inputs are regenerated from the seeds by `acts` / `weights` below (CPU generators).
"""
import gzip
import importlib
import io
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))

ACT_CASES = [dict(shape=(64, 1024), seed=11), dict(shape=(3, 7, 512), seed=12), dict(shape=(200, 128), seed=13)]
GEMM_CASES = [dict(M=96, N=384, K=512, seed=21), dict(M=16, N=256, K=1024, seed=22)]


def acts(shape, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(*shape, generator=g) * torch.exp(torch.randn(*shape[:-1], 1, generator=g))
    x[..., 3] *= 30.0
    return x.to(torch.bfloat16)


def weights(N, K, seed):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(N, K, generator=g) * 0.02).to(torch.bfloat16)


def main(out):
    sys.path.insert(0, ROOT)
    from oracle import ref_harness as rh
    rh.activate('cuda')
    kernel = importlib.import_module('llmc.compression.quantization.kernel')
    rec = {'act_quant': [], 'gemm': [], 'torch': torch.__version__, 'device': torch.cuda.get_device_name(0)}
    for c in ACT_CASES:
        y, s = kernel.act_quant(acts(c['shape'], c['seed']).cuda(), 128)
        rec['act_quant'].append({'case': c, 'codes': y.view(torch.uint8).cpu(), 'scales': s.cpu()})
    for c in GEMM_CASES:
        x = acts((c['M'], c['K']), c['seed']).cuda()
        w = weights(c['N'], c['K'], c['seed'] + 100).cuda()
        xq, xs = kernel.act_quant(x, 128)
        wq, ws = kernel.weight_cast_to_fp8(w, 128)
        cc = kernel.fp8_gemm(xq, xs, wq, ws)                    # torch.get_default_dtype() = fp32
        old = torch.get_default_dtype()
        torch.set_default_dtype(torch.bfloat16)
        try:
            back = kernel.weight_cast_to_bf16(wq, ws, 128)
        finally:
            torch.set_default_dtype(old)
        rec['gemm'].append({'case': c, 'xq': xq.view(torch.uint8).cpu(), 'xs': xs.cpu(), 'wq': wq.view(torch.uint8).cpu(),
                            'ws': ws.float().cpu(), 'c': cc.float().cpu(), 'w_back': back.cpu()})
    buf = io.BytesIO()
    torch.save(rec, buf)
    with gzip.open(out, 'wb') as f:
        f.write(buf.getvalue())
    print('wrote', out)


if __name__ == '__main__':
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(HERE, 'fp8_infer_golden.pt.gz'))
