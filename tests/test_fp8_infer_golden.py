"""tests/golden/fp8_infer_golden.pt.gz — outputs of the reference's own Triton kernels (kernel.py) produced on the B200 box
by tests/golden/make_fp8_infer_golden.py — as the comparand of
  * oracle/fp8_infer_oracle.py (CPU, IEEE divisions) in the `-m "not gpu"` suite, and
  * the CUDA kernels (csrc/fp8_infer.cu) in the `-m gpu` suite, without needing oracle/_ref.
Triton lowers fp32 `/` to div.full.f32 (<= 2 ulp, not the IEEE quotient): against the IEEE oracle the scales may sit one
fp32 ulp apart and codes on an exact e4m3 rounding tie one code step apart; the CUDA act_quant issues the same
instruction and is bit-exact. fp8_gemm contracts identical codes and scales: 1e-5."""
import gzip
import io
import os

import pytest
import torch

from oracle import fp8_infer_oracle as fo
from tests.golden.make_fp8_infer_golden import acts

F8 = torch.float8_e4m3fn
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'fp8_infer_golden.pt.gz')


def _load():
    with gzip.open(GOLDEN, 'rb') as f:
        return torch.load(io.BytesIO(f.read()), weights_only=False)


G = _load()


def _ulp(a, b):
    return int((a.contiguous().view(torch.int32).long() - b.contiguous().view(torch.int32).long()).abs().max())


@pytest.mark.parametrize('rec', G['act_quant'], ids=[str(tuple(r['case']['shape'])) for r in G['act_quant']])
def test_oracle_act_quant_vs_triton_golden(rec):
    x = acts(rec['case']['shape'], rec['case']['seed'])
    y, s = fo.act_quant(x, 128)
    assert s.shape == rec['scales'].shape and y.shape == rec['codes'].shape
    assert _ulp(s, rec['scales']) <= 1
    d = (y.view(torch.uint8).to(torch.int16) - rec['codes'].to(torch.int16)).abs()
    frac = float((d > 0).float().mean())
    print(f"[act_quant {rec['case']['shape']}] oracle vs Triton: scale ulp {_ulp(s, rec['scales'])}, code mismatch {frac:.2e}")
    assert int(d.max()) <= 1 and frac < 5e-3


@pytest.mark.parametrize('rec', G['gemm'], ids=[f"{r['case']['M']}x{r['case']['N']}x{r['case']['K']}" for r in G['gemm']])
def test_oracle_fp8_gemm_and_dequant_vs_triton_golden(rec):
    xq, wq = rec['xq'].view(F8), rec['wq'].view(F8)
    c = fo.fp8_gemm(xq, rec['xs'], wq, rec['ws'], out_dtype=torch.float32)
    rel = float((c - rec['c']).norm() / rec['c'].norm())
    print(f"[fp8_gemm {rec['case']}] oracle vs Triton: relative Frobenius error {rel:.2e}")
    assert rel < 1e-5
    # weight_cast_to_bf16 (kernel.py:115-138): one fp32 product per element, rounded once
    N, K = wq.shape
    s_full = rec['ws'].repeat_interleave(128, 0)[:N].repeat_interleave(128, 1)[:, :K]
    back = (wq.float() * s_full).to(torch.bfloat16)
    assert torch.equal(back.view(torch.int16), rec['w_back'].view(torch.int16))


@pytest.mark.gpu
def test_cuda_kernels_vs_triton_golden():
    from lightcompress_b200 import fp8_infer as fi
    dev = 'cuda:0'
    for rec in G['act_quant']:
        x = acts(rec['case']['shape'], rec['case']['seed']).to(dev)
        y, s = fi.act_quant(x, 128)
        assert torch.equal(s.cpu().view(torch.int32), rec['scales'].view(torch.int32)), 'act_quant scales are not bit-exact'
        assert torch.equal(y.view(torch.uint8).cpu(), rec['codes']), 'act_quant codes are not bit-exact'
    for rec in G['gemm']:
        xq, wq = rec['xq'].view(F8).to(dev), rec['wq'].view(F8).to(dev)
        c = fi.fp8_gemm(xq, rec['xs'].to(dev), wq, rec['ws'].to(dev), out_dtype=torch.float32).cpu()
        assert float((c - rec['c']).norm() / rec['c'].norm()) < 1e-5
        assert float((c - rec['c']).abs().max() / rec['c'].abs().max()) < 1e-5
        back = fi.weight_cast_to_bf16(wq, rec['ws'].to(dev), 128).cpu()
        assert torch.equal(back.view(torch.int16), rec['w_back'].view(torch.int16))
