"""GPU parity of the real-quant packers (lcb_pack_rows, lcb_awq_gemm_pack) against outputs of the UNMODIFIED reference
classmethods VllmRealQuantLinear.pack / AutoawqRealQuantLinear.gemm_pack (tests/golden/pack_golden.pt.gz). Bit-exact."""
import gzip
import io
import os

import pytest
import torch

from lightcompress_b200 import pack as P
from lightcompress_b200 import quant as Q
from tests.golden.make_golden import make_input

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'
PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'pack_golden.pt.gz')
with gzip.open(PATH, 'rb') as f:
    GOLD = torch.load(io.BytesIO(f.read()), weights_only=False)


@pytest.mark.parametrize('rec', GOLD['vllm'], ids=[f"b{r['case']['bit']}-{r['case']['shape'][1]}" for r in GOLD['vllm']])
def test_vllm_pack_matches_reference(rec):
    c = rec['case']
    x = make_input(c['shape'], getattr(torch, c['dtype']), c['seed']).to(DEV)
    q = (Q.IntegerQuantizer(c['bit'], True, 'per_group', group_size=c['gs']) if c['gs']
         else Q.IntegerQuantizer(c['bit'], True, 'per_channel'))
    codes, scales, _ = q.real_quant_weight_dynamic(x)
    assert codes.dtype == rec['codes'].dtype and torch.equal(codes.cpu(), rec['codes'])
    packed, s16 = P.vllm_pack(codes, scales, c['bit'])
    assert packed.dtype == torch.int32 and torch.equal(packed.cpu(), rec['packed'])
    assert s16.dtype == torch.float16 and torch.equal(s16.cpu().view(torch.int16), rec['scales16'].view(torch.int16))


@pytest.mark.parametrize('rec', GOLD['awq'], ids=[f"{r['case']['dtype']}-g{r['case']['gs']}" for r in GOLD['awq']])
def test_awq_gemm_pack_matches_reference(rec):
    c = rec['case']
    x = make_input(c['shape'], getattr(torch, c['dtype']), c['seed']).to(DEV)
    q = Q.IntegerQuantizer(4, False, 'per_group', group_size=c['gs'])
    _, scales, zeros = q.real_quant_weight_dynamic(x)
    assert torch.equal(zeros.cpu(), rec['zeros'])
    qw, s16, qz = P.awq_gemm_pack(x, scales, zeros, 4, c['gs'])
    assert torch.equal(qw.cpu(), rec['qweight'])
    assert torch.equal(qz.cpu(), rec['qzeros'])
    assert torch.equal(s16.cpu().view(torch.int16), rec['scales16'].view(torch.int16))


def test_pack_large_roundtrip_property():
    # Llama-3-8B q_proj size: unpacking the words gives back code + 8 for every element
    R, C = 4096, 4096
    codes = torch.randint(-8, 8, (R, C), dtype=torch.int32, device=DEV)
    packed, _ = P.vllm_pack(codes, torch.ones(1, device=DEV), 4)
    shifts = torch.arange(8, device=DEV, dtype=torch.int32) * 4
    un = ((packed.unsqueeze(-1) >> shifts) & 0xF).reshape(R, C)
    assert torch.equal(un, codes + 8)


def test_pack_rejects_bad_input():
    with pytest.raises(NotImplementedError):
        P.awq_gemm_pack(torch.zeros(32, 128, device=DEV), torch.ones(32, 1, device=DEV), torch.zeros(32, 1, device=DEV), 8, 128)
    with pytest.raises(RuntimeError):
        P.vllm_pack(torch.zeros(4, 8, dtype=torch.int32), torch.ones(4, 1), 4)
