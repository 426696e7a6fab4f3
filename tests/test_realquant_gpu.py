"""Real-quant deploy through the plug-ins: `deploy('vllm_quant' | 'autoawq_quant' | ...)` replaces every
block linear by the backend's real-quant module — codes, scales, zero-points quantized and packed on the GPU — and the
buffers are bit-equal to what the UNMODIFIED reference pipeline deploys (base_bq.py:933-962, module_utils.py:855-1158:
numpy packing on the host, a Python loop per input channel). BASELINE configs[4] (DeepSeek-V2-Lite MoE, FP8 E4M3
per-block) runs on a synthetic ModuleList-of-experts block: the reference cannot (its FP8 rounding lives in qtorch, which
is not installable offline), so that case is checked against the C oracle instead."""
import copy

import pytest
import torch
import torch.nn as nn

from lightcompress_b200 import algo
from oracle import quant_oracle as qo
from oracle import ref_harness as rh

pytestmark = pytest.mark.gpu
DEV = torch.device('cuda:0')


def _tiny_llama(seed=0):
    from transformers import LlamaConfig, LlamaForCausalLM
    torch.manual_seed(seed)
    cfg = LlamaConfig(vocab_size=512, hidden_size=256, intermediate_size=512, num_hidden_layers=2, num_attention_heads=4,
                      num_key_value_heads=4, max_position_embeddings=256, initializer_range=0.06)
    return LlamaForCausalLM(cfg).to(torch.bfloat16).eval()


def _buffers(model):
    out = {}
    for i, b in enumerate(model.model.layers):
        for n, m in b.named_modules():
            if hasattr(m, 'in_features') and not isinstance(m, nn.Linear):
                for k, v in m.named_buffers():
                    out[f'{i}.{n}.{k}'] = v.detach().to(DEV)
    return out


CASES = [
    ('vllm_quant', {'bit': 8, 'symmetric': True, 'granularity': 'per_channel'}),
    ('vllm_quant', {'bit': 4, 'symmetric': True, 'granularity': 'per_group', 'group_size': 128, 'need_pack': True}),
    ('lightllm_quant', {'bit': 8, 'symmetric': True, 'granularity': 'per_channel'}),
    ('autoawq_quant', {'bit': 4, 'symmetric': False, 'granularity': 'per_group', 'group_size': 128,
                       'pack_version': 'gemm_pack'}),
]


@pytest.mark.skipif(not rh.available(), reason='oracle/_ref missing (python oracle/build_ref.py)')
@pytest.mark.parametrize('fmt,wcfg', CASES, ids=[f'{f}-w{w["bit"]}-{w["granularity"]}' for f, w in CASES])
def test_real_quant_deploy_equals_reference_pipeline(fmt, wcfg):
    model = _tiny_llama(seed=11)
    q = {'method': 'RTN', 'weight': dict(wcfg)}
    calib = [{'input_ids': torch.randint(0, 512, (1, 32), generator=torch.Generator().manual_seed(1))}]
    ref = copy.deepcopy(model)
    rh.run_pipeline(ref, 'Llama', copy.deepcopy(q), calib, device='cuda', deploy=fmt)
    got = copy.deepcopy(model).to(DEV)
    a = algo.ALGO_REGISTRY['RTN'](algo.adapters.LlamaAdapter(got), copy.deepcopy(q), None, None, {})
    a.run_block_loop()
    a.deploy(fmt)
    br, bg = _buffers(ref), _buffers(got)
    assert len(br) >= 14 * 2 and br.keys() == bg.keys(), (sorted(br)[:6], sorted(bg)[:6])
    for k in br:
        assert br[k].dtype == bg[k].dtype and br[k].shape == bg[k].shape, (k, br[k].dtype, bg[k].dtype, br[k].shape, bg[k].shape)
        assert torch.equal(br[k], bg[k]), k
    m_ref, m_got = ref.model.layers[0].mlp.up_proj, got.model.layers[0].mlp.up_proj
    assert type(m_ref).__name__ == type(m_got).__name__
    for attr in ('weight_shape', 'weight_dtype', 'scales_shape', 'scales_dtype', 'zeros_shape', 'zeros_dtype'):
        assert getattr(m_ref, attr) == getattr(m_got, attr), attr


class _Expert(nn.Module):
    def __init__(self, hidden, inter):
        super().__init__()
        self.gate_proj = nn.Linear(hidden, inter, bias=False)
        self.up_proj = nn.Linear(hidden, inter, bias=False)
        self.down_proj = nn.Linear(inter, hidden, bias=False)


class _MoEBlock(nn.Module):
    """The linears of a DeepSeek-V2-Lite MoE block that matter for shapes (models/deepseekv2.py:67-190): attention with the
    576-row kv_a_proj (4.5 x 128: ragged per-block tiles), routed experts in a ModuleList, shared experts."""

    def __init__(self, hidden=512, inter=352, n_experts=6):
        super().__init__()
        self.self_attn = nn.ModuleDict({'q_proj': nn.Linear(hidden, 768, bias=False),
                                        'kv_a_proj_with_mqa': nn.Linear(hidden, 576, bias=False),
                                        'o_proj': nn.Linear(hidden, hidden, bias=False)})
        self.mlp = nn.ModuleDict({'experts': nn.ModuleList([_Expert(hidden, inter) for _ in range(n_experts)]),
                                  'shared_experts': _Expert(hidden, 2 * inter)})


class _MoEAdapter(object):
    block_name_prefix = 'model.layers'

    def __init__(self, blocks):
        self.blocks = blocks

    def get_blocks(self):
        return self.blocks

    def get_block_linears(self, block):
        return {n: m for n, m in block.named_modules() if isinstance(m, nn.Linear)}

    def get_subsets_in_block(self, block):
        return []


def test_fp8_per_block_moe_block_deploys_through_rtn():
    torch.manual_seed(0)
    blocks = nn.ModuleList([_MoEBlock(), _MoEBlock()]).to(torch.bfloat16).to(DEV)
    originals = {n: p.detach().clone() for n, p in blocks.named_parameters()}
    q = {'weight': {'quant_type': 'float-quant', 'bit': 'e4m3', 'symmetric': True, 'granularity': 'per_block', 'block_size': 128,
                    'use_qtorch': True},
         'act': {'quant_type': 'float-quant', 'bit': 'e4m3', 'symmetric': True, 'granularity': 'per_group', 'group_size': 128,
                 'use_qtorch': True}}
    a = algo.ALGO_REGISTRY['RTN'](_MoEAdapter(blocks), q, None, None, {})
    a.run_block_loop()
    a.deploy('vllm_quant')
    n_checked = 0
    for bi, block in enumerate(blocks):
        for name, m in block.named_modules():
            if not isinstance(m, algo.VllmRealQuantLinear):
                assert not isinstance(m, nn.Linear), name      # every linear was replaced
                continue
            w = originals[f'{bi}.{name}.weight']
            R, C = w.shape
            assert m.weight.dtype == torch.float8_e4m3fn and tuple(m.weight.shape) == (R, C)
            assert m.weight_scale_inv.dtype == torch.float32
            assert tuple(m.weight_scale_inv.shape) == (-(-R // 128), -(-C // 128))
            codes, scales, _ = qo.fakequant(w.cpu(), R, C, 128, 128, kind=1, sym=True, qmin=-448.0, qmax=448.0,
                                            out_mode=qo.OUT_FP8, math_dtype=torch.float32)
            assert torch.equal(m.weight.view(torch.uint8).cpu(), codes), name
            assert torch.equal(m.weight_scale_inv.cpu().reshape(-1), scales), name
            n_checked += 1
    assert n_checked == 2 * (3 + 3 * 6 + 3)
