"""End-to-end perplexity parity (north star: "end-to-end perplexity within 0.01").

No dataset or checkpoint is reachable offline, so the model is a random-init 2-block Llama (initializer_range 0.06, which
gives it opinions) and the "text" is random tokens labelled by the float model's own greedy next-token prediction: the
float model scores ln ppl ~3.8 on it, 4-bit RTN ~3.98, GPTQ ~3.96, so the quantization effect (~0.18) is two orders of
magnitude above the bar. What is compared is the SAME model quantized two ways — by the CUDA plug-ins (ALGO_REGISTRY on the GPU) and by a block
loop composed from the CPU oracles (C oracle fake-quant, gptq_oracle Hessian / factor / column update) — and then scored
by the same evaluation forward. "Within 0.01" is stated for models whose perplexity is ~6; on this model (perplexity ~
vocabulary size) the equivalent bar is relative: |ln ppl_gpu - ln ppl_oracle| <= 0.01 / 6.
"""
import copy
import math

import pytest
import torch

from lightcompress_b200 import algo
from oracle import gptq_oracle as go
from oracle import quant_oracle as qo

pytestmark = pytest.mark.gpu
DEV = torch.device('cuda:0')
W4 = {'bit': 4, 'symmetric': False, 'granularity': 'per_group', 'group_size': 128}
SPEC = dict(kind=0, sym=False, qmin=0.0, qmax=15.0)
BAR = 0.01 / 6


def tiny_llama(seed=0):
    from transformers import LlamaConfig, LlamaForCausalLM
    torch.manual_seed(seed)
    cfg = LlamaConfig(vocab_size=512, hidden_size=256, intermediate_size=512, num_hidden_layers=2, num_attention_heads=4,
                      num_key_value_heads=4, max_position_embeddings=256, initializer_range=0.06)
    return LlamaForCausalLM(cfg).to(torch.bfloat16).eval()


def tokens(n, seq, seed):
    g = torch.Generator().manual_seed(seed)
    return [torch.randint(0, 512, (1, seq), generator=g) for _ in range(n)]


@torch.no_grad()
def teacher_labels(model, batches):
    model = model.to(DEV)
    return [model(input_ids=ids.to(DEV)).logits.float()[0, :-1].argmax(-1) for ids in batches]


@torch.no_grad()
def log_ppl(model, batches, labels):
    """mean cross-entropy (fp32 logits) against the float model's greedy labels, on the GPU."""
    model = model.to(DEV)
    tot, n = 0.0, 0
    for ids, tgt in zip(batches, labels):
        logits = model(input_ids=ids.to(DEV)).logits.float()
        tot += float(torch.nn.functional.cross_entropy(logits[0, :-1], tgt, reduction='sum'))
        n += tgt.numel()
    return tot / n


def block_linears(block):
    return {n: m for n, m in block.named_modules() if isinstance(m, torch.nn.Linear)}


@torch.no_grad()
def oracle_rtn(model):
    for block in model.model.layers:
        for m in block_linears(block).values():
            w = m.weight.data
            q, _, _ = qo.fakequant(w, w.shape[0], w.shape[1], 1, 128, **SPEC)
            m.weight.data = q
    return model


@torch.no_grad()
def oracle_gptq(model, calib):
    """GPTQ block loop on the CPU from the oracle pieces, true_sequential off: every Hessian of a block comes from one
    forward of the float block (gptq.py:247-295), then per linear: factor (:128-176), column update (:199-244), and the
    deploy-time fake-quant of the compensated weight with the stored group qparams (:398-452)."""
    emb = model.model.embed_tokens
    rot = model.model.rotary_emb
    hidden = [emb(ids) for ids in calib]
    pos = torch.arange(calib[0].shape[1]).unsqueeze(0)
    for block in model.model.layers:
        lin = block_linears(block)
        H = {n: torch.zeros(m.in_features, m.in_features) for n, m in lin.items()}
        ns = {n: 0 for n in lin}

        def hook(m, x, y, name):
            ns[name] = go.add_batch(H[name], ns[name], x[0])     # H updated in place
        handles = [m.register_forward_hook(lambda m, x, y, name=n: hook(m, x, y, name)) for n, m in lin.items()]
        outs = []
        for h in hidden:
            o = block(h, position_ids=pos, position_embeddings=rot(h, pos))
            outs.append(o[0] if isinstance(o, tuple) else o)
        for hd in handles:
            hd.remove()
        for n, m in lin.items():
            Wf, Hinv = go.process_hessian(H[n].clone(), m.weight.data.clone(), 0.01)
            tmp, _, sc, zp = go.weight_transform(Wf, Hinv, 128, 128, SPEC)
            R, C = tmp.shape
            s = sc.t().reshape(R, C // 128, 1)
            z = zp.t().reshape(R, C // 128, 1)
            t = tmp.reshape(R, C // 128, 128)
            q = (torch.clamp(torch.round(t / s) + z, 0.0, 15.0) - z) * s      # fake_quant_weight_static, fp32 (quant.py:699-717)
            m.weight.data = q.reshape(R, C).to(torch.bfloat16)
        hidden = outs
    return model


def run_plugin(method, model, calib, special=None):
    m = copy.deepcopy(model).to(DEV)
    ad = algo.adapters.LlamaAdapter(m)
    inp = None
    if method != 'RTN':
        inp = ad.collect_first_block_input([{'input_ids': ids.to(DEV)} for ids in calib])
    cfg = {'weight': dict(W4)}
    if special:
        cfg['special'] = special
    a = algo.ALGO_REGISTRY[method](ad, cfg, inp, None, {'calib': {'seq_len': 128}})
    a.run_block_loop()
    a.deploy('fake_quant')
    return m


def test_rtn_perplexity_equals_oracle_pipeline():
    model = tiny_llama()
    ev = tokens(8, 128, 100)
    got = run_plugin('RTN', model, None)
    ref = oracle_rtn(copy.deepcopy(model))
    for a, b in zip(got.model.layers, ref.model.layers):
        for (n, x), (_, y) in zip(block_linears(a).items(), block_linears(b).items()):
            assert torch.equal(x.weight.data.cpu().view(torch.int16), y.weight.data.view(torch.int16)), n
    lab = teacher_labels(copy.deepcopy(model), ev)
    assert log_ppl(got, ev, lab) == log_ppl(ref, ev, lab)
    assert log_ppl(got, ev, lab) - log_ppl(copy.deepcopy(model), ev, lab) > 0.05     # 4-bit RTN costs something here


def test_gptq_perplexity_within_bar_of_oracle_pipeline():
    model = tiny_llama(seed=2)
    calib, ev = tokens(16, 128, 101), tokens(48, 128, 102)     # 6k evaluation tokens
    sp = {'actorder': False, 'static_groups': False, 'percdamp': 0.01, 'blocksize': 128, 'true_sequential': False}
    got = run_plugin('GPTQ', model, calib, sp)
    ref = oracle_gptq(copy.deepcopy(model).float().to(torch.bfloat16), calib)
    lab = teacher_labels(copy.deepcopy(model), ev)
    lp_float, lp_got, lp_ref = log_ppl(copy.deepcopy(model), ev, lab), log_ppl(got, ev, lab), log_ppl(ref, ev, lab)
    print('log-ppl float, gpu plug-in, oracle pipeline:', lp_float, lp_got, lp_ref)
    assert abs(lp_got - lp_ref) <= BAR, (lp_got, lp_ref, lp_float)
    # and the quantization did something measurable, so the bar is not vacuous
    assert lp_got - lp_float > 20 * BAR and math.isfinite(lp_got)
    same = [float((a.weight.data.cpu() == b.weight.data.cpu()).float().mean())
            for x, y in zip(got.model.layers, ref.model.layers)
            for a, b in zip(block_linears(x).values(), block_linears(y).values())]
    # q / k / v of block 0 see bit-identical inputs: the same grid points up to Hessian-rounding flips. Every later linear's
    # input already differs between the CPU and the cuBLAS / cuDNN forward, and GPTQ's error feedback amplifies that into
    # other (equally good) grid points — which is why the bar is on perplexity
    assert min(same[:3]) > 0.99 and min(same) > 0.5, same
