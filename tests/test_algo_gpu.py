"""GPU tests of the algorithm layer: GPTQLayer vs the oracle, the four ALGO_REGISTRY plug-ins end to end on a tiny
random-init Llama, and patch_reference on stand-in classes with the reference's attribute contract."""
import copy

import pytest
import torch
import torch.nn as nn

from lightcompress_b200 import algo
from lightcompress_b200 import quant as Q
from oracle import gptq_oracle as go

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


def _rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm())


def test_gptq_layer_matches_oracle():
    g = torch.Generator().manual_seed(0)
    R, C, T = 96, 384, 512
    W = (torch.randn(R, C, generator=g) * 0.02).to(torch.bfloat16)
    xs = [(torch.randn(1, T, C, generator=g) * torch.exp(0.5 * torch.randn(C, generator=g))).to(torch.bfloat16) for _ in range(3)]
    H = torch.zeros(C, C)
    n = 0
    for x in xs:
        n = go.add_batch(H, n, x)
    Wf, Hinv = go.process_hessian(H, W, 0.01)
    tmp_o, loss_o, sc_o, zp_o = go.weight_transform(Wf, Hinv, 128, 128, dict(kind=0, sym=False, qmin=0.0, qmax=15.0))
    wq = Q.IntegerQuantizer(4, False, 'per_group', group_size=128)
    layer = algo.GPTQLayer(W.to(DEV), wq, blocksize=128, percdamp=0.01)
    for x in xs:
        layer.add_batch(x.to(DEV))
    tmp = layer.quantize()
    assert tmp.dtype == torch.float32 and tmp.shape == W.shape
    assert _rel(tmp.cpu(), tmp_o) < 2e-3
    assert _rel(layer.scales.cpu(), sc_o) < 2e-3
    s, z = layer.buf_qparams()
    assert s.shape == (R * C // 128, 1) and z.shape == s.shape
    # deploy-time fake quant with the stored qparams lands on the quantization grid
    args = {'scales': s, 'zeros': z, 'qmax': wq.qmax, 'qmin': wq.qmin}
    fq = wq.fake_quant_weight_static(tmp, args)
    codes = torch.round(fq.reshape(-1, 128) / s + z)
    assert float(codes.min()) >= 0 and float(codes.max()) <= 15


def _tiny_llama(dtype=torch.bfloat16, seed=0):
    from transformers import LlamaConfig, LlamaForCausalLM
    torch.manual_seed(seed)
    cfg = LlamaConfig(vocab_size=512, hidden_size=256, intermediate_size=512, num_hidden_layers=2, num_attention_heads=4,
                      num_key_value_heads=4, max_position_embeddings=256)
    m = LlamaForCausalLM(cfg).to(dtype).to(DEV).eval()
    return m


def _calib(n=4, seq=128, vocab=512, seed=1):
    g = torch.Generator().manual_seed(seed)
    return [{'input_ids': torch.randint(0, vocab, (1, seq), generator=g).to(DEV)} for _ in range(n)]


def _run(method, quant_cfg, model, data):
    ad = algo.adapters.LlamaAdapter(model)
    inp = ad.collect_first_block_input(data)
    a = algo.ALGO_REGISTRY[method](ad, quant_cfg, inp, None, {'calib': {'seq_len': 128}})
    a.run_block_loop()
    a.deploy('fake_quant')
    return a


W4 = {'bit': 4, 'symmetric': False, 'granularity': 'per_group', 'group_size': 128}


@torch.no_grad()
def _logits(model, data):
    return torch.cat([model(**b).logits.float() for b in data])


@pytest.mark.parametrize('method,cfg', [
    ('RTN', {'weight': W4}),
    ('GPTQ', {'weight': W4, 'special': {'actorder': False, 'static_groups': False, 'percdamp': 0.01, 'blocksize': 128,
                                        'true_sequential': True}}),
    ('Awq', {'weight': W4, 'special': {'trans': True, 'trans_version': 'v2', 'weight_clip': True, 'clip_sym': False}}),
    ('SmoothQuant', {'weight': {'bit': 8, 'symmetric': True, 'granularity': 'per_channel'},
                     'act': {'bit': 8, 'symmetric': True, 'granularity': 'per_token'}, 'special': {'alpha': 0.8}}),
])
def test_block_loop_end_to_end(method, cfg):
    import lightcompress_b200.algo.adapters  # noqa: F401
    model = _tiny_llama()
    data = _calib()
    ref = _logits(model, data)
    q = copy.deepcopy(model)
    a = _run(method, copy.deepcopy(cfg), q, data)
    out = _logits(q, data)
    assert torch.isfinite(out).all()
    err = float((out - ref).pow(2).mean() / ref.pow(2).mean())
    assert err < 0.2, err
    # every block linear is on a 4-bit grid: at most 16 distinct values per group of 128 (w-only methods)
    if method in ('RTN', 'Awq'):
        for blk in a.blocks:
            for n, m in a.model.get_block_linears(blk).items():
                g = m.weight.data.float().reshape(-1, 128).sort(dim=1).values
                distinct = 1 + (g[:, 1:] != g[:, :-1]).sum(dim=1)
                assert int(distinct.max()) <= 16, n
    if method == 'Awq':
        assert len(a.best_index) == 2 * 4   # 4 subsets per block searched (v->o has matching shapes here)


def test_gptq_shared_hessian_equals_per_linear():
    # q/k/v and gate/up see the same input: one shared HessianState per subset must give what one per linear gives
    model = _tiny_llama(seed=5)
    data = _calib(n=4, seed=6)
    outs = {}
    for share in (True, False):
        q = copy.deepcopy(model)
        cfg = {'weight': W4, 'special': {'actorder': False, 'static_groups': False, 'percdamp': 0.01, 'blocksize': 128,
                                         'true_sequential': True, 'share_hessian': share}}
        a = _run('GPTQ', cfg, q, data)
        outs[share] = torch.cat([m.weight.data.float().flatten() for blk in a.blocks
                                 for m in a.model.get_block_linears(blk).values()])
    rel = float((outs[True] - outs[False]).norm() / outs[False].norm())
    assert rel < 1e-3, rel     # split-K order of the Hessian atomics differs run to run: not bit-equal, same grid points
    assert float((outs[True] == outs[False]).float().mean()) > 0.99


def test_gptq_graph_replayed_forward_equals_eager(monkeypatch):
    # the recorded-and-replayed block forward (CUDA graph, hooks included) must do what the per-sample eager loop does:
    # same Hessian sample count, same quantized weights (up to the run-to-run order of the Hessian reduce-adds)
    from lightcompress_b200.algo import blockwise as bw
    from lightcompress_b200.algo import gptq as lcgptq
    model = _tiny_llama(seed=7)
    data = _calib(n=6, seed=8)
    used, counts = [], []
    orig = bw.BaseBlockwiseQuantization._block_forward_graph
    monkeypatch.setattr(bw.BaseBlockwiseQuantization, '_block_forward_graph',
                        lambda self, *a, **k: (used.append(1), orig(self, *a, **k))[1])
    fin = lcgptq.HessianState.finalize
    monkeypatch.setattr(lcgptq.HessianState, 'finalize',
                        lambda self: (counts.append(self.nsamples) if self.H is None else None, fin(self))[1])
    outs = {}
    for graph in (True, False):
        q = copy.deepcopy(model)
        cfg = {'weight': W4, 'special': {'actorder': False, 'static_groups': False, 'percdamp': 0.01, 'blocksize': 128,
                                         'true_sequential': True, 'cuda_graph': graph, 'forward_batch': 1}}
        n_used = len(used)
        a = _run('GPTQ', cfg, q, data)
        assert (len(used) > n_used) == graph
        outs[graph] = torch.cat([m.weight.data.float().flatten() for blk in a.blocks
                                 for m in a.model.get_block_linears(blk).values()])
    assert counts and all(c == 6 for c in counts), counts
    rel = float((outs[True] - outs[False]).norm() / outs[False].norm())
    assert rel < 1e-3, rel
    assert float((outs[True] == outs[False]).float().mean()) > 0.99


def test_gptq_beats_rtn_on_calibration_data():
    model = _tiny_llama(seed=3)
    data = _calib(n=8, seed=4)
    ref = _logits(model, data)
    errs = {}
    for method, cfg in (('RTN', {'weight': W4}),
                        ('GPTQ', {'weight': W4, 'special': {'actorder': False, 'static_groups': False, 'percdamp': 0.01,
                                                            'blocksize': 128, 'true_sequential': True}})):
        q = copy.deepcopy(model)
        _run(method, copy.deepcopy(cfg), q, data)
        errs[method] = float((_logits(q, data) - ref).pow(2).mean())
    assert errs['GPTQ'] < errs['RTN'], errs


# ---- patch_reference on stand-ins that follow the reference's attribute contract (gptq.py / awq.py) ------------------------
class _RefLikeGPTQ(object):
    """Attributes as gptq.py:29-56, 104-176, 297-309; the three hot methods raise unless patched."""

    def __init__(self, wquantizer, blocksize=128):
        self.wquantizer, self.blocksize = wquantizer, blocksize
        self.static_groups = self.actorder = self.owq = False
        self.percdamp = 0.01
        self.layers_cache, self.groups, self.qparams = {}, [], {}

    def layer_init(self, layer, name):
        C = layer.weight.shape[1]
        self.layers_cache[name] = {'H': torch.zeros((C, C), device=layer.weight.device), 'nsamples': 0, 'columns': C}
        self.columns, self.n_out, self.n_nonout = C, 0, C

    def add_batch(self, layer, name, inp, out):
        raise AssertionError('reference path taken')

    def process_hessian_and_weights(self, layer, name):   # the torch part of gptq.py:128-176
        W = layer.weight.data.clone().float()
        H = self.layers_cache[name]['H']
        dead = torch.diag(H) == 0
        H[dead, dead] = 1
        W[:, dead] = 0
        d = torch.arange(self.columns, device=H.device)
        H[d, d] += self.percdamp * torch.mean(torch.diag(H))
        H = torch.linalg.cholesky(torch.cholesky_inverse(torch.linalg.cholesky(H)), upper=True)
        return W, H

    def weight_transform(self, W, Hinv, Losses, tmp):
        raise AssertionError('reference path taken')


def test_patch_reference_gptq_contract():
    reg = {'GPTQ': _RefLikeGPTQ}
    assert algo.patch_reference(reg) == ['GPTQ'] and algo.patch_reference(reg) == []
    g = torch.Generator().manual_seed(5)
    R, C = 64, 256
    layer = nn.Linear(C, R, bias=False).to(DEV)
    layer.weight.data = (torch.randn(R, C, generator=g) * 0.02).to(DEV)
    xs = [torch.randn(1, 300, C, generator=g).to(torch.bfloat16) for _ in range(2)]
    wq = Q.IntegerQuantizer(4, False, 'per_group', group_size=128)
    obj = _RefLikeGPTQ(wq)
    obj.layer_init(layer, 'l')
    for x in xs:
        obj.add_batch(layer, 'l', x.to(DEV), None)
    assert obj.layers_cache['l']['nsamples'] == 2
    W, Hinv = obj.process_hessian_and_weights(layer, 'l')
    Losses, tmp = torch.zeros_like(W), torch.zeros_like(W)
    obj.weight_transform(W, Hinv, Losses, tmp)
    # oracle on the same inputs
    H = torch.zeros(C, C)
    n = 0
    for x in xs:
        n = go.add_batch(H, n, x)
    Wf, Hinv_o = go.process_hessian(H, layer.weight.data.cpu(), 0.01)
    tmp_o, loss_o, sc_o, zp_o = go.weight_transform(Wf, Hinv_o, 128, 128, dict(kind=0, sym=False, qmin=0.0, qmax=15.0))
    assert _rel(tmp.cpu(), tmp_o) < 2e-3
    assert len(obj.groups) == C // 128 and obj.groups[0]['scale'].shape == (R, 1) and obj.groups[1]['zero'].shape == (R, 1)
    assert _rel(torch.stack([g_['scale'].reshape(-1) for g_ in obj.groups]).cpu(), sc_o) < 2e-3


def test_batched_block_forward_equals_per_sample():
    """`forward_batch` samples per block forward (GPTQ: the hooks only accumulate sums over tokens): same outputs per
    sample, same Hessian and batch count as one forward per sample."""
    model = _tiny_llama(seed=9)
    data = _calib(n=6, seed=10)
    cfg = {'weight': W4, 'special': {'actorder': False, 'static_groups': False, 'percdamp': 0.01, 'blocksize': 128,
                                     'true_sequential': False}}
    res = {}
    for fb in (1, 4):
        c = copy.deepcopy(cfg)
        c['special'].update(forward_batch=fb, cuda_graph=False)
        ad = algo.adapters.LlamaAdapter(copy.deepcopy(model))
        inp = ad.collect_first_block_input(data)
        a = algo.ALGO_REGISTRY['GPTQ'](ad, c, inp, None, {'calib': {'seq_len': 128}})
        blk = a.blocks[0].to(DEV)
        a.block_idx = 0
        a.block_init(blk)
        handles = a.register_hooks(a.model.get_block_linears(blk), {})
        out = a.block_forward(blk)
        for h in handles:
            h.remove()
        g = a.layers_cache['mlp.down_proj']
        res[fb] = (torch.cat(out), g.hessian.S.clone(), g.hessian.nsamples)
    assert res[1][2] == res[4][2] == 6
    assert res[1][0].shape == res[4][0].shape
    assert float((res[1][0].float() - res[4][0].float()).abs().max()) <= 2e-2 * float(res[1][0].float().abs().max())
    S1, S4 = res[1][1].triu(), res[4][1].triu()
    assert float((S1 - S4).abs().max() / S1.abs().max()) < 1e-2      # bf16 activations of two GEMM batchings


def test_child_output_reuse_changes_nothing():
    """The shipped GPTQ recipe (act-order, true_sequential, quant_out) makes five passes over a block; norms / attention
    whose parameters did not change since an earlier pass are looked up instead of recomputed (_ChildCache). Same deployed
    weights and same block outputs, bit for bit, as with the lookup disabled."""
    model = _tiny_llama(seed=12)
    data = _calib(n=8, seed=13)
    outs = {}
    for reuse in (True, 'direct', 'budget0', False):
        q = copy.deepcopy(model)
        cfg = {'weight': W4, 'quant_out': True,
               'special': {'actorder': True, 'static_groups': False, 'percdamp': 0.01, 'blocksize': 128,
                           'true_sequential': True, 'forward_batch': 4,
                           'reuse_child_outputs': True if reuse == 'budget0' else reuse,
                           'reuse_budget_gb': 0 if reuse == 'budget0' else 64}}
        a = _run('GPTQ', cfg, q, data)
        outs[reuse] = (torch.cat([m.weight.data.float().flatten() for blk in a.blocks
                                  for m in a.model.get_block_linears(blk).values()]),
                       torch.cat([t.float().flatten() for t in a.input['data']]))
    for mode in (True, 'direct', 'budget0'):     # every level / the block's direct children only / memory budget exceeded
        assert torch.equal(outs[mode][0], outs[False][0]), mode
        assert torch.equal(outs[mode][1], outs[False][1]), mode


def test_block_loop_prefetches_blocks_that_start_on_the_host():
    """SURVEY §8f N4 (base_bq.py:397): a model whose blocks still live in host memory is calibrated block by block while
    the next block is copied in on a side stream by a worker thread — same weights as the model that starts on the GPU,
    with and without the prefetch, and every block ends on the device."""
    import copy
    from transformers import LlamaConfig, LlamaForCausalLM
    torch.manual_seed(3)
    cfg = LlamaConfig(vocab_size=256, hidden_size=256, intermediate_size=512, num_hidden_layers=4, num_attention_heads=4,
                      num_key_value_heads=2, max_position_embeddings=256)
    base = LlamaForCausalLM(cfg).to(torch.bfloat16).eval()
    g = torch.Generator().manual_seed(1)
    calib = [torch.randint(0, 256, (1, 64), generator=g) for _ in range(8)]
    q = {'weight': {'bit': 4, 'symmetric': False, 'granularity': 'per_group', 'group_size': 128},
         'special': {'actorder': False, 'percdamp': 0.01, 'blocksize': 128, 'true_sequential': True}, 'quant_out': True}

    def run(host_blocks, prefetch):
        m = copy.deepcopy(base).to(DEV)
        ad = algo.adapters.LlamaAdapter(m)
        inp = ad.collect_first_block_input([{'input_ids': ids.to(DEV)} for ids in calib])
        if host_blocks:
            for b in m.model.layers:
                b.to('cpu')
        cfgq = copy.deepcopy(q)
        cfgq['special']['prefetch_blocks'] = prefetch
        a = algo.ALGO_REGISTRY['GPTQ'](ad, cfgq, inp, None, {'calib': {'seq_len': 64}})
        a.run_block_loop()
        assert all(p.is_cuda for b in m.model.layers for p in b.parameters())
        return [p.detach().float().cpu() for b in m.model.layers for p in b.parameters()]

    ref = run(False, True)
    for host_blocks, prefetch in ((True, True), (True, False)):
        got = run(host_blocks, prefetch)
        # the Hessian's split-K reductions are not run-to-run deterministic: equal up to GPTQ's own reproducibility
        same = sum(float((a == b).float().mean()) for a, b in zip(ref, got)) / len(ref)
        assert same > 0.99, same
