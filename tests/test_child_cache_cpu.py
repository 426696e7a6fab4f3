"""CPU: host logic of the block loop's module-output reuse (_ChildCache, algo/blockwise.py) on a toy block — full passes
and passes that stop at a hooked layer (true_sequential), with layers "quantized" in between: the outputs and the collected
layer inputs equal a plain recomputation bit for bit, at every reuse level and when the memory budget is exceeded."""
import copy

import pytest
import torch
import torch.nn as nn

from lightcompress_b200.algo import blockwise as bw


class _Inner(nn.Module):
    def __init__(self):
        super().__init__()
        self.f1, self.act, self.f2 = nn.Linear(8, 16), nn.ReLU(), nn.Linear(16, 8)

    def forward(self, x):
        return self.f2(self.act(self.f1(x)) * 2.0)


class _Block(nn.Module):
    def __init__(self):
        super().__init__()
        self.ln, self.attn, self.ln2, self.mlp = nn.LayerNorm(8), _Inner(), nn.LayerNorm(8), _Inner()

    def forward(self, x):
        x = x + self.attn(self.ln(x))
        return x + self.mlp(self.ln2(x))


def _plugin(special):
    o = object.__new__(bw.BaseBlockwiseQuantization)
    o.special, o.forward_batch, o._child_cache, o._hooked_modules = special, 2, False, set()
    o.input = {'kwargs': [{}]}
    return o


def _sequential_passes(block, data, special):
    """P1..P4 stop at attn.f1 / attn.f2 / mlp.f1 / mlp.f2 (collecting that layer's input), the layer is then changed;
    P5 runs the whole block."""
    b, o, feats, looked_up = copy.deepcopy(block), _plugin(special), [], 0
    for target in ('attn.f1', 'attn.f2', 'mlp.f1', 'mlp.f2'):
        m, got = b.get_submodule(target), []

        def pre(mod, x, got=got):
            got.append(x[0].clone())
            raise bw._StopForward()
        h = m.register_forward_pre_hook(pre)
        o._hooked_modules = {m}
        with torch.no_grad():
            o._block_forward_batched(b, data, True)
        h.remove()
        o._hooked_modules = set()
        feats.append(torch.cat(got))
        m.weight.data = m.weight.data * 1.1
    with torch.no_grad():
        feats.append(torch.cat(o._block_forward_batched(b, data, False)))
    return feats, o


@pytest.mark.parametrize('special', [{'reuse_child_outputs': True}, {'reuse_child_outputs': 'direct'},
                                     {'reuse_child_outputs': True, 'reuse_budget_gb': 0}])
def test_reuse_equals_recompute(special):
    torch.manual_seed(0)
    block = _Block()
    data = [torch.randn(1, 4, 8) for _ in range(6)]
    ref, _ = _sequential_passes(block, data, {'reuse_child_outputs': False})
    got, o = _sequential_passes(block, data, special)
    for a, b in zip(ref, got):
        assert torch.equal(a, b)
    cache = o._child_cache
    assert cache is not False and cache.order[''] == ['ln', 'attn', 'ln2', 'mlp']
    if special.get('reuse_child_outputs') is True and 'reuse_budget_gb' not in special:
        assert cache.deep and 'mlp.f1' in cache.out and len(cache.out['attn']) == 3      # 3 chunks of 2 samples
    else:
        assert not cache.deep and not any('.' in k for k in cache.out)


def test_a_module_whose_own_tensor_changed_is_not_entered():
    """own_signature: when a parent's OWN parameter changes, what flows between its children may change: no child reuse."""
    class P(nn.Module):
        def __init__(self):
            super().__init__()
            self.scale = nn.Parameter(torch.ones(8))
            self.a, self.b = nn.Linear(8, 8), nn.Linear(8, 8)

        def forward(self, x):
            return self.b(self.a(x) * self.scale)

    class B(nn.Module):
        def __init__(self):
            super().__init__()
            self.p = P()

        def forward(self, x):
            return self.p(x)
    torch.manual_seed(1)
    blk, data = B(), [torch.randn(1, 3, 8) for _ in range(4)]
    o = _plugin({'reuse_child_outputs': True})
    with torch.no_grad():
        o._block_forward_batched(blk, data, False)
        blk.p.scale.data = blk.p.scale.data * 3.0          # `b` is "clean" but its input changed through p.scale
        got = torch.cat(o._block_forward_batched(blk, data, False))
        ref = torch.cat([blk(x) for x in data])
    assert torch.equal(got, ref)
