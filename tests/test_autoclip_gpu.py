"""GPU parity of the AutoClip path against the pinned oracle (oracle/autoclip_oracle.py, itself bit-exact against the
UNMODIFIED reference AutoClipper: tests/test_autoclip_oracle.py).

 * clamped fake-quant (fake_quant_weight_clipped -> lcb_fakequant with clip_ratio): bit-exact against the reference's
   golden q_w (auto_clip.py:258-260) and against the oracle on random tensors;
 * lcb_autoclip_search: the selected (best_max, best_min) must equal the oracle's on every (channel, group): the kernel
   follows the reference's arithmetic op for op — every product x*w rounded to the tensor dtype, fp32 sum, one rounding,
   rounded difference / square, fp32 mean — so the only freedom left is the ORDER of the fp32 sum over a group's rounded
   products (torch CPU: vectorised cascade; here: the tensor pipe), which can move an output by one ulp when the sum sits
   on a rounding boundary. A disagreement is accepted only on an exact tie of the oracle's own error curve."""
import gzip
import io
import os

import pytest
import torch

from lightcompress_b200 import ops
from lightcompress_b200 import quant as Q
from oracle import autoclip_oracle as aco
from oracle import quant_oracle as qo
from tests.golden.make_autoclip_golden import inputs

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'
PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'autoclip_golden.pt.gz')
with gzip.open(PATH, 'rb') as f:
    GOLD = torch.load(io.BytesIO(f.read()), weights_only=False)
IDS = [f"oc{r['case']['oc']}-C{r['case']['C']}-{r['case']['dtype']}-sym{int(r['case']['clip_sym'])}" for r in GOLD]


def spec_of(c):
    qmin, qmax = (-(2.0 ** (c['bit'] - 1)), 2.0 ** (c['bit'] - 1) - 1) if c['sym'] else (0.0, 2.0 ** c['bit'] - 1)
    return dict(kind=0, sym=c['sym'], qmin=qmin, qmax=qmax)


@pytest.mark.parametrize('rec', GOLD, ids=IDS)
def test_clipped_fake_quant_matches_reference_golden(rec):
    c = rec['case']
    W, _ = inputs(c)
    wq = Q.IntegerQuantizer(c['bit'], c['sym'], 'per_group', group_size=c['gs'])
    for i_s, ref in rec['q_w'].items():
        got = wq.fake_quant_weight_clipped(W.to(DEV), 1 - i_s / 20, c['clip_sym']).cpu()
        assert torch.equal(got.view(torch.int16), ref.view(torch.int16)), f'i_s={i_s}'


@pytest.mark.parametrize('dtype', [torch.bfloat16, torch.float16])
@pytest.mark.parametrize('gran,gs,cols', [('per_group', 128, 1024), ('per_group', 64, 512), ('per_channel', 0, 4096),
                                          ('per_channel', 0, 14336), ('per_channel', 0, 333)])
@pytest.mark.parametrize('sym,clip_sym', [(False, False), (True, True), (False, True)])
def test_clipped_fake_quant_matches_oracle(dtype, gran, gs, cols, sym, clip_sym):
    g = torch.Generator().manual_seed(cols + gs)
    W = (torch.randn(24, cols, generator=g) * 0.02)
    W[3, :64] = W[3, :64].abs() + 0.01
    W[5, :64] = -W[5, :64].abs() - 0.01
    W = W.to(dtype)
    kw = {'group_size': gs} if gran == 'per_group' else {}
    wq = Q.IntegerQuantizer(4, sym, gran, **kw)
    spec = dict(kind=0, sym=sym, qmin=-8.0 if sym else 0.0, qmax=7.0 if sym else 15.0)
    for ratio in (0.95, 0.7, 0.55):
        got = wq.fake_quant_weight_clipped(W.to(DEV), ratio, clip_sym).cpu()
        ref, _, _ = qo.fakequant(W, 24, cols, 1, gs if gs else cols, clip=(-ratio if clip_sym else ratio), **spec)
        assert torch.equal(got.view(torch.int16), ref.view(torch.int16)), ratio


def _check_search(W, x, c, tok):
    spec = spec_of(c)
    bmax_o, bmin_o, errs = aco.auto_clip_layer(W, x, spec, c['gs'], c['clip_sym'], n_sample_token=tok)
    wq = Q.IntegerQuantizer(c['bit'], c['sym'], 'per_group', group_size=c['gs'])
    bmax, bmin = ops.autoclip_search(W.to(DEV), x.reshape(-1, x.shape[-1]).to(DEV), wq, c['clip_sym'])
    bmax, bmin = bmax.cpu(), bmin.cpu()
    assert bmax.shape == bmax_o.shape and bmin.shape == bmin_o.shape
    same = (bmax.view(torch.int16) == bmax_o.view(torch.int16)) & (bmin.view(torch.int16) == bmin_o.view(torch.int16))
    agree = float(same.float().mean())
    # disagreements must be near-ties of the oracle's own error curve
    errs = errs.float()                                    # [steps, oc, ng]
    org_max = (W.reshape(W.shape[0], -1, c['gs']).abs() if c['clip_sym'] else W.reshape(W.shape[0], -1, c['gs'])).amax(-1)
    ratio = (bmax.squeeze(-1).float() / org_max.float()).clamp(0.5, 1.0)
    step = torch.round((1 - ratio) * 20).long().clamp(0, errs.shape[0] - 1)
    e_sel = errs.gather(0, step.unsqueeze(0)).squeeze(0)
    e_min = errs.min(0).values
    bad = (~same.squeeze(-1)) & (e_sel > 1.005 * e_min + 1e-12)      # one ulp of a bf16 / fp16 error mean
    return agree, int(bad.sum()), same.numel()


@pytest.mark.parametrize('rec', GOLD, ids=IDS)
def test_autoclip_search_matches_oracle(rec):
    c = rec['case']
    W, x = inputs(c)
    agree, bad, n = _check_search(W, x, c, c['tok'])
    print(f'[autoclip {c}] identical (best_max, best_min) on {agree * n:.0f} of {n} (channel, group) pairs')
    # the arithmetic is the reference's op for op; only the ORDER of the fp32 sum over a group's rounded products is not
    # pinned (torch CPU: vectorised cascade), which can move an output by one ulp when the sum sits on a rounding boundary
    assert agree >= 0.999 and bad == 0, f'agreement {agree:.5f}, {bad} picks are not exact ties of the oracle error curve'


def test_autoclip_search_ragged_tokens_and_llama_shape():
    c = dict(oc=256, C=1024, tok=75, dtype='bfloat16', seed=31, bit=4, sym=False, gs=128, clip_sym=False)
    W, x = inputs(c)
    agree, bad, n = _check_search(W, x, c, c['tok'])    # 75 tokens: padded to 80 rows, mean over 75
    print(f'[autoclip ragged] identical on {agree * n:.0f} of {n}')
    assert agree >= 0.999 and bad == 0, (agree, bad)


def test_autoclip_rejects_unsupported():
    wq = Q.IntegerQuantizer(4, False, 'per_group', group_size=192)
    W = torch.randn(64, 384, device=DEV).to(torch.bfloat16)
    x = torch.randn(16, 384, device=DEV).to(torch.bfloat16)
    with pytest.raises(NotImplementedError):
        ops.autoclip_search(W, x, wq, False)            # groups are tiles of <= 128 or k * 128 columns: no kernel, no fallback
    with pytest.raises(RuntimeError):
        ops.autoclip_search(W.cpu(), x.cpu(), Q.IntegerQuantizer(4, False, 'per_group', group_size=128), False)


class _RefLikeAutoClipper(object):
    """Attribute contract of auto_clip.py:21-42; the hot method raises unless patched."""

    def __init__(self, wquantizer, clip_sym):
        self.wquantizer, self.aquantizer = wquantizer, None
        self.clip_version, self.clip_sym, self.w_only, self.padding_mask = 'v1', clip_sym, True, None

    def auto_clip_layer(self, block_idx, layer_name, w, inputs, n_grid=20, max_shrink=0.5, n_sample_token=512, eps=0.0):
        raise AssertionError('reference path taken')


def test_patch_autoclipper_contract():
    from lightcompress_b200.algo import patch_autoclipper
    assert patch_autoclipper(_RefLikeAutoClipper) and not patch_autoclipper(_RefLikeAutoClipper)
    c = GOLD[0]['case']
    W, x = inputs(c)
    wq = Q.IntegerQuantizer(c['bit'], c['sym'], 'per_group', group_size=c['gs'])
    ac = _RefLikeAutoClipper(wq, c['clip_sym'])
    bmax, bmin = ac.auto_clip_layer(0, 'fc', W.to(DEV), [x.to(DEV)], n_sample_token=c['tok'])
    assert bmax.shape == GOLD[0]['best_max'].shape and bmax.dtype == W.dtype
    same = (bmax.cpu().view(torch.int16) == GOLD[0]['best_max'].view(torch.int16))
    assert float(same.float().mean()) >= 0.999


def test_row_shards_equal_the_full_search():
    """every (channel, group) is searched independently, so out-channel shards (auto_clip_layer_sharded deals
    them over the ranks) give exactly the rows of the full search."""
    c = dict(oc=256, C=512, tok=64, dtype='bfloat16', seed=41, bit=4, sym=False, gs=128, clip_sym=False)
    W, x = inputs(c)
    wq = Q.IntegerQuantizer(4, False, 'per_group', group_size=128)
    xf = x.reshape(-1, x.shape[-1]).to(DEV)
    full_max, full_min = ops.autoclip_search(W.to(DEV), xf, wq, False)
    for r0 in range(0, 256, 64):
        mx, mn = ops.autoclip_search(W[r0:r0 + 64].contiguous().to(DEV), xf, wq, False)
        assert torch.equal(mx, full_max[r0:r0 + 64]) and torch.equal(mn, full_min[r0:r0 + 64])
