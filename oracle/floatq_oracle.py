"""TEST INFRASTRUCTURE ONLY — never imported by lightcompress_b200/ or by a timed region.

CPU restatement of FloatQuantizer WITHOUT qtorch (per-element scales): get_float_qparams (quant.py:1005-1027) and
quant / dequant (quant.py:1061-1076), one statement per torch op of the reference, each op = fp32 compute + one rounding
to the chain's dtype (the tensor dtype when e_bits < 5, fp32 when e_bits >= 5 or for fp32 tensors — `maxval.to(float32)`,
quant.py:1016-1017, promotes everything downstream). The reference's 0-dim CUDA operands `e_bits` / `m_bits` are cast to
the op's dtype when TensorIterator loads them; Python scalars enter as fp32.

Pinned by tests/golden/floatq_golden.pt.gz — outputs of the UNMODIFIED reference class, which only runs on a CUDA device
(`.cuda()` inside get_float_qparams), generated on the B200 box by tests/golden/make_floatq_golden.py. The CPU's libm
log2f / powf and the GPU's libdevice routines may differ in the last bit of an fp32 result; after the chain's roundings that
shows on a handful of elements per million (tests/test_floatq_oracle.py states the bound), which is why the CUDA kernel is
pinned against the reference directly (tests/test_floatq_gpu.py) and this restatement is the `-m "not gpu"` cross-check.
"""
import torch


def _rd(t, dt):
    return t.to(dt).float()


def fake_quant(x, unit, e_bits, m_bits):
    """x: tensor whose last dimension is split into units of `unit` elements (per_channel / per_token: unit = last dim).
    Returns (fake-quantized tensor in x.dtype, xc, scales) with xc / scales in the chain's dtype, shaped [n_units, unit]."""
    org = x.dtype
    dt = torch.float32 if (e_bits >= 5 or org == torch.float32) else org
    t = x.reshape(-1, unit).float()                      # exact image of the tensor
    mx, mn = t.amax(dim=-1, keepdim=True), t.amin(dim=-1, keepdim=True)
    maxval = torch.maximum(mx, -mn)
    two_e = torch.tensor(2.0).pow(torch.tensor(float(e_bits)))
    c = torch.log2(2.0 - torch.tensor(2.0).pow(torch.tensor(-float(m_bits))))
    c = _rd(c, dt)                                       # loaded as an operand of a `dt` op
    bias = _rd(two_e - _rd(torch.log2(maxval), dt), dt)
    bias = _rd(bias + c, dt)
    bias = _rd(bias - 1.0, dt)
    xc = torch.minimum(torch.maximum(t, -maxval), maxval)
    ls = torch.floor(_rd(_rd(torch.log2(xc.abs()), dt) + bias, dt))
    ls = torch.clamp(ls, min=1.0)
    e = _rd(_rd(ls - float(m_bits), dt) - bias, dt)
    s = _rd(torch.pow(torch.tensor(2.0), e), dt)
    s = torch.where(s == 0, torch.ones_like(s), s)
    q = torch.round(_rd(_rd(xc / s, dt) + 0.0, dt))
    y = _rd(_rd(q - 0.0, dt) * s, dt)
    return y.to(org).reshape(x.shape), xc.to(dt), s.to(dt)
