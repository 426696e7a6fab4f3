"""TEST INFRASTRUCTURE — CPU restatement of calib_algo='mse' (BaseQuantizer.get_mse_range, quant.py:145-203).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this. Pinned to outputs of the unmodified
reference (tests/golden/mse_golden.pt.gz, made by tests/golden/make_mse_golden.py): tests/test_mse_oracle.py.

The search, per quantization unit (row of the reshaped tensor), all in fp32:
    for i in range(int(maxshrink * mse_grid)):  p = 1 - i / mse_grid
        (xmin, xmax) = p * (min, max);  err = sum |quant_dequant(x; qparams(xmin, xmax)) - x| ** 2.4
        if err < best: best = err; keep (xmin, xmax)
One property of the reference is part of the contract: `best_min_val` / `best_max_val` ALIAS the min / max the next
candidate is derived from (quant.py:164, 193-195 write through views of `min_val`), so after an improvement at step i the
following candidates shrink the already-shrunk range: xmin_{j} = p_j * xmin_best, not p_j * min.
"""
import torch


def qparams(mn, mx, kind, sym, qmin, qmax, round_zp=True):
    """get_qparams (quant.py:545-559) on fp32 ranges. kind 0 = integer, else float (sym only)."""
    if sym or kind != 0:
        a = torch.max(mx.abs(), mn.abs()).clamp(min=1e-5)
        return a / qmax, torch.zeros_like(a)
    s = (mx - mn).clamp(min=1e-5) / (qmax - qmin)
    if round_zp:
        z = (qmin - torch.round(mn / s)).clamp(qmin, qmax)
    else:
        z = qmin - (mn / s)
    return s, z


def quant_dequant(x, s, z, kind, qmin, qmax, round_zp=True):
    """IntegerQuantizer.quant_dequant (quant.py:699-717) / FloatQuantizer with use_qtorch (:1061-1081), fp32."""
    if kind == 0:
        if round_zp:
            q = torch.clamp(torch.round(x / s) + z, qmin, qmax)
        else:
            q = torch.clamp(torch.round(x / s.clamp_min(1e-9) + z), qmin, qmax)
        return (q - z) * s
    s = torch.where(s == 0, torch.ones_like(s), s)
    dt = torch.float8_e4m3fn if kind == 1 else torch.float8_e5m2
    return ((x / s + 0.0).float().to(dt).float()) * s


def get_mse_range(t, kind, sym, qmin, qmax, round_zp=True, maxshrink=0.8, mse_grid=100, norm=2.4):
    """t: [units, unit_len] any float dtype. Returns (min, max) fp32 [units, 1]."""
    t = t.float()
    mn = t.amin(dim=-1, keepdim=True)
    mx = t.amax(dim=-1, keepdim=True)
    best = torch.full([t.shape[0]], float('inf'))
    for i in range(int(maxshrink * mse_grid)):
        p = 1 - i / mse_grid
        xmin, xmax = p * mn, p * mx
        s, z = qparams(xmin, xmax, kind, sym, qmin, qmax, round_zp)
        q = quant_dequant(t, s, z, kind, qmin, qmax, round_zp)
        err = (q - t).abs_().pow_(norm).sum(1)
        better = err < best
        if bool(better.any()):
            best[better] = err[better]
            mn[better] = xmin[better]      # aliasing of the reference: the next candidates start from the kept range
            mx[better] = xmax[better]
    return mn, mx


def fake_quant(x2d_units, kind, sym, qmin, qmax, round_zp=True, **kw):
    """fake_quant_weight_dynamic with calib_algo='mse' on a [units, unit_len] view: fp32 qparams -> the quant-dequant
    promotes to fp32 -> result cast back to the tensor dtype. Returns (fake, scales, zeros, min, max)."""
    mn, mx = get_mse_range(x2d_units, kind, sym, qmin, qmax, round_zp, **kw)
    s, z = qparams(mn, mx, kind, sym, qmin, qmax, round_zp)
    out = quant_dequant(x2d_units.float() if kind == 0 else x2d_units, s, z, kind, qmin, qmax, round_zp)
    return out.to(x2d_units.dtype), s, z, mn, mx
