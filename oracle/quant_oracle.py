"""ctypes front-end of oracle/lc_oracle.c operating on CPU torch tensors (torch = memory only).

`OracleQuantizer` restates the call contract of the reference quantizers
(/root/reference/llmc/compression/quantization/quant.py:661-1221) on top of lco_fakequant so that the
parity tests read like reference calls. TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).
"""
import ctypes

import torch

from . import build as _b

_DT = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}
OUT_FAKE, OUT_I8, OUT_U8, OUT_I32, OUT_FP8, OUT_NONE, OUT_CODE_T = range(7)


class Spec(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int32) for n in
                ('dtype', 'math_dtype', 'qparam_dtype', 'kind', 'sym', 'round_zp', 'out_mode', 'qp_math_plus1')] + \
               [(n, ctypes.c_float) for n in ('qmin', 'qmax', 'output_scale_factor', 'clip_ratio')]


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(_b.build())
        vp, i64 = ctypes.c_void_p, ctypes.c_int64
        _lib.lco_fakequant.restype = ctypes.c_int
        _lib.lco_fakequant.argtypes = [vp, vp, vp, vp, i64, i64, i64, i64, ctypes.POINTER(Spec), ctypes.c_int]
        _lib.lco_check_markstein.restype = i64
        _lib.lco_check_markstein.argtypes = [vp, i64, vp, i64]
        _lib.lco_check_magic_round.restype = i64
        _lib.lco_check_magic_round.argtypes = [vp, i64]
        _lib.lco_round_fp8.restype = ctypes.c_float
        _lib.lco_round_fp8.argtypes = [ctypes.c_float, ctypes.c_int]
        _lib.lco_encode_fp8.restype = ctypes.c_int
        _lib.lco_encode_fp8.argtypes = [ctypes.c_float, ctypes.c_int]
    return _lib


def fakequant(x, rows, cols, ur, uc, *, kind=0, sym=True, round_zp=True, qmin=-8.0, qmax=7.0, out_mode=OUT_FAKE,
              math_dtype=None, qp_math=None, dynamic=True, scales=None, zeros=None, osf=1.0, clip=1.0, out_dtype=None):
    """Run lco_fakequant. Returns (out, scales, zeros)."""
    assert x.device.type == 'cpu'
    x = x.contiguous()
    md = math_dtype or x.dtype
    sp = Spec()
    qd = qp_math or md
    sp.dtype, sp.math_dtype, sp.qparam_dtype = _DT[x.dtype], _DT[md], _DT[qd]
    sp.qp_math_plus1 = _DT[qd] + 1
    sp.kind, sp.sym, sp.round_zp, sp.out_mode = kind, int(sym), int(round_zp), out_mode
    sp.qmin, sp.qmax, sp.output_scale_factor, sp.clip_ratio = qmin, qmax, osf, clip
    nunits = (-(-rows // min(ur, max(rows, 1)))) * (-(-cols // min(uc, max(cols, 1)))) if rows and cols else 0
    od = {OUT_FAKE: x.dtype, OUT_CODE_T: x.dtype, OUT_I8: torch.int8, OUT_U8: torch.uint8, OUT_I32: torch.int32,
          OUT_FP8: torch.uint8}.get(out_mode)
    out = torch.empty(x.shape, dtype=out_dtype or od) if out_mode != OUT_NONE else None
    if dynamic:
        scales = torch.empty(nunits, dtype=qd)
        zeros = torch.empty(nunits, dtype=qd) if (not sym and kind == 0) else None
    else:
        scales = scales.to(qd).contiguous()
        zeros = zeros.to(qd).contiguous() if (zeros is not None and not sym and kind == 0) else None
    st = lib().lco_fakequant(x.data_ptr(), out.data_ptr() if out is not None else None, scales.data_ptr(),
                             zeros.data_ptr() if zeros is not None else None, rows, cols, ur, uc, sp, int(dynamic))
    if st == -2:
        raise ValueError('group size does not divide the last dimension')
    assert st == 0, st
    return out, scales, zeros
