"""ctypes binding of include/lcb200.h — the only way Python reaches the CUDA kernels.

There is NO fallback: if liblcb200.so is missing or a call fails this module raises. Status codes are
mapped onto the exception types the reference raises for the same condition.
"""
import ctypes
import os

from . import _build

LCB_F32, LCB_F16, LCB_BF16 = 0, 1, 2
KIND_INT, KIND_FP8_E4M3, KIND_FP8_E5M2 = 0, 1, 2
OUT_FAKE, OUT_I8, OUT_U8, OUT_I32, OUT_FP8, OUT_NONE, OUT_CODE_T, OUT_DEQUANT = range(8)

_c_i64, _c_i32, _c_vp, _c_f = ctypes.c_int64, ctypes.c_int32, ctypes.c_void_p, ctypes.c_float


class QuantSpec(ctypes.Structure):
    _fields_ = [
        ('dtype', _c_i32), ('math_dtype', _c_i32), ('qparam_dtype', _c_i32), ('kind', _c_i32),
        ('sym', _c_i32), ('round_zp', _c_i32), ('out_mode', _c_i32), ('qp_math_plus1', _c_i32),
        ('qmin', _c_f), ('qmax', _c_f), ('output_scale_factor', _c_f), ('clip_ratio', _c_f),
    ]


# name -> (restype, argtypes); must list every symbol include/lcb200.h declares
SIGNATURES = {
    'lcb_abi_version': (ctypes.c_int, []),
    'lcb_last_error': (ctypes.c_char_p, []),
    'lcb_fakequant': (ctypes.c_int, [_c_vp, _c_vp, _c_vp, _c_vp, _c_i64, _c_i64, _c_i64, _c_i64,
                                     ctypes.POINTER(QuantSpec), ctypes.c_int, _c_vp, _c_vp]),
    'lcb_fakequant_workspace_bytes': (_c_i64, []),
    'lcb_fakequant_plan_create': (_c_vp, [_c_i64, _c_i64, _c_i64, _c_i64, ctypes.POINTER(QuantSpec), ctypes.c_int]),
    'lcb_fakequant_plan_run': (ctypes.c_int, [_c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_vp]),
    'lcb_fakequant_plan_destroy': (None, [_c_vp]),
    'lcb_minmax': (ctypes.c_int, [_c_vp, _c_vp, _c_vp, _c_i64, _c_i64, _c_i64, _c_i64, _c_i32, _c_i32, _c_i32, _c_vp, _c_vp]),
    'lcb_qparams_from_range': (ctypes.c_int, [_c_vp, _c_vp, _c_vp, _c_vp, _c_i64, ctypes.POINTER(QuantSpec), _c_vp]),
    'lcb_mse_range': (ctypes.c_int, [_c_vp, _c_vp, _c_vp, _c_i64, _c_i64, ctypes.POINTER(QuantSpec), _c_i32,
                                     ctypes.c_double, ctypes.c_float, _c_vp]),
    'lcb_fakequant_multi': (ctypes.c_int, [_c_vp, _c_vp, _c_vp, _c_vp, _c_i32, _c_i64, ctypes.POINTER(QuantSpec), _c_vp]),
    'lcb_fakequant_block_fp8_multi': (ctypes.c_int, [_c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_i32, ctypes.POINTER(QuantSpec), _c_vp]),
    'lcb_fp8_block_dequant': (ctypes.c_int, [_c_vp, _c_vp, _c_vp, _c_i64, _c_i64, _c_i64, _c_i32, _c_i32, _c_vp]),
    'lcb_act_quant_fp8': (ctypes.c_int, [_c_vp, _c_vp, _c_vp, _c_i64, _c_i64, _c_i32, _c_vp]),
    'lcb_fp8_gemm': (ctypes.c_int, [_c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_i64, _c_i64, _c_i64, _c_i32, _c_i32, _c_vp]),
    'lcb_fp8_gemm_ws': (ctypes.c_int, [_c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_i64, _c_i64, _c_i64, _c_i32, _c_i32, _c_vp, _c_vp]),
    'lcb_fp8_gemm_workspace_bytes': (_c_i64, [_c_i64, _c_i64, _c_i64]),
    'lcb_hessian_accum': (ctypes.c_int, [_c_vp, _c_vp, _c_i64, _c_i64, _c_i32, _c_vp, _c_vp]),
    'lcb_hessian_workspace_bytes': (_c_i64, [_c_i64, _c_i64, _c_i32]),
    'lcb_floatquant_elem': (ctypes.c_int, [_c_vp, _c_vp, _c_vp, _c_vp, _c_i64, _c_i64, _c_i32, _c_i32, _c_i32, _c_vp]),
    'lcb_floatquant_elem_static': (ctypes.c_int, [_c_vp, _c_vp, _c_vp, _c_i64, _c_i32, _c_i32, _c_vp]),
    'lcb_sqdiff_sum': (ctypes.c_int, [_c_vp, _c_vp, _c_vp, _c_i64, _c_i32, _c_vp]),
    'lcb_hessian_accum_f32_noalloc': (ctypes.c_int, [_c_vp, _c_vp, _c_i64, _c_i64, _c_vp]),
    'lcb_hessian_finalize': (ctypes.c_int, [_c_vp, _c_i64, ctypes.c_double, _c_vp]),
    'lcb_hessian_packed_elems': (_c_i64, [_c_i64]),
    'lcb_hessian_pack': (ctypes.c_int, [_c_vp, _c_vp, _c_i64, ctypes.c_double, _c_vp]),
    'lcb_hessian_unpack_finalize': (ctypes.c_int, [_c_vp, _c_vp, _c_i64, _c_vp]),
    'lcb_gptq_weight_transform': (ctypes.c_int, [_c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_i64,
                                                 _c_i64, _c_i64, _c_i32, _c_i32, _c_i32,
                                                 ctypes.POINTER(QuantSpec), _c_vp, _c_vp]),
    'lcb_gptq_workspace_bytes': (_c_i64, [_c_i64, _c_i64, _c_i32]),
    'lcb_cholinv_upper': (ctypes.c_int, [_c_vp, _c_i64, _c_i64, _c_f, _c_i32, _c_vp, _c_vp, _c_vp]),
    'lcb_cholinv_workspace_bytes': (_c_i64, [_c_i64, _c_i32]),
    'lcb_scale_cols': (ctypes.c_int, [_c_vp, _c_vp, _c_vp, _c_i64, _c_i64, _c_i32, _c_i32, _c_vp]),
    'lcb_linear_mse': (ctypes.c_int, [_c_vp, _c_vp, _c_vp, _c_vp, _c_i64, _c_i64, _c_i64, _c_i32, _c_vp]),
    'lcb_linear': (ctypes.c_int, [_c_vp, _c_vp, _c_vp, _c_vp, _c_i64, _c_i64, _c_i64, _c_i32, _c_vp]),
    'lcb_awq_candidates': (ctypes.c_int, [_c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_i64, _c_i64, _c_i64, _c_i32,
                                          _c_i32, ctypes.POINTER(QuantSpec), _c_vp, _c_vp]),
    'lcb_awq_workspace_bytes': (_c_i64, [_c_i64, _c_i64, _c_i64, _c_i32]),
    'lcb_autoclip_search': (ctypes.c_int, [_c_vp, _c_vp, _c_vp, _c_vp, _c_i64, _c_i64, _c_i64, _c_i64, _c_i32, _c_i32,
                                           _c_i32, _c_i32, ctypes.POINTER(QuantSpec), _c_vp, _c_vp]),
    'lcb_autoclip_workspace_bytes': (_c_i64, [_c_i64, _c_i64, _c_i64, _c_i64]),
    'lcb_autoclip_init': (ctypes.c_int, [_c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_i64, _c_i64, _c_i32, _c_i32, _c_i32, _c_vp]),
    'lcb_autoclip_step': (ctypes.c_int, [_c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_i64, _c_i64, _c_i64, _c_i64,
                                         _c_i32, _c_f, _c_i32, _c_i32, _c_i32, _c_vp]),
    'lcb_pack_rows': (ctypes.c_int, [_c_vp, _c_vp, _c_i64, _c_i64, _c_i32, _c_i32, _c_vp]),
    'lcb_awq_gemm_pack': (ctypes.c_int, [_c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_i64, _c_i64, _c_i32, _c_i32, _c_vp]),
    'lcb_col_absstat': (ctypes.c_int, [_c_vp, _c_vp, _c_i64, _c_i64, _c_i32, _c_i32, _c_vp]),
}

_lib = None


class LcbError(RuntimeError):
    pass


def lib_path():
    return _build.LIB


def load():
    """dlopen liblcb200.so and bind every symbol. Raises if the extension was not built."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise LcbError(
            f'{path} is missing: the CUDA extension was not built. Run `python -m lightcompress_b200._build` '
            '(or __graft_entry__.build()). There is no CPU fallback.')
    lib = ctypes.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    if lib.lcb_abi_version() != 1:
        raise LcbError('liblcb200.so ABI version mismatch')
    _lib = lib
    return lib


_EXC = {-1: AssertionError, -2: ValueError, -3: NotImplementedError, -4: RuntimeError, -5: ValueError}


def check(status, what):
    if status == 0:
        return
    detail = load().lcb_last_error().decode() if status == -4 else ''
    raise _EXC.get(status, RuntimeError)(f'{what} failed with lcb status {status} {detail}')
