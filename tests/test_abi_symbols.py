"""CPU: liblcb200.so loads and exports every symbol include/lcb200.h declares (no compute calls)."""
import ctypes
import os
import re

from lightcompress_b200 import _build, _cabi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, 'include', 'lcb200.h')).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return sorted(set(re.findall(r'\b(lcb_[a-z0-9_]+)\s*\(', src)))


def test_header_symbols_are_exported_and_bound():
    _build.build()
    lib = ctypes.CDLL(_build.LIB)
    syms = declared_symbols()
    assert len(syms) >= 12
    for s in syms:
        assert hasattr(lib, s), f'{s} declared in lcb200.h but not exported'
        assert s in _cabi.SIGNATURES, f'{s} has no ctypes signature'
    assert set(_cabi.SIGNATURES) == set(syms)


def test_abi_version_and_struct_layout():
    lib = _cabi.load()
    assert lib.lcb_abi_version() == 1
    assert ctypes.sizeof(_cabi.QuantSpec) == 48
    assert lib.lcb_fakequant_workspace_bytes() > 0


def test_sass_is_sm100a_only():
    import subprocess
    out = subprocess.run(['cuobjdump', '-lelf', _build.LIB], capture_output=True, text=True).stdout
    archs = set(re.findall(r'sm_(\d+a?)', out))
    assert archs == {'100a'}, archs


def test_every_entry_point_cites_the_reference_interface_it_replaces():
    """The boundary rule: each declaration in include/lcb200.h sits under a comment that cites the reference file:line it
    replaces, or says that it has no counterpart in the reference."""
    src = open(os.path.join(ROOT, 'include', 'lcb200.h')).read()
    comments = [(m.end(), m.group(1)) for m in re.finditer(r'/\*(.*?)\*/', src, flags=re.S)]
    blank = re.sub(r'/\*.*?\*/', lambda m: ' ' * len(m.group(0)), src, flags=re.S)
    missing = []
    for m in re.finditer(r'\b(lcb_[a-z0-9_]+)\s*\(', blank):
        before = [c for end, c in comments if end <= m.start()]
        text = before[-1] if before else ''
        if not (re.search(r'\.py:\d+', text) or 'no counterpart in the reference' in text.lower()):
            missing.append(m.group(1))
    assert not missing, missing


def test_new_entry_points_validate_their_arguments_before_touching_the_device():
    """Argument checks of the round-2 entries return a status without a CUDA call (safe on a machine without a GPU)."""
    lib = _cabi.load()
    vp = ctypes.c_void_p
    one = ctypes.create_string_buffer(64)
    p = ctypes.cast(one, vp)
    assert lib.lcb_floatquant_elem(None, p, None, None, 4, 8, 4, 3, 2, None) == -1          # no input
    assert lib.lcb_floatquant_elem(p, None, None, None, 4, 8, 4, 3, 2, None) == -1          # nothing to produce
    assert lib.lcb_floatquant_elem(p, p, None, None, 4, 8, 0, 3, 2, None) == -1             # e_bits out of range
    assert lib.lcb_floatquant_elem(p, p, None, None, 4, 8, 4, 3, 7, None) == -3             # unknown dtype
    assert lib.lcb_floatquant_elem(p, p, None, None, 0, 8, 4, 3, 2, None) == 0              # empty tensor
    assert lib.lcb_floatquant_elem_static(p, None, p, 8, 2, 2, None) == -1
    assert lib.lcb_floatquant_elem_static(p, p, p, 0, 2, 2, None) == 0
    assert lib.lcb_sqdiff_sum(None, p, p, 8, 2, None) == -1 and lib.lcb_sqdiff_sum(p, p, p, 0, 2, None) == 0
    assert lib.lcb_hessian_accum_f32_noalloc(None, p, 8, 8, None) == -1
    assert lib.lcb_hessian_accum_f32_noalloc(p, p, 0, 8, None) == 0
    assert lib.lcb_hessian_accum_f32_noalloc(p, p, 1025, 4096, None) == -2                   # beyond the static scratch
    assert b'LCB_HESSIAN_NOALLOC_MAX_ELEMS' in lib.lcb_last_error()
    assert lib.lcb_fp8_gemm_ws(None, p, p, p, p, None, 8, 8, 128, 2, 0, None, None) == -1
    assert lib.lcb_fp8_gemm_ws(p, p, p, p, p, None, 0, 8, 128, 2, 0, None, None) == 0        # M == 0
    assert lib.lcb_fakequant_block_fp8_multi(None, None, None, None, None, 0, None, None) == -1
