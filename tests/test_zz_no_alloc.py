"""GPU: lcb_hessian_accum_f32_noalloc (scratch in a static __device__ array, no allocation inside the call) against the
existing lcb_hessian_accum entry with a caller workspace, both through the C ABI. Collected last on purpose.
Tolerance: the two entries launch the same kernels; the fp32 TMA reduce-adds of split-K tiles may
land in a different order, which test_hessian_gpu.py bounds by 1e-6 of the matrix scale — same bar here, plus the 2e-5
bar against an fp64 product."""
import ctypes

import pytest
import torch

from lightcompress_b200 import _cabi

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'
MAX_ELEMS = 4 << 20     # LCB_HESSIAN_NOALLOC_MAX_ELEMS (include/lcb200.h)
LCB_F32 = 0


def _acts(T, C, seed):
    g = torch.Generator().manual_seed(seed)
    gain = torch.exp(torch.randn(C, generator=g))
    return (torch.randn(T, C, generator=g) * gain).to(DEV)


def _rel(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream(torch.device(DEV)).cuda_stream)


def _accum_ws(lib, S, x):
    T, C = x.shape
    ws = torch.empty(int(lib.lcb_hessian_workspace_bytes(T, C, LCB_F32)), dtype=torch.uint8, device=DEV)
    assert lib.lcb_hessian_accum(S.data_ptr(), x.data_ptr(), T, C, LCB_F32, ws.data_ptr(), _stream()) == 0
    torch.cuda.synchronize()


def _accum_noalloc(lib, S, x):
    T, C = x.shape
    st = lib.lcb_hessian_accum_f32_noalloc(S.data_ptr(), x.data_ptr(), T, C, _stream())
    assert st == 0, (st, lib.lcb_last_error().decode())
    torch.cuda.synchronize()


def _upper_blocks(S):
    """The contract of lcb_hessian_accum: only 128x128 blocks (i, j >= i) are written."""
    C = S.shape[0]
    blk = torch.arange(C, device=S.device) // 128
    return torch.where(blk[None, :] >= blk[:, None], S, torch.zeros_like(S))


@pytest.mark.parametrize('shapes', [((768, 1024), (300, 1024)),      # second call smaller: stale scratch beyond it
                                    ((1024, 4096), (64, 4096)),     # first call exactly at the limit
                                    ((96, 576), (2048, 576)),
                                    ((40, 100), (200, 100))])       # C % 8 != 0: SIMT path, scratch unused
def test_noalloc_matches_workspace_entry_two_calls_in_a_row(shapes):
    lib = _cabi.load()
    with torch.cuda.device(DEV):
        C = shapes[0][1]
        xs = [_acts(T, C, 31 + i) for i, (T, _) in enumerate(shapes)]
        S_ref = torch.zeros(C, C, device=DEV)
        S_new = torch.zeros(C, C, device=DEV)
        exact = torch.zeros(C, C, device=DEV, dtype=torch.float64)
        for i, x in enumerate(xs):
            _accum_ws(lib, S_ref, x)
            torch.cuda.synchronize()
            free0 = torch.cuda.mem_get_info()[0]
            tmem0 = torch.cuda.memory_allocated()
            _accum_noalloc(lib, S_new, x)
            if i > 0:   # the first call of a process may load the module (lazy loading); after that nothing may move
                assert torch.cuda.memory_allocated() == tmem0
                assert torch.cuda.mem_get_info()[0] >= free0, 'device memory shrank across an allocation-free call'
            exact += x.double().t() @ x.double()
            a, b = _upper_blocks(S_new), _upper_blocks(S_ref)
            assert _rel(a, b) < 1e-6, f'call {i}: no-alloc entry differs from lcb_hessian_accum'
            assert _rel(a, _upper_blocks(exact)) < 2e-5, f'call {i}: no-alloc entry differs from the fp64 product'


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    lib = _cabi.load()
    with torch.cuda.device(DEV):
        T, C = 1025, 4096
        assert T * C > MAX_ELEMS
        x = torch.ones(T, C, device=DEV)
        S = torch.zeros(C, C, device=DEV)
        st = lib.lcb_hessian_accum_f32_noalloc(S.data_ptr(), x.data_ptr(), T, C, _stream())
        assert st == -2     # LCB_E_SHAPE
        msg = lib.lcb_last_error().decode()
        assert 'LCB_HESSIAN_NOALLOC_MAX_ELEMS' in msg and '4194304' in msg
        with pytest.raises(ValueError):
            _cabi.check(st, 'lcb_hessian_accum_f32_noalloc')
        torch.cuda.synchronize()
        assert float(S.abs().max()) == 0.0, 'a rejected call must launch nothing'
        # the entry still works afterwards
        x2 = _acts(128, 256, 7)
        S1 = torch.zeros(256, 256, device=DEV)
        S2 = torch.zeros(256, 256, device=DEV)
        _accum_ws(lib, S1, x2)
        _accum_noalloc(lib, S2, x2)
        assert _rel(_upper_blocks(S2), _upper_blocks(S1)) < 1e-6
        assert lib.lcb_hessian_accum_f32_noalloc(S1.data_ptr(), x2.data_ptr(), 0, 256, _stream()) == 0
        assert lib.lcb_hessian_accum_f32_noalloc(None, x2.data_ptr(), 8, 256, _stream()) == -1
