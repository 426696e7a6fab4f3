"""Thin torch-facing wrappers of the non-quantizer C-ABI entry points (include/lcb200.h).

torch is used for device memory and the current stream only. Every function fails loudly when the extension is
missing or the tensors are not on a CUDA device: there is no CPU path.
"""
import torch

from . import _cabi
from .quant import _DT, _on_device, _stream


def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError(f'lightcompress_b200: tensor on {t.device}; the sm_100a extension is the only implementation')


def hessian_accum(S, x):
    """S += x^T x on tensor cores (upper 128x128 blocks only; see lcb_hessian_accum).

    S: fp32 [C, C] contiguous, zero-initialised before the first call. x: [..., C] f32 / f16 / bf16, any number of
    tokens. Replaces the matmul of GPTQ.add_batch (gptq.py:283-290).
    """
    _need_cuda(S, x)
    assert S.dtype == torch.float32 and S.is_contiguous() and S.dim() == 2 and S.shape[0] == S.shape[1]
    C = S.shape[0]
    x2 = x.reshape(-1, x.shape[-1])
    assert x2.shape[1] == C, f'input has {x2.shape[1]} channels, S is {C}x{C}'
    if not x2.is_contiguous():
        x2 = x2.contiguous()
    lib = _cabi.load()
    nbytes = int(lib.lcb_hessian_workspace_bytes(x2.shape[0], C, _DT[x2.dtype]))
    ws = torch.empty(nbytes, dtype=torch.uint8, device=S.device) if nbytes else None
    with _on_device(S.device):
        st = lib.lcb_hessian_accum(S.data_ptr(), x2.data_ptr(), x2.shape[0], C, _DT[x2.dtype],
                                   ws.data_ptr() if ws is not None else None, _stream(S.device.index))
    _cabi.check(st, 'lcb_hessian_accum')
    return S


def hessian_finalize(S, nsamples):
    """In place: H = (2 / nsamples) * sym(S) — the reference's running-average Hessian (gptq.py:283-290)."""
    _need_cuda(S)
    lib = _cabi.load()
    with _on_device(S.device):
        st = lib.lcb_hessian_finalize(S.data_ptr(), S.shape[0], 2.0 / float(nsamples), _stream(S.device.index))
    _cabi.check(st, 'lcb_hessian_finalize')
    return S


def col_absstat(x, out, mean=False):
    """out[c] = max(out[c], max_r |x[r, c]|) (or += sum_r |x[r, c]| when mean). x: [..., C]; out: fp32 [C]."""
    _need_cuda(x, out)
    x2 = x.reshape(-1, x.shape[-1])
    if not x2.is_contiguous():
        x2 = x2.contiguous()
    assert out.dtype == torch.float32 and out.is_contiguous() and out.numel() == x2.shape[1]
    lib = _cabi.load()
    with _on_device(x.device):
        st = lib.lcb_col_absstat(x2.data_ptr(), out.data_ptr(), x2.shape[0], x2.shape[1], _DT[x2.dtype],
                                 1 if mean else 0, _stream(x.device.index))
    _cabi.check(st, 'lcb_col_absstat')
    return out
