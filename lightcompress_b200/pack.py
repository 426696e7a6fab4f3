"""Real-quant export packers on the GPU, mirroring the reference classmethods:

    vllm_pack(weight_codes, scales, bit)                   VllmRealQuantLinear.pack        module_utils.py:931-955
    awq_gemm_pack(weight, scales, zeros, bit, group_size)  AutoawqRealQuantLinear.gemm_pack module_utils.py:1099-1158

The reference packs with numpy on the host, or with a Python loop over input channels; here each is one kernel
(csrc/pack.cu). No CPU path: tensors must live on a CUDA device.
"""
import torch

from . import _cabi
from .quant import _DT, _on_device, _stream


def _need_cuda(*ts):
    for t in ts:
        if not t.is_cuda:
            raise RuntimeError(f'lightcompress_b200.pack: tensor on {t.device}; the sm_100a extension is the only implementation')


def vllm_pack(weight, scales, bit):
    """weight: integer codes [R, C] (int8 / int32, signed: the symmetric real_quant output). Returns (int32 [R, ceil(C /
    (32 / bit))], scales.to(float16))."""
    _need_cuda(weight)
    if weight.dtype not in (torch.int8, torch.int32):
        raise TypeError(f'vllm_pack: codes must be int8 or int32, got {weight.dtype}')
    w = weight.contiguous()
    R, C = w.shape
    pf = 32 // int(bit)
    out = torch.empty((R, -(-C // pf)), dtype=torch.int32, device=w.device)
    lib = _cabi.load()
    with _on_device(w.device):
        st = lib.lcb_pack_rows(w.data_ptr(), out.data_ptr(), R, C, w.element_size(), int(bit), _stream(w.device.index))
    _cabi.check(st, 'lcb_pack_rows')
    return out, scales.to(torch.float16)


def awq_gemm_pack(weight, scales, zeros, bit, group_size):
    """weight [out, in] (the float weight of the layer), scales / zeros [out, in / group] as real_quant_weight_dynamic
    returns them. Returns (qweight int32 [in, out / 8], scales fp16 [in / group, out], qzeros int32 [in / group, out / 8])."""
    if int(bit) != 4:
        raise NotImplementedError('Only 4-bit are supported for now.')   # module_utils.py:1134
    assert scales is not None and zeros is not None
    _need_cuda(weight, scales, zeros)
    sc = scales.t().contiguous().to(torch.float16)
    zp = zeros.t().contiguous().to(torch.int32)
    w = weight.contiguous()
    R, C = w.shape
    G = C // int(group_size)
    if sc.shape != (G, R) or zp.shape != (G, R):
        raise ValueError(f'awq_gemm_pack: scales / zeros must be [{R}, {G}], got {tuple(scales.shape)} / {tuple(zeros.shape)}')
    qweight = torch.empty((C, R // 8), dtype=torch.int32, device=w.device)
    qzeros = torch.empty((G, R // 8), dtype=torch.int32, device=w.device)
    lib = _cabi.load()
    with _on_device(w.device):
        st = lib.lcb_awq_gemm_pack(w.data_ptr(), sc.data_ptr(), zp.data_ptr(), qweight.data_ptr(), qzeros.data_ptr(), R, C,
                                   int(group_size), _DT[w.dtype], _stream(w.device.index))
    _cabi.check(st, 'lcb_awq_gemm_pack')
    return qweight, sc, qzeros
