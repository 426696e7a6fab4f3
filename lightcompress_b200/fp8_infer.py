"""Block-wise FP8 inference step: the mirror of the reference's Triton module
`llmc/compression/quantization/kernel.py` — `act_quant` (:31-54), `fp8_gemm` (:218-242) — and of
`block_wise_fp8_forward_func` / `LlmcFp8Linear.forward` (module_utils.py:41-46, 244-261), same names, argument
meaning, return shapes / dtypes and assertion behaviour. `weight_cast_to_fp8` / `weight_cast_to_bf16` live in
`quant.py`. Everything runs in liblcb200.so (csrc/fp8_infer.cu); there is no CPU path.
"""
import torch

from . import _cabi
from .quant import _DT, _on_device, _require_cuda, _stream, weight_cast_to_bf16, weight_cast_to_fp8  # noqa: F401

_F8 = torch.float8_e4m3fn


def act_quant(x, block_size=128):
    """kernel.py:31-54 → (e4m3 codes shaped like x, fp32 scales [..., x.size(-1) // block_size])."""
    assert x.is_contiguous(), 'Input tensor must be contiguous'
    assert (
        x.size(-1) % block_size == 0
    ), f'Last dimension size must be divisible by block_size (block_size={block_size})'
    _require_cuda(x, 'act_quant')
    lib = _cabi.load()
    y = torch.empty(x.shape, dtype=torch.uint8, device=x.device)
    s = torch.empty(*x.size()[:-1], x.size(-1) // block_size, dtype=torch.float32, device=x.device)
    with _on_device(x.device):
        st = lib.lcb_act_quant_fp8(x.data_ptr(), y.data_ptr(), s.data_ptr(), s.numel(), block_size, _DT[x.dtype],
                                   _stream(x.device.index))
    _cabi.check(st, 'lcb_act_quant_fp8')
    return y.view(_F8), s


_WS = {}


def _zero_workspace(device, nbytes):
    """The split-K workspace of lcb_fp8_gemm_ws: zero on entry, left zero by the call — one buffer per (device, stream)
    serves every call ordered on that stream (calls on different streams may overlap on the GPU and get their own)."""
    if torch.cuda.is_current_stream_capturing():     # a graph owns its buffers: zeroed by a captured memset, not cached
        return torch.zeros(nbytes, dtype=torch.uint8, device=device)
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    ws = _WS.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = _WS[key] = torch.zeros(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
    return ws


def fp8_gemm(a, a_s, b, b_s, bias=None, out_dtype=None):
    """kernel.py:218-242: c[..., N] = (a * a_s) @ (b * b_s)^T with 128-deep K blocks, in torch.get_default_dtype()
    (or `out_dtype`). `bias` (not a reference argument) fuses the `y += bias` of block_wise_fp8_forward_func."""
    assert a.is_contiguous() and b.is_contiguous(), 'Input tensors must be contiguous'
    assert (
        a_s.is_contiguous() and b_s.is_contiguous()
    ), 'Scaling factor tensors must be contiguous'
    if not a.is_cuda:
        raise RuntimeError(f'fp8_gemm: tensor is on {a.device}; lightcompress_b200 has no CPU path')
    if a.dtype != _F8 or b.dtype != _F8:
        raise TypeError(f'fp8_gemm: operands must be float8_e4m3fn, got {a.dtype} / {b.dtype}')
    if a_s.dtype != torch.float32 or b_s.dtype != torch.float32:
        raise TypeError('fp8_gemm: scaling factors must be float32')
    K = a.size(-1)
    M = a.numel() // K
    N = b.size(0)
    assert b.size(1) == K
    nkb = (K + 127) // 128
    assert a_s.numel() == M * nkb and b_s.numel() == ((N + 127) // 128) * nkb, 'scale shapes do not match 128-blocks'
    dt = out_dtype or torch.get_default_dtype()
    if dt not in _DT:
        raise TypeError(f'fp8_gemm: unsupported output dtype {dt}')
    c = torch.empty(*a.size()[:-1], N, dtype=dt, device=a.device)
    lib = _cabi.load()
    if bias is not None:
        bias = bias.contiguous()
        if bias.dtype not in _DT or bias.numel() != N:
            raise TypeError('fp8_gemm: bias must be a float tensor of N elements')
    # few output tiles (decode shapes): split-K over the idle SMs through an fp32 workspace (lcb_fp8_gemm_ws)
    nws = int(lib.lcb_fp8_gemm_workspace_bytes(M, N, K)) if M <= 256 else 0
    ws = _zero_workspace(a.device, nws) if nws else None
    with _on_device(a.device):
        st = lib.lcb_fp8_gemm_ws(a.data_ptr(), a_s.data_ptr(), b.data_ptr(), b_s.data_ptr(), c.data_ptr(),
                                 bias.data_ptr() if bias is not None else None, M, N, K, _DT[dt],
                                 _DT[bias.dtype] if bias is not None else 0, ws.data_ptr() if ws is not None else None,
                                 _stream(a.device.index))
    _cabi.check(st, 'lcb_fp8_gemm')
    return c


def block_wise_fp8_forward_func(x, w, w_scale, block_size, bias):
    """module_utils.py:41-46: act_quant → fp8_gemm → bf16 (+ bias). Two launches; the bias add rides the GEMM epilogue."""
    assert block_size == 128, 'the reference GEMM kernel fixes the K / N scale block at 128 (kernel.py:143, 196)'
    xq, scale = act_quant(x.contiguous(), block_size)
    return fp8_gemm(xq, scale, w, w_scale, bias=bias, out_dtype=torch.bfloat16)


class LlmcFp8Linear(torch.nn.Module):
    """module_utils.py:223-288: block-scaled e4m3 weight + `weight_scale_inv`; forward = block_wise_fp8_forward_func."""

    def __init__(self, in_features, out_features, bias, block_size):
        super().__init__()
        self.block_size = block_size
        self.in_features = in_features
        self.out_features = out_features
        if bias:
            self.bias = torch.nn.Parameter(torch.empty(out_features))
        else:
            self.register_parameter('bias', None)
        self.weight = torch.nn.Parameter(torch.empty(out_features, in_features, dtype=_F8), requires_grad=False)
        so = (out_features + block_size - 1) // block_size
        si = (in_features + block_size - 1) // block_size
        self.weight_scale_inv = torch.nn.Parameter(torch.empty(so, si, dtype=torch.float32), requires_grad=False)

    @torch.no_grad()
    def forward(self, x):
        if self.weight.dtype == _F8:
            return block_wise_fp8_forward_func(x, self.weight, self.weight_scale_inv, self.block_size, self.bias)
        return torch.nn.functional.linear(x, self.weight, self.bias)

    @classmethod
    @torch.no_grad()
    def new(cls, module, block_size=128):
        """From a float nn.Linear: weight_cast_to_fp8 (quant.py:34-43) supplies codes and scales."""
        m = cls(module.in_features, module.out_features, module.bias is not None, block_size)
        codes, scales = weight_cast_to_fp8(module.weight.data, block_size)
        m.weight = torch.nn.Parameter(codes, requires_grad=False)
        m.weight_scale_inv = torch.nn.Parameter(scales.float().reshape(m.weight_scale_inv.shape), requires_grad=False)
        if module.bias is not None:
            m.bias = torch.nn.Parameter(module.bias.data.clone(), requires_grad=False)
        return m

    def extra_repr(self):
        return (f'in_features={self.in_features}, out_features={self.out_features}, bias={self.bias is not None}, '
                f'block_size={self.block_size}')
