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


def hessian_pack(S, count):
    """The maintained half of S (upper 128-blocks) + this rank's batch count as one contiguous fp32 buffer: the payload of
    the single all-reduce of a token-sharded Hessian (lcb_hessian_pack)."""
    _need_cuda(S)
    lib = _cabi.load()
    C = S.shape[0]
    packed = torch.empty(int(lib.lcb_hessian_packed_elems(C)), dtype=torch.float32, device=S.device)
    with _on_device(S.device):
        st = lib.lcb_hessian_pack(S.data_ptr(), packed.data_ptr(), C, float(count), _stream(S.device.index))
    _cabi.check(st, 'lcb_hessian_pack')
    return packed


def hessian_unpack_finalize(S, packed):
    """In place: S <- the all-reduced sums, then H = (2 / total count) * sym(S), the count read on the device."""
    _need_cuda(S, packed)
    lib = _cabi.load()
    with _on_device(S.device):
        st = lib.lcb_hessian_unpack_finalize(S.data_ptr(), packed.data_ptr(), S.shape[0], _stream(S.device.index))
    _cabi.check(st, 'lcb_hessian_unpack_finalize')
    return S


def cholinv_upper(H, percdamp=0.0, splits=3, check=True):
    """In place: H <- upper Cholesky factor of inverse(H + percdamp * mean(diag H) * I) — the damping plus
    `cholesky -> cholesky_inverse -> cholesky(upper=True)` of GPTQ.process_hessian_and_weights (gptq.py:165-174) as one
    blocked reverse Cholesky + triangular inverse on tcgen05 (lcb_cholinv_upper). H: fp32 [n, n] contiguous, SPD.
    Raises torch.linalg.LinAlgError like the reference when a pivot is not positive (check=True costs one host sync)."""
    _need_cuda(H)
    assert H.dtype == torch.float32 and H.dim() == 2 and H.shape[0] == H.shape[1] and H.is_contiguous()
    n = H.shape[0]
    lib = _cabi.load()
    ws = torch.empty(int(lib.lcb_cholinv_workspace_bytes(n, int(splits))), dtype=torch.uint8, device=H.device)
    info = torch.zeros(1, dtype=torch.int32, device=H.device)
    with _on_device(H.device):
        st = lib.lcb_cholinv_upper(H.data_ptr(), n, n, float(percdamp), int(splits), ws.data_ptr(), info.data_ptr(),
                                   _stream(H.device.index))
    _cabi.check(st, 'lcb_cholinv_upper')
    if check:
        bad = int(info.item())
        if bad:
            raise torch.linalg.LinAlgError(
                f'linalg.cholesky: The factorization could not be completed because the input is not positive-definite '
                f'(the leading minor of order {bad} is not positive-definite).')
    return H


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


def gptq_weight_transform(W, Hinv, blocksize, group_size, qmin, qmax, sym, round_zp=True, n_nonout=None,
                          static_scales=None, static_zeros=None, col_group=None):
    """GPTQ.weight_transform (gptq.py:199-244) on the GPU. W fp32 [R, C] is updated in place like the reference's W
    (trailing columns receive the lazy updates); Hinv is the upper Cholesky factor of the inverse damped Hessian.

    group_size > 0 and no static_scales: per-group qparams are searched on the fly (gptq.py:215-223) and returned as
    [ceil(n_nonout / group_size), R]. With static_scales [G, R] (+ static_zeros, col_group int32 [C] or None = col //
    group_size, or group 0 when group_size <= 0) they are inputs. Returns (tmp, Losses, scales, zeros|None).
    """
    _need_cuda(W, Hinv, static_scales, static_zeros, col_group)
    assert W.dtype == torch.float32 and W.is_contiguous() and Hinv.dtype == torch.float32
    Hinv = Hinv.contiguous()   # torch.linalg.cholesky(upper=True) returns a transposed view
    R, C = W.shape
    n_nonout = C if n_nonout is None else int(n_nonout)
    lib = _cabi.load()
    spec = _cabi.QuantSpec()
    spec.dtype = spec.math_dtype = spec.qparam_dtype = _cabi.LCB_F32
    spec.kind, spec.sym, spec.round_zp, spec.out_mode = _cabi.KIND_INT, int(bool(sym)), int(bool(round_zp)), _cabi.OUT_FAKE
    spec.qmin, spec.qmax, spec.output_scale_factor, spec.clip_ratio = float(qmin), float(qmax), 1.0, 1.0
    static = static_scales is not None
    if static:
        gs = static_scales.to(torch.float32).contiguous()
        gz = static_zeros.to(torch.float32).contiguous() if (static_zeros is not None and not sym) else None
        if not sym and gz is None:
            raise ValueError('asymmetric static groups need static_zeros')
        cg = col_group.to(torch.int32).contiguous() if col_group is not None else None
    else:
        if group_size <= 0:
            raise ValueError('per-channel GPTQ needs the layer qparams (static_scales)')
        G = -(-n_nonout // group_size)
        gs = torch.empty((G, R), dtype=torch.float32, device=W.device)
        gz = None if sym else torch.empty((G, R), dtype=torch.float32, device=W.device)
        cg = None
    tmp = torch.zeros_like(W)
    losses = torch.zeros_like(W)
    ws = torch.empty(int(lib.lcb_gptq_workspace_bytes(R, C, blocksize)), dtype=torch.uint8, device=W.device)
    with _on_device(W.device):
        st = lib.lcb_gptq_weight_transform(
            W.data_ptr(), Hinv.data_ptr(), tmp.data_ptr(), losses.data_ptr(), gs.data_ptr(),
            gz.data_ptr() if gz is not None else None, cg.data_ptr() if cg is not None else None, R, C, n_nonout,
            int(blocksize), int(group_size), 1 if static else 0, spec, ws.data_ptr(), _stream(W.device.index))
    _cabi.check(st, 'lcb_gptq_weight_transform')
    return tmp, losses, gs, gz


def scale_cols(x, scales, div=False, out=None):
    """x * scales[None, :] (Awq.scaling_weight, awq.py:41-47) or x / scales[None, :] (scaling_input, base_bq.py:877-889),
    in the tensor dtype with one rounding, IEEE division. x: [..., C] contiguous; scales: [C] of the same dtype."""
    _need_cuda(x, scales)
    x = x if x.is_contiguous() else x.contiguous()
    s = scales.reshape(-1).to(x.dtype).contiguous()
    C = x.shape[-1]
    assert s.numel() == C
    out = torch.empty_like(x) if out is None else out
    lib = _cabi.load()
    with _on_device(x.device):
        st = lib.lcb_scale_cols(x.data_ptr(), s.data_ptr(), out.data_ptr(), x.numel() // C, C, _DT[x.dtype],
                                1 if div else 0, _stream(x.device.index))
    _cabi.check(st, 'lcb_scale_cols')
    return out


def linear_mse(x, w, org_out, loss=None):
    """sum((org_out - F.linear(x, w))^2) with the reference's roundings (awq.py:134-145) without materialising the
    output. Returns a 1-element float64 device tensor (accumulated into `loss` when given)."""
    _need_cuda(x, w, org_out)
    x2 = x.reshape(-1, x.shape[-1])
    o2 = org_out.reshape(-1, org_out.shape[-1])
    assert x2.is_contiguous() and w.is_contiguous() and o2.is_contiguous()
    assert x2.dtype == w.dtype == o2.dtype and x2.shape[1] == w.shape[1] and o2.shape == (x2.shape[0], w.shape[0])
    if loss is None:
        loss = torch.zeros(1, dtype=torch.float64, device=x.device)
    lib = _cabi.load()
    with _on_device(x.device):
        st = lib.lcb_linear_mse(x2.data_ptr(), w.data_ptr(), o2.data_ptr(), loss.data_ptr(), x2.shape[0], w.shape[0],
                                x2.shape[1], _DT[x2.dtype], _stream(x.device.index))
    _cabi.check(st, 'lcb_linear_mse')
    return loss


def sqdiff_mean(a, b):
    """`(a - b).float().pow(2).mean()` (calculate_loss, awq.py:134-145) in one pass (lcb_sqdiff_sum): the difference in
    the tensor dtype, squares and sum in fp32 / fp64. Returns a 0-dim float64 device tensor (no host sync)."""
    _need_cuda(a, b)
    assert a.shape == b.shape and a.dtype == b.dtype
    a2 = a if a.is_contiguous() else a.contiguous()
    b2 = b if b.is_contiguous() else b.contiguous()
    out = torch.zeros(1, dtype=torch.float64, device=a.device)
    lib = _cabi.load()
    with _on_device(a.device):
        st = lib.lcb_sqdiff_sum(a2.data_ptr(), b2.data_ptr(), out.data_ptr(), a2.numel(), _DT[a2.dtype], _stream(a.device.index))
    _cabi.check(st, 'lcb_sqdiff_sum')
    return out[0] / float(max(a2.numel(), 1))


def linear(x, w, bias=None):
    """torch.nn.functional.linear(x, w, bias) for 16-bit tensors on the tcgen05 contraction (lcb_linear): fp32
    accumulation, one rounding. x [..., C], w [R, C] of the same dtype; bias None or [R]."""
    _need_cuda(x, w)
    C = x.shape[-1]
    x2 = x.reshape(-1, C)
    if not x2.is_contiguous():
        x2 = x2.contiguous()
    w = w if w.is_contiguous() else w.contiguous()
    assert x2.dtype == w.dtype and w.shape[1] == C
    R = w.shape[0]
    out = torch.empty((x2.shape[0], R), dtype=x.dtype, device=x.device)
    b = None
    if bias is not None:
        b = bias.to(x.dtype).contiguous()
    lib = _cabi.load()
    with _on_device(x.device):
        st = lib.lcb_linear(x2.data_ptr(), w.data_ptr(), b.data_ptr() if b is not None else None, out.data_ptr(),
                            x2.shape[0], R, C, _DT[x.dtype], _stream(x.device.index))
    _cabi.check(st, 'lcb_linear')
    return out.reshape(*x.shape[:-1], R)


def awq_candidates(W, x, org_out, cand_scales, wquantizer):
    """Losses (mean squared error, float64 [n_cand]) of every AWQ scale candidate for a subset whose inspect module is
    the linear itself (awq.py:202-248). cand_scales: [n_cand, C] in the tensor dtype."""
    _need_cuda(W, x, org_out, cand_scales)
    x2 = x.reshape(-1, x.shape[-1]).contiguous()
    o2 = org_out.reshape(-1, org_out.shape[-1]).contiguous()
    W = W.contiguous()
    cs = cand_scales.to(W.dtype).contiguous()
    R, C = W.shape
    T = x2.shape[0]
    n = cs.shape[0]
    g = wquantizer.granularity
    if g == 'per_group':
        gs = wquantizer.group_size
    elif g == 'per_channel':
        gs = 0
    else:
        raise NotImplementedError(f'awq_candidates: weight granularity {g}')
    qmin, qmax = wquantizer._qrange()
    spec = wquantizer._spec(W.dtype, W.dtype, W.dtype, _cabi.OUT_FAKE, qmin, qmax, 1.0, 1.0, _DT[W.dtype] + 1)
    lib = _cabi.load()
    loss = torch.zeros(n, dtype=torch.float64, device=W.device)
    ws = torch.empty(int(lib.lcb_awq_workspace_bytes(T, R, C, n)), dtype=torch.uint8, device=W.device)
    with _on_device(W.device):
        st = lib.lcb_awq_candidates(W.data_ptr(), x2.data_ptr(), o2.data_ptr(), cs.data_ptr(), loss.data_ptr(), T, R, C, n,
                                    gs, spec, ws.data_ptr(), _stream(W.device.index))
    _cabi.check(st, 'lcb_awq_candidates')
    return loss / float(T * R)


def _autoclip_geometry(W, wquantizer):
    g = wquantizer.granularity
    if g == 'per_group':
        gs = wquantizer.group_size
    elif g == 'per_channel':
        gs = 0
    else:
        raise NotImplementedError(f'autoclip: weight granularity {g}')
    return gs, (W.shape[1] // gs if gs > 0 else 1)


def autoclip_pad_tokens(x2):
    """[tokens, C] -> ([T_rows, C] contiguous with T_rows = tokens rounded up to 8, zero rows appended; tokens)."""
    tokens, C = x2.shape
    T_rows = -(-tokens // 8) * 8
    if T_rows != tokens:   # zero rows add nothing to the squared error; the mean still divides by `tokens`
        xp = torch.zeros((T_rows, C), dtype=x2.dtype, device=x2.device)
        xp[:tokens] = x2
        x2 = xp
    return x2.contiguous(), tokens


class AutoClipState(object):
    """Buffers of one taken-apart AutoClip search (lcb_autoclip_init / lcb_autoclip_step): the host composes the weights
    of every shrink step itself (clip_version 'v2', w+a)."""

    def __init__(self, W, wquantizer, clip_sym, T_rows, tokens):
        _need_cuda(W)
        if W.dtype not in (torch.bfloat16, torch.float16):
            raise TypeError(f'autoclip: W is {W.dtype}')
        self.W = W.contiguous()
        self.R, self.C = W.shape
        self.gs, self.ng = _autoclip_geometry(W, wquantizer)
        self.clip_sym, self.T_rows, self.tokens = bool(clip_sym), int(T_rows), int(tokens)
        dev, nu = W.device, self.R * self.ng
        self.best_max = torch.empty((self.R, self.ng, 1), dtype=W.dtype, device=dev)
        self.best_min = torch.empty((self.R, self.ng, 1), dtype=W.dtype, device=dev)
        self.org_max = torch.empty(nu, dtype=torch.float32, device=dev)
        self.org_min = torch.empty(nu, dtype=torch.float32, device=dev)
        self.min_err = torch.empty(nu, dtype=torch.float32, device=dev)
        self.org = torch.empty((self.ng, self.R, self.T_rows), dtype=W.dtype, device=dev)
        lib = _cabi.load()
        with _on_device(dev):
            st = lib.lcb_autoclip_init(self.W.data_ptr(), self.best_max.data_ptr(), self.best_min.data_ptr(),
                                       self.org_max.data_ptr(), self.org_min.data_ptr(), self.min_err.data_ptr(), self.R, self.C,
                                       self.gs, 1 if clip_sym else 0, _DT[W.dtype], _stream(dev.index))
        _cabi.check(st, 'lcb_autoclip_init')

    def _step(self, Wq, x, ratio, is_org):
        _need_cuda(Wq, x)
        if Wq.dtype != self.W.dtype or x.dtype != self.W.dtype or tuple(Wq.shape) != (self.R, self.C) \
                or tuple(x.shape) != (self.T_rows, self.C):
            raise ValueError(f'autoclip step: Wq {tuple(Wq.shape)} {Wq.dtype}, x {tuple(x.shape)} {x.dtype}')
        Wq, x = Wq.contiguous(), x.contiguous()
        lib = _cabi.load()
        with _on_device(Wq.device):
            st = lib.lcb_autoclip_step(Wq.data_ptr(), x.data_ptr(), self.org.data_ptr(), self.org_max.data_ptr(),
                                       self.org_min.data_ptr(), self.min_err.data_ptr(), self.best_max.data_ptr(),
                                       self.best_min.data_ptr(), self.R, self.C, self.T_rows, self.tokens, self.gs, float(ratio),
                                       1 if self.clip_sym else 0, 1 if is_org else 0, _DT[Wq.dtype], _stream(Wq.device.index))
        _cabi.check(st, 'lcb_autoclip_step')

    def original(self, x):
        """org_out = (x * w).sum(-1) on the plain inputs (auto_clip.py:151)."""
        self._step(self.W, x, 1.0, True)

    def candidate(self, Wq, xq, ratio):
        """One shrink step: Wq = this step's fake-quantized weights, xq = the (fake-quantized) inputs (auto_clip.py:154-179)."""
        self._step(Wq, xq, ratio, False)


def autoclip_search(W, x, wquantizer, clip_sym, n_grid=20, max_shrink=0.5):
    """AutoClip v1 search of one linear (auto_clip.py:84-191): returns (best_max_val, best_min_val), each [R, n_group, 1]
    in W.dtype. W [R, C] bf16 / f16; x [tokens, C] = the sampled tokens (auto_clip.py:131-133) in the same dtype."""
    _need_cuda(W, x)
    W = W.contiguous()
    R, C = W.shape
    x2 = x.reshape(-1, C)
    if x2.dtype != W.dtype:
        raise TypeError(f'autoclip_search: x is {x2.dtype}, W is {W.dtype}')
    x2, tokens = autoclip_pad_tokens(x2)
    T_rows = x2.shape[0]
    gs, ng = _autoclip_geometry(W, wquantizer)
    qmin, qmax = wquantizer._qrange()
    spec = wquantizer._spec(W.dtype, W.dtype, W.dtype, _cabi.OUT_FAKE, qmin, qmax, 1.0, 1.0, _DT[W.dtype] + 1)
    lib = _cabi.load()
    best_max = torch.empty((R, ng, 1), dtype=W.dtype, device=W.device)
    best_min = torch.empty((R, ng, 1), dtype=W.dtype, device=W.device)
    ws = torch.empty(int(lib.lcb_autoclip_workspace_bytes(R, C, T_rows, gs)), dtype=torch.uint8, device=W.device)
    with _on_device(W.device):
        st = lib.lcb_autoclip_search(W.data_ptr(), x2.data_ptr(), best_max.data_ptr(), best_min.data_ptr(), R, C, T_rows,
                                     tokens, gs, int(n_grid), int(max_shrink * n_grid), 1 if clip_sym else 0, spec,
                                     ws.data_ptr(), _stream(W.device.index))
    _cabi.check(st, 'lcb_autoclip_search')
    return best_max, best_min
