"""CPU restatement of the reference's GPTQ path — TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Each function follows /root/reference/llmc/compression/quantization/gptq.py line by line, on CPU torch tensors
(torch = memory + IEEE fp32 ops; the reference itself is eager torch, so this IS its arithmetic):
  add_batch                     gptq.py:254-295   (world_size 1: the all-reduce is the identity)
  process_hessian               gptq.py:128-176   (dead columns, damping, Cholesky inverse factor)
  weight_transform              gptq.py:199-244   (+ search_column_qparams :359-366)
Parity pinning: tests/golden/make_algo_golden.py drives the UNMODIFIED reference class methods
(object.__new__(GPTQ), SURVEY.md §8c recipe 2) on seeded inputs and stores their outputs in tests/golden/algo_golden.pt.gz;
tests/test_algo_oracle_golden.py checks this restatement against them.
"""
import math

import torch

from . import quant_oracle as qo


def add_batch(H, nsamples, inp):
    """gptq.py:254-295 for a Linear layer. H fp32 [C, C] (updated in place), inp [b, T, C] or [T, C]. Returns nsamples."""
    if inp.dim() == 2:
        inp = inp.unsqueeze(0)
    tmp = inp.shape[0]
    x = inp.reshape(-1, inp.shape[-1]).t()
    H *= nsamples / (nsamples + tmp)
    nsamples += tmp
    chunk = math.sqrt(2 / nsamples) * x.float()
    H += chunk.matmul(chunk.t())
    return nsamples


def process_hessian(H, W, percdamp):
    """gptq.py:139-176 without actorder / owq. Returns (W, Hinv_upper); H is consumed."""
    W = W.clone().float()
    dead = torch.diag(H) == 0
    H[dead, dead] = 1
    W[:, dead] = 0
    damp = percdamp * torch.mean(torch.diag(H))
    diag = torch.arange(H.shape[0])
    H[diag, diag] += damp
    H = torch.linalg.cholesky(H)
    H = torch.cholesky_inverse(H)
    H = torch.linalg.cholesky(H, upper=True)
    return W, H


def _qparams_cols(Wcols, spec):
    """IntegerQuantizer.get_tensor_qparams on an fp32 [R, g] slice with per-row units (quant.py:690-697)."""
    R, g = Wcols.shape
    _, s, z = qo.fakequant(Wcols.contiguous(), R, g, 1, g, out_mode=qo.OUT_NONE, **spec)
    if z is None:
        z = torch.zeros(R)
    return s.float().reshape(R, 1), z.float().reshape(R, 1)


def _quant_dequant_col(w, scale, zero, spec):
    """IntegerQuantizer.quant_dequant of one fp32 column [R, 1] with [R, 1] qparams (quant.py:699-717)."""
    R = w.shape[0]
    out, _, _ = qo.fakequant(w.reshape(R, 1).contiguous(), R, 1, 1, 1, dynamic=False, scales=scale.reshape(-1),
                             zeros=zero.reshape(-1), **spec)
    return out.reshape(R)


def weight_transform(W, Hinv, blocksize, group_size, spec, static_groups=None):
    """gptq.py:199-244. W fp32 [R, C] (mutated like the reference), Hinv upper factor. spec: kind/sym/qmin/qmax kwargs
    of quant_oracle.fakequant. group_size <= 0: per_channel (qparams from the whole row, searched once up front as
    `search_layer_qparams` does with the collected buffers). Returns (tmp, Losses, scales [C/g, R], zeros [C/g, R])."""
    R, C = W.shape
    Losses = torch.zeros_like(W)
    tmp = torch.zeros_like(W)
    scales, zeros = [], []
    if group_size <= 0:
        scale, zero = _qparams_cols(W, spec)
    for i1 in range(0, C, blocksize):
        i2 = min(i1 + blocksize, C)
        count = i2 - i1
        W1, Hinv1 = W[:, i1:i2].clone(), Hinv[i1:i2, i1:i2]
        tmp1, Err1, Losses1 = torch.zeros_like(W1), torch.zeros_like(W1), torch.zeros_like(W1)
        for i in range(count):
            w, d = W1[:, i], Hinv1[i, i]
            if group_size > 0:
                if static_groups is not None:
                    scale, zero = static_groups[(i1 + i) // group_size]
                elif (i1 + i) % group_size == 0:
                    scale, zero = _qparams_cols(W[:, (i1 + i):min(i1 + i + group_size, C)], spec)
                    scales.append(scale.reshape(-1).clone())
                    zeros.append(zero.reshape(-1).clone())
            q = _quant_dequant_col(w, scale, zero, spec)
            tmp1[:, i] = w
            Losses1[:, i] = ((w - q) ** 2) / (2 * d ** 2)
            err1 = (w - q) / d
            W1[:, i:] -= err1.unsqueeze(1).matmul(Hinv1[i, i:].unsqueeze(0))
            Err1[:, i] = err1
        tmp[:, i1:i2], Losses[:, i1:i2] = tmp1, Losses1
        W[:, i2:] -= Err1.matmul(Hinv[i1:i2, i2:])
    sc = torch.stack(scales) if scales else scale.reshape(1, -1)
    zp = torch.stack(zeros) if zeros else zero.reshape(1, -1)
    return tmp, Losses, sc, zp
