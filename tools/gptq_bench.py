"""GPTQ calibration seconds per Llama-3-8B transformer block on one B200 (BASELINE configs[1]).

Per linear: Hessian accumulation over the calibration set (128 batches x 2048 tokens, gptq_w_only.yml), the damped
Cholesky-inverse factor (lcb_cholinv_upper; the torch / cuSOLVER chain is timed beside it) and the blocked column update
(lcb_gptq_weight_transform). Synthetic activations / random weights of the named shapes.

    python tools/gptq_bench.py [--quick] [--batches N]
"""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lightcompress_b200 import ops  # noqa: E402

# (name, R = out_features, C = in_features, how many linears of the block share this shape)
LLAMA3_8B = [('q/o_proj', 4096, 4096, 2), ('k/v_proj', 1024, 4096, 2), ('gate/up_proj', 14336, 4096, 2),
             ('down_proj', 4096, 14336, 1)]
# distinct layer INPUTS among the linears of each row: q/k/v share one, o has its own, gate/up share one, down has its own.
# algo.GPTQ (share_hessian, the default) accumulates and factorises one Hessian per distinct input.
N_INPUTS = {'q/o_proj': 2, 'k/v_proj': 0, 'gate/up_proj': 1, 'down_proj': 1}


def ev_time(fn, rep=1):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(rep):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e-3 / rep


def main():
    quick = '--quick' in sys.argv
    nb = 128
    if '--batches' in sys.argv:
        nb = int(sys.argv[sys.argv.index('--batches') + 1])
    run(torch.device('cuda:0'), nb, quick, emit=True)


def run(dev, nb=128, quick=False, emit=False):
    """Returns {'hessian_s', 'cholesky_s', 'update_s', 'calib_s_per_block', 'batches'} for one Llama-3-8B block."""
    T = 2048
    total = {'hessian_s': 0.0, 'cholesky_s': 0.0, 'update_s': 0.0}
    shared = {'hessian_s': 0.0, 'cholesky_s': 0.0}
    rows = []
    for name, R, C, mult in LLAMA3_8B:
        g = torch.Generator(device=dev).manual_seed(C + R)
        gain = torch.exp(0.5 * torch.randn(C, device=dev, generator=g))
        xs = [(torch.randn(T, C, device=dev, generator=g) * gain).to(torch.bfloat16) for _ in range(4)]
        S = torch.zeros(C, C, device=dev)
        ops.hessian_accum(S, xs[0])
        t_h = ev_time(lambda: [ops.hessian_accum(S, x) for x in xs], 3) / len(xs)      # per 2048-token batch
        x4 = torch.cat(xs, 0)
        t_h4 = ev_time(lambda: ops.hessian_accum(S, x4), 3) / 4                        # 4 batches per call
        S.zero_()
        for x in xs:
            ops.hessian_accum(S, x)
        ops.hessian_finalize(S, len(xs))
        H = S
        W = (torch.randn(R, C, device=dev, generator=g) * 0.02)

        def chol():
            Hd = H.clone()
            d = torch.arange(C, device=dev)
            Hd[d, d] += 0.01 * torch.mean(torch.diag(Hd))
            L = torch.linalg.cholesky(Hd)
            Hi = torch.cholesky_inverse(L)
            return torch.linalg.cholesky(Hi, upper=True)
        t_c_torch = ev_time(chol, 1)                                      # the reference's cuSOLVER chain, for comparison
        Hw = H.clone()
        ops.cholinv_upper(Hw, 0.01, check=False)
        t_c = ev_time(lambda: ops.cholinv_upper(Hw.copy_(H), 0.01, check=False), 2)
        Hinv = Hw
        ops.gptq_weight_transform(W.clone(), Hinv, 128, 128, 0.0, 15.0, False)   # warm-up (attributes, allocator)
        Wc = W.clone()
        t0 = time.perf_counter()
        t_u = ev_time(lambda: ops.gptq_weight_transform(Wc, Hinv, 128, 128, 0.0, 15.0, False), 1)
        wall = time.perf_counter() - t0
        flops_h = 2.0 * T * C * C
        row = {'linear': name, 'R': R, 'C': C, 'x': mult, 'hessian_ms_per_batch': round(t_h * 1e3, 4),
               'hessian_ms_per_batch_4x': round(t_h4 * 1e3, 4), 'hessian_gemm_equiv_TFLOPs': round(flops_h / t_h4 / 1e12, 1),
               'cholinv_ms': round(t_c * 1e3, 2), 'torch_cholesky_chain_ms': round(t_c_torch * 1e3, 2), 'update_ms': round(t_u * 1e3, 3), 'update_wall_ms': round(wall * 1e3, 3),
               'update_gemm_equiv_TFLOPs': round(float(R) * C * C / t_u / 1e12, 2)}
        rows.append(row)
        if emit:
            print(json.dumps(row), flush=True)
        total['hessian_s'] += mult * nb * t_h4
        total['cholesky_s'] += mult * t_c
        total['update_s'] += mult * t_u
        shared['hessian_s'] += N_INPUTS[name] * nb * t_h4
        shared['cholesky_s'] += N_INPUTS[name] * t_c
        del xs, x4, S, H, W, Wc, Hinv, Hw
        torch.cuda.empty_cache()
        if quick and name.startswith('gate'):
            break
    total['calib_s_per_block'] = sum(total.values())
    total['batches'] = nb
    # one Hessian + one factor per distinct input (what algo.GPTQ runs); the per-linear figures above follow the reference
    total['shared_inputs'] = {'hessian_s': shared['hessian_s'], 'cholesky_s': shared['cholesky_s'],
                              'calib_s_per_block': shared['hessian_s'] + shared['cholesky_s'] + total['update_s']}
    if emit:
        print(json.dumps({'llama3_8b_block': total}), flush=True)
    return total


if __name__ == '__main__':
    main()
