"""Size sweep of the fused fake-quant kernels against a plain device copy (the measured HBM ceiling).

    python tools/fq_sweep.py [--quick]

Prints one JSON line per (case, size): algorithmic GB/s of our kernel, of `b.copy_(a)` on the same buffers, and the
ratio. Timing: CUDA events around REP back-to-back launches on the current stream, after warm-up; buffers are
rotated so that every launch reads data that is not L2 resident (sum of live buffers > 126 MB for the small sizes).
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lightcompress_b200.quant import FloatQuantizer, IntegerQuantizer  # noqa: E402


def time_calls(fn, bufs, rep):
    for b in bufs[:3]:
        fn(b)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(rep):
        fn(bufs[i % len(bufs)])
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e-3 / rep


def main():
    quick = '--quick' in sys.argv
    dev = torch.device('cuda:0')
    cases = [
        ('int4 asym g128 bf16 (w4a16 weights)', lambda: IntegerQuantizer(4, False, 'per_group', group_size=128), 'w', 4096),
        ('int8 sym per_token bf16 C=4096 (w8a8 acts)', lambda: IntegerQuantizer(8, True, 'per_token'), 'a', 4096),
        ('int8 sym per_token bf16 C=14336', lambda: IntegerQuantizer(8, True, 'per_token'), 'a', 14336),
        ('int8 sym per_channel bf16 C=4096', lambda: IntegerQuantizer(8, True, 'per_channel'), 'w', 4096),
        ('fp8 e4m3 per_block128 bf16', lambda: FloatQuantizer('e4m3', True, 'per_block', block_size=128, use_qtorch=True), 'w', 2048),
        ('fp8 e4m3 per_group128 bf16 (acts)', lambda: FloatQuantizer('e4m3', True, 'per_group', group_size=128, use_qtorch=True), 'a', 2048),
        ('int4 sym g128 fp16', lambda: IntegerQuantizer(4, True, 'per_group', group_size=128), 'w16', 4096),
    ]
    sizes = [2 ** 22, 2 ** 24, 2 ** 26, 2 ** 28] if quick else [2 ** 20, 2 ** 22, 2 ** 24, 2 ** 26, 2 ** 28, 2 ** 30]
    only = [a[7:] for a in sys.argv if a.startswith('--only=')]
    for name, mk, kind, cols in cases:
        if only and not any(o in name for o in only):
            continue
        q = mk()
        for numel in sizes:
            rows = max(1, numel // cols)
            dt = torch.float16 if kind == 'w16' else torch.bfloat16
            nb = max(2, min(8, (400 << 20) // (rows * cols * 2)))
            bufs = [(torch.randn(rows, cols, device=dev, dtype=torch.float32) * 0.02).to(dt) for _ in range(nb)]
            outs = [torch.empty_like(b) for b in bufs[:2]]
            fn = (q.fake_quant_act_dynamic if kind == 'a' else q.fake_quant_weight_dynamic)
            rep = max(4, min(50, (8 << 30) // (rows * cols * 4)))
            t = time_calls(fn, bufs, rep)
            tc = time_calls(lambda b: outs[0].copy_(b), bufs, rep)
            alg = rows * cols * 4
            print(json.dumps({'case': name, 'numel': rows * cols, 'us': round(t * 1e6, 2), 'GBps': round(alg / t / 1e9, 1),
                              'copy_us': round(tc * 1e6, 2), 'copy_GBps': round(alg / tc / 1e9, 1),
                              'vs_copy': round(tc / t, 3)}), flush=True)
            del bufs, outs
            torch.cuda.empty_cache()


if __name__ == '__main__':
    main()
