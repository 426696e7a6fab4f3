"""bench.py — headline benchmark of the B200-native LLMC quantization hot path.

Metric (BASELINE.json): fake-quant HBM GB/s on Llama-3-8B shapes (+ GPTQ calibration s/block as extra keys).
A "step" = one pass of the w4a16 group-128 fake-quant (IntegerQuantizer(4, asym, per_group 128), the quantization
of BASELINE configs[1]) over the 7 linear weights of one Llama-3-8B transformer block (218.1 M bf16 elements,
436 MB in + 436 MB out: larger than the 126 MB L2, so no flush is needed between iterations), issued as ONE
multi-tensor launch (fake_quant_weights_dynamic -> lcb_fakequant_multi), the way the block loop deploys a block.
`extras.per_tensor_calls_GBps` is the same work through 7 reference-API calls (fake_quant_weight_dynamic).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

N > 1 (torchrun): independent linear layers / blocks shard naturally — every rank processes its own block, no
data-path collective; value = total bytes of all ranks / max-over-ranks time ("scaling": "weak").
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

LLAMA3_8B_BLOCK = [('q_proj', 4096, 4096), ('k_proj', 1024, 4096), ('v_proj', 1024, 4096), ('o_proj', 4096, 4096),
                   ('gate_proj', 14336, 4096), ('up_proj', 14336, 4096), ('down_proj', 4096, 14336)]
METRIC = 'fake-quant HBM GB/s (% of peak); GPTQ/AWQ calib s/block, Llama-3-8B'
WORKLOAD = 'Llama-3-8B block (7 linears, 218.1M bf16 params) w4a16 asym group-128 fake_quant_weight_dynamic'


def peaks():
    """(HBM GB/s, bf16 TFLOP/s sustained, source, bf16 TFLOP/s burst). Loops shorter than ~1 s are held to the BURST figure."""
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        d = json.load(open(p))
        burst = float(d.get('bf16_tflops', 1600.0))
        return float(d['hbm_gbs']), float(d.get('bf16_tflops_sustained', burst)), 'measured', burst
    return 6650.0, 1590.0, 'fallback', 1590.0


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', f'--id={self.index}', f'--query-gpu={self.Q}',
                                          '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([f.strip() for f in line.split(',')])

    def stop(self):
        if not self.proc:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = sorted(int(r[0]) for r in self.rows if len(r) >= 6 and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) >= 6 and r[1].isdigit()]
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 6 and r[2 + i] == 'Active' for r in self.rows)]
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'reasons': reasons, 'samples': len(sm)}


def make_weights(torch, dev, seed):
    g = torch.Generator(device=dev).manual_seed(seed)
    ws = []
    for _, r, c in LLAMA3_8B_BLOCK:
        w = torch.randn((r, c), generator=g, device=dev, dtype=torch.float32) * 0.02
        ws.append(w.to(torch.bfloat16))
    return ws


def _host_cores():
    return len(os.sched_getaffinity(0)) if hasattr(os, 'sched_getaffinity') else os.cpu_count()


def cpu_block_weights(torch):
    g = torch.Generator().manual_seed(0)
    return [(torch.randn(r, c, generator=g) * 0.02).to(torch.bfloat16) for _, r, c in LLAMA3_8B_BLOCK]


def cpu_reference_step(torch, ws, q):
    """One step of the workload on the host: fake_quant_weight_dynamic over the 7 weights of the block."""
    t0 = time.perf_counter()
    for w in ws:
        q(w)
    return time.perf_counter() - t0


def cpu_arm():
    """(callable(weight) -> fake-quantized weight, kind, description). The reference's OWN quantizer
    (llmc/compression/quantization/quant.py, shipped unmodified under oracle/_ref) on CPU tensors with every host core;
    the C/OpenMP port (oracle/lc_oracle.c) only when that copy is absent."""
    import ctypes
    import torch
    cores = _host_cores()
    try:   # torchrun exports OMP_NUM_THREADS=1: the CPU arm is entitled to every host core it can use
        ctypes.CDLL('libgomp.so.1').omp_set_num_threads(cores)
    except OSError:
        pass
    torch.set_num_threads(cores)
    from oracle import ref_harness
    if ref_harness.available():
        refq = ref_harness.load_quant_only()
        q = refq.IntegerQuantizer(4, False, 'per_group', group_size=128)
        return q.fake_quant_weight_dynamic, 'reference', 'IntegerQuantizer.fake_quant_weight_dynamic (quant.py:833-869) on CPU tensors'
    from oracle import quant_oracle as qo
    kw = dict(kind=0, sym=False, qmin=0.0, qmax=15.0)
    return (lambda w: qo.fakequant(w, w.shape[0], w.shape[1], 1, 128, **kw)[0]), 'port', 'oracle/lc_oracle.c (OpenMP)'


def cpu_baseline(steps=3):
    """The CPU arm on the WHOLE workload (the 7 weights of one Llama-3-8B block), best of `steps`."""
    import torch
    fn, kind, what = cpu_arm()
    ws = cpu_block_weights(torch)
    cpu_reference_step(torch, ws[:2], fn)
    best = min(cpu_reference_step(torch, ws, fn) for _ in range(steps))
    nbytes = sum(w.numel() for w in ws) * 4
    return {'value': nbytes / best / 1e9, 'unit': 'GB/s', 'cores': _host_cores(), 'kind': kind,
            'sample': f'the whole step: 7 weights of one Llama-3-8B block (218.1M bf16 elements), best of {steps}; {what}'}, best


def extras(torch, dev, tf_peak, tf_burst, hbm_peak):
    """Tensor-path numbers of the same hot path (short runs): Hessian, GPTQ update, AWQ candidate at Llama-3-8B shapes."""
    from lightcompress_b200 import ops
    from lightcompress_b200.quant import IntegerQuantizer
    res = {}

    def ev(fn, rep):
        fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(rep):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) * 1e-3 / rep

    C, T = 4096, 8192
    g = torch.Generator(device=dev).manual_seed(0)
    x = torch.randn(T, C, device=dev, generator=g).to(torch.bfloat16)
    S = torch.zeros(C, C, device=dev)
    t = ev(lambda: ops.hessian_accum(S, x), 20)
    res['hessian'] = {'C': C, 'T': T, 'ms': t * 1e3, 'gemm_equiv_TFLOPs': 2.0 * T * C * C / t / 1e12,
                      'hw_syrk_TFLOPs': 2.0 * T * C * C / t / 1e12 * 33 / 64,
                      'frac_of_bf16_burst_peak_hw': 2.0 * T * C * C / t / 1e12 * 33 / 64 / tf_burst}
    ops.hessian_finalize(S, 20)
    Hw = S.clone()
    ops.cholinv_upper(Hw, 0.01, check=False)
    t = ev(lambda: ops.cholinv_upper(Hw.copy_(S), 0.01, check=False), 3)
    res['cholinv_ms_C4096'] = t * 1e3      # damping + cholesky -> cholesky_inverse -> cholesky(upper) of gptq.py:165-174
    Hinv = Hw
    W = torch.randn(C, C, device=dev, generator=g) * 0.02
    t = ev(lambda: ops.gptq_weight_transform(W.clone(), Hinv, 128, 128, 0.0, 15.0, False), 2)
    res['gptq_update_q_proj_ms'] = t * 1e3
    Wb = W.to(torch.bfloat16)
    org = torch.nn.functional.linear(x, Wb)
    cand = (torch.rand(4, C, device=dev, generator=g) + 0.5).to(torch.bfloat16)
    wq = IntegerQuantizer(4, False, 'per_group', group_size=128)
    t = ev(lambda: ops.awq_candidates(Wb, x, org, cand, wq), 3) / 4
    # block-wise FP8 inference step: act_quant (HBM bound) and the fp8 GEMM with fp32 block scales
    from lightcompress_b200 import fp8_infer as fi
    xa = torch.randn(8192, 7168, device=dev, generator=g).to(torch.bfloat16)
    ta = ev(lambda: fi.act_quant(xa), 20)
    a8, as8 = fi.act_quant(xa)
    b8 = (torch.randn(2048, 7168, device=dev, generator=g) * 50).clamp(-448, 448).to(torch.float8_e4m3fn)
    bs8 = torch.rand(16, 56, device=dev, generator=g) * 0.01 + 1e-3
    tg = ev(lambda: fi.fp8_gemm(a8, as8, b8, bs8, out_dtype=torch.bfloat16), 10)
    aq_bytes = xa.numel() * 3 + xa.numel() // 128 * 4
    res['fp8_act_quant'] = {'shape': [8192, 7168], 'ms': ta * 1e3, 'GBps': aq_bytes / ta / 1e9}
    res['fp8_gemm'] = {'M': 8192, 'N': 2048, 'K': 7168, 'ms': tg * 1e3, 'TFLOPs': 2.0 * 8192 * 2048 * 7168 / tg / 1e12}
    del x, S, W, Wb, org, cand, Hinv, Hw, xa, a8, as8, b8, bs8
    torch.cuda.empty_cache()
    sys.path.insert(0, os.path.join(ROOT, 'tools'))
    import gptq_bench   # GPTQ calibration of one Llama-3-8B block: 128 x 2048 tokens, 7 linears (BASELINE configs[1])
    res['gptq_calib_llama3_8b_block'] = gptq_bench.run(dev, nb=128)
    res['awq_candidate'] = {'T': T, 'R': C, 'C': C, 'ms': t * 1e3, 'TFLOPs': 2.0 * T * C * C / t / 1e12,
                            'frac_of_bf16_burst_peak': 2.0 * T * C * C / t / 1e12 / tf_burst}
    res.update(hbm_extras(torch, dev, hbm_peak))
    import gc
    gc.collect()
    torch.cuda.empty_cache()
    # END TO END through ALGO_REGISTRY on one Llama-3-8B-sized decoder block (hooked forwards included), in a process of its
    # own: what is timed is a calibration run, whose allocator / library state is not that of a process that has just run
    # every other extra (in-process the same block measured 0.54 s against 0.35 s on its own)
    r = subprocess.run([sys.executable, os.path.join(ROOT, 'tools', 'block_bench.py'), 'gptq', 'awq'], capture_output=True,
                       text=True, timeout=900)
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith('{')]
    bb = json.loads(lines[-1]) if r.returncode == 0 and lines else {'error': (r.stderr or r.stdout)[-400:]}
    if 'gptq_s_per_block' in bb:   # 2 T sum C^2 = 160.6 TFLOP of Hessian work per block (GEMM convention)
        bb['gptq_frac_of_bf16_sustained'] = 160.6 / bb['gptq_s_per_block'] / tf_peak
    res['calib_e2e'] = bb
    return res


def nested_block_bench(world):
    """tools/block_bench.py under its own torchrun on `world` GPUs of this node; returns its JSON record."""
    env = {k: v for k, v in os.environ.items()
           if not (k.startswith('TORCHELASTIC') or k in ('RANK', 'LOCAL_RANK', 'WORLD_SIZE', 'LOCAL_WORLD_SIZE', 'GROUP_RANK',
                                                         'GROUP_WORLD_SIZE', 'ROLE_RANK', 'ROLE_WORLD_SIZE', 'ROLE_NAME',
                                                         'MASTER_ADDR', 'MASTER_PORT', 'TORCH_NCCL_ASYNC_ERROR_HANDLING'))}
    port = int(os.environ.get('MASTER_PORT', '29500')) + 37
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', f'--nproc-per-node={world}', '--master-addr', '127.0.0.1',
           '--master-port', str(port), os.path.join(ROOT, 'tools', 'block_bench.py'), 'gptq', 'awq']
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env)
    except subprocess.TimeoutExpired:
        return {'error': 'timeout'}
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith('{')]
    return json.loads(lines[-1]) if r.returncode == 0 and lines else {'error': (r.stderr or r.stdout)[-400:]}


def hbm_extras(torch, dev, hbm_peak):
    """The other HBM-bound call sites of the path at the reference's own call sizes, through the reference API, one call
    per tensor. Inputs rotate through a pool larger than the 126 MB L2 and every output stays alive until the clock
    stops (a recycled output block would be rewritten in L2 and never reach HBM)."""
    from lightcompress_b200.quant import FloatQuantizer, IntegerQuantizer
    res = {}

    def timed_calls(fn, xs, n_calls, graph=False):
        keep = []
        for i in range(n_calls):     # one untimed pass: afterwards the caching allocator owns every block the timed pass
            keep.append(fn(xs[i % len(xs)]))   # needs (n_calls outputs alive at once), no cudaMalloc inside the clock
        keep.clear()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if graph:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for i in range(n_calls):
                    keep.append(fn(xs[i % len(xs)]))
            g.replay()
            torch.cuda.synchronize()
            a.record()
            g.replay()
            b.record()
        else:
            a.record()
            for i in range(n_calls):
                keep.append(fn(xs[i % len(xs)]))
            b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) * 1e-3 / n_calls

    # configs[2]: SmoothQuant w8a8 — per-token dynamic int8 fake-quant of the activation in front of every linear
    # (module_utils.py:799-813: 224 calls per Llama-3-8B forward)
    q8 = IntegerQuantizer(8, True, 'per_token')
    gen = torch.Generator(device=dev).manual_seed(11)
    act = {}
    for shape in ((1, 2048, 4096), (8, 2048, 4096), (8, 2048, 14336)):
        numel = shape[0] * shape[1] * shape[2]
        pool = max(2, min(16, (300 << 20) // (numel * 2) + 1))
        xs = [torch.randn(shape, device=dev, generator=gen, dtype=torch.float32).to(torch.bfloat16) for _ in range(pool)]
        n_calls = max(8, min(224, (6 << 30) // (numel * 2)))
        rec = {}
        for mode in ('eager', 'graph'):
            t = timed_calls(q8.fake_quant_act_dynamic, xs, n_calls, graph=(mode == 'graph'))
            rec[mode] = {'us_per_call': t * 1e6, 'GBps': numel * 4 / t / 1e9, 'frac': numel * 4 / t / 1e9 / hbm_peak}
        rec['calls'] = n_calls
        act['x'.join(map(str, shape))] = rec
        del xs
        torch.cuda.empty_cache()
    res['act_per_token_int8'] = act
    # configs[4]: DeepSeek-V2-Lite MoE block, FP8 E4M3 per-block (128 x 128) real quant of the 66 x 3 expert matrices
    f8 = FloatQuantizer('e4m3', True, 'per_block', block_size=128, use_qtorch=True)
    ws = []
    for _ in range(66):
        for r, c in ((1408, 2048), (1408, 2048), (2048, 1408)):
            ws.append((torch.randn(r, c, device=dev, generator=gen, dtype=torch.float32) * 0.02).to(torch.bfloat16))
    nbytes = sum(w.numel() for w in ws) * 3 + sum(-(-w.shape[0] // 128) * -(-w.shape[1] // 128) for w in ws) * 4

    def ev(fn, rep=3):
        outs = [fn() for _ in range(rep)]     # untimed: the allocator now owns `rep` generations of outputs
        del outs
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        outs = [fn() for _ in range(rep)]
        b.record()
        torch.cuda.synchronize()
        del outs
        return a.elapsed_time(b) * 1e-3 / rep
    t_multi = ev(lambda: f8.real_quant_weights_dynamic(ws))
    t_each = ev(lambda: [f8.real_quant_weight_dynamic(w) for w in ws])
    res['fp8_per_block_deepseek_v2_lite_moe_block'] = {
        'matrices': len(ws), 'bytes': nbytes,
        'multi_launch': {'ms': t_multi * 1e3, 'GBps': nbytes / t_multi / 1e9, 'frac': nbytes / t_multi / 1e9 / hbm_peak},
        'per_tensor_calls': {'ms': t_each * 1e3, 'GBps': nbytes / t_each / 1e9, 'frac': nbytes / t_each / 1e9 / hbm_peak}}
    del ws
    torch.cuda.empty_cache()
    return res


def sharded_extras(torch, dist, dev, world, rank):
    """The sharded tensor paths of SURVEY.md §8e, every rank participating (world = 1 gives the single-GPU value of the
    same item, so that the driver's N = 1 record carries the denominator of every speed-up).
    Hessian: the 128 x 2048 calibration tokens of one linear input are split over the ranks, every rank accumulates its
    shard into its own S (tcgen05 SYRK); ONE all-reduce of the maintained half of S with the batch count riding in the
    same buffer, scale applied on the device — checked against the single-rank Hessian of the same data.
    AWQ: the (candidate, row tile) units of the 20-candidate search of a linear are dealt over the ranks, one
    all-reduce(SUM) of 20 partial losses. Times are device times, max over ranks."""
    from lightcompress_b200 import ops
    from lightcompress_b200.algo import awq as lcawq
    from lightcompress_b200.algo import dist as lcdist
    from lightcompress_b200.quant import IntegerQuantizer
    res = {}

    def timed(fn, rep=3):
        best = None
        for _ in range(rep):
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            out = fn()
            b.record()
            torch.cuda.synchronize()
            t = torch.tensor([a.elapsed_time(b)], device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            best = float(t.item()) if best is None else min(best, float(t.item()))
        return best, out

    T, batches = 2048, 128
    for C in (4096, 8192, 14336):   # Llama-3-8B hidden / Llama-3-70B hidden / Llama-3-8B intermediate (down_proj input)
        mine = list(range(rank, batches, world))

        def batch(i):    # batch i is the same tensor on whichever rank owns it: two distinct patterns keep memory small
            g = torch.Generator(device=dev).manual_seed(1000 + (i & 1))
            return torch.randn(T, C, device=dev, generator=g).to(torch.bfloat16)
        xs = [batch(0), batch(1)]

        def hess():
            S = torch.zeros(C, C, device=dev)
            for i in mine:
                ops.hessian_accum(S, xs[i & 1])
            return lcdist.finalize_sharded_hessian(S, len(mine))
        hess()
        ms, H = timed(hess)
        rec = {'ms': ms, 'batches_total': batches, 'batches_per_rank': len(mine),
               'gemm_equiv_TFLOPs_aggregate': 2.0 * T * batches * C * C / (ms * 1e-3) / 1e12,
               'allreduce_payload_MB': ops.hessian_pack(H, 0).numel() * 4 / 1e6 if world > 1 else 0.0}
        if world > 1 and C == 4096:   # hardware check: the N-rank H equals the 1-rank H of the same 128 batches
            S1 = torch.zeros(C, C, device=dev)
            for i in range(batches):
                ops.hessian_accum(S1, xs[i & 1])
            H1 = ops.hessian_finalize(S1, batches)
            rec['max_rel_diff_vs_single_rank'] = float((H - H1).abs().max() / H1.abs().max())
            # two fp32 summation orders over 262 144 tokens: the bar the tests hold H to against an fp64 product
            assert rec['max_rel_diff_vs_single_rank'] < 2e-5, rec
            del S1, H1
        res[f'hessian_C{C}'] = rec
        del H, xs
        torch.cuda.empty_cache()
    wq = IntegerQuantizer(4, False, 'per_group', group_size=128)
    # T = 128 samples x 512 tokens, the calibration set of awq_w_only.yml
    for name, R, C, T in (('llama3_8b_o_proj', 4096, 4096, 65536), ('llama3_70b_o_proj', 8192, 8192, 65536)):
        g = torch.Generator(device=dev).manual_seed(7)      # identical data on every rank
        x = (torch.randn(T, C, device=dev, generator=g) * torch.exp(torch.randn(C, device=dev, generator=g))).to(torch.bfloat16)
        W = (torch.randn(R, C, device=dev, generator=g) * 0.02).to(torch.bfloat16)
        org = torch.nn.functional.linear(x, W)

        def awq():
            return lcawq.awq_search_scale(x, W, wq, 'v2', org_out=org)
        awq()
        ms, (sc, best_i, losses) = timed(awq)
        n_tiles, units = lcdist.shard_units(20)
        res[f'awq_20_candidates_{name}'] = {'ms': ms, 'units_per_rank': len(units), 'row_tiles': n_tiles,
                                            'best_index': int(best_i), 'T': T, 'R': R, 'C': C,
                                            'TFLOPs_aggregate': 20 * 2.0 * T * R * C / (ms * 1e-3) / 1e12}
        del x, W, org
        torch.cuda.empty_cache()
    # AutoClip v1 of one linear, output channels dealt over the ranks, one all_gather of the ranges
    from lightcompress_b200.algo.autoclip import auto_clip_layer_sharded
    g = torch.Generator(device=dev).manual_seed(8)
    xc = torch.randn(512, 4096, device=dev, generator=g).to(torch.bfloat16)
    Wc = (torch.randn(14336, 4096, device=dev, generator=g) * 0.02).to(torch.bfloat16)
    auto_clip_layer_sharded(Wc, xc, wq, False, 512)
    ms, (bmx, bmn) = timed(lambda: auto_clip_layer_sharded(Wc, xc, wq, False, 512))
    res['autoclip_gate_proj_14336x4096'] = {'ms': ms, 'rows_per_rank': 14336 // world if 14336 % (world * 16) == 0 else 14336,
                                            'tokens': 512, 'checksum': float(bmx.float().sum())}
    del xc, Wc, bmx, bmn
    torch.cuda.empty_cache()
    # configs[4]: the 66 x 3 expert matrices of a DeepSeek-V2-Lite MoE block, FP8 E4M3 per-block real quant, expert e on rank
    # e mod world (rank-local results: a checkpoint writer takes them from the rank that owns the expert)
    from lightcompress_b200.algo import rtn_real_quantize_sharded
    from lightcompress_b200.quant import FloatQuantizer
    f8 = FloatQuantizer('e4m3', True, 'per_block', block_size=128, use_qtorch=True)
    g = torch.Generator(device=dev).manual_seed(9)
    weights = {}
    for e in range(66):
        for nm, (r, c) in (('gate_proj', (1408, 2048)), ('up_proj', (1408, 2048)), ('down_proj', (2048, 1408))):
            weights[f'experts.{e}.{nm}'] = (torch.randn(r, c, device=dev, generator=g, dtype=torch.float32) * 0.02).to(torch.bfloat16)
    nbytes = sum(w.numel() for w in weights.values()) * 3
    rtn_real_quantize_sharded(weights, f8, gather=False)
    ms, out = timed(lambda: rtn_real_quantize_sharded(weights, f8, gather=False))
    res['fp8_expert_sharding_deepseek_v2_lite_block'] = {'ms': ms, 'matrices_total': len(weights), 'matrices_this_rank': len(out),
                                                         'GBps_aggregate': nbytes / (ms * 1e-3) / 1e9}
    return res


def run_reference(args):
    """`--impl reference`: the reference's own CPU implementation of the path on the host cores — every step is the
    whole workload (7 weights, 872 MB of algorithmic traffic), rank 0 only."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    import torch
    fn, kind, what = cpu_arm()
    ws = cpu_block_weights(torch)
    nbytes = sum(w.numel() for w in ws) * 4
    for _ in range(max(1, min(args.warmup, 3))):
        cpu_reference_step(torch, ws, fn)
    times = [cpu_reference_step(torch, ws, fn) for _ in range(args.steps)]
    total = sum(times)
    v = nbytes * args.steps / total / 1e9
    cb = {'value': v, 'unit': 'GB/s', 'cores': _host_cores(), 'kind': kind,
          'sample': f'every step = the whole workload (7 weights, 218.1M bf16 elements); {what}; '
                    f'best step {nbytes / min(times) / 1e9:.2f} GB/s'}
    print(json.dumps({'impl': 'reference', 'metric': METRIC, 'value': v, 'unit': 'GB/s', 'n_gpus': args.gpus,
                      'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': total / args.steps * 1e3,
                      'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'bf16',
                      'data': 'synthetic', 'config': {'workload': WORKLOAD, 'device': f'host CPU, {cb["cores"]} threads'},
                      'cpu_baseline': cb,
                      'e2e': {'value': v, 'unit': 'GB/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='ours')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-extras', action='store_true')
    ap.add_argument('--sharded-only', action='store_true', help='of the extras, run only the `sharded` block')
    args = ap.parse_args()
    if args.impl == 'reference':
        return run_reference(args)

    if os.environ.get('LCB_BENCH_TRACE'):   # debugging aid: dump every thread's stack if the run stalls
        import faulthandler
        faulthandler.dump_traceback_later(int(os.environ['LCB_BENCH_TRACE']), exit=True, file=sys.stderr)
    import torch
    import torch.distributed as dist
    from lightcompress_b200 import _cabi
    from lightcompress_b200.quant import IntegerQuantizer

    _cabi.load()  # fail loudly when the extension is missing
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)

    ws = make_weights(torch, dev, seed=rank)
    numel = sum(w.numel() for w in ws)
    alg_bytes = numel * 4  # 2 B in + 2 B out per element
    q = IntegerQuantizer(4, False, 'per_group', group_size=128)

    def step():
        return q.fake_quant_weights_dynamic(ws)

    def step_per_tensor():
        return [q.fake_quant_weight_dynamic(w) for w in ws]

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local)
    sampler.start()   # nvidia-smi needs ~100 ms per sample: it runs from the warm-up to the end of the e2e leg
    for _ in range(max(args.warmup, 3)):
        step()
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    sync()
    ms = e0.elapsed_time(e1) / args.steps
    t = torch.tensor([ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    value = alg_bytes * world / (ms * 1e-3) / 1e9

    # end to end through the public API with HOST buffers: pinned H2D of the weights, fake-quant, D2H of the result
    hin = [torch.empty(w.shape, dtype=w.dtype).pin_memory() for w in ws]
    hout = [torch.empty(w.shape, dtype=w.dtype).pin_memory() for w in ws]
    for h, w in zip(hin, ws):
        h.copy_(w)

    # the package's host-buffer entry point (lightcompress_b200.host.HostFakeQuant): row chunks flow pinned host -> H2D ->
    # fused kernel -> D2H -> pinned host on three streams (PCIe is full duplex), every step moves all 436 MB each way
    from lightcompress_b200.host import HostFakeQuant
    pipe = HostFakeQuant(q, dev)

    e2e_steps = max(2, min(args.steps, 5))
    pipe.run(hin, hout)
    pipe.synchronize()
    sync()
    e0.record()
    pipe.begin()
    for _ in range(e2e_steps):
        pipe.run(hin, hout)
    pipe.end()
    e1.record()
    sync()
    pipe.synchronize()
    clocks = sampler.stop()
    t = torch.tensor([e0.elapsed_time(e1) / e2e_steps], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_ms = float(t.item())
    # the platform ceiling of that leg: the same bytes as plain pinned copies both ways at once, no kernel
    sys.path.insert(0, os.path.join(ROOT, 'tools'))
    import memcpy_ceiling
    sync()
    ceil_s, ceil_n = memcpy_ceiling.measure(dev, total_mb=(numel * 2) >> 20, chunk_mb=16, reps=3, barrier=sync)
    t = torch.tensor([ceil_s], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ceiling = 2 * ceil_n * world / float(t.item()) / 1e9
    e2e = {'value': alg_bytes * world / (e2e_ms * 1e-3) / 1e9, 'unit': 'GB/s', 'h2d_bytes_per_step': numel * 2,
           'raw_copy_ceiling_GBps': ceiling, 'frac_of_raw_copy_ceiling': alg_bytes * world / (e2e_ms * 1e-3) / 1e9 / ceiling,
           'd2h_bytes_per_step': numel * 2, 'ms_per_step': e2e_ms, 'api': 'lightcompress_b200.host.HostFakeQuant.run',
           'checked': bool(torch.equal(hout[0].view(torch.int16), q.fake_quant_weight_dynamic(ws[0]).cpu().view(torch.int16)))}

    hbm_peak, tf_peak, which, tf_burst = peaks()
    per_gpu = alg_bytes / (ms * 1e-3) / 1e9
    roofline = {'bound': 'hbm', 'kernel': 'fq_group_multi_kernel<bf16,NV=2,int,CHAIN_BF16_SMALL>', 'achieved': per_gpu,
                'peak': hbm_peak, 'peak_source': f'{which} (MEASURED_PEAKS.json hbm_gbs)', 'unit': 'GB/s',
                'frac': per_gpu / hbm_peak, 'frac_of_nominal_8TBs': per_gpu / 8000.0, 'traffic': None,
                'algorithmic_bytes_per_launch': 'numel*4 = 872.4 MB (bf16 in + bf16 out of 218.1M elements); 1 launch per step'}
    tr = os.path.join(ROOT, 'profiles', 'fakequant_traffic.json')
    if os.path.exists(tr):
        roofline['traffic'] = json.load(open(tr)).get('dram_bytes_per_launch')

    out = {'metric': METRIC, 'value': value, 'unit': 'GB/s', 'n_gpus': world, 'steps': args.steps,
           'warmup': max(args.warmup, 3), 'ms_per_step': ms, 'higher_is_better': True, 'scaling': 'weak',
           'vs_baseline': None, 'dtype': 'bf16', 'data': 'synthetic',
           'config': {'workload': WORKLOAD, 'l2': 'inputs (436 MB) + outputs (436 MB) per step exceed the 126 MB L2',
                      'parallelism': f'layers/blocks sharded x{world}, no collective'},
           'clocks': clocks, 'e2e': e2e, 'gpu_launches': args.steps, 'roofline': roofline}

    if rank == 0 and world == 1 and not args.no_extras and not args.sharded_only:   # single-GPU side numbers; at N > 1 every rank leaves together
        for _ in range(3):
            step_per_tensor()
        sync()
        e0.record()
        for _ in range(args.steps):
            step_per_tensor()
        e1.record()
        sync()
        pt_ms = e0.elapsed_time(e1) / args.steps
        out['extras'] = {'per_tensor_calls_GBps': alg_bytes / (pt_ms * 1e-3) / 1e9, 'per_tensor_calls_ms': pt_ms,
                         'per_tensor_calls_frac': alg_bytes / (pt_ms * 1e-3) / 1e9 / hbm_peak}
        # the same 7 reference-API calls recorded into a CUDA graph (what a deploy loop over many blocks would replay)
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            keep = step_per_tensor()
        gr.replay()
        sync()
        e0.record()
        for _ in range(args.steps):
            gr.replay()
        e1.record()
        sync()
        g_ms = e0.elapsed_time(e1) / args.steps
        out['extras'].update({'per_tensor_calls_graph_GBps': alg_bytes / (g_ms * 1e-3) / 1e9, 'per_tensor_calls_graph_ms': g_ms})
        del gr, keep
        del ws, hin, hout
        torch.cuda.empty_cache()
        out['extras'].update(extras(torch, dev, tf_peak, tf_burst, hbm_peak))
    if not args.no_extras:
        if world > 1:
            del ws, hin, hout
        torch.cuda.empty_cache()
        out['sharded'] = sharded_extras(torch, dist, dev, world, rank)
    if rank == 0 and world == 1 and not args.no_cpu_baseline and not args.sharded_only:
        out['cpu_baseline'] = cpu_baseline()[0]
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
        if rank == 0 and not args.no_extras:
            # END-TO-END s / block at N ranks (GPTQ: calibration samples sharded, one all-reduce of the packed Hessian per
            # distinct input; AWQ: candidate x row-tile units and AutoClip channels sharded), as a job of its own on the
            # same N GPUs: this rank launches it after the group is gone, the other ranks have left
            torch.cuda.empty_cache()
            out['sharded']['calib_e2e'] = nested_block_bench(world)
    if rank == 0:
        print(json.dumps(out), flush=True)


if __name__ == '__main__':
    main()
