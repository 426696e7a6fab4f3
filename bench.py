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
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d['hbm_gbs']), float(d.get('bf16_tflops_sustained', d.get('bf16_tflops', 1400.0))), 'measured'
    return 6650.0, 1590.0, 'fallback'


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


def cpu_baseline(steps=3, sample=('q_proj', 4096, 4096)):
    """The oracle port (oracle/lc_oracle.c, OpenMP over all host cores) on a bounded sample of the workload."""
    import ctypes
    import torch
    from oracle import quant_oracle as qo
    cores = len(os.sched_getaffinity(0)) if hasattr(os, 'sched_getaffinity') else os.cpu_count()
    try:   # torchrun exports OMP_NUM_THREADS=1: the CPU arm is entitled to every host core it can use
        ctypes.CDLL('libgomp.so.1').omp_set_num_threads(cores)
    except OSError:
        pass
    _, r, c = sample
    x = (torch.randn(r, c, generator=torch.Generator().manual_seed(0)) * 0.02).to(torch.bfloat16)
    kw = dict(kind=0, sym=False, qmin=0.0, qmax=15.0)
    qo.fakequant(x, r, c, 1, 128, **kw)
    best = 1e30
    for _ in range(steps):
        t0 = time.perf_counter()
        qo.fakequant(x, r, c, 1, 128, **kw)
        best = min(best, time.perf_counter() - t0)
    return {'value': r * c * 4 / best / 1e9, 'unit': 'GB/s', 'cores': cores, 'kind': 'port',
            'sample': f'{sample[0]} [{r},{c}] bf16 w4 asym g128, best of {steps} (oracle/lc_oracle.c, OpenMP)'}, best


def extras(torch, dev, tf_peak):
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
                      'hw_syrk_TFLOPs': 2.0 * T * C * C / t / 1e12 * 33 / 64, 'frac_of_bf16_peak_hw': 2.0 * T * C * C / t / 1e12 * 33 / 64 / tf_peak}
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
                            'frac_of_bf16_peak': 2.0 * T * C * C / t / 1e12 / tf_peak}
    return res


def sharded_extras(torch, dist, dev, world, rank):
    """N > 1, every rank participates: the sharded tensor paths of SURVEY.md §8e at Llama-3-8B shapes.
    Hessian: the 128 x 2048 calibration tokens of one linear input are split over the ranks, every rank accumulates its
    shard into its own S (tcgen05 SYRK), ONE all-reduce of S over NCCL, then the finalize kernel — checked against the
    batch count. AWQ: the 20 grid candidates of a 4096 x 4096 linear are split over the ranks (n = rank mod world), the
    argmin is reduced. Times are device times, max over ranks."""
    from lightcompress_b200 import ops
    from lightcompress_b200.algo import dist as lcdist
    from lightcompress_b200.quant import IntegerQuantizer
    res = {}

    def timed(fn):
        dist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        out = fn()
        b.record()
        torch.cuda.synchronize()
        t = torch.tensor([a.elapsed_time(b)], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), out

    T, batches = 2048, 128
    for C in (4096, 14336):
        g = torch.Generator(device=dev).manual_seed(100 + rank)
        xs = [torch.randn(T, C, device=dev, generator=g).to(torch.bfloat16) for _ in range(2)]
        mine = len(range(rank, batches, world))

        def hess():
            S = torch.zeros(C, C, device=dev)
            for i in range(mine):
                ops.hessian_accum(S, xs[i & 1])
            n = lcdist.allreduce_hessian(S, mine)
            return ops.hessian_finalize(S, n), n
        hess()
        ms, (H, n) = timed(hess)
        res[f'hessian_C{C}'] = {'ms': ms, 'batches_total': int(n), 'batches_per_rank': mine,
                                'gemm_equiv_TFLOPs_aggregate': 2.0 * T * batches * C * C / (ms * 1e-3) / 1e12}
        assert int(n) == batches
        del H, xs
        torch.cuda.empty_cache()
    C, T = 4096, 8192
    g = torch.Generator(device=dev).manual_seed(7)      # identical data on every rank
    x = torch.randn(T, C, device=dev, generator=g).to(torch.bfloat16)
    W = (torch.randn(C, C, device=dev, generator=g) * 0.02).to(torch.bfloat16)
    org = torch.nn.functional.linear(x, W)
    cand = (torch.rand(20, C, device=dev, generator=g) + 0.5).to(torch.bfloat16)
    wq = IntegerQuantizer(4, False, 'per_group', group_size=128)
    idx = lcdist.shard_candidates(20)

    def awq():
        losses = ops.awq_candidates(W, x, org, cand[idx], wq)
        return lcdist.argmin_across_ranks([float(v) for v in losses.tolist()], idx, dev)
    awq()
    ms, (best_l, best_i) = timed(awq)
    res['awq_20_candidates'] = {'ms': ms, 'candidates_per_rank': len(idx), 'best_index': int(best_i), 'T': T, 'R': C, 'C': C}
    return res


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    total = 0.0
    for _ in range(args.warmup):
        cpu_baseline(1)
    vals = []
    for _ in range(args.steps):
        cb, t = cpu_baseline(1)
        vals.append(cb['value'])
        total += t
    v = sum(vals) / len(vals)
    cb['value'] = v
    print(json.dumps({'impl': 'reference', 'metric': METRIC, 'value': v, 'unit': 'GB/s', 'n_gpus': args.gpus,
                      'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': total / args.steps * 1e3,
                      'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'bf16',
                      'data': 'synthetic', 'config': {'workload': WORKLOAD, 'sample': cb['sample']},
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
    e2e = {'value': alg_bytes * world / (e2e_ms * 1e-3) / 1e9, 'unit': 'GB/s', 'h2d_bytes_per_step': numel * 2,
           'd2h_bytes_per_step': numel * 2, 'ms_per_step': e2e_ms, 'api': 'lightcompress_b200.host.HostFakeQuant.run',
           'checked': bool(torch.equal(hout[0].view(torch.int16), q.fake_quant_weight_dynamic(ws[0]).cpu().view(torch.int16)))}

    hbm_peak, tf_peak, which = peaks()
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

    if rank == 0 and world == 1 and not args.no_extras:   # single-GPU side numbers; at N > 1 every rank leaves together
        for _ in range(3):
            step_per_tensor()
        sync()
        e0.record()
        for _ in range(args.steps):
            step_per_tensor()
        e1.record()
        sync()
        pt_ms = e0.elapsed_time(e1) / args.steps
        out['extras'] = {'per_tensor_calls_GBps': alg_bytes / (pt_ms * 1e-3) / 1e9, 'per_tensor_calls_ms': pt_ms}
        del ws, hin, hout
        torch.cuda.empty_cache()
        out['extras'].update(extras(torch, dev, tf_peak))
    if world > 1 and not args.no_extras:
        del ws, hin, hout
        torch.cuda.empty_cache()
        out['sharded'] = sharded_extras(torch, dist, dev, world, rank)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        out['cpu_baseline'] = cpu_baseline()[0]
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
