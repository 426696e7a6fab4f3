"""Platform ceiling of the host-buffer leg (bench.py `e2e`): plain pinned cudaMemcpyAsync host -> device and device -> host
running CONCURRENTLY on two streams (one cudaMemcpyAsync per chunk), no kernel — what HostFakeQuant could reach at best.

    python tools/memcpy_ceiling.py [--mb 436] [--chunk-mb 16]            one GPU
    torchrun --nproc-per-node N tools/memcpy_ceiling.py                  N GPUs at once (shared host memory / PCIe roots)

Prints one JSON line per rank 0: per-direction GB/s of this rank (max-over-ranks time) and the aggregate over the ranks,
in the `e2e` unit (algorithmic bytes = H2D + D2H bytes)."""
import json
import os
import sys

import torch


def measure(dev, total_mb=436, chunk_mb=16, reps=5, both=True, barrier=None):
    """`barrier` (a callable) runs after the buffers exist and the copies are warm, right before the clock starts: with
    several ranks the pinned allocations (hundreds of ms, serialised by the OS) would otherwise stagger the ranks and the
    copies of different GPUs would not contend."""
    n = total_mb << 20
    c = chunk_mb << 20
    hin = torch.empty(n, dtype=torch.uint8).pin_memory()
    hout = torch.empty(n, dtype=torch.uint8).pin_memory()
    din = torch.empty(n, dtype=torch.uint8, device=dev)
    dout = torch.empty(n, dtype=torch.uint8, device=dev)
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def once():
        for o in range(0, n, c):
            with torch.cuda.stream(s1):
                din[o:o + c].copy_(hin[o:o + c], non_blocking=True)
            if both:
                with torch.cuda.stream(s2):
                    hout[o:o + c].copy_(dout[o:o + c], non_blocking=True)
    once()
    torch.cuda.synchronize(dev)
    if barrier is not None:
        barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    cur = torch.cuda.current_stream(dev)
    e0.record(cur)
    s1.wait_event(e0)
    s2.wait_event(e0)
    for _ in range(reps):
        once()
    cur.wait_stream(s1)
    cur.wait_stream(s2)
    e1.record(cur)
    torch.cuda.synchronize(dev)
    return e0.elapsed_time(e1) * 1e-3 / reps, n


def main():
    argv = sys.argv[1:]
    mb = int(argv[argv.index('--mb') + 1]) if '--mb' in argv else 436
    chunk = int(argv[argv.index('--chunk-mb') + 1]) if '--chunk-mb' in argv else 16
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group('nccl', device_id=dev)
    res = {}
    for name, both in (('h2d_only', False), ('h2d_and_d2h', True)):
        def bar():
            if world > 1:
                dist.barrier()
                torch.cuda.synchronize(dev)
        t, n = measure(dev, mb, chunk, both=both, barrier=bar)
        tt = torch.tensor([t], device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t = float(tt.item())
        moved = n * (2 if both else 1)
        res[name] = {'ms': t * 1e3, 'GBps_per_rank': moved / t / 1e9, 'GBps_aggregate': moved * world / t / 1e9}
    if rank == 0:
        print(json.dumps({'n_gpus': world, 'mb': mb, 'chunk_mb': chunk, **res}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
