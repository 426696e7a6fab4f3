"""Result-only collectives of the sharded paths. Works on any initialised torch.distributed backend
(NCCL over NVLink on the B200 box, gloo in the CPU tests); every helper is the identity when no group is initialised."""
import torch
import torch.distributed as dist


def world():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def allreduce_hessian(S, nsamples):
    """Sum the un-normalised Hessians and the batch counts of all ranks (ONE collective per linear; replaces the
    per-batch all-reduce + `/= world_size` of gptq.py:292-295). Returns the global batch count."""
    rank, ws = world()
    if ws == 1:
        return nsamples
    n = torch.tensor([float(nsamples)], dtype=torch.float64, device=S.device)
    dist.all_reduce(S, op=dist.ReduceOp.SUM)
    dist.all_reduce(n, op=dist.ReduceOp.SUM)
    return float(n.item())


def finalize_sharded_hessian(S, nsamples):
    """H = (2 / total batches) * sym(sum over ranks of S), in place, for a token-sharded Hessian on CUDA devices: the
    maintained half of S (upper 128-blocks) and the batch count travel in ONE all-reduce (ops.hessian_pack), the scale is
    applied with the count read on the device (no second collective, no host sync). Returns S (now H)."""
    from .. import ops
    rank, ws = world()
    if ws == 1:
        return ops.hessian_finalize(S, nsamples)
    packed = ops.hessian_pack(S, nsamples)
    dist.all_reduce(packed, op=dist.ReduceOp.SUM)
    return ops.hessian_unpack_finalize(S, packed)


def shard_units(n_cand, n_tiles=None):
    """Work units (candidate, row tile) of this rank for a grid search whose loss is a SUM over output rows: with
    n_tiles = world / gcd(n_cand, world) row tiles the n_cand * n_tiles units divide evenly over the ranks (20 candidates
    on 8 ranks: 2 tiles, 5 units each — candidate-only sharding gives 3 / 2 / 2 ...). Returns (n_tiles, [(cand, tile)])."""
    import math
    rank, ws = world()
    if n_tiles is None:
        n_tiles = ws // math.gcd(n_cand, ws)
    units = [(u // n_tiles, u % n_tiles) for u in range(rank, n_cand * n_tiles, ws)]
    return n_tiles, units


def reduce_sum(t):
    rank, ws = world()
    if ws > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t


def shard_candidates(n_cand):
    """Candidate indices of this rank: n = rank (mod world)."""
    rank, ws = world()
    return list(range(rank, n_cand, ws))


def argmin_across_ranks(losses, indices, device):
    """Global (loss, index) minimum over the candidates each rank evaluated. Ties -> lowest index, which mimics the strict
    `<` of the sequential search (awq.py:245). losses / indices: python lists of this rank."""
    rank, ws = world()
    best_l, best_i = float('inf'), -1
    for l, i in zip(losses, indices):
        if l < best_l or (l == best_l and i < best_i):
            best_l, best_i = l, i
    if ws == 1:
        if best_i < 0:
            raise RuntimeError('argmin_across_ranks: every candidate loss is NaN / inf (the reference ends with '
                               'best_scales = None, awq.py:245-248, and fails on its next use)')
        return best_l, best_i
    t = torch.tensor([best_l, float(best_i if best_i >= 0 else 1 << 30)], dtype=torch.float64, device=device)
    gathered = [torch.zeros_like(t) for _ in range(ws)]
    dist.all_gather(gathered, t)
    for g in gathered:
        l, i = float(g[0]), int(g[1])
        if l < best_l or (l == best_l and i < best_i):
            best_l, best_i = l, i
    if best_i < 0 or best_i >= (1 << 30):
        raise RuntimeError('argmin_across_ranks: every candidate loss on every rank is NaN / inf')
    return best_l, best_i


def allreduce_max(t):
    rank, ws = world()
    if ws > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t
