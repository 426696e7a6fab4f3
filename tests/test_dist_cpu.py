"""Host logic of the sharded paths on gloo, world_size 2 (CPU): result-only collectives of SURVEY.md §8e."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _plain(o):
    if isinstance(o, torch.Tensor):
        return ('__t__', o.numpy().copy())
    if isinstance(o, dict):
        return {k: _plain(v) for k, v in o.items()}
    if isinstance(o, (list, tuple)):
        return type(o)(_plain(v) for v in o)
    return o


def _unplain(o):
    if isinstance(o, tuple) and len(o) == 2 and isinstance(o[0], str) and o[0] == '__t__':
        return torch.from_numpy(o[1])
    if isinstance(o, dict):
        return {k: _unplain(v) for k, v in o.items()}
    if isinstance(o, (list, tuple)):
        return type(o)(_unplain(v) for v in o)
    return o


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from lightcompress_b200.algo import dist as lcdist
    out = {}
    # token-sharded Hessian: S_r = X_r^T X_r, one all-reduce, global batch count
    g = torch.Generator().manual_seed(0)
    X = torch.randn(4, 64, 16, generator=g)
    mine = X[rank::world]
    S = sum(x.t() @ x for x in mine)
    n = lcdist.allreduce_hessian(S, mine.shape[0])
    out['n'] = n
    out['H'] = S * (2.0 / n)
    # candidate sharding + MINLOC with ties -> lowest index
    cands = lcdist.shard_candidates(20)
    out['cands'] = cands
    losses_all = [abs(i - 7) * 0.5 + (0.0 if i != 11 else -2.0) for i in range(20)]   # min at 7 (0.0) and 11 (0.0): tie
    out['best'] = lcdist.argmin_across_ranks([losses_all[i] for i in cands], cands, torch.device('cpu'))
    t = torch.tensor([float(rank), 5.0 - rank])
    out['max'] = lcdist.allreduce_max(t)
    # (candidate, row tile) work units: every unit on exactly one rank, partial sums all-reduced
    nt, units = lcdist.shard_units(20)
    out['units'] = (nt, units)
    sums = torch.zeros(20, dtype=torch.float64)
    for n, tile in units:
        sums[n] += (n + 1) * 10.0 + tile
    out['sums'] = lcdist.reduce_sum(sums)
    nt3, units3 = lcdist.shard_units(3)     # 3 candidates on 2 ranks: 2 row tiles, 3 units each
    out['units3'] = (nt3, units3)
    # AutoClip: output channels dealt over the ranks, results gathered (the search itself is injected)
    from lightcompress_b200.algo.autoclip import auto_clip_layer_sharded
    w = torch.arange(64 * 8, dtype=torch.float32).reshape(64, 8)
    seen = []

    def fake_search(w_, x_, q_, sym_, n_sample_token=None):
        seen.append(tuple(w_.shape))
        return w_.amax(1).reshape(-1, 1, 1), w_.amin(1).reshape(-1, 1, 1)
    out['clip'] = auto_clip_layer_sharded(w, None, None, False, search=fake_search)
    out['clip_rows'] = seen
    # expert sharding: weight i on rank i mod world, results gathered on every rank (fp8 codes travel as bytes)
    from lightcompress_b200.algo import rtn_real_quantize_sharded
    weights = {f'experts.{i}.w': torch.full((4, 8), float(i + 1)) for i in range(5)}
    calls = []

    def fake_quantize(w):
        calls.append(float(w[0, 0]))
        return (w / 8).to(torch.float8_e4m3fn), w.amax(dim=1, keepdim=True), None
    res = rtn_real_quantize_sharded(weights, None, quantize=fake_quantize)
    out['expert_calls'] = calls
    out['experts'] = {n: (c.float(), s, z) for n, (c, s, z) in res.items()}
    out['local_only'] = sorted(rtn_real_quantize_sharded(weights, None, quantize=fake_quantize, gather=False))
    q.put((rank, _plain(out)))      # numpy, not torch storages: no fd hand-over that races with this process exiting
    dist.barrier()
    dist.destroy_process_group()


def test_result_only_collectives_world2():
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = {r: _unplain(o) for r, o in (q.get(timeout=120) for _ in ps)}
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    g = torch.Generator().manual_seed(0)
    X = torch.randn(4, 64, 16, generator=g)
    H = sum(x.t() @ x for x in X) * (2.0 / 4)
    for r in range(2):
        assert res[r]['n'] == 4.0
        assert torch.allclose(res[r]['H'], H, rtol=1e-6, atol=1e-6)
        assert res[r]['cands'] == list(range(r, 20, 2))
        assert res[r]['best'] == (0.0, 7)          # tie between 7 (rank 1) and 11 (rank 1)... lowest index wins
        assert torch.equal(res[r]['max'], torch.tensor([1.0, 5.0]))
        assert res[r]['units'] == (1, [(n, 0) for n in range(r, 20, 2)])
        assert torch.equal(res[r]['sums'], torch.tensor([(n + 1) * 10.0 for n in range(20)], dtype=torch.float64))
        assert res[r]['units3'][0] == 2 and len(res[r]['units3'][1]) == 3
        w = torch.arange(64 * 8, dtype=torch.float32).reshape(64, 8)
        assert res[r]['clip_rows'] == [(32, 8)]                      # half the rows per rank
        assert torch.equal(res[r]['clip'][0].reshape(-1), w.amax(1)) and torch.equal(res[r]['clip'][1].reshape(-1), w.amin(1))
        assert res[r]['expert_calls'][:3 - r] == [float(i + 1) for i in range(r, 5, 2)]     # only its own experts
        assert sorted(res[r]['experts']) == [f'experts.{i}.w' for i in range(5)]
        for i in range(5):
            c, s_, z = res[r]['experts'][f'experts.{i}.w']
            assert torch.equal(c, torch.full((4, 8), (i + 1) / 8.0)) and z is None
            assert torch.equal(s_, torch.full((4, 1), float(i + 1)))
        assert res[r]['local_only'] == [f'experts.{i}.w' for i in range(r, 5, 2)]


def test_single_process_is_identity():
    from lightcompress_b200.algo import dist as lcdist
    S = torch.eye(3)
    assert lcdist.allreduce_hessian(S, 5) == 5 and torch.equal(S, torch.eye(3))
    assert lcdist.shard_candidates(4) == [0, 1, 2, 3]
    assert lcdist.shard_units(3) == (1, [(0, 0), (1, 0), (2, 0)])
    assert lcdist.argmin_across_ranks([3.0, 1.0, 1.0], [0, 1, 2], torch.device('cpu')) == (1.0, 1)
