"""CPU, world_size 2 over gloo: the sample-sharding host logic (row ownership, the single
all-reduce of per-rank sums, division by the global count) reproduces the single-process result.
The local evaluator is the oracle -- the kernels themselves are exercised by the -m gpu tests."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT, PKG


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    import sys
    for p in (ROOT, PKG):
        if p not in sys.path:
            sys.path.insert(0, p)
    from nfcore import sharding
    from oracle import planar_torch as pt
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    rs = np.random.RandomState(0)
    K, D, N = 6, 2, 1001                      # odd N: ranks own 501 and 500 rows
    z0 = rs.randn(N, D)
    params = np.concatenate([0.1 + 0.05 * rs.randn(2 * K * D), 0.1 * np.ones(K)])
    lo, hi = sharding.shard_bounds(N, rank, world)
    p = torch.tensor(params, requires_grad=True)
    lq, lp, _ = pt.script_terms(p, torch.tensor(z0[lo:hi]), pt.p1, 1e-7, K)
    F_sum = torch.sum(lq) - torch.sum(lp)      # what the kernel returns: a SUM over local rows
    g_sum, = torch.autograd.grad(F_sum, p)
    sums = torch.cat([F_sum.detach().reshape(1), g_sum])
    n_global = sharding.reduce_sums(sums, hi - lo)
    q.put((rank, lo, hi, n_global, (sums / n_global).numpy()))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharded_sums_match_single_process():
    from nfcore import sharding
    from oracle import planar_torch as pt
    world, port = 2, _free_port()
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert [(r[1], r[2]) for r in res] == [(0, 501), (501, 1001)]
    assert all(r[3] == 1001 for r in res)
    rs = np.random.RandomState(0)
    K, D, N = 6, 2, 1001
    z0 = rs.randn(N, D)
    params = np.concatenate([0.1 + 0.05 * rs.randn(2 * K * D), 0.1 * np.ones(K)])
    F_ref, g_ref = pt.script_value_and_grad(params, z0, pt.p1, 1e-7, K)
    for r in res:
        assert abs(r[4][0] - F_ref) <= 1e-12 * max(1, abs(F_ref))
        np.testing.assert_allclose(r[4][1:], g_ref, rtol=1e-10, atol=1e-13)
    np.testing.assert_array_equal(res[0][4], res[1][4])       # every rank holds the same reduced vector


def test_shard_bounds_cover_all_rows():
    from nfcore import sharding
    for n in (0, 1, 7, 8, 1 << 20, (1 << 26) + 3):
        for world in (1, 2, 4, 8):
            b = [sharding.shard_bounds(n, r, world) for r in range(world)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1


def _count_worker(rank, world, port, q):
    import sys
    for p in (ROOT, PKG):
        if p not in sys.path:
            sys.path.insert(0, p)
    from nfcore import sharding
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    out = []
    for N in (9, 10, 9, 1, 0):                 # rank 0 owns 5 rows for N = 9 and for N = 10: its shard size alone says nothing
        lo, hi = sharding.shard_bounds(N, rank, world)
        sums = torch.full((3,), float(hi - lo), dtype=torch.float32)
        n = sharding.reduce_with_count(sums, hi - lo)
        out.append((N, n, sums.tolist()))
    q.put((rank, out))
    dist.barrier()
    dist.destroy_process_group()


def test_global_count_travels_with_every_exchange():
    """The global sample count is part of every exchange: a caller that changes N between calls gets the right
    normalisation on every rank and no rank skips a collective (a cache keyed on the local shard size would)."""
    world, port = 2, _free_port()
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    procs = [ctx.Process(target=_count_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, out in res:
        for N, n, sums in out:
            assert n == N
            assert sums == [float(N)] * 3
