"""Under torchrun (NCCL): the sample-sharded objective equals the single-GPU objective.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/dist_check.py
"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'vi-normflows_b200')]
import numpy as np, torch, torch.distributed as dist
from nfcore import sharding
from nfcore.objective import SharedFlowObjective

rank, world = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
torch.cuda.set_device(int(os.environ['LOCAL_RANK']))
dist.init_process_group('nccl', device_id=torch.device('cuda', int(os.environ['LOCAL_RANK'])))
rs = np.random.RandomState(0)
ok = True
for (K, D, N, target) in ((32, 2, 200003, 'p1'), (32, 2, 200003 * world, 'p1'), (64, 2, 333333 * world, 'p4'),
                          (8, 16, 5001, ('gaussdiag', np.zeros(16), np.zeros(16) - 0.2))):
    z0 = rs.randn(N, D)
    params = np.concatenate([(0.1 + 0.02 * rs.randn(2 * K * D)) / np.sqrt(D / 2), 0.1 * np.ones(K)])
    for dtype, tol in (('float64', 1e-11), ('float32', 1e-5)):
        full = SharedFlowObjective(target, K, D, dtype=dtype)
        F1, g1 = full.value_and_grad(params, z0)
        lo, hi = sharding.shard_bounds(N, rank, world)
        sh = SharedFlowObjective(target, K, D, dtype=dtype, process_group=dist.group.WORLD)
        F2, g2 = sh.value_and_grad(params, z0[lo:hi])
        eF = abs(F1 - F2) / max(1, abs(F1)); eg = np.abs(g1 - g2).max() / np.abs(g1).max()
        good = eF <= tol and eg <= tol
        ok &= good
        if rank == 0:
            print('K=%d D=%d N=%d %s world=%d: |dF|=%.2e |dg|=%.2e %s' % (K, D, N, dtype, world, eF, eg, 'OK' if good else 'FAIL'), flush=True)
# the device optimisation loop, sample-sharded: every rank trains on its shard, all end with the parameters one GPU finds
K, D, N, T = 8, 2, 4096 * world + 12, 40
eps_all = rs.randn(T, N, D).astype(np.float32)
init = np.concatenate([0.1 * np.ones(2 * K * D), 0.1 * np.ones(K)])
solo = SharedFlowObjective('p1', K, D, dtype='float32')
for optimizer in ('adam', 'rmsprop'):
    x1, tr1 = solo.train(init, eps_all, optimizer=optimizer, step_size=1e-3)
    lo, hi = sharding.shard_bounds(N, rank, world)
    sh = SharedFlowObjective('p1', K, D, dtype='float32', process_group=dist.group.WORLD)
    x2, tr2 = sh.train(init, np.ascontiguousarray(eps_all[:, lo:hi]), optimizer=optimizer, step_size=1e-3)
    ex, et = np.abs(x1 - x2).max(), np.abs(tr1 - tr2).max()
    allx = [None] * world
    dist.all_gather_object(allx, x2.tobytes())
    same = all(a == allx[0] for a in allx)
    good = ex <= 2e-5 and et <= 2e-5 and same and np.abs(x1 - init).max() > 1e-3
    ok &= good
    if rank == 0:
        print('sharded device loop %s, %d steps, world=%d: |dx|=%.2e |dF trace|=%.2e identical on all ranks=%s %s'
              % (optimizer, T, world, ex, et, same, 'OK' if good else 'FAIL'), flush=True)
# the NVLink one-shot exchange against NCCL, repeated so that both slot sets and the sequence logic are exercised
from nfcore.peer import SmallAllReduce
ar = SmallAllReduce(dist.group.WORLD)
g = torch.Generator(device='cuda').manual_seed(100 + rank)
for it in range(50):
    v = torch.randn(195, device='cuda', generator=g)
    ref = v.clone(); dist.all_reduce(ref)
    out = ar(v.clone())
    torch.cuda.synchronize()
    err = float((out - ref).abs().max())
    if err > 1e-5:
        ok = False
        print('rank %d it %d peer all-reduce mismatch %.3e' % (rank, it, err), flush=True)
ar.check()
# the in-place form that carries the sample count (what the objective uses when the kernel did not do the exchange)
for it in range(6):
    n = 195 + it
    v = torch.randn(n + 3, device='cuda', generator=g)
    ref = v[:n].clone(); dist.all_reduce(ref)
    ar.counted(v, n, 1000 + rank + it)
    torch.cuda.synchronize()
    tail = v[n:].cpu().numpy()
    cnt = ar.global_count(tail)
    if float((v[:n] - ref).abs().max()) > 1e-5 or cnt != sum(1000 + r + it for r in range(world)):
        ok = False
        print('rank %d counted exchange mismatch (count %d)' % (rank, cnt), flush=True)
if rank == 0:
    print('peer all-reduce (50 rounds) %s' % ('OK' if ok else 'FAIL'), flush=True)
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
