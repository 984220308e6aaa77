"""BASELINE config 5: K = 64 planar layers, D in {2, 16, 64, 256}, 2^26 Monte-Carlo samples in total, sharded over the
ranks of one box (strong scaling), one exchange of the 3+2KD+2K result sums per step.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 --master-port P tools/scale_c5.py [--log2n 26] [--steps 3]
    python tools/scale_c5.py                      (one GPU)

Per D one line.  `ms_per_step` is device time: CUDA events around `steps` enqueued steps of (objective launches +
exchange), max over ranks -- what an on-device training loop sees.  `ms_per_call` is the same step through the
NumPy-facing call (host parameter algebra, upload of the layer constants, read-back of F and the gradient, one
synchronisation per step).  `exchange_share` compares with the same launches without the exchange.  Targets as SURVEY.md section 8d: D = 2 the potential U1, D > 2 a
diagonal Gaussian through log_mvn math.  Run it at G = 1, 2, 4, 8 and divide: tools/scale_c5_table.py does that.
"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'vi-normflows_b200')]
import numpy as np, torch, torch.distributed as dist
from nfcore import sharding
from nfcore.objective import SharedFlowObjective

ap = argparse.ArgumentParser()
ap.add_argument('--log2n', type=int, default=26)
ap.add_argument('--steps', type=int, default=3)
ap.add_argument('--dims', default='2,16,64,256')
args = ap.parse_args()
world, rank, local = int(os.environ.get('WORLD_SIZE', '1')), int(os.environ.get('RANK', '0')), int(os.environ.get('LOCAL_RANK', '0'))
torch.cuda.set_device(local)
dev = torch.device('cuda', local)
if world > 1:
    dist.init_process_group('nccl', device_id=dev)
group = dist.group.WORLD if world > 1 else None
K, N = 64, 1 << args.log2n
for D in [int(x) for x in args.dims.split(',')]:
    rs = np.random.RandomState(5)
    W = (0.1 + 0.01 * rs.randn(K, D)) / np.sqrt(D / 2); U = (0.1 + 0.01 * rs.randn(K, D)) / np.sqrt(D / 2); B = 0.1 * np.ones(K)
    params = np.concatenate([W.ravel(), U.ravel(), B])
    target = 'p1' if D == 2 else ('gaussdiag', rs.uniform(-1, 1, D), rs.uniform(-1, 1, D))
    lo, hi = sharding.shard_bounds(N, rank, world)
    n_loc = hi - lo
    gen = torch.Generator(device=dev).manual_seed(4000 + rank)
    eps = torch.empty(n_loc, D, dtype=torch.float32, device=dev)
    step_rows = max(1, (1 << 28) // D)
    for a in range(0, n_loc, step_rows):                       # generated in pieces: randn's own temporaries stay small
        eps[a:a + step_rows].normal_(generator=gen)
    obj = SharedFlowObjective(target, K, D, dtype='float32', process_group=group)
    solo = SharedFlowObjective(target, K, D, dtype='float32')             # same launches, no exchange
    for o in (obj, solo):
        o.value_and_grad(params, eps)
    times = {}
    for name, o in (('sharded', obj), ('local', solo)):
        for mode in ('device', 'call'):
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(args.steps):
                if mode == 'call':
                    F, g = o.value_and_grad(params, eps)
                else:
                    o._launch(params, eps, True)
                    if o.group is not None and not o._exchanged:
                        o._exchange_only(eps.shape[0])
            e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / args.steps], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            times[name, mode] = float(t.item())
    if rank == 0:
        exch = 'kernel last block (NVLink peer)' if obj._small_ar is not None and D <= 2 else ('peer kernel' if obj._small_ar is not None else ('NCCL' if world > 1 else 'none'))
        print(json.dumps({'config': 'C5', 'K': K, 'D': D, 'log2_samples_total': args.log2n, 'n_gpus': world, 'ms_per_step': times['sharded', 'device'],
                          'samples_per_s': N / times['sharded', 'device'] * 1e3, 'ms_per_call': times['sharded', 'call'],
                          'ms_local_kernels_only': times['local', 'device'],
                          'exchange_share': max(0.0, 1 - times['local', 'device'] / times['sharded', 'device']), 'exchange': exch,
                          'result_floats': 3 + 2 * K * D + 2 * K, 'F': float(F)}), flush=True)
    del eps, obj, solo
    torch.cuda.empty_cache()
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
