"""Where does the host-buffer call spend its time?  (tuning aid)"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'vi-normflows_b200')]
import numpy as np, torch
import bench
from nfcore import semantics as sem
from nfcore.objective import SharedFlowObjective

dev = torch.device('cuda:0')
N = 1 << 20
host = torch.randn(N, 2).pin_memory()
d = torch.empty(N, 2, device=dev)
for _ in range(3): d.copy_(host, non_blocking=True)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(20): d.copy_(host, non_blocking=True)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 20
print('H2D 8 MiB pinned: %.1f us  (%.1f GB/s)' % (dt * 1e6, 8.388608e6 / dt / 1e9))
W, U, B = bench.params_np()
params = sem.pack(W, U, B)
obj = SharedFlowObjective('p1', 32, 2, dtype='float32')
for n in (1024, 1 << 18, 1 << 20):
    h = host[:n]
    for _ in range(3): obj.value_and_grad(params, h)
    t0 = time.perf_counter()
    for _ in range(50): obj.value_and_grad(params, h)
    dt = (time.perf_counter() - t0) / 50
    print('value_and_grad N=%8d: %.1f us/call -> %.3e samples/s' % (n, dt * 1e6, n / dt))
