"""Host-side stages of the NumPy-facing objective call (tuning aid): python tools/e2e_stages.py  [NF_SESSION_TIMING=1 prints the C side]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'vi-normflows_b200')]
import numpy as np, torch
import bench
from nfcore import semantics as sem
from nfcore.objective import SharedFlowObjective
W, U, B = bench.params_np()
params = sem.pack(W, U, B)
obj = SharedFlowObjective('p1', 32, 2, dtype='float32')
host = torch.randn(1 << 20, 2).pin_memory()
for n in (1024, 1 << 20):
    h = host[:n]
    for _ in range(5): obj.value_and_grad(params, h)
    T = [0.0] * 4
    reps = 50
    for _ in range(reps):
        t0 = time.perf_counter()
        lc64, p64 = obj._constants(params)
        t1 = time.perf_counter()
        res, nn, p64 = obj._host_eval(params, h, True)
        t2 = time.perf_counter()
        g = obj._chain(p64, res, nn)
        t3 = time.perf_counter()
        T[0] += t1 - t0; T[1] += t2 - t1; T[2] += t3 - t2
    print('N=%d  _constants %.1f us | _host_eval (incl. its own _constants) %.1f us | _chain %.1f us' % (n, T[0] / reps * 1e6, T[1] / reps * 1e6, T[2] / reps * 1e6))
