"""Per-call time of the host-buffer objective call over the first 600 calls (PCIe link ramp; tuning aid)."""
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
host = [torch.randn(1 << 20, 2).pin_memory() for _ in range(4)]
ts = []
for i in range(600):
    t0 = time.perf_counter(); obj.value_and_grad(params, host[i % 4]); ts.append(time.perf_counter() - t0)
    if i == 299: time.sleep(0.5)          # an idle gap in the middle
ts = np.array(ts) * 1e6
for a in range(0, 600, 20):
    print('calls %3d-%3d: mean %.0f us  min %.0f  max %.0f' % (a, a + 19, ts[a:a + 20].mean(), ts[a:a + 20].min(), ts[a:a + 20].max()))
