"""Wide-D free energy end to end from host samples (SharedFlowObjective.value_and_grad on a NumPy array, i.e. the
host-buffer session): python tools/gram_e2e.py D logN [K].  NF_NO_GRAM=1 forces the direct kernel."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'vi-normflows_b200'))
import numpy as np
import torch
from nfcore.objective import SharedFlowObjective

D, logN = int(sys.argv[1]), int(sys.argv[2])
K = int(sys.argv[3]) if len(sys.argv) > 3 else 64
N = 1 << logN
rs = np.random.RandomState(0)
z0 = torch.from_numpy(rs.randn(N, D).astype(np.float32)).pin_memory().numpy()
sc = 1.0 / np.sqrt(D / 2)
params = np.concatenate([sc * 0.5 * rs.randn(2 * K * D), 0.3 * rs.randn(K)])
mu, ls = 0.2 * rs.randn(D), 0.3 * rs.randn(D) - 4.0 / D
obj = SharedFlowObjective(('gaussdiag', mu, ls), K, D, dtype='float32')
for _ in range(2):
    F, g = obj.value_and_grad(params, z0)
t0 = time.perf_counter()
n = 5
for _ in range(n):
    F, g = obj.value_and_grad(params, z0)
dt = (time.perf_counter() - t0) / n
print('D=%d K=%d N=2^%d  %.2f ms per value_and_grad from host samples (%.1f MB of samples)  %.3e samples/s  F=%.6f' % (
    D, K, logN, dt * 1e3, N * D * 4 / 1e6, N / dt, F))
