"""Where does a flow-VAE gradient step (BASELINE C4) spend its time?  (tuning aid)"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'vi-normflows_b200'), os.path.join(ROOT, 'tests')]
import numpy as np, torch
from test_gpu_vae import _mnist_closures
from normflows import optimization
K, D, Bsz = 16, 40, 4096
enc, dec, encode, decode, unpack, logp = _mnist_closures(K, D)
X = (np.random.RandomState(2).rand(Bsz, 784) < 0.13).astype(np.float32)
p = np.random.RandomState(3).randn(enc.D + dec.D) * 0.05
F = optimization.make_free_energy(logp, X, D, K, encode, decode, 10000, Bsz, dtype=torch.float32)
objective, gradient = optimization.gradient_create(F, D, Bsz, unpack, dtype=torch.float32)
for _ in range(3): gradient(p, 0)
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(5): gradient(p, 0)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by='cuda_time_total', row_limit=22, max_name_column_width=60))
print(prof.key_averages().table(sort_by='cpu_time_total', row_limit=12, max_name_column_width=60))
