"""BASELINE config 4 (flow-VAE, batch 4096, D = 40, K = 16, H = 64, L = 3): time per optimisation step.
    host loop   : autograd.misc.optimizers.adam(gradient, ...) -- parameters and gradient cross PCIe every step (round 1)
    device loop : optimization.optimize(...) -- parameters, Adam moments and the update stay on the device
    device loop + device draws : the same with NORMFLOWS_DEVICE_RNG (z0 drawn on the device instead of 164 k host randn)
"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'vi-normflows_b200'), os.path.join(ROOT, 'tests')]
import numpy as np, torch
from test_gpu_vae import _mnist_closures
from autograd.misc.optimizers import adam
from normflows import optimization
K, D, Bsz, T = 16, 40, 4096, 60
enc, dec, encode, decode, unpack, logp = _mnist_closures(K, D)
X = (np.random.RandomState(2).rand(4 * Bsz, 784) < 0.13).astype(np.float32)
p0 = np.random.RandomState(3).randn(enc.D + dec.D) * 0.05
optimization.results = type(optimization.results)('/tmp')

def host_loop():
    F = optimization.make_free_energy(logp, X, D, K, encode, decode, T, Bsz, dtype=torch.float32)
    _, gradient = optimization.gradient_create(F, D, 4 * Bsz, unpack, dtype=torch.float32)
    adam(gradient, p0, step_size=1e-3, num_iters=T)

def device_loop():
    optimization.optimize(logp, X, D, K, 4 * Bsz, p0, unpack, encode, decode, T, Bsz, 1e-3, verbose=False)

for name, fn, env in (('host loop', host_loop, None), ('device loop', device_loop, None), ('device loop + device draws', device_loop, '11')):
    if env:
        os.environ['NORMFLOWS_DEVICE_RNG'] = env
    else:
        os.environ.pop('NORMFLOWS_DEVICE_RNG', None)
    fn()                                    # warm-up (allocator, attribute set-up)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    fn()
    torch.cuda.synchronize()
    print('%-28s %.2f ms per step (%d steps, batch %d)' % (name, (time.perf_counter() - t0) / T * 1e3, T, Bsz), flush=True)
