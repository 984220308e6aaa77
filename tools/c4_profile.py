"""BASELINE config 4, device loop with device draws: where the host time of a step goes (cProfile) -- tuning aid.
    python tools/c4_profile.py            prints ms/step and the top cumulative entries"""
import cProfile, io, os, pstats, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'vi-normflows_b200'), os.path.join(ROOT, 'tests')]
os.environ['NORMFLOWS_DEVICE_RNG'] = '11'
import numpy as np, torch
from test_gpu_vae import _mnist_closures
from normflows import optimization
K, D, Bsz, T = 16, 40, 4096, int(sys.argv[1]) if len(sys.argv) > 1 else 60
enc, dec, encode, decode, unpack, logp = _mnist_closures(K, D)
X = (np.random.RandomState(2).rand(4 * Bsz, 784) < 0.13).astype(np.float32)
p0 = np.random.RandomState(3).randn(enc.D + dec.D) * 0.05
optimization.results = type(optimization.results)('/tmp')
def run():
    optimization.optimize(logp, X, D, K, 4 * Bsz, p0, unpack, encode, decode, T, Bsz, 1e-3, verbose=False)
run(); torch.cuda.synchronize()
t0 = time.perf_counter(); run(); torch.cuda.synchronize()
print('%.2f ms per step' % ((time.perf_counter() - t0) / T * 1e3))
if os.environ.get('C4_NOPROF'):
    sys.exit(0)
pr = cProfile.Profile(); pr.enable(); run(); torch.cuda.synchronize(); pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats('cumulative').print_stats(45); print(s.getvalue()[:9000])
