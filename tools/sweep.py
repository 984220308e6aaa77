"""Quick device-resident throughput probe of the headline configuration (tuning aid)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'vi-normflows_b200')]
import numpy as np, torch
from nfcore import ops, semantics as sem
import bench

def run(tag, K=32, N=1 << 20, steps=100, check=None):
    dev = torch.device('cuda:0')
    rs = np.random.RandomState(7)
    W = 0.1 + 0.01 * rs.randn(K, 2); U = 0.1 + 0.01 * rs.randn(K, 2); B = 0.1 * np.ones(K)
    S, c = sem.forward(W, U)
    a = [torch.as_tensor(x, dtype=torch.float32).to(dev).contiguous() for x in (W, S, B, c)]
    g = torch.Generator(device=dev).manual_seed(5)
    eps = [torch.randn(N, 2, device=dev, generator=g) for _ in range(16)]
    for i in range(5):
        out = ops.planar_elbo(eps[i % 16], *a, 'p1')
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        ops.planar_elbo(eps[i % 16], *a, 'p1', out=out)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    ref = ops.planar_elbo(eps[0], *a, 'p1')
    sig = [float(ref[0][0]) / N, float(ref[1].abs().sum()), float(ref[4].abs().sum())]
    print('%-16s K=%d N=2^%d  %8.1f us/step  %.3e samples/s  F=%.7f |gw|=%.5e |gc|=%.5e' % (tag, K, N.bit_length() - 1, ms * 1e3, N / ms * 1e3, *sig), flush=True)

if __name__ == '__main__':
    for cfg in os.environ.get('SWEEP', '4,0;8,0;8,8;8,4;6,4;4,4;4,2').split(';'):
        os.environ['NF_ELBO_CFG'] = cfg
        run('cfg=' + cfg)
        run('cfg=' + cfg, N=1 << 24, steps=10)
