"""Quick device-resident throughput probe of the headline configuration (tuning aid).

    python tools/sweep.py                 every packed-pair configuration in SWEEP (one subprocess each: the library
                                          reads NF_PAIR_CFG once per process), at 2^20 and 2^24 samples per step
"""
import os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'vi-normflows_b200')]


def run(tag, K=32, N=1 << 20, steps=100, target='p1'):
    import numpy as np, torch
    from nfcore import ops, semantics as sem
    dev = torch.device('cuda:0')
    rs = np.random.RandomState(7)
    W = 0.1 + 0.01 * rs.randn(K, 2); U = 0.1 + 0.01 * rs.randn(K, 2); B = 0.1 * np.ones(K)
    S, c = sem.forward(W, U)
    a = [torch.as_tensor(x, dtype=torch.float32).to(dev).contiguous() for x in (W, S, B, c)]
    g = torch.Generator(device=dev).manual_seed(5)
    nbuf = 16 if N <= 1 << 22 else 2
    eps = [torch.randn(N, 2, device=dev, generator=g) for _ in range(nbuf)]
    for i in range(5):
        out = ops.planar_elbo(eps[i % nbuf], *a, target)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        ops.planar_elbo(eps[i % nbuf], *a, target, out=out)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    ref = ops.planar_elbo(eps[0], *a, target)
    sig = [float(ref[0][0]) / N, float(ref[1].abs().sum()), float(ref[4].abs().sum())]
    print('%-16s K=%d N=2^%d  %8.1f us/step  %.3e samples/s  F=%.7f |gw|=%.5e |gc|=%.5e' % (tag, K, N.bit_length() - 1, ms * 1e3, N / ms * 1e3, *sig), flush=True)


if __name__ == '__main__':
    K = os.environ.get('SWEEP_K', '32')
    for cfg in os.environ.get('SWEEP', 'default;16,6,4,0;14,8,4,2;16,8,4,2;12,8,4,2;off').split(';'):
        for lg, steps in ((20, 100), (24, 10)):
            subprocess.run([sys.executable, os.path.join(ROOT, 'tools', 'one.py'), cfg, K, str(lg), str(steps)])
