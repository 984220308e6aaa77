"""Why does bench.py's 20-step figure differ from tools/sweep.py's 100-step one?  Varies the number of rotating
sample buffers (L2-resident or not) and the number of timed steps."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'vi-normflows_b200')]
import numpy as np, torch
from nfcore import ops, semantics as sem
K, N = 32, 1 << 20
dev = torch.device('cuda:0')
rs = np.random.RandomState(7)
W = 0.1 + 0.01 * rs.randn(K, 2); U = 0.1 + 0.01 * rs.randn(K, 2); B = 0.1 * np.ones(K)
S, c = sem.forward(W, U)
a = [torch.as_tensor(x, dtype=torch.float32).to(dev).contiguous() for x in (W, S, B, c)]
eps = [torch.randn(N, 2, device=dev) for _ in range(32)]
out = ops.planar_elbo(eps[0], *a, 'p1')
for nbuf in (4, 16, 32):
    for steps in (20, 100, 400):
        for i in range(3):
            ops.planar_elbo(eps[i % nbuf], *a, 'p1', out=out)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            ops.planar_elbo(eps[i % nbuf], *a, 'p1', out=out)
        e1.record(); torch.cuda.synchronize()
        print('buffers %2d steps %3d: %.1f us/step' % (nbuf, steps, e0.elapsed_time(e1) / steps * 1e3), flush=True)
