"""Back-to-back launches vs isolated launches of the headline kernel (tuning aid)."""
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
for i in range(200):
    ops.planar_elbo(eps[i % 32], *a, 'p1', out=out)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(200):
    ops.planar_elbo(eps[i % 32], *a, 'p1', out=out)
e1.record(); torch.cuda.synchronize()
print('back to back: %.1f us/step' % (e0.elapsed_time(e1) / 200 * 1e3))
ts = []
for i in range(200):
    torch.cuda.synchronize()
    e0.record()
    ops.planar_elbo(eps[i % 32], *a, 'p1', out=out)
    e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1) * 1e3)
print('isolated (sync between): median %.1f us, min %.1f' % (np.median(ts), min(ts)))
# isolated but GPU kept warm by a long dummy kernel right before? -> skip
