"""Wide-D free energy, one shape (for timing and ncu): python tools/gram_probe.py D logN [K] [steps]
Prints us/step of nf_planar_elbo_fwd_bwd_f32 (whatever path the library picks; NF_NO_GRAM=1 forces the
direct kernel, NF_GRAM_CHUNK=<samples> sets the chunk)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'vi-normflows_b200'))
import torch
from nfcore import ops

D, logN = int(sys.argv[1]), int(sys.argv[2])
K = int(sys.argv[3]) if len(sys.argv) > 3 else 64
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 10
N = 1 << logN
dev = 'cuda:0'
torch.manual_seed(0)
nbuf = 4
eps = [torch.randn(N, D, device=dev) for _ in range(nbuf)]
sc = 0.5 / (D / 2) ** 0.5
w, s = sc * torch.randn(K, D, device=dev), sc * torch.randn(K, D, device=dev)
b, c = 0.3 * torch.randn(K, device=dev), 0.3 * torch.randn(K, device=dev)
tp = torch.cat([0.2 * torch.randn(D, device=dev), 0.3 * torch.randn(D, device=dev) - 4.0 / D])
out = ops.planar_elbo(eps[0], w, s, b, c, 'gaussdiag', tp)
for i in range(3):
    ops.planar_elbo(eps[i % nbuf], w, s, b, c, 'gaussdiag', tp, out=out)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter()
e0.record()
for i in range(steps):
    ops.planar_elbo(eps[i % nbuf], w, s, b, c, 'gaussdiag', tp, out=out)
e1.record()
t_host = (time.perf_counter() - t0) / steps * 1e6
torch.cuda.synchronize()
us = e0.elapsed_time(e1) * 1e3 / steps
print('D=%d K=%d N=2^%d  %.1f us/step (host enqueue %.1f us/step)  %.3e samples/s  F=%s' % (
    D, K, logN, us, t_host, N / us * 1e6, out[0].cpu().numpy()))
