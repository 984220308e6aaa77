"""Matrix-product path vs direct kernel over a grid of (D, K, N), one process: python tools/gram_crossover.py
For each shape: us per call with NF_GRAM=1, with NF_NO_GRAM=1, and which one the library's cost model picks."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'vi-normflows_b200'))
import torch
from nfcore import ops

dev = 'cuda:0'


def timed(fn, steps):
    for i in range(2): fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps): fn(i)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / steps


def setenv(gram):
    os.environ.pop('NF_GRAM', None); os.environ.pop('NF_NO_GRAM', None)
    if gram is True: os.environ['NF_GRAM'] = '1'
    if gram is False: os.environ['NF_NO_GRAM'] = '1'


print('%5s %4s %6s %10s %10s %10s  %s' % ('D', 'K', 'N', 'gram us', 'direct us', 'default us', 'model picked'))
bad = 0
for D in (32, 64, 128, 256):
    for K in (16, 32, 64):
        for logN in (10, 12, 13, 14, 15, 16, 17):
            N = 1 << logN
            if N * D * 4 > (256 << 20): continue
            torch.manual_seed(0)
            eps = [torch.randn(N, D, device=dev) for _ in range(2)]
            sc = 0.5 / (D / 2) ** 0.5
            w, s = sc * torch.randn(K, D, device=dev), sc * torch.randn(K, D, device=dev)
            b, c = 0.3 * torch.randn(K, device=dev), 0.3 * torch.randn(K, device=dev)
            tp = torch.cat([0.2 * torch.randn(D, device=dev), 0.3 * torch.randn(D, device=dev) - 4.0 / D])
            res = {}
            for name, g in (('gram', True), ('direct', False), ('default', None)):
                setenv(g)
                out = ops.planar_elbo(eps[0], w, s, b, c, 'gaussdiag', tp)
                res[name] = timed(lambda i: ops.planar_elbo(eps[i % 2], w, s, b, c, 'gaussdiag', tp, out=out), 8)
            pick = 'gram' if abs(res['default'] - res['gram']) < abs(res['default'] - res['direct']) else 'direct'
            best = 'gram' if res['gram'] < res['direct'] else 'direct'
            flag = '' if pick == best or abs(res['gram'] - res['direct']) < 0.1 * min(res['gram'], res['direct']) else '   <-- wrong pick'
            bad += bool(flag)
            print('%5d %4d   2^%-2d %10.1f %10.1f %10.1f  %s%s' % (D, K, logN, res['gram'], res['direct'], res['default'], pick, flag), flush=True)
print('wrong picks (beyond 10 %%): %d' % bad)
