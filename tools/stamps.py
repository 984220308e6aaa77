"""Phase timestamps of the packed-pair kernel (tuning aid): where the time of one launch goes.
    NF_PAIR_CFG=16,6,4 NF_PAIR_FLAGS=2 python tools/stamps.py [logN]
"""
import ctypes, os, sys
os.environ.setdefault('NF_PAIR_CFG', '16,6,4')
os.environ['NF_PAIR_FLAGS'] = str(int(os.environ.get('NF_PAIR_FLAGS', '0')) | 2)
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'vi-normflows_b200')]
import numpy as np, torch
from nfcore import ops, semantics as sem, _lib

lg = int(sys.argv[1]) if len(sys.argv) > 1 else 20
N, K = (1 << lg) if lg < 64 else lg, 32       # lg >= 64: the sample count itself
dev = torch.device('cuda:0')
rs = np.random.RandomState(7)
W = 0.1 + 0.01 * rs.randn(K, 2); U = 0.1 + 0.01 * rs.randn(K, 2); B = 0.1 * np.ones(K)
S, c = sem.forward(W, U)
a = [torch.as_tensor(x, dtype=torch.float32).to(dev).contiguous() for x in (W, S, B, c)]
eps = [torch.randn(N, 2, device=dev) for _ in range(8)]
out = ops.planar_elbo(eps[0], *a, 'p1')
lib = _lib.load()
lib.nf_debug_pair_stamps.restype = ctypes.c_int
lib.nf_debug_pair_stamps.argtypes = [ctypes.c_void_p]
for rep in range(3):
    for i in range(8):
        ops.planar_elbo(eps[i], *a, 'p1', out=out)
    torch.cuda.synchronize()
    st = np.zeros((256, 8), dtype=np.uint64)
    lib.nf_debug_pair_stamps(st.ctypes.data)
    full = st.astype(np.int64)
    st = st[:148].astype(np.int64)
    t0 = st[:, 0].min()
    rel = (st[:, :5] - t0) / 1e3
    last = int(np.argmax(st[:, 4]))
    for b in range(2):
        wt = (full[200 + 4 * b:200 + 4 * b + 4].reshape(-1)[:32] - st[:, 0].min()) / 1e3
        wt = wt[wt > 0]
        print('   block %d: warps finish at ' % b + ' '.join('%.1f' % x for x in wt))
    ph = (full[216:232].reshape(16, 2, 4) - t0) / 1e3          # block 0: [warp][round][start, forward done, target done, reverse done]
    for rr in range(2):
        v = ph[:, rr]
        if (v[:, 3] > 0).all():
            print('   block 0 round %d: start %.1f..%.1f | forward %.2f..%.2f us | target %.2f..%.2f | reverse %.2f..%.2f  (per warp: fwd/tgt/rev)'
                  % (rr, v[:, 0].min(), v[:, 0].max(), (v[:, 1] - v[:, 0]).min(), (v[:, 1] - v[:, 0]).max(), (v[:, 2] - v[:, 1]).min(),
                     (v[:, 2] - v[:, 1]).max(), (v[:, 3] - v[:, 2]).min(), (v[:, 3] - v[:, 2]).max()))
            print('      warps 0,1,4,5 start/fwd/rev: ' + ' '.join('%.1f/%.1f/%.1f' % (v[w, 0], v[w, 1] - v[w, 0], v[w, 3] - v[w, 2]) for w in (0, 1, 4, 5)))
    print('   last block %d: ticket %.2f | rows loaded and staged %.2f | vector built %.2f | outputs written %.2f us'
          % (last, rel[last, 3], (st[last, 6] - t0) / 1e3, (st[last, 7] - t0) / 1e3, rel[last, 4]))
    print('N=2^%d  entry %.1f..%.1f us | set-up done %.1f..%.1f | samples done %.1f..%.1f (median %.1f) | ticket %.1f..%.1f | finalize done %.1f (block %d; its ticket at %.1f)'
          % (lg, rel[:, 0].min(), rel[:, 0].max(), rel[:, 1].min(), rel[:, 1].max(), rel[:, 2].min(), rel[:, 2].max(), np.median(rel[:, 2]),
             rel[:, 3].min(), rel[:, 3].max(), rel[last, 4], last, rel[last, 3]))
