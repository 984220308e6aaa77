"""Device-timed throughput of every BASELINE.json configuration on one GPU (secondary table;
bench.py is the contract benchmark).  Prints one line per case and writes profiles-style text.
    python tools/bench_configs.py [--quick]
"""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'vi-normflows_b200')]
import numpy as np, torch
from nfcore import ops, semantics as sem
from nfcore.objective import SharedFlowObjective

dev = torch.device('cuda:0')
SMS, FMAX = 148, 1.965e9
P_FP32, P_MUFU = SMS * 128 * 2 * FMAX, SMS * 16 * FMAX


def roofline(K, D, target, bytes_per=4):
    ct, mt = {'u': (64, 10), 'mix': (2 * (6 * D + 10), 5), 'gauss': (6 * D, 1)}[target]
    flops = K * (14 * D + 25) + 2 * D + ct
    mufu = 4 * K + mt
    return 1.0 / max(flops / P_FP32, mufu / P_MUFU, bytes_per * D / 6550.7e9), flops, mufu


def timed(fn, steps, warm=3):
    for i in range(warm): fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps): fn(i)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps * 1e-3


def shared_case(name, K, D, N, target, tname, tparams=None, steps=50, dtype=torch.float32):
    rs = np.random.RandomState(7)
    W = (0.1 + 0.01 * rs.randn(K, D)) / np.sqrt(D / 2); U = (0.1 + 0.01 * rs.randn(K, D)) / np.sqrt(D / 2); B = 0.1 * np.ones(K)
    S, c = sem.forward(W, U)
    a = [torch.as_tensor(x).to(dev, dtype).contiguous() for x in (W, S, B, c)]
    nbuf = max(2, min(16, int((300 << 20) // (N * D * 4))))
    eps = [torch.randn(N, D, device=dev, dtype=dtype) for _ in range(nbuf)]
    tp = None if tparams is None else torch.as_tensor(tparams).to(dev, dtype)
    out = ops.planar_elbo(eps[0], *a, tname, tp)
    dt = timed(lambda i: ops.planar_elbo(eps[i % nbuf], *a, tname, tp, out=out), steps)
    roof, fl, mu = roofline(K, D, target)
    v = N / dt
    note, extra = '', {}
    if D >= 32 and dtype == torch.float32:
        # what runs is the matrix-product chain.  Its own bounds per sample: HBM traffic of the chain as it is
        # decomposed (every [N][D] and [KP][N] intermediate written once and read by its consumers: 32 D + 44 KP bytes),
        # and the tensor time of five products at three TF32 passes (5 * 2 K D * 3 flop at the nominal dense TF32 rate)
        KP = 16 if K <= 16 else (32 if K <= 32 else 64)
        hbm_bytes = 32 * D + 44 * KP
        t_hbm = hbm_bytes / 6455.9e9
        t_tc = 30.0 * K * D / 1.1e15
        extra = {'chain_hbm_bytes_per_sample': hbm_bytes, 'frac_chain_hbm': t_hbm * v, 'frac_tf32x3': t_tc * v}
        note = ('  [matrix-product path: %.2f of the HBM time of its own chain (%d B/sample), %.2f of the 3-pass TF32 tensor time; '
                'the roofline column is the CUDA-core bound of the direct algorithm]' % (t_hbm * v, hbm_bytes, t_tc * v))
    print('%-34s K=%-3d D=%-4d N=2^%-2d %9.1f us/step %10.3e samples/s  roofline %.3e  frac %.3f%s' %
          (name, K, D, N.bit_length() - 1, dt * 1e6, v, roof, v / roof, note), flush=True)
    r = {'case': name, 'K': K, 'D': D, 'N': N, 'us_per_step': dt * 1e6, 'samples_per_s': v, 'roofline': roof, 'frac': v / roof}
    r.update(extra)
    return r


def main():
    quick = '--quick' in sys.argv
    res = []
    # C1: latency-bound, through the NumPy-facing API, float64
    obj = SharedFlowObjective(('gaussdiag', np.array([1.0, -1.0]), np.log(np.array([0.5, 2.0]))), 8, 2, dtype='float64')
    rs = np.random.RandomState(0)
    params = np.concatenate([0.1 * np.ones(32), 0.1 * np.ones(8)])
    for _ in range(5): obj.value_and_grad(params, rs.randn(100, 2))
    t0 = time.perf_counter()
    for _ in range(300): obj.value_and_grad(params, rs.randn(100, 2))
    dt = (time.perf_counter() - t0) / 300
    print('%-34s K=8   D=2    N=100   %9.1f us/step  %.0f objective+gradient calls/s (host API, float64)' % ('C1 gaussian, NumPy API', dt * 1e6, 1 / dt), flush=True)
    res.append({'case': 'C1', 'us_per_call': dt * 1e6})
    eps_all = rs.randn(2000, 100, 2)
    obj.train(params, eps_all[:50], optimizer='adam', step_size=1e-3)
    t0 = time.perf_counter()
    obj.train(params, eps_all, optimizer='adam', step_size=1e-3)
    dt2 = (time.perf_counter() - t0) / 2000
    print('%-34s K=8   D=2    N=100   %9.1f us/step  %.0f Adam steps/s (nf_script_train_f64: loop on the device, incl. upload of all draws)' % ('C1 gaussian, device loop', dt2 * 1e6, 1 / dt2), flush=True)
    res.append({'case': 'C1 device loop', 'us_per_step': dt2 * 1e6})
    # C2
    for K in (16, 32):
        for t in ('p1', 'p2', 'p3', 'p4'):
            res.append(shared_case('C2 %s' % t, K, 2, 1 << 20, 'u', t))
    # C3
    mix = np.concatenate([[0.3, 0.7], [-1.5, 0.0, 1.5, 0.5], [1.0, 0.5, 0.3, 2.0]])
    res.append(shared_case('C3 mixture', 32, 2, 1 << 24, 'mix', 'mixdiag', mix, steps=10))
    # C5 (single-GPU slice of the sweep)
    res.append(shared_case('C5 U1', 64, 2, 1 << 23, 'u', 'p1', steps=10))
    for D, N in ((16, 1 << 21), (64, 1 << 19), (128, 1 << 18), (256, 1 << 17), (512, 1 << 16)):
        rs = np.random.RandomState(5)
        tp = np.concatenate([rs.uniform(-1, 1, D), rs.uniform(-1, 1, D)])
        res.append(shared_case('C5 gaussdiag', 64, D, N if not quick else N >> 2, 'gauss', 'gaussdiag', tp, steps=5))
    # C4: flow-VAE step through optimization.gradient (MLP + amortised flow + Bernoulli), float32
    sys.path.insert(0, os.path.join(ROOT, 'tests'))
    from test_gpu_vae import _mnist_closures
    from normflows import optimization
    K, D, Bsz = 16, 40, 4096
    enc, dec, encode, decode, unpack, logp = _mnist_closures(K, D)
    X = (np.random.RandomState(2).rand(Bsz, 784) < 0.13).astype(np.float32)
    p = np.random.RandomState(3).randn(enc.D + dec.D) * 0.05
    F = optimization.make_free_energy(logp, X, D, K, encode, decode, 10000, Bsz, dtype=torch.float32)
    objective, gradient = optimization.gradient_create(F, D, Bsz, unpack, dtype=torch.float32)
    for _ in range(3): gradient(p, 0)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(20): gradient(p, 0)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 20
    print('%-34s K=16  D=40   B=4096 %9.1f us/step %10.3e images/s (host API incl. H2D of X and D2H of the gradient)' % ('C4 flow-VAE gradient', dt * 1e6, Bsz / dt), flush=True)
    res.append({'case': 'C4', 'us_per_step': dt * 1e6, 'images_per_s': Bsz / dt})
    print(json.dumps(res))


if __name__ == '__main__':
    main()
