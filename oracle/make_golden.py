"""Generate tests/golden/*.npz by running the UNMODIFIED reference (build container only).

    python -m oracle.make_golden            # from the repo root

TEST INFRASTRUCTURE.  Needs /root/reference; the GPU box only ever sees the
committed .npz files.  Forward values ("F_ref", "zK_ref", ...) come from the
reference's own functions executed under ``oracle.refshim``.  Gradients come
from two independent sources that the CPU tests compare with each other:
"g_fd" = central differences of the reference's own forward, and "g_torch" =
torch float64 autograd over ``oracle.planar_torch`` (the stand-in for HIPS
autograd 1.3, which is not installable here).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import refshim, planar_torch as pt  # noqa: E402

OUT = os.path.join(ROOT, 'tests', 'golden')
EPS = 1e-7


def central_diff(f, x, h):
    g = np.zeros_like(x)
    for i in range(x.size):
        xp = x.copy(); xp[i] += h
        xm = x.copy(); xm[i] -= h
        g[i] = (f(xp) - f(xm)) / (2 * h)
    return g


def script_params(K, D, kind, seed):
    rs = np.random.RandomState(seed)
    if kind == 'init':        # near Final (master).ipynb cell 16 init (0.1*ones) -- SURVEY §8d C2
        W = 0.1 + 0.01 * rs.randn(K, D)
        U = 0.1 + 0.01 * rs.randn(K, D)
        B = 0.1 * np.ones(K)
    elif kind == 'wide':      # spread-out parameters: exercises u_hat, big tanh arguments
        W = 0.6 * rs.randn(K, D)
        U = 0.6 * rs.randn(K, D)
        B = 0.5 * rs.randn(K)
    else:
        raise ValueError(kind)
    return np.concatenate([W.ravel(), U.ravel(), B.ravel()])


def gen_script(ns, dist):
    import torch
    mix_mu = np.array([[-1.5, 0.0], [1.5, 0.5]])
    mix_sig = np.array([np.diag([1.0, 0.5]), np.diag([0.3, 2.0])])
    mix_pi = np.array([0.3])
    g_mu = np.array([1.0, -1.0]); g_ls = np.log(np.array([0.5, 2.0]))
    rs16 = np.random.RandomState(5)
    g16_mu = rs16.uniform(-1, 1, 16) * 0.2; g16_ls = rs16.uniform(-1, 1, 16)
    cases = [
        # name, target-name, K, D, N, params-kind
        ('script_p1_K8', 'p1', 8, 2, 100, 'init'),
        ('script_p1_K32', 'p1', 32, 2, 2048, 'init'),
        ('script_p1_K32_wide', 'p1', 32, 2, 1024, 'wide'),
        ('script_p2_K16', 'p2', 16, 2, 2048, 'init'),
        ('script_p2_K16_wide', 'p2', 16, 2, 1024, 'wide'),
        ('script_p3_K16', 'p3', 16, 2, 2048, 'init'),
        ('script_p3_K16_wide', 'p3', 16, 2, 1024, 'wide'),
        ('script_p4_K32', 'p4', 32, 2, 2048, 'init'),
        ('script_p4_K8_wide', 'p4', 8, 2, 1024, 'wide'),
        ('script_trial1_K8', 'trial1', 8, 2, 512, 'init'),
        ('script_gmm1d_K4', 'gmm1d', 4, 1, 256, 'wide'),
        ('script_mix2d_K32', 'mix', 32, 2, 1024, 'init'),
        ('script_mix2d_K8_wide', 'mix', 8, 2, 512, 'wide'),
        ('script_gauss2d_K8', 'gauss', 8, 2, 100, 'init'),
        ('script_gauss16d_K8', 'gauss16', 8, 16, 256, 'wide'),
    ]
    for i, (name, tname, K, D, N, kind) in enumerate(cases):
        z0 = np.random.RandomState(1000 + i).randn(N, D)
        params = script_params(K, D, kind, 7 + i)
        if D == 16:
            params[:2 * K * D] *= 1 / np.sqrt(D / 2)
        tparams = np.zeros(0)
        if tname in ('p1', 'p2', 'p3', 'p4', 'trial1'):
            ref_t, tor_t = ns[tname], pt.TARGETS[tname]
        elif tname == 'gmm1d':
            ref_t, tor_t = ns['gmm'], pt.make_gmm1d()
            tparams = np.array([0.3, 0.7, -1.0, 3.0, 1.0, 1.0])
        elif tname == 'mix':
            ref_t = lambda z: dist.prob_gm(z, mix_mu, mix_sig, mix_pi)
            tor_t = lambda z: pt.prob_gm(z, torch.tensor(mix_mu), torch.tensor(mix_sig), torch.tensor(mix_pi))
            tparams = np.concatenate([mix_pi, [1 - mix_pi.sum()], mix_mu.ravel(),
                                      np.array([np.diag(s) for s in mix_sig]).ravel()])
        elif tname == 'gauss':
            ref_t = lambda z: np.exp(dist.log_mvn(z, g_mu, g_ls))
            tor_t = lambda z: torch.exp(pt.log_mvn(z, g_mu, g_ls))
            tparams = np.concatenate([g_mu, g_ls])
        elif tname == 'gauss16':
            ref_t = lambda z: np.exp(dist.log_mvn(z, g16_mu, g16_ls))
            tor_t = lambda z: torch.exp(pt.log_mvn(z, g16_mu, g16_ls))
            tparams = np.concatenate([g16_mu, g16_ls])
        f = refshim.reference_script_objective(ns, ref_t, z0, EPS, K)
        F_ref = f(params)
        # other_metrics (get_data.py:80-95): mean log q_K and mean log p on the same z0
        om = ns['gradient_create'](ref_t, EPS, D, N, K)[3]
        with refshim.PinnedNormal(z0):
            mlq, mlp = om(params, 0)
        g_fd = central_diff(f, params, 1e-6)
        F_t, g_t = pt.script_value_and_grad(params, z0, tor_t, EPS, K)
        rel = np.abs(g_fd - g_t).max() / (np.abs(g_t).max() + 1e-300)
        print(f'{name:26s} F_ref={F_ref:+.12e}  |F_t-F_ref|={abs(F_t - F_ref):.1e}  fd-vs-torch rel={rel:.1e}')
        np.savez_compressed(os.path.join(OUT, name + '.npz'), target=tname, K=K, D=D, N=N, eps=EPS,
                            z0=z0, params=params, target_params=tparams, F_ref=F_ref,
                            mean_log_q=float(mlq), mean_log_p=float(mlp), g_fd=g_fd, g_torch=g_t, F_torch=F_t)


def gen_gmm_kat(dist):
    # normflows/normflows/test_gmm.py:6-11 -- the only fixture-like inputs in the reference
    mus = np.array([[0], [4]]); sig = np.array([[1], [1]]); pi = np.array([0.3])
    xx = np.linspace(-3, 8, 100).reshape(-1, 1)
    yy = dist.prob_gm(xx, mus, sig, pi)
    np.savez_compressed(os.path.join(OUT, 'gmm_kat.npz'), xx=xx, mus=mus, sig=sig, pi=pi, yy=yy)
    print('gmm_kat first values', yy[:3, 0], 'integral', np.trapezoid(yy[:, 0], xx[:, 0]))


def gen_planar_flow(flows):
    for D in (1, 2, 40):
        rs = np.random.RandomState(30 + D)
        N = 64
        z, w, u, b = rs.randn(N, D), 0.7 * rs.randn(N, D), 0.7 * rs.randn(N, D), rs.randn(N)
        out = flows.planar_flow(z, w, u, b)
        uhat = flows._get_uhat(u, w)
        np.savez_compressed(os.path.join(OUT, f'planar_flow_D{D}.npz'), z=z, w=w, u=u, b=b, out=out, uhat=uhat)
        print(f'planar_flow D={D} out[0]={out[0, :2]}')


def gen_amortised(optimization, dist):
    """optimization.py:74-85 on explicit encoder outputs (no MLP): K=4, N=48, D=8 and K=16, N=32, D=40."""
    import torch
    for (K, N, D, seed) in ((4, 48, 8, 50), (16, 32, 40, 51), (3, 40, 1, 52), (8, 64, 2, 53)):
        rs = np.random.RandomState(seed)
        z0 = rs.randn(N, D)
        mu0 = 0.3 * rs.randn(N, D); ls0 = 0.3 * rs.randn(N, D)
        sc = 1.0 / np.sqrt(D)
        W = sc * rs.randn(K, N, D); U = sc * rs.randn(K, N, D); B = 0.3 * rs.randn(K, N)
        Xb = np.zeros((N, 3))

        # drive the reference's own F with identity-like closures so that F = first - beta*mean(log p(z)) - third
        def encode(phi, X):
            return mu0, ls0, W, U, B

        seen = {}

        def decode(theta, zk):
            seen['zK'] = np.array(zk)
            return np.zeros((zk.shape[0], 3))

        def logp(X, Z, logits):
            return dist.log_std_norm(Z)
        F = refshim.capture_optimize_F(optimization, logp, Xb, D, K, N, np.zeros(2), lambda p: (p[:1], p[1:]),
                                       encode, decode, 400, N, 1e-3)
        t = 50
        state = optimization.rs.get_state()
        optimization.rs = _FixedRandn(z0)
        try:
            F_ref = float(F(np.zeros(2), t))
        finally:
            optimization.rs = np.random.RandomState(0)
            optimization.rs.set_state(state)
        # torch oracle: value and grads wrt the encoder outputs
        tz0 = torch.tensor(z0)
        leaves = [torch.tensor(a, requires_grad=True) for a in (mu0, ls0, W, U, B)]
        zk, rsum = pt.amortised_flow_terms(tz0, *leaves)
        beta_t = pt.beta_schedule(t, 400)
        F_t = torch.mean(pt.logq0_axis0(tz0, D)) - torch.mean(pt.log_std_norm(zk)) * beta_t - torch.mean(rsum)
        grads = torch.autograd.grad(F_t, leaves)
        print(f'amortised K={K} N={N} D={D}: F_ref={F_ref:+.12e} |F_t-F_ref|={abs(float(F_t) - F_ref):.1e} '
              f'|zK diff|={np.abs(zk.detach().numpy() - seen["zK"]).max():.1e}')
        np.savez_compressed(os.path.join(OUT, f'amortised_K{K}_D{D}.npz'), K=K, N=N, D=D, t=t, max_iter=400,
                            z0=z0, mu0=mu0, ls0=ls0, W=W, U=U, B=B, F_ref=F_ref, zK_ref=seen['zK'],
                            running_sum=rsum.detach().numpy(), beta_t=beta_t,
                            g_mu0=grads[0].numpy(), g_ls0=grads[1].numpy(), g_W=grads[2].numpy(),
                            g_U=grads[3].numpy(), g_B=grads[4].numpy())


class _FixedRandn:
    def __init__(self, z0):
        self.z0 = z0

    def randn(self, *shape):
        assert tuple(shape) == self.z0.shape, (shape, self.z0.shape)
        return self.z0


def gen_mlp(nn_models):
    rs = np.random.RandomState(60)
    arch = {'width': 16, 'hidden_layers': 3, 'input_dim': 12, 'output_dim': 7, 'activation_fn_type': 'relu',
            'activation_fn_params': '', 'activation_fn': lambda x: np.maximum(x, 0)}
    ff = nn_models.Feedforward(arch)
    w = 0.5 * rs.randn(1, ff.D)
    x = rs.randn(12, 9)
    out = ff.forward(w, x)
    np.savez_compressed(os.path.join(OUT, 'mlp_forward.npz'), weights=w, x=x, out=out, dims=np.array([12, 16, 3, 7]))
    print('mlp forward', out.shape, ff.D)


def gen_mnist_F(optimization):
    """The real thing: src/learning_mnist.py's encode/decode/logp through optimization.optimize's F
    (K=4, D=40, H=64, L=3 as in the driver) on a synthetic binarised batch."""
    import torch
    drv = refshim.load_mnist_driver()
    rs = np.random.RandomState(2)
    Bsz = 24
    X = (rs.rand(Bsz, 784) < 0.13).astype(int)
    unpack_params, encode, decode = drv.make_unpack_params()
    params = np.random.RandomState(3).randn(drv.encoder.D + drv.decoder.D) * 0.05
    z0 = np.random.RandomState(0).randn(Bsz, drv.dim_z)
    max_iter, t = 40, 3
    np.random.seed(11)       # make_batch_iter permutes with global np.random (utils.py:50)
    F = refshim.capture_optimize_F(optimization, drv.logp, X, drv.dim_z, drv.K, Bsz, params, unpack_params,
                                   encode, decode, max_iter, Bsz, 1e-3)
    # one batch per epoch == a permutation of X: recover it to feed the oracle the same rows
    np.random.seed(11)
    from normflows import utils
    perm_batch = utils.make_batch_iter(X, Bsz, max_iter)(t)
    optimization.rs = _FixedRandn(z0)
    try:
        F_ref = float(F(params, t))
    finally:
        optimization.rs = np.random.RandomState(0)
    enc_dims = (784, 64, 3, 2 * 40 + 2 * 40 * 4 + 4)
    dec_dims = (40, 64, 3, 784)
    p = torch.tensor(params, requires_grad=True)
    F_t = pt.mnist_F(p, torch.tensor(z0), torch.tensor(perm_batch, dtype=torch.float64), t, max_iter,
                     enc_dims, dec_dims, 40, 4)
    g, = torch.autograd.grad(F_t, p)
    # finite differences of the reference F on a few coordinates
    idx = np.random.RandomState(4).choice(params.size, 24, replace=False)
    g_fd = np.zeros(idx.size)
    optimization.rs = _FixedRandn(z0)
    try:
        for j, i in enumerate(idx):
            h = 1e-5
            pp = params.copy(); pp[i] += h
            pm = params.copy(); pm[i] -= h
            g_fd[j] = (float(F(pp, t)) - float(F(pm, t))) / (2 * h)
    finally:
        optimization.rs = np.random.RandomState(0)
    rel = np.abs(g_fd - g.numpy()[idx]).max() / np.abs(g.numpy()[idx]).max()
    print(f'mnist F: F_ref={F_ref:+.12e} |F_t-F_ref|={abs(float(F_t) - F_ref):.1e} fd-vs-torch rel={rel:.1e}')
    np.savez_compressed(os.path.join(OUT, 'mnist_F_K4_D40.npz'), X=perm_batch.astype(np.uint8), z0=z0,
                        params=params.astype(np.float64), t=t, max_iter=max_iter, F_ref=F_ref, F_torch=float(F_t),
                        g_torch=g.numpy().astype(np.float32), fd_idx=idx, g_fd=g_fd,
                        enc_dims=np.array(enc_dims), dec_dims=np.array(dec_dims), K=4, D=40)


def main():
    if not refshim.available():
        raise SystemExit('reference not found at ' + refshim.REFERENCE_ROOT)
    os.makedirs(OUT, exist_ok=True)
    ns = refshim.load_script_namespace('get_data.py', 142)
    from normflows import flows, distributions, optimization, nn_models
    gen_gmm_kat(distributions)
    gen_planar_flow(flows)
    gen_mlp(nn_models)
    gen_script(ns, distributions)
    gen_amortised(optimization, distributions)
    gen_mnist_F(optimization)


if __name__ == '__main__':
    main()
