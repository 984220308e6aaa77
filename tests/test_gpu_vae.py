"""GPU parity for the flow-VAE path (BASELINE config 4): Feedforward.forward (nn_models.py:41-84),
log_bern_mult (distributions.py:86-89) and the amortised free energy F of optimization.py:66-92
driven through closures written exactly like src/learning_mnist.py:57-99."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def _rel(a, ref):
    return float(np.abs(np.asarray(a) - ref).max() / max(np.abs(ref).max(), 1e-300))


def _arch(d_in, h, l, d_out, act='relu', out_act=None):
    from normflows import transformations
    a = {'width': h, 'hidden_layers': l, 'input_dim': d_in, 'output_dim': d_out, 'activation_fn_type': act,
         'activation_fn_params': '', 'activation_fn': transformations.relu}
    if out_act is not None:
        a['output_activation_fn'] = out_act
    return a


def test_feedforward_forward_matches_reference():
    from normflows import nn_models
    g = np.load(os.path.join(GOLDEN, 'mlp_forward.npz'))
    d = [int(v) for v in g['dims']]
    ff = nn_models.Feedforward(_arch(*d))
    assert ff.D == g['weights'].shape[1]
    out = ff.forward(g['weights'], g['x'])
    assert out.shape == g['out'].shape
    assert _rel(out, g['out']) <= 1e-10
    out32 = ff.forward(g['weights'].astype(np.float32), g['x'].astype(np.float32))
    assert _rel(out32, g['out']) <= 1e-4


@pytest.mark.parametrize('hidden,out_act,dims,N', [
    ('relu', None, (37, 24, 3, 19), 150), ('relu', 'sigmoid', (37, 24, 3, 19), 150), ('rbf', None, (37, 24, 3, 19), 150),
    # shapes the tensor-core (tcgen05, TF32x3) path takes: every dimension a multiple of 4, batch >= 128
    ('relu', None, (784, 64, 3, 1376), 384), ('relu', 'sigmoid', (40, 64, 3, 784), 256), ('relu', None, (64, 64, 1, 64), 1000)])
def test_mlp_backward_matches_autograd(hidden, out_act, dims, N):
    from nfcore import ops
    torch.manual_seed(0)
    dev = 'cuda:0'
    P = (dims[0] * dims[1] + dims[1]) + (dims[3] * dims[1] + dims[3]) + (dims[2] - 1) * (dims[1] ** 2 + dims[1])
    w = (0.3 * torch.randn(P, dtype=torch.float64)).requires_grad_(True)
    x = torch.randn(N, dims[0], dtype=torch.float64, requires_grad=True)
    g_out = torch.randn(N, dims[3], dtype=torch.float64)

    def ref(wv, xv):       # same layout walk as nn_models.py:59-80
        i = 0
        a = xv
        act = (lambda v: torch.relu(v)) if hidden == 'relu' else (lambda v: torch.exp(-v * v))
        kin = dims[0]
        for _ in range(dims[2]):
            W = wv[i:i + dims[1] * kin].reshape(dims[1], kin); i += dims[1] * kin
            b = wv[i:i + dims[1]]; i += dims[1]
            a = act(a @ W.T + b)
            kin = dims[1]
        W = wv[i:i + dims[3] * kin].reshape(dims[3], kin); i += dims[3] * kin
        o = a @ W.T + wv[i:]
        return torch.sigmoid(o) if out_act == 'sigmoid' else o
    o_ref = ref(w, x)
    gw_ref, gx_ref = torch.autograd.grad((o_ref * g_out).sum(), (w, x))
    for dt, tol in ((torch.float64, 1e-10), (torch.float32, 1e-4)):
        wd, xd, gd = (t.detach().to(dev, dt).contiguous() for t in (w, x, g_out))
        out, ws = ops.mlp_forward(wd, xd, dims, hidden, out_act)
        assert _rel(out.cpu().double().numpy(), o_ref.detach().numpy()) <= tol
        g_w, g_x = ops.mlp_backward(wd, xd, dims, hidden, out_act, ws, out, gd, need_gx=True)
        assert _rel(g_w.cpu().double().numpy(), gw_ref.numpy()) <= tol
        assert _rel(g_x.cpu().double().numpy(), gx_ref.numpy()) <= tol


@pytest.mark.parametrize('M,N,K,act', [(128, 64, 32, 0), (300, 100, 40, 1), (4096, 64, 784, 2), (64, 784, 4096, 0)])
def test_tensor_core_gemm_is_float32_accurate(M, N, K, act):
    """nf_gemm_tn_tf32x3 (TMA + tcgen05 TF32 with the hi/lo split) against a float64 matmul."""
    import ctypes
    import nfcore
    lib = nfcore.load()
    torch.manual_seed(M + N + K)
    dev = 'cuda:0'
    A = torch.randn(M, K, device=dev)
    B = torch.randn(N, K, device=dev) / K ** 0.5
    bias = torch.randn(N, device=dev)
    C = torch.full((M, N), float('nan'), device=dev)
    CT = torch.full((N, M), float('nan'), device=dev)
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    rc = lib.nf_gemm_tn_tf32x3(p(A), K, p(B), K, M, N, K, p(C), N, p(CT), M, p(bias), act,
                               ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    torch.cuda.synchronize()
    ref = A.double() @ B.double().T + bias.double()
    ref = torch.relu(ref) if act == 1 else (torch.sigmoid(ref) if act == 2 else ref)
    assert _rel(C.double().cpu().numpy(), ref.cpu().numpy()) <= 5e-5
    assert torch.equal(C, CT.T)
    # a single TF32 pass would be ~5e-4: the split really is in effect
    assert _rel(C.double().cpu().numpy(), ref.cpu().numpy()) < 1e-4


@pytest.mark.parametrize('M,N,K,act,out', [(128 * 700, 64, 64, 0, 'CT'), (128 * 650 + 37, 64, 96, 1, 'C'), (128 * 300, 200, 64, 0, 'C'),
                                           (128 * 600, 40, 256, 2, 'CT'), (128 * 640 + 5, 64, 64, 0, 'C'), (128 * 333 + 77, 100, 32, 0, 'CT')])
def test_persistent_gemm_matches_per_tile_kernel(M, N, K, act, out, monkeypatch):
    """Short contraction, thousands of tiles, one output (C or C^T): the persistent kernel (tile loop per CTA, double-buffered
    TMEM accumulator, bulk tensor stores).  Same MMAs in the same order as the per-tile kernel, so the two must agree bit
    for bit, and both with float64; rows / columns past M / N must stay untouched."""
    import ctypes
    import nfcore
    lib = nfcore.load()
    torch.manual_seed(M % 1000 + N + K)
    dev = 'cuda:0'
    A = torch.randn(M, K, device=dev)
    B = torch.randn(N, K, device=dev) / K ** 0.5
    bias = torch.randn(N, device=dev)
    p = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
    pad = 8                                   # guard rows / columns around the output: must keep their fill value
    ldo = ((N if out == 'C' else M) + pad + 3) // 4 * 4
    outs = []
    for persist in ('1', '0'):
        monkeypatch.setenv('NF_GEMM_PERSIST', persist)
        buf = torch.full(((M if out == 'C' else N) + pad, ldo), -7.0, device=dev)
        C, CT = (buf, None) if out == 'C' else (None, buf)
        rc = lib.nf_gemm_tn_tf32x3(p(A), K, p(B), K, M, N, K, p(C), ldo if out == 'C' else 0, p(CT), ldo if out == 'CT' else 0, p(bias), act,
                                   ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
        assert rc == 0
        torch.cuda.synchronize()
        outs.append(buf)
    assert torch.equal(outs[0], outs[1])
    got = outs[0][:M, :N] if out == 'C' else outs[0][:N, :M].T
    guard = torch.cat([outs[0][(M if out == 'C' else N):].reshape(-1), outs[0][:, (N if out == 'C' else M):].reshape(-1)])
    assert bool((guard == -7.0).all())
    ref = A.double() @ B.double().T + bias.double()
    ref = torch.relu(ref) if act == 1 else (torch.sigmoid(ref) if act == 2 else ref)
    assert _rel(got.double().cpu().numpy(), ref.cpu().numpy()) <= 5e-5


def test_bernoulli_ll_matches_oracle():
    from nfcore import ops
    from oracle import planar_np as pn
    rs = np.random.RandomState(1)
    X = (rs.rand(300, 784) < 0.13).astype(np.float64)
    # keep p a float32-representable distance from 0 and 1: 1 - p cancels catastrophically in float32
    # next to saturation, which is a property of the input precision, not of the kernel
    p = 0.001 + 0.998 * rs.rand(300, 784)
    ref = pn.log_bern_mult(X, p)
    g_ref = X / (1e-7 + p) - (1 - X) / (1e-7 + 1 - p)
    for dt, tol in ((torch.float64, 1e-12), (torch.float32, 1e-5)):
        out, g = ops.bernoulli_ll(torch.as_tensor(X).to('cuda:0', dt), torch.as_tensor(p).to('cuda:0', dt), 1e-7, -0.5)
        assert abs(float(out) - ref) <= tol * abs(ref)
        assert _rel(g.cpu().double().numpy(), -0.5 * g_ref) <= (1e-10 if dt == torch.float64 else 1e-4)


def _mnist_closures(K, dim_z):
    """encode / decode / logp written as in src/learning_mnist.py:57-99 (same slicing and reshapes)."""
    from normflows import nn_models, transformations, distributions
    dim_x = 784
    encoder = nn_models.Feedforward(_arch(dim_x, 64, 3, 2 * dim_z + 2 * dim_z * K + K))
    decoder = nn_models.Feedforward(_arch(dim_z, 64, 3, dim_x, out_act=transformations.sigmoid))

    def encode(weights, X):
        N = X.shape[0]
        phi = encoder.forward(weights.reshape(1, -1), X.T)[0]
        mu0 = phi[:dim_z].reshape(N, dim_z)
        log_sigma_diag0 = phi[dim_z:2 * dim_z].reshape(N, dim_z)
        W = phi[2 * dim_z:2 * dim_z + K * dim_z].reshape(K, N, dim_z)
        U = phi[2 * dim_z + K * dim_z:2 * dim_z + 2 * K * dim_z].reshape(K, N, dim_z)
        b = phi[-K:].reshape(K, N)
        return mu0, log_sigma_diag0, W, U, b

    def decode(weights, Z):
        logits = decoder.forward(weights.reshape(1, -1), Z.T)[0]
        return logits.T

    def unpack_params(params):
        return params[:encoder.D], params[encoder.D:]

    def logp(X, Z, logits):
        return distributions.log_bern_mult(X, logits) + distributions.log_std_norm(Z)
    return encoder, decoder, encode, decode, unpack_params, logp


class _FixedRandn:
    def __init__(self, z0):
        self.z0 = z0

    def randn(self, *shape):
        assert tuple(shape) == self.z0.shape
        return self.z0


@pytest.mark.parametrize('dtype,tol', [(torch.float64, 1e-10), (torch.float32, 1e-4)])
def test_mnist_free_energy_and_gradient(dtype, tol, monkeypatch):
    from normflows import optimization
    from oracle import planar_torch as pt
    g = np.load(os.path.join(GOLDEN, 'mnist_F_K4_D40.npz'))
    K, D = int(g['K']), int(g['D'])
    encoder, decoder, encode, decode, unpack_params, logp = _mnist_closures(K, D)
    assert encoder.D + decoder.D == g['params'].size
    X = g['X'].astype(np.float64)
    t, max_iter = int(g['t']), int(g['max_iter'])
    monkeypatch.setattr(optimization, 'rs', _FixedRandn(g['z0']))
    F = optimization.make_free_energy(logp, X, D, K, encode, decode, max_iter, X.shape[0], dtype=dtype,
                                      batch_iter=lambda tt: X)
    objective, gradient = optimization.gradient_create(F, D, X.shape[0], unpack_params, dtype=dtype)
    val = objective(g['params'], t)
    assert abs(val - float(g['F_ref'])) <= tol * abs(float(g['F_ref'])), (val, float(g['F_ref']))
    grad = gradient(g['params'], t)
    # oracle gradient (torch float64 autograd over the op-for-op restatement)
    p = torch.tensor(g['params'], requires_grad=True)
    Fo = pt.mnist_F(p, torch.tensor(g['z0']), torch.tensor(X), t, max_iter, tuple(int(v) for v in g['enc_dims']),
                    tuple(int(v) for v in g['dec_dims']), D, K)
    g_ref, = torch.autograd.grad(Fo, p)
    assert _rel(grad, g_ref.numpy()) <= tol
    # and the finite differences of the REFERENCE's own F on a few coordinates
    assert np.abs(grad[g['fd_idx']] - g['g_fd']).max() <= max(tol, 5e-5) * np.abs(g_ref.numpy()).max()


@pytest.mark.parametrize('conditioned', [False, True])
def test_mnist_free_energy_at_baseline_size(conditioned, monkeypatch):
    """BASELINE config 4 at full size (batch 4096, 784 pixels, D = 40, K = 16, H = 64, L = 3; SURVEY.md section 8d C4
    inputs: Bernoulli(0.13) pixels, parameters 0.05 N(0,1) as src/learning_mnist.py:83-86, z0 from RandomState(0)): F and
    its gradient over all 209 904 parameters against oracle.planar_torch.mnist_F (torch float64 autograd).
    float64: 1e-10 on F, 1e-7 on the gradient, both parameter sets.  float32: 1e-4 on F, both sets.
    With the reference's initialisation (conditioned=False) the batch is ill-conditioned twice over: some of its 65 536
    (sample, layer) pairs land within 1e-3 of the log|1 + h'c| singularity, and per-sample directions w with
    ||w||^2 near 0 blow up u_hat = u + (m(w.u) - w.u) w / ||w||^2 (flows.py:39); single samples then dominate the
    gradient (max |g| about 5e6) and float32 cannot hold it.  The float32 gradient is therefore compared on the
    conditioned set: the same draw with the encoder's output layer scaled by 0.15 and its W / U bias blocks raised
    by 0.2 / 0.1, which puts (sum w)(sum u) near +8 and ||w||^2 near 1.2 for every sample (asserted through the oracle)."""
    from normflows import optimization
    from oracle import planar_torch as pt
    K, D, Bsz = 16, 40, 4096
    encoder, decoder, encode, decode, unpack_params, logp = _mnist_closures(K, D)
    X = (np.random.RandomState(2).rand(Bsz, 784) < 0.13).astype(np.float64)
    params = np.random.RandomState(3).randn(encoder.D + decoder.D) * 0.05
    assert params.size == 148000 + 61904
    d_out = 2 * D + 2 * D * K + K
    if conditioned:
        params[encoder.D - (64 + 1) * d_out:encoder.D] *= 0.15       # W_out (d_out x 64) and b_out of the encoder
        bias = params[encoder.D - d_out:encoder.D]
        bias[2 * D:2 * D + K * D] += 0.2                              # rows that become W (learning_mnist.py:65)
        bias[2 * D + K * D:2 * D + 2 * K * D] += 0.1                  # rows that become U (:66)
    z0 = np.random.RandomState(0).randn(Bsz, D)
    t, max_iter = 0, 10000
    enc_dims, dec_dims = (784, 64, 3, d_out), (D, 64, 3, 784)
    p = torch.tensor(params, requires_grad=True)
    Fo = pt.mnist_F(p, torch.tensor(z0), torch.tensor(X), t, max_iter, enc_dims, dec_dims, D, K)
    g_ref, = torch.autograd.grad(Fo, p)
    Fo = float(Fo.detach())
    with torch.no_grad():
        enc = pt.mnist_encode(p[:encoder.D], torch.tensor(X), enc_dims, D, K)
        near = pt.amortised_min_det_factor(torch.tensor(z0), *enc)
        w2 = float((enc[2] ** 2).sum(-1).min())
    assert (near > 0.9 and w2 > 1.0) if conditioned else (near < 1e-3), (near, w2)
    for dtype in (torch.float64, torch.float32):
        monkeypatch.setattr(optimization, 'rs', _FixedRandn(z0))
        F = optimization.make_free_energy(logp, X, D, K, encode, decode, max_iter, Bsz, dtype=dtype, batch_iter=lambda tt: X)
        objective, gradient = optimization.gradient_create(F, D, Bsz, unpack_params, dtype=dtype)
        val = objective(params, t)
        grad = gradient(params, t)
        eF, eg = abs(val - Fo) / abs(Fo), _rel(grad, g_ref.numpy())
        print('C4 B=4096 conditioned=%s %s: F err %.2e, grad err %.2e, min|1+h\'c| %.2e, min||w||^2 %.2e' % (conditioned, dtype, eF, eg, near, w2))
        if dtype == torch.float64:
            assert eF <= 1e-10 and eg <= 1e-7, (eF, eg)
        else:
            assert eF <= 1e-4, eF
            if conditioned:
                assert eg <= 1e-4, eg


def test_optimize_runs_and_decreases_free_energy(tmp_path, monkeypatch):
    """A short Adam run through the drop-in optimize(): HEAD signature, callback cadence, return value."""
    from normflows import optimization
    K, D = 2, 8
    encoder, decoder, encode, decode, unpack_params, logp = _mnist_closures(K, D)
    rs = np.random.RandomState(2)
    X = (rs.rand(256, 784) < 0.13).astype(int)
    init = np.random.RandomState(3).randn(encoder.D + decoder.D) * 0.05
    monkeypatch.setattr(optimization, 'results', tmp_path)
    np.random.seed(0)
    optimization.rs = np.random.RandomState(0)
    F = optimization.make_free_energy(logp, X, D, K, encode, decode, 60, 64)
    obj, _ = optimization.gradient_create(F, D, 256, unpack_params)
    before = np.mean([obj(init, 59) for _ in range(4)])
    np.random.seed(0)
    optimization.rs = np.random.RandomState(0)
    phi, theta = optimization.optimize(logp, X, D, K, 256, init, unpack_params, encode, decode, 60, 64, 1e-3, verbose=True)
    assert phi.shape == (encoder.D,) and theta.shape == (decoder.D,)
    after = np.mean([obj(np.concatenate([phi, theta]), 59) for _ in range(4)])
    assert np.isfinite(after) and after < before
    assert (tmp_path / 'free_energy.txt').read_text().startswith(f'\n{K} flows: ')


def test_device_adam_loop_equals_host_adam_loop(tmp_path, monkeypatch):
    """optimize() keeps parameters, Adam moments and the update on the device (nf_adam_step).  In float64 its result
    after 25 iterations equals autograd.misc.optimizers.adam(gradient, init, ...) driven from the host on the same
    mini-batches and the same z0 draws (optimization.py:118), to 1e-10."""
    from autograd.misc.optimizers import adam
    from normflows import optimization
    K, D, T, Bsz = 2, 8, 25, 64
    encoder, decoder, encode, decode, unpack_params, logp = _mnist_closures(K, D)
    X = (np.random.RandomState(2).rand(256, 784) < 0.13).astype(int)
    init = np.random.RandomState(3).randn(encoder.D + decoder.D) * 0.05
    monkeypatch.setattr(optimization, 'results', tmp_path)
    monkeypatch.setenv('NORMFLOWS_PRECISION', 'fp64')
    np.random.seed(0)
    optimization.rs = np.random.RandomState(0)
    F = optimization.make_free_energy(logp, X, D, K, encode, decode, T, Bsz)
    _, gradient = optimization.gradient_create(F, D, 256, unpack_params)
    x_host = adam(gradient, init, step_size=1e-3, num_iters=T)
    np.random.seed(0)
    optimization.rs = np.random.RandomState(0)
    phi, theta = optimization.optimize(logp, X, D, K, 256, init, unpack_params, encode, decode, T, Bsz, 1e-3, verbose=False)
    x_dev = np.concatenate([phi, theta])
    assert np.abs(x_host - init).max() > 1e-3
    assert np.abs(x_dev - x_host).max() <= 1e-10 * max(1.0, np.abs(x_host).max()), np.abs(x_dev - x_host).max()


def test_free_energy_with_device_side_draws():
    """make_free_energy(device_rng=seed): z0 drawn on the device (opt-in, not the reference's stream).  Same seed, same
    value; the value is a plausible free energy of the same model (close to the host-draw value at the same parameters)."""
    from normflows import optimization
    K, D, Bsz = 4, 8, 512
    encoder, decoder, encode, decode, unpack_params, logp = _mnist_closures(K, D)
    X = (np.random.RandomState(2).rand(Bsz, 784) < 0.13).astype(np.float64)
    params = np.random.RandomState(3).randn(encoder.D + decoder.D) * 0.05
    vals = []
    for rep in range(2):
        F = optimization.make_free_energy(logp, X, D, K, encode, decode, 100, Bsz, dtype=torch.float64, batch_iter=lambda tt: X, device_rng=7)
        objective, gradient = optimization.gradient_create(F, D, Bsz, unpack_params, dtype=torch.float64)
        vals.append(objective(params, 50))
        g = gradient(params, 50)
        assert np.isfinite(g).all() and np.abs(g).max() > 0
    assert vals[0] == vals[1]
    optimization.rs = np.random.RandomState(0)
    F = optimization.make_free_energy(logp, X, D, K, encode, decode, 100, Bsz, dtype=torch.float64, batch_iter=lambda tt: X)
    host = optimization.gradient_create(F, D, Bsz, unpack_params, dtype=torch.float64)[0](params, 50)
    assert abs(vals[0] - host) <= 0.05 * abs(host)
