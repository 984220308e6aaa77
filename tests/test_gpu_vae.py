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


@pytest.mark.parametrize('dtype,tol', [(torch.float64, 1e-10), (torch.float32, 2e-4)])
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
