"""CPU: the oracle restatements reproduce the committed golden vectors (which were
produced by running the reference itself, oracle/make_golden.py)."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, golden_script_cases
from oracle import planar_np as pn, planar_torch as pt


def np_target(name, tp, D):
    if name in pn.TARGETS:
        return pn.TARGETS[name]
    if name == 'gmm1d':
        return lambda z: pn.gmm1d(z)
    if name == 'mix':
        G = tp.size // (1 + 2 * D)
        pi, mu, var = tp[:G], tp[G:G + G * D].reshape(G, D), tp[G + G * D:].reshape(G, D)
        sig = np.array([np.diag(v) for v in var])
        return lambda z: pn.prob_gm(z, mu, sig, pi[:-1])
    if name in ('gauss', 'gauss16'):
        return lambda z: np.exp(pn.log_mvn(z, tp[:D], tp[D:]))
    raise KeyError(name)


def torch_target(name, tp, D):
    if name in pt.TARGETS:
        return pt.TARGETS[name]
    if name == 'gmm1d':
        return pt.make_gmm1d()
    if name == 'mix':
        G = tp.size // (1 + 2 * D)
        pi, mu, var = tp[:G], tp[G:G + G * D].reshape(G, D), tp[G + G * D:].reshape(G, D)
        sig = torch.tensor(np.array([np.diag(v) for v in var]))
        return lambda z: pt.prob_gm(z, torch.tensor(mu), sig, torch.tensor(pi[:-1]))
    if name in ('gauss', 'gauss16'):
        return lambda z: torch.exp(pt.log_mvn(z, tp[:D], tp[D:]))
    raise KeyError(name)


@pytest.mark.parametrize('case', golden_script_cases())
def test_script_objective_matches_reference(case):
    g = np.load(os.path.join(GOLDEN, case + '.npz'))
    K, D = int(g['K']), int(g['D'])
    tgt = np_target(str(g['target']), g['target_params'], D)
    F = pn.script_objective(g['params'], g['z0'], tgt, float(g['eps']), K)
    assert abs(F - float(g['F_ref'])) <= 1e-13 * max(1.0, abs(float(g['F_ref'])))
    lq, lp, _ = pn.script_terms(g['params'], g['z0'], tgt, float(g['eps']), K)
    assert abs(np.mean(lq) - float(g['mean_log_q'])) <= 1e-12 * max(1.0, abs(float(g['mean_log_q'])))
    assert abs(np.mean(lp) - float(g['mean_log_p'])) <= 1e-12 * max(1.0, abs(float(g['mean_log_p'])))


@pytest.mark.parametrize('case', golden_script_cases())
def test_torch_gradient_matches_reference_finite_differences(case):
    g = np.load(os.path.join(GOLDEN, case + '.npz'))
    K, D = int(g['K']), int(g['D'])
    tgt = torch_target(str(g['target']), g['target_params'], D)
    F, grad = pt.script_value_and_grad(g['params'], g['z0'], tgt, float(g['eps']), K)
    assert abs(F - float(g['F_ref'])) <= 1e-13 * max(1.0, abs(float(g['F_ref'])))
    np.testing.assert_allclose(grad, g['g_torch'], rtol=1e-12, atol=1e-14)
    # central differences of the REFERENCE forward (h=1e-6): agreement limited by the fd error,
    # larger for the "wide" parameter sets that sit near the log|1+h'c| singularity
    scale = np.abs(grad).max()
    tol = 2e-4 if 'wide' in case else 1e-6
    assert np.abs(g['g_fd'] - grad).max() <= tol * scale


def test_gmm_known_answer():
    # normflows/normflows/test_gmm.py:6-11 inputs; SURVEY.md §4 lists the first values
    g = np.load(os.path.join(GOLDEN, 'gmm_kat.npz'))
    yy = pn.prob_gm(g['xx'], g['mus'], g['sig'], g['pi'])
    np.testing.assert_allclose(yy, g['yy'], rtol=1e-14)
    np.testing.assert_allclose(yy[:3, 0], [0.00132955, 0.00184412, 0.00252646], rtol=1e-5)


@pytest.mark.parametrize('D', [1, 2, 40])
def test_planar_flow_matches_reference(D):
    g = np.load(os.path.join(GOLDEN, f'planar_flow_D{D}.npz'))
    np.testing.assert_allclose(pn.get_uhat(g['u'], g['w']), g['uhat'], rtol=1e-13)
    np.testing.assert_allclose(pn.planar_flow(g['z'], g['w'], g['u'], g['b']), g['out'], rtol=1e-13, atol=1e-14)
    out_t = pt.planar_flow(*(torch.tensor(g[k]) for k in ('z', 'w', 'u', 'b'))).numpy()
    np.testing.assert_allclose(out_t, g['out'], rtol=1e-12, atol=1e-13)


def test_mlp_forward_matches_reference():
    g = np.load(os.path.join(GOLDEN, 'mlp_forward.npz'))
    d = [int(v) for v in g['dims']]
    np.testing.assert_allclose(pn.mlp_forward(g['weights'], g['x'], *d), g['out'], rtol=1e-13, atol=1e-13)
    out_t = pt.mlp_forward(torch.tensor(g['weights']), torch.tensor(g['x']), *d).numpy()
    np.testing.assert_allclose(out_t, g['out'][0], rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize('name', ['amortised_K4_D8', 'amortised_K16_D40', 'amortised_K3_D1', 'amortised_K8_D2'])
def test_amortised_flow_matches_reference(name):
    g = np.load(os.path.join(GOLDEN, name + '.npz'))
    zk, rs = pn.amortised_flow_terms(g['z0'], g['mu0'], g['ls0'], g['W'], g['U'], g['B'])
    np.testing.assert_allclose(zk, g['zK_ref'], rtol=1e-12, atol=1e-13)
    D = int(g['D'])
    F = np.mean(pn.logq0_axis0(g['z0'], D)) - np.mean(pn.log_std_norm(zk)) * float(g['beta_t']) - np.mean(rs)
    assert abs(F - float(g['F_ref'])) <= 1e-12 * abs(float(g['F_ref']))
    assert abs(pn.beta_schedule(int(g['t']), int(g['max_iter'])) - float(g['beta_t'])) < 1e-15


def test_mnist_F_matches_reference():
    g = np.load(os.path.join(GOLDEN, 'mnist_F_K4_D40.npz'))
    enc, dec = tuple(int(v) for v in g['enc_dims']), tuple(int(v) for v in g['dec_dims'])
    p = torch.tensor(g['params'], requires_grad=True)
    F = pt.mnist_F(p, torch.tensor(g['z0']), torch.tensor(g['X'].astype(np.float64)), int(g['t']),
                   int(g['max_iter']), enc, dec, int(g['D']), int(g['K']))
    assert abs(float(F) - float(g['F_ref'])) <= 1e-11 * abs(float(g['F_ref']))
    gr, = torch.autograd.grad(F, p)
    idx = g['fd_idx']
    np.testing.assert_allclose(gr.numpy()[idx], g['g_fd'], rtol=2e-5, atol=1e-6)
    # numpy restatement of encode/decode path gives the same F
    X = g['X'].astype(np.float64)
    n_phi = pn.mlp_param_count(*enc)
    phi, theta = g['params'][:n_phi], g['params'][n_phi:]
    enc_out = pn.mnist_encode(phi, X, enc, int(g['D']), int(g['K']))
    Fn = pn.amortised_F(g['z0'], enc_out, lambda th, z: pn.mnist_decode(th, z, dec), theta, X, pn.mnist_logp,
                        int(g['t']), int(g['max_iter']))
    assert abs(Fn - float(g['F_ref'])) <= 1e-11 * abs(float(g['F_ref']))
