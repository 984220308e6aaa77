"""GPU tests of the evaluation surface (csrc/nf_eval.cu): target densities at given points and samples pushed through
the shared-parameter flow -- what visualise_flow (Final (master).ipynb cell 16) computes with NumPy -- against the
oracle, and the distributions.py density helpers on CUDA tensors against the reference's own known answer
(test_gmm.py:6-11 -> tests/golden/gmm_kat.npz)."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def _rel(a, ref):
    return float(np.abs(np.asarray(a) - ref).max() / max(np.abs(ref).max(), 1e-300))


@pytest.mark.parametrize('name', ['p1', 'p2', 'p3', 'p4', 'trial1'])
def test_potential_density_on_a_grid(name):
    from normflows import objectives
    from oracle import planar_np as pn
    xs, ys = np.linspace(-5, 5, 201), np.linspace(-4.5, 4.5, 151)
    X, Y = np.meshgrid(xs, ys)
    ref = getattr(pn, name)(np.stack([X.ravel(), Y.ravel()], axis=1)).reshape(Y.shape)
    got = objectives.density_grid(name, xs, ys)
    assert got.shape == (151, 201)
    assert _rel(got, ref) <= 1e-10
    got32 = objectives.density_grid(name, xs, ys, dtype='float32')
    assert got32.dtype == np.float32 and _rel(got32, ref) <= 1e-4


@pytest.mark.parametrize('D', [1, 2, 3, 4, 7, 40, 256])
def test_gaussian_and_mixture_densities(D):
    from normflows import objectives
    from oracle import planar_torch as pt
    rs = np.random.RandomState(D)
    N, G = 1000 + D, 3
    Z = rs.randn(N, D)
    mu, ls = 0.5 * rs.randn(D), 0.4 * rs.randn(D) - (3.0 / D if D > 4 else 0)
    ref = torch.exp(pt.log_mvn(torch.tensor(Z), mu, ls)).numpy()
    assert _rel(objectives.target_density(('gaussdiag', mu, ls), Z), ref) <= 1e-10
    if ref.max() > 1e-30:             # a 256-dimensional density underflows float32 (1e-179): float64 only there
        assert _rel(objectives.target_density(('gaussdiag', mu, ls), Z, dtype='float32'), ref) <= 1e-4
    pi = np.array([0.2, 0.5]); mmu = 0.5 * rs.randn(G, D); var = 0.8 + rs.rand(G, D)
    sig = torch.tensor(np.array([np.diag(v) for v in var])) if D > 1 else torch.tensor(var)
    ref = pt.prob_gm(torch.tensor(Z), torch.tensor(mmu), sig, torch.tensor(pi)).numpy().ravel()
    assert _rel(objectives.target_density(('mixdiag', pi, mmu, var), Z), ref) <= 1e-10
    if 1 < D <= 4:
        A = rs.randn(G, D, D)
        cov = np.einsum('gij,gkj->gik', A, A) + 0.5 * np.eye(D)
        ref = pt.prob_gm(torch.tensor(Z), torch.tensor(mmu), torch.tensor(cov), torch.tensor(pi)).numpy().ravel()
        assert _rel(objectives.target_density(('mixdiag', pi, mmu, cov), Z), ref) <= 1e-10


@pytest.mark.parametrize('D,K', [(1, 4), (2, 32), (3, 5), (4, 64), (16, 8), (40, 16), (300, 3)])
def test_flow_samples_match_oracle(D, K):
    from normflows import objectives
    from oracle import planar_torch as pt
    rs = np.random.RandomState(10 * D + K)
    N = 2000 + D
    z0 = rs.randn(N, D)
    params = np.concatenate([(0.3 * rs.randn(2 * K * D)) / np.sqrt(max(D, 2) / 2), 0.2 * rs.randn(K)])
    lq_ref, _, zK_ref = pt.script_terms(torch.tensor(params), torch.tensor(z0), lambda z: torch.ones(z.shape[0]), 1e-7, K)
    zK, lq = objectives.flow_samples(params, z0, K, D)
    assert zK.shape == (N, D) and lq.shape == (N,)
    assert _rel(zK, zK_ref.numpy()) <= 1e-10 and _rel(lq, lq_ref.numpy()) <= 1e-10
    zK32, lq32 = objectives.flow_samples(params, z0, K, D, dtype='float32')
    assert _rel(zK32, zK_ref.numpy()) <= 1e-4 and _rel(lq32, lq_ref.numpy()) <= 1e-4


def test_distribution_helpers_on_cuda_tensors():
    """prob_gm / mvn / log_mvn given CUDA tensors run on the GPU; prob_gm on the inputs of the reference's test_gmm.py
    reproduces the values the reference itself computed (tests/golden/gmm_kat.npz, made by oracle/make_golden.py)."""
    from normflows import distributions
    from oracle import planar_np as pn
    g = np.load(os.path.join(GOLDEN, 'gmm_kat.npz'))
    x = torch.as_tensor(g['xx']).cuda()
    out = distributions.prob_gm(x, g['mus'], g['sig'], g['pi'])
    assert out.is_cuda and out.shape == (x.shape[0], 1)
    assert _rel(out.cpu().numpy(), g['yy']) <= 1e-12
    rs = np.random.RandomState(0)
    Z = rs.randn(500, 3)
    mu = rs.randn(3)
    A = rs.randn(3, 3); cov = A @ A.T + np.eye(3)
    assert _rel(distributions.mvn(torch.as_tensor(Z).cuda(), mu, cov).cpu().numpy(), pn.mvn(Z, mu, cov)) <= 1e-10
    ls = 0.3 * rs.randn(3)
    assert _rel(distributions.log_mvn(torch.as_tensor(Z).cuda(), mu, ls).cpu().numpy(), pn.log_mvn(Z, mu, ls)) <= 1e-10
