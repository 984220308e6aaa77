"""CPU check of the algebra behind the wide-D "Gram" path (csrc/nf_elbo_gram.cu).

For shared layer constants (w_k, s_k, b_k, c_k) the planar chain of get_data.py:104-109
    a_k = w_k.z_{k-1} + b_k ; t_k = tanh(a_k) ; z_k = z_{k-1} + s_k t_k
only ever moves z inside span{s_j}, so with M[k][j] = w_k.s_j
    a_k   = (z_0 W^T)[k] + b_k + sum_{j<k} M[k][j] t_j             (K x K recurrence per sample)
    z_K   = z_0 + T S
and the reverse pass is the mirror image: with gz_K = dF/dz_K and P = gz_K S^T
    gz_k.s_k = P[k] + sum_{j>k} delta_j M[j][k]
    g_s = T^T gz_K + triu(T^T Delta, 1) W          g_w = Delta^T z_0 + triu(T^T Delta, 1)^T S
Everything of size D is a matrix product; the sequential part is O(K^2) scalars per sample.
This file restates both forms in NumPy and checks them against torch autograd (float64)
through the oracle's target; the CUDA path is compared with the oracle in test_gpu_flows.py.
"""
import numpy as np
import pytest
import torch

from oracle import planar_torch as pt


def _target_and_grad(zK, mu, ls, eps):
    """log(eps + N(z; mu, diag(exp ls))) and d(-that)/dz  (distributions.py:35-54 through exp)."""
    D = zK.shape[1]
    iv = np.exp(-ls)
    df = zK - mu
    q = np.sum(df * df * iv, axis=1)
    dens = np.exp(-0.5 * (D * np.log(2 * np.pi) + ls.sum()) - 0.5 * q)
    lp = np.log(eps + dens)
    gz = (dens / (eps + dens))[:, None] * df * iv
    return lp, gz


def direct_form(z0, w, s, b, c, mu, ls, eps):
    N, D = z0.shape
    K = w.shape[0]
    z = z0.copy()
    T = np.zeros((N, K))
    ld = np.zeros(N)
    for k in range(K):
        t = np.tanh(z @ w[k] + b[k])
        T[:, k] = t
        ld += np.log(eps + np.abs(1 + (1 - t * t) * c[k]))
        z = z + np.outer(t, s[k])
    lq = -0.5 * np.sum(np.log(2 * np.pi) + z0 ** 2, axis=1) - ld
    lp, gz = _target_and_grad(z, mu, ls, eps)
    g_w, g_s = np.zeros_like(w), np.zeros_like(s)
    g_b, g_c = np.zeros(K), np.zeros(K)
    for k in range(K - 1, -1, -1):
        t = T[:, k]
        z = z - np.outer(t, s[k])
        g_s[k] = t @ gz
        hp = 1 - t * t
        x = 1 + hp * c[k]
        gx = -np.sign(x) / (eps + np.abs(x))
        g_c[k] = np.sum(gx * hp)
        ga = (gz @ s[k] - 2 * c[k] * gx * t) * hp
        g_b[k] = ga.sum()
        g_w[k] = ga @ z
        gz = gz + np.outer(ga, w[k])
    return np.array([np.sum(lq - lp), lq.sum(), lp.sum()]), g_w, g_s, g_b, g_c


def gram_form(z0, w, s, b, c, mu, ls, eps):
    N, D = z0.shape
    K = w.shape[0]
    M = w @ s.T                                   # M[k][j] = w_k . s_j
    A = z0 @ w.T + b                              # product 1
    T = np.zeros((N, K))
    ld = np.zeros(N)
    for j in range(K):                            # forward recurrence, column form
        t = np.tanh(A[:, j])
        T[:, j] = t
        ld += np.log(eps + np.abs(1 + (1 - t * t) * c[j]))
        A[:, j + 1:] += np.outer(t, M[j + 1:, j])
    zK = z0 + T @ s                               # product 2
    lq = -0.5 * np.sum(np.log(2 * np.pi) + z0 ** 2, axis=1) - ld
    lp, gzK = _target_and_grad(zK, mu, ls, eps)
    P = gzK @ s.T                                 # product 3
    Dl = np.zeros((N, K))
    g_b, g_c = np.zeros(K), np.zeros(K)
    for k in range(K - 1, -1, -1):                # reverse recurrence, column form
        t = T[:, k]
        hp = 1 - t * t
        x = 1 + hp * c[k]
        gx = -np.sign(x) / (eps + np.abs(x))
        g_c[k] = np.sum(gx * hp)
        ga = (P[:, k] - 2 * c[k] * gx * t) * hp
        Dl[:, k] = ga
        g_b[k] = ga.sum()
        P[:, :k] += np.outer(ga, M[k, :k])
    U = np.triu(T.T @ Dl, 1)                      # product 6: U[k][j] = sum_n t_k delta_j, j > k
    g_s = T.T @ gzK + U @ w                       # product 4
    g_w = Dl.T @ z0 + U.T @ s                     # product 5
    return np.array([np.sum(lq - lp), lq.sum(), lp.sum()]), g_w, g_s, g_b, g_c


@pytest.mark.parametrize('D,K,N', [(8, 5, 64), (64, 16, 200), (256, 64, 50)])
def test_gram_form_equals_direct_form_and_autograd(D, K, N):
    rs = np.random.RandomState(7 * D + K)
    z0 = rs.randn(N, D)
    sc = 1.0 / np.sqrt(D / 2)
    w, s = sc * 0.5 * rs.randn(K, D), sc * 0.5 * rs.randn(K, D)
    b, c = 0.3 * rs.randn(K), 0.4 * rs.randn(K)
    mu, ls = 0.2 * rs.randn(D), 0.3 * rs.randn(D) - 4.0 / D
    eps = 1e-7
    Fd, *gd = direct_form(z0, w, s, b, c, mu, ls, eps)
    Fg, *gg = gram_form(z0, w, s, b, c, mu, ls, eps)
    np.testing.assert_allclose(Fg, Fd, rtol=1e-12, atol=1e-9)
    for a, r in zip(gg, gd):
        np.testing.assert_allclose(a, r, rtol=1e-9, atol=1e-9 * max(1.0, np.abs(r).max()))
    # the same quantity through autograd on the oracle's target
    tw, ts_, tb, tc = (torch.tensor(v, requires_grad=True) for v in (w, s, b, c))
    z = torch.tensor(z0)
    ldt = 0.
    for k in range(K):
        t = torch.tanh(z @ tw[k] + tb[k])
        ldt = ldt + torch.log(eps + torch.abs(1 + (1 - t * t) * tc[k]))
        z = z + torch.outer(t, ts_[k])
    lq = -0.5 * torch.sum(np.log(2 * np.pi) + torch.tensor(z0) ** 2, 1) - ldt
    lp = torch.log(eps + torch.exp(pt.log_mvn(z, mu, ls)))
    F = torch.sum(lq - lp.reshape(-1))
    F.backward()
    assert abs(float(F.detach()) - Fg[0]) <= 1e-10 * max(1.0, abs(float(F.detach())))
    for a, r in zip(gg, (tw.grad, ts_.grad, tb.grad, tc.grad)):
        r = r.numpy()
        np.testing.assert_allclose(a, r, rtol=1e-8, atol=1e-9 * max(1.0, np.abs(r).max()))
