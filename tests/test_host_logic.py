"""CPU: host-side algebra (semantics chain rule, optimisers, target resolution) and the
C-ABI surface (library loads, exports every symbol include/nfcore.h declares)."""
import ctypes
import os

import numpy as np
import pytest
import torch

import nfcore
from nfcore import semantics as sem
from oracle import planar_torch as pt


@pytest.mark.parametrize('semantics', ['script', 'paper'])
def test_semantics_chain_rule(semantics):
    rs = np.random.RandomState(0)
    K, D = 5, 3
    W, U = rs.randn(K, D), rs.randn(K, D)
    g_w, g_s, g_b, g_c = rs.randn(K, D), rs.randn(K, D), rs.randn(K), rs.randn(K)
    Wt = torch.tensor(W, requires_grad=True); Ut = torch.tensor(U, requires_grad=True)
    wu = (Wt * Ut).sum(1)
    alpha = pt.m(wu) - wu
    n = torch.linalg.norm(Wt, dim=1)
    if semantics == 'script':
        S = Ut + alpha[:, None] * Wt / n[:, None]; c = wu
    else:
        S = Ut + alpha[:, None] * Wt / (n ** 2)[:, None]; c = (Wt * S).sum(1)
    S_np, c_np = sem.forward(W, U, semantics)
    np.testing.assert_allclose(S_np, S.detach().numpy(), rtol=1e-13)
    np.testing.assert_allclose(c_np, c.detach().numpy(), rtol=1e-13)
    L = (torch.tensor(g_w) * Wt).sum() + (torch.tensor(g_s) * S).sum() + (torch.tensor(g_c) * c).sum()
    dW_t, dU_t = torch.autograd.grad(L, (Wt, Ut))
    dW, dU, dB = sem.backward(W, U, g_w, g_s, g_b, g_c, semantics)
    np.testing.assert_allclose(dW, dW_t.numpy(), rtol=1e-11, atol=1e-12)
    np.testing.assert_allclose(dU, dU_t.numpy(), rtol=1e-11, atol=1e-12)
    np.testing.assert_allclose(dB, g_b)


def test_optimizers_closed_form():
    from autograd.misc.optimizers import adam, rmsprop, sgd
    g = lambda x, i: 2 * x
    x0 = np.array([1.0, -2.0])
    calls = []
    x = adam(g, x0, callback=lambda x, i, gg: calls.append((i, x.copy())), num_iters=3, step_size=0.1)
    # first adam step moves every coordinate by exactly step_size (bias-corrected m/sqrt(v) = sign)
    np.testing.assert_allclose(calls[1][1], x0 - 0.1 * np.sign(x0), rtol=1e-6)
    assert [c[0] for c in calls] == [0, 1, 2] and x.shape == x0.shape
    x = rmsprop(g, x0, None, 1, 0.1)            # positional form used at get_data.py:140
    gg = 2 * x0
    np.testing.assert_allclose(x, x0 - 0.1 * gg / (np.sqrt(0.9 + 0.1 * gg ** 2) + 1e-8))
    x = sgd(g, x0, None, 2, 0.1)                # positional form used at experimentation.py:109
    v1 = -0.1 * 2 * x0; x1 = x0 + 0.1 * v1
    v2 = 0.9 * v1 - 0.1 * 2 * x1
    np.testing.assert_allclose(x, x1 + 0.1 * v2)


def test_library_loads_and_exports_header_symbols():
    assert os.path.isfile(nfcore.LIB_PATH), 'run __graft_entry__.build() first'
    lib = ctypes.CDLL(nfcore.LIB_PATH)
    declared = nfcore.declared_symbols()
    assert 'nf_planar_elbo_fwd_bwd_f32' in declared and 'nf_planar_elbo_fwd_bwd_f64' in declared
    for name in declared:
        assert hasattr(lib, name), 'libnfcore.so does not export ' + name
    nfcore.load()
    from nfcore import _lib
    assert set(_lib.PROTOTYPES) == set(declared)
    assert lib.nf_built_for_sm() == 100
    # every status of the header has its own explanation
    lib.nf_status_string.restype = ctypes.c_char_p
    texts = [lib.nf_status_string(c) for c in range(0, -8, -1)]
    assert len(set(texts)) == 8 and b'unknown' not in b''.join(texts) and lib.nf_status_string(-99) == b'unknown status'
    assert b'did not arrive' in lib.nf_status_string(-7)            # NF_ERR_PEER


def test_no_cpu_fallback():
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    from nfcore.objective import SharedFlowObjective
    with pytest.raises(nfcore.NFError):
        SharedFlowObjective('p1', 4, 2)


def test_resolve_target():
    from nfcore.objective import resolve_target
    assert resolve_target('p1', 2) == ('p1', None)
    def p3(z): return z
    assert resolve_target(p3, 2)[0] == 'p3'
    name, tp = resolve_target('gmm', 1)
    np.testing.assert_allclose(tp, [0.3, 0.7, -1, 3, 1, 1])
    name, tp = resolve_target(('mixdiag', [0.3], [[0, 1], [2, 3]], [[1, 1], [2, 2]]), 2)
    np.testing.assert_allclose(tp[:2], [0.3, 0.7])
    name, tp2 = resolve_target(('mixdiag', [0.3], [[0, 1], [2, 3]], [np.diag([1., 1.]), np.diag([2., 2.])]), 2)
    np.testing.assert_allclose(tp2, tp)
    # non-diagonal covariance matrices (distributions.py:23-25) are a target of their own, up to four dimensions
    name, tp3 = resolve_target(('mixdiag', [0.3], [[0, 1], [2, 3]], [np.array([[1., .2], [.2, 1.]]), np.diag([2., 2.])]), 2)
    assert name == 'mixfull' and tp3.size == 2 * (1 + 2 + 4)
    np.testing.assert_allclose(tp3[:2], [0.3, 0.7])
    np.testing.assert_allclose(tp3[6:10], [1., .2, .2, 1.])
    with pytest.raises(nfcore.NFError):
        resolve_target(('mixdiag', [0.3], np.zeros((2, 5)), np.tile(np.eye(5) + 0.1, (2, 1, 1))), 5)
    with pytest.raises(nfcore.NFError):
        resolve_target(lambda z: z, 2)
    with pytest.raises(nfcore.NFError):
        resolve_target('p1', 3)


@pytest.mark.parametrize('semantics', ['script', 'paper'])
@pytest.mark.parametrize('K,D', [(1, 1), (8, 2), (5, 40)])
def test_c_host_algebra_matches_numpy_algebra(semantics, K, D):
    """nf_script_constants_host / nf_script_chain_host (plain C in libnfcore, no GPU needed) against the NumPy
    statement of the same map and chain rule in nfcore/semantics.py (which the oracle tests pin to get_data.py:106-108)."""
    import ctypes
    from nfcore import _lib, semantics as sem
    lib = _lib.load()
    rs = np.random.RandomState(K * 100 + D)
    W, U, B = rs.randn(K, D), rs.randn(K, D), rs.randn(K)
    p = sem.pack(W, U, B)
    sid = sem.SEMANTICS.index(semantics)
    lc = np.empty(2 * K * D + 2 * K)
    assert lib.nf_script_constants_host(ctypes.c_void_p(p.ctypes.data), K, D, sid, ctypes.c_void_p(lc.ctypes.data)) == 0
    S, c = sem.forward(W, U, semantics)
    np.testing.assert_allclose(lc, np.concatenate([W.ravel(), S.ravel(), B, c]), rtol=1e-13, atol=1e-14)
    g = rs.randn(2 * K * D + 2 * K)
    out = np.empty(2 * K * D + K)
    assert lib.nf_script_chain_host(ctypes.c_void_p(p.ctypes.data), ctypes.c_void_p(g.ctypes.data), K, D, sid,
                                    ctypes.c_double(0.25), ctypes.c_void_p(out.ctypes.data)) == 0
    dW, dU, dB = sem.backward(W, U, g[:K * D].reshape(K, D), g[K * D:2 * K * D].reshape(K, D), g[2 * K * D:2 * K * D + K],
                              g[2 * K * D + K:], semantics)
    np.testing.assert_allclose(out, 0.25 * sem.pack(dW, dU, dB), rtol=1e-12, atol=1e-13)
    assert lib.nf_script_constants_host(None, K, D, sid, ctypes.c_void_p(lc.ctypes.data)) != 0


def test_philox_restatement_known_answers():
    """oracle/philox_np.py against the published Philox-4x32-10 known-answer vectors (Random123 kat_vectors)."""
    from oracle import philox_np as ph
    kat = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
           ((0xffffffff,) * 4, (0xffffffff, 0xffffffff), (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0), (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, want in kat:
        got = tuple(int(x[0]) for x in ph.philox4x32_10(*[[c] for c in ctr], *key))
        assert got == want
    z = ph.normal(5, 0, 7, 2)
    assert z.shape == (7, 2) and np.array_equal(ph.normal(5, 0, 3, 2), z[:3])
