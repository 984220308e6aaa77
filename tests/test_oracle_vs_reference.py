"""CPU, build container only: the oracle against the UNMODIFIED reference run under the
import shim, on fresh random inputs (skipped where /root/reference does not exist)."""
import numpy as np
import pytest

from oracle import refshim, planar_np as pn

pytestmark = pytest.mark.skipif(not refshim.available(), reason='reference checkout not present')


@pytest.fixture(scope='module')
def ref():
    ns = refshim.load_script_namespace('get_data.py', 142)
    from normflows import flows, distributions
    yield ns, flows, distributions
    refshim.uninstall()


@pytest.mark.parametrize('tname', ['p1', 'p2', 'p3', 'p4', 'trial1'])
def test_potentials_and_objective(ref, tname):
    ns, _, _ = ref
    rs = np.random.RandomState(123)
    z = 2.5 * rs.randn(500, 2)
    np.testing.assert_allclose(pn.TARGETS[tname](z), ns[tname](z), rtol=1e-13, atol=1e-300)
    K = 12
    params = 0.5 * rs.randn(2 * K * 2 + K)
    z0 = rs.randn(300, 2)
    f = refshim.reference_script_objective(ns, ns[tname], z0, 1e-7, K)
    assert abs(pn.script_objective(params, z0, pn.TARGETS[tname], 1e-7, K) - f(params)) < 1e-12


def test_distributions(ref):
    _, flows, dist = ref
    rs = np.random.RandomState(5)
    Z = rs.randn(50, 3)
    mu = rs.randn(3); ls = 0.3 * rs.randn(3)
    np.testing.assert_allclose(pn.log_mvn(Z, mu, ls), dist.log_mvn(Z, mu, ls), rtol=1e-13)
    Z1 = rs.randn(20, 1)
    np.testing.assert_allclose(pn.log_mvn(Z1, np.array([0.2]), np.array([0.3])),
                               dist.log_mvn(Z1, np.array([0.2]), np.array([0.3])), rtol=1e-13)
    np.testing.assert_allclose(pn.log_std_norm(Z), dist.log_std_norm(Z), rtol=1e-14)
    X = (rs.rand(7, 11) < 0.3).astype(float); p = rs.rand(7, 11)
    assert abs(pn.log_bern_mult(X, p) - dist.log_bern_mult(X, p)) < 1e-12
    w, u, b = rs.randn(50, 3), rs.randn(50, 3), rs.randn(50)
    np.testing.assert_allclose(pn.planar_flow(Z, w, u, b), flows.planar_flow(Z, w, u, b), rtol=1e-13)
