"""The reference's one HEAD-signature driver, src/learning_mnist.py, executed UNMODIFIED against this package
(`normflows` here is vi-normflows_b200/normflows): its module-level code, `make_unpack_params`, `get_init_params`,
`logp` and `run_optimization`.  Only where the reference checkout exists (the build container); `mlxtend` -- the
MNIST loader the driver imports, not installed here -- is stubbed, the data are synthetic Bernoulli pixels.

CPU part: the driver's definitions execute and bind to `optimization.optimize`'s HEAD signature
(optimization.py:40-42).  GPU part (needs both a device and the checkout): 20 iterations of `run_optimization`.
"""
import os
import sys
import types

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, 'vi-normflows_b200')
DRIVER = '/root/reference/src/learning_mnist.py'

pytestmark = pytest.mark.skipif(not os.path.isfile(DRIVER), reason='reference checkout not present')


def _load_driver():
    # this package's normflows, not the reference's (oracle.refshim may have installed that one earlier in the session)
    for name in [m for m in sys.modules if m == 'normflows' or m.startswith('normflows.')]:
        f = getattr(sys.modules[name], '__file__', '') or ''
        if not f.startswith(PKG):
            del sys.modules[name]
    if PKG not in sys.path:
        sys.path.insert(0, PKG)
    stub, data = types.ModuleType('mlxtend'), types.ModuleType('mlxtend.data')
    data.loadlocal_mnist = lambda **kw: (_ for _ in ()).throw(RuntimeError('no MNIST files in the test'))
    stub.data = data
    saved = {k: sys.modules.get(k) for k in ('mlxtend', 'mlxtend.data')}
    sys.modules['mlxtend'], sys.modules['mlxtend.data'] = stub, data
    try:
        ns = {'__name__': 'learning_mnist_under_test'}
        with open(DRIVER) as fh:
            exec(compile(fh.read(), DRIVER, 'exec'), ns)
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    import normflows
    assert normflows.__file__.startswith(PKG)
    return ns


def test_driver_definitions_bind_to_this_package():
    ns = _load_driver()
    dim_z, K, dim_x, width = ns['dim_z'], ns['K'], ns['dim_x'], ns['width']
    n_enc_out = 2 * dim_z + 2 * dim_z * K + K
    assert ns['encoder'].D == dim_x * width + width + 2 * (width * width + width) + width * n_enc_out + n_enc_out
    assert ns['decoder'].D == dim_z * width + width + 2 * (width * width + width) + width * dim_x + dim_x
    unpack_params, encode, decode = ns['make_unpack_params']()
    np.random.seed(0)
    init = ns['get_init_params']()
    assert init.shape == (ns['encoder'].D + ns['decoder'].D,)
    phi, theta = unpack_params(init)
    assert phi.shape == (ns['encoder'].D,) and theta.shape == (ns['decoder'].D,)
    from normflows import optimization
    a = optimization._parse_optimize_args((ns['logp'], None, dim_z, K, 10, init, unpack_params, encode, decode, 20, 8, 1e-3), {})
    assert a['encode'] is encode and a['decode'] is decode and a['max_iter'] == 20 and a['step_size'] == 1e-3


@pytest.mark.gpu
def test_driver_runs_twenty_iterations(tmp_path, monkeypatch):
    torch = pytest.importorskip('torch')
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    ns = _load_driver()
    monkeypatch.chdir(tmp_path)
    rs = np.random.RandomState(3)
    X = (rs.rand(128, ns['dim_x']) < 0.13).astype(int)
    unpack_params, encode, decode = ns['make_unpack_params']()
    np.random.seed(1)
    init = ns['get_init_params']()
    phi, theta = ns['run_optimization'](X, init, unpack_params, encode, decode, max_iter=20, batch_size=16, N=None, step_size=1e-3)
    assert phi.shape == (ns['encoder'].D,) and theta.shape == (ns['decoder'].D,)
    assert np.all(np.isfinite(phi)) and np.all(np.isfinite(theta))
    assert np.abs(np.concatenate([phi, theta]) - init).max() > 0      # Adam moved the parameters
