"""Import shim that runs the UNMODIFIED reference in the build container.

TEST INFRASTRUCTURE.  ``/root/reference`` exists only in the build container,
so this module is used by ``oracle/make_golden.py`` (which writes the committed
fixtures under tests/golden/) and by CPU tests that skip when the reference is
absent.  Nothing that runs on the GPU box imports it.

The reference needs HIPS ``autograd`` (environment.yml:9), ``matplotlib``,
``mlxtend`` and ``statsmodels``; none is installed and none can be.  The shim
provides:
  * ``autograd.numpy`` -> numpy, ``autograd.numpy.random`` -> numpy.random,
    ``autograd.scipy`` -> scipy  (forward values are therefore the reference's
    own arithmetic, executed by NumPy);
  * ``autograd.grad`` -> raises (no gradients from the reference here);
  * ``autograd.misc.optimizers`` -> stubs that raise;
  * empty ``matplotlib.pyplot``, ``mlxtend.data.loadlocal_mnist``,
    ``statsmodels.api``.
"""
import os
import sys
import types

REFERENCE_ROOT = os.environ.get('NF_REFERENCE_ROOT', '/root/reference')


def available():
    return os.path.isfile(os.path.join(REFERENCE_ROOT, 'get_data.py'))


def _raise(*a, **k):
    raise RuntimeError('autograd is not available: the reference cannot produce gradients here')


class _Anything(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith('__'):
            raise AttributeError(name)
        return lambda *a, **k: None


_saved = {}


def install():
    """Put the stand-in modules into sys.modules and the reference package on sys.path."""
    import numpy
    import scipy
    names = ['autograd', 'autograd.numpy', 'autograd.numpy.random', 'autograd.scipy',
             'autograd.misc', 'autograd.misc.optimizers', 'matplotlib', 'matplotlib.pyplot',
             'mlxtend', 'mlxtend.data', 'statsmodels', 'statsmodels.api', 'normflows']
    for n in list(sys.modules):
        if n in names or n.startswith('normflows.'):
            _saved[n] = sys.modules.pop(n)
    ag = types.ModuleType('autograd')
    ag.numpy = numpy
    ag.scipy = scipy
    ag.grad = lambda f, *a, **k: _raise
    misc = types.ModuleType('autograd.misc')
    opt = types.ModuleType('autograd.misc.optimizers')
    opt.adam = opt.rmsprop = opt.sgd = _raise
    misc.optimizers = opt
    ag.misc = misc
    sys.modules.update({'autograd': ag, 'autograd.numpy': numpy, 'autograd.numpy.random': numpy.random,
                        'autograd.scipy': scipy, 'autograd.misc': misc, 'autograd.misc.optimizers': opt})
    mpl = _Anything('matplotlib')
    plt = _Anything('matplotlib.pyplot')
    mpl.pyplot = plt
    mlx = types.ModuleType('mlxtend')
    mlxd = types.ModuleType('mlxtend.data')
    mlxd.loadlocal_mnist = _raise
    mlx.data = mlxd
    sm = types.ModuleType('statsmodels')
    sma = _Anything('statsmodels.api')
    sm.api = sma
    sys.modules.update({'matplotlib': mpl, 'matplotlib.pyplot': plt, 'mlxtend': mlx, 'mlxtend.data': mlxd,
                        'statsmodels': sm, 'statsmodels.api': sma})
    pkg = os.path.join(REFERENCE_ROOT, 'normflows')
    if pkg not in sys.path:
        sys.path.insert(0, pkg)


def uninstall():
    """Remove the stand-ins and the reference package again (so the product's own
    ``normflows`` / ``autograd`` compat modules can be imported in the same process)."""
    pkg = os.path.join(REFERENCE_ROOT, 'normflows')
    while pkg in sys.path:
        sys.path.remove(pkg)
    for n in list(sys.modules):
        if n.split('.')[0] in ('autograd', 'matplotlib', 'mlxtend', 'statsmodels', 'normflows'):
            del sys.modules[n]
    sys.modules.update(_saved)
    _saved.clear()


def load_script_namespace(script='get_data.py', last_line=142):
    """exec the definitions (not the training run at the bottom) of a top-level
    reference script: get_data.py lines 1-142, experimentation.py lines 1-74."""
    install()
    with open(os.path.join(REFERENCE_ROOT, script)) as f:
        src = ''.join(f.readlines()[:last_line])
    ns = {'__name__': 'reference_' + script.replace('.py', '')}
    exec(compile(src, os.path.join(REFERENCE_ROOT, script), 'exec'), ns)
    return ns


class PinnedNormal:
    """Context manager that makes ``np.random.multivariate_normal`` (get_data.py:100)
    return a preset array once, so the reference objective runs on chosen z0."""

    def __init__(self, z0):
        self.z0 = z0

    def __enter__(self):
        import numpy
        self._orig = numpy.random.multivariate_normal
        z0 = self.z0
        numpy.random.multivariate_normal = lambda mean, cov, size=None: z0
        return self

    def __exit__(self, *exc):
        import numpy
        numpy.random.multivariate_normal = self._orig
        return False


def reference_script_objective(ns, target, z0, eps, K):
    """Returns f(params)->F evaluating the reference's own ``variational_objective``
    (get_data.py:97-113) on pinned z0."""
    N, D = z0.shape
    objective = ns['gradient_create'](target, eps, D, N, K)[0]

    def f(params):
        with PinnedNormal(z0):
            return float(objective(params, 0))
    return f


def load_mnist_driver():
    """Import src/learning_mnist.py (module level builds encoder/decoder; main() is not run)."""
    install()
    import importlib.util
    path = os.path.join(REFERENCE_ROOT, 'src', 'learning_mnist.py')
    spec = importlib.util.spec_from_file_location('reference_learning_mnist', path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def capture_optimize_F(optimization, logp, X, D, K, N, init_params, unpack_params, encode, decode,
                       max_iter, batch_size, step_size):
    """Reach ``optimize``'s private objective (optimization.py:66-92) by replacing
    ``grad``/``adam``/``tqdm`` in the reference module for the duration of the call."""
    captured = {}

    def fake_grad(f):
        captured['objective'] = f
        return f

    class _Bar:
        def __init__(self, *a, **k): pass
        def update(self, *a): pass
        def close(self): pass
        write = staticmethod(lambda *a, **k: None)

    def fake_adam(gradient, init, step_size=None, callback=None, num_iters=None):
        return init

    saved = (optimization.grad, optimization.adam, optimization.tqdm)
    optimization.grad, optimization.adam, optimization.tqdm = fake_grad, fake_adam, _Bar
    try:
        optimization.optimize(logp, X, D, K, N, init_params, unpack_params, encode, decode,
                              max_iter, batch_size, step_size, verbose=False)
    finally:
        optimization.grad, optimization.adam, optimization.tqdm = saved
    return captured['objective']
