"""Minimal provider of the ``autograd`` names the reference's drivers import.

HIPS autograd 1.3 (environment.yml:9) is neither vendored by the reference nor
installable here.  The drivers only need: ``autograd.numpy`` (NumPy itself),
``autograd.numpy.random``, ``autograd.scipy``, ``autograd.misc.optimizers`` and
``grad``.  Gradients of the free energy come from the CUDA kernels; ``grad`` is
therefore only able to return the gradient function of an objective built by this
package (it carries a ``.gradient`` attribute) and raises for anything else -- there
is no tape and no CPU fallback.
"""
import numpy  # noqa: F401
import scipy  # noqa: F401

from . import misc  # noqa: F401


def grad(fun, argnum=0):
    g = getattr(fun, 'gradient', None)
    if g is None:
        raise NotImplementedError(
            'autograd.grad: no tape here -- only objectives created by normflows '
            '(optimization.gradient_create / objectives.gradient_create) carry a CUDA gradient')
    return g
