"""adam / rmsprop / sgd with the call signatures and update rules of autograd 1.3's
``autograd.misc.optimizers`` (restated from its published algorithm; the module is a
third-party dependency of the reference, pinned at environment.yml:9, not vendored).

Call sites in the reference: optimization.py:118 (adam, keywords), get_data.py:140
(rmsprop, positional), experimentation.py:109 (sgd, positional).  ``grad(x, i)`` is
called once per iteration and ``callback(x, i, g)`` runs BEFORE the update.  Works on
flat float64 NumPy vectors (what every reference driver passes).
"""
import numpy as np


def sgd(grad, x, callback=None, num_iters=200, step_size=0.1, mass=0.9):
    x = np.array(x, dtype=np.float64)
    velocity = np.zeros(len(x))
    for i in range(num_iters):
        g = grad(x, i)
        if callback:
            callback(x, i, g)
        velocity = mass * velocity - (1.0 - mass) * g
        x = x + step_size * velocity
    return x


def rmsprop(grad, x, callback=None, num_iters=100, step_size=0.1, gamma=0.9, eps=10 ** -8):
    x = np.array(x, dtype=np.float64)
    avg_sq_grad = np.ones(len(x))
    for i in range(num_iters):
        g = grad(x, i)
        if callback:
            callback(x, i, g)
        avg_sq_grad = avg_sq_grad * gamma + g ** 2 * (1 - gamma)
        x = x - step_size * g / (np.sqrt(avg_sq_grad) + eps)
    return x


def adam(grad, x, callback=None, num_iters=100, step_size=0.001, b1=0.9, b2=0.999, eps=10 ** -8):
    x = np.array(x, dtype=np.float64)
    m = np.zeros(len(x))
    v = np.zeros(len(x))
    for i in range(num_iters):
        g = grad(x, i)
        if callback:
            callback(x, i, g)
        m = (1 - b1) * g + b1 * m
        v = (1 - b2) * (g ** 2) + b2 * v
        mhat = m / (1 - b1 ** (i + 1))
        vhat = v / (1 - b2 ** (i + 1))
        x = x - step_size * mhat / (np.sqrt(vhat) + eps)
    return x
