"""``autograd.numpy.random`` -> numpy.random (same global state, same RandomState class)."""
from numpy.random import *  # noqa: F401,F403
from numpy.random import RandomState, seed, randn, rand, permutation, multivariate_normal, normal, randint, binomial, choice  # noqa: F401
