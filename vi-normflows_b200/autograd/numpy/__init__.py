"""``autograd.numpy`` -> NumPy (the drivers use it as a plain array library)."""
from numpy import *  # noqa: F401,F403
from numpy import random, linalg, fft  # noqa: F401
import numpy as _np

__all__ = [n for n in dir(_np) if not n.startswith('_')]


def __getattr__(name):
    return getattr(_np, name)
