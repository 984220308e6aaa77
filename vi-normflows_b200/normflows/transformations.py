"""normflows/normflows/transformations.py: elementwise helpers (host-side NumPy; the
MLP kernels fuse their own ReLU / sigmoid)."""
import numpy as np

eps = 1e-7                                    # transformations.py:4


def sigmoid(x):                               # :7-8
    return 1 / (1 + np.exp(-x))


def logit(x):                                 # :11-12
    return np.log(eps + x / (1 - x))


def affine(Z, slope, intercept):              # :15-18
    return np.matmul(slope, Z.T).T + intercept


relu = lambda x: np.maximum(x, 0)             # :21  # noqa: E731
