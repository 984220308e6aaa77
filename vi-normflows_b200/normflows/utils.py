"""normflows/normflows/utils.py: mini-batch schedule and the (plot-only) reconstruction dump."""
import numpy as np
import numpy.random as npr

from .config import figs, rs, figname
from .distributions import make_samples_z, sample_from_pz
from .transformations import affine
from . import plotting


def clear_figs():
    [f.unlink() for f in figs.glob('*')]


def get_samples_from_params(phi, theta, X, K):       # utils.py:15-22 (stale helper, kept verbatim in behaviour)
    N, D = X.shape
    Zhat = make_samples_z(*phi, K)
    ZK = Zhat[K, :, :]                               # raises IndexError like the reference
    mu_z, log_sigma_diag_pz, logit_pi, A, B = theta
    Xhat = affine(ZK, A, B) + rs.randn(N, D)
    return Xhat, ZK


def compare_reconstruction(phi, theta, x_true, encode, decode, K, t):    # utils.py:25-38
    eps = 1e-7
    mu0, log_sigma_diag0, W, U, b = encode(phi, x_true)
    K = W.shape[0]
    z = sample_from_pz(mu0, log_sigma_diag0, W, U, b, K)
    logits = decode(theta, z)
    xhat = npr.binomial(1, logits + eps)
    if plotting.plt is not None:
        plotting.plot_mnist(x_true.reshape(28, 28), xhat.reshape(28, 28))
        plotting.plt.savefig(figname.format(K, t))
        plotting.plt.close()
    return xhat


def make_batch_schedule(N, batch_size, max_iter):
    """The index schedule of utils.py:41-52: per epoch one global-RNG permutation split into batches.
    Consumes np.random exactly like the reference's make_batch_iter."""
    n_batches = int(np.ceil(N / batch_size))
    n_epochs = int(np.ceil(max_iter / n_batches))
    idx = np.arange(N)
    sched = []
    for i in range(n_epochs):
        sched.append(np.array_split(np.random.permutation(idx), n_batches))

    def batch_indices(t):
        return sched[int(np.floor(t / n_batches))][t % n_batches]
    return batch_indices


def make_batch_iter(X, batch_size, max_iter):        # utils.py:41-60 (global np.random.permutation)
    N, D = X.shape
    n_batches = int(np.ceil(N / batch_size))
    n_epochs = int(np.ceil(max_iter / n_batches))
    idx = np.arange(N)
    batch_sched = []
    for i in range(n_epochs):
        idx_shuffled = np.random.permutation(idx)
        batch_sched.append(np.array_split(idx_shuffled, n_batches))

    def get_batch(t):
        epoch = int(np.floor(t / n_batches))
        batch = t % n_batches
        idx_batch = batch_sched[epoch][batch]
        return X[idx_batch].reshape(len(idx_batch), D)
    return get_batch
