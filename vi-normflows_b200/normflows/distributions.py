"""normflows/normflows/distributions.py: densities and sampling.

Given CUDA tensors, ``mvn`` / ``prob_gm`` / ``log_mvn`` evaluate on the GPU (nf_target_density_*: the same fused
target code the objectives run, csrc/nf_targets.cuh) and return CUDA tensors.  Given NumPy arrays they are the
reference's host-side formulas: small conveniences the drivers apply to a few hundred points they hold on the host
(test_gmm.py:6-11, the plotting paths) -- by design not a compute path; the objectives never go through them.
``sample_from_pz`` pushes its samples through the K planar layers on the GPU.
"""
import numpy as np
import torch

from .transformations import sigmoid
from .flows import planar_flow


def _is_t(*xs):
    return any(isinstance(x, torch.Tensor) for x in xs)


def _gpu_density(Z, kind, *parts):
    """CUDA tensor Z (N, D) -> (N,) density through nf_target_density_* (kind: 'mixdiag' | 'mixfull' | 'gaussdiag')."""
    from nfcore import ops
    tp = torch.cat([torch.as_tensor(np.asarray(p.detach().cpu() if isinstance(p, torch.Tensor) else p, dtype=np.float64).ravel())
                    for p in parts]).to(Z.device, Z.dtype)
    return ops.target_density(Z, kind, tp)


def mvn(Z, mu, sigma_diag):                           # distributions.py:10-32
    N, D = Z.shape
    if isinstance(Z, torch.Tensor) and Z.is_cuda:
        sg = np.asarray(sigma_diag.detach().cpu() if isinstance(sigma_diag, torch.Tensor) else sigma_diag, dtype=np.float64)
        if len(sg) == 1:                              # a scalar variance (:20-22): D must be 1
            return _gpu_density(Z, 'mixdiag', [1.0], mu, sg.reshape(-1)[:1]).reshape(N, 1)
        if D <= 4:                                    # a full matrix (:23-25)
            return _gpu_density(Z, 'mixfull', [1.0], mu, sg.reshape(D, D)).reshape(N, 1)
    mu = mu.reshape(-1, D)
    if len(sigma_diag) == 1:
        siginv = 1 / sigma_diag
        det_sigma = sigma_diag
    else:                                             # a full matrix, as the reference treats it (:23-25)
        siginv = np.linalg.inv(sigma_diag)
        det_sigma = np.linalg.det(sigma_diag)
    zdiff = Z - mu
    const = (2 * np.pi) ** (-D / 2) * det_sigma ** (-1 / 2)
    prod = np.dot(zdiff, siginv).reshape(N, D)
    return const * np.exp(-0.5 * np.sum(prod * zdiff, axis=1)).reshape(N, 1)


def log_mvn(Z, mu, log_sigma_diag):                   # distributions.py:35-54
    N, D = Z.shape
    if isinstance(Z, torch.Tensor) and Z.is_cuda and not (isinstance(mu, torch.Tensor) and mu.requires_grad):
        return torch.log(_gpu_density(Z, 'gaussdiag', mu, log_sigma_diag))
    mu = mu.reshape(-1, D)
    logdet_sigma = np.sum(log_sigma_diag)
    if len(log_sigma_diag) == 1:
        siginv = np.exp(1 - log_sigma_diag)           # kept as in the reference (:46-47)
    else:
        siginv = np.linalg.inv(np.diag(np.exp(log_sigma_diag)))
    const = -D / 2 * np.log(2 * np.pi) - 0.5 * logdet_sigma
    zdiff = Z - mu
    return const - 0.5 * np.sum(np.matmul(zdiff, siginv) * zdiff, axis=1)


def log_std_norm(x):                                  # distributions.py:57
    if _is_t(x):                                      # inside the GPU objective: stays on the device / on the tape
        return -x.shape[1] / 2 * np.log(2 * np.pi) - 0.5 * torch.sum(x * x, dim=1)
    return -x.shape[1] / 2 * np.log(2 * np.pi) - 0.5 * np.sum(x * x, axis=1)


def prob_gm(Z, mu, sigma_diag, pi):                   # distributions.py:60-69
    G = pi.shape[0] + 1
    if isinstance(Z, torch.Tensor) and Z.is_cuda:
        N, D = Z.shape
        tonp = lambda a: np.asarray(a.detach().cpu() if isinstance(a, torch.Tensor) else a, dtype=np.float64)
        pi_all = np.concatenate([tonp(pi).ravel(), [1 - tonp(pi).sum()]])
        sg = tonp(sigma_diag)
        if D == 1:
            return _gpu_density(Z, 'mixdiag', pi_all, tonp(mu), sg.reshape(G)).reshape(N, 1)
        if D <= 4:
            return _gpu_density(Z, 'mixfull', pi_all, tonp(mu), sg.reshape(G, D, D)).reshape(N, 1)
    prob = 0
    for g in range(G):
        pi_g = 1 - np.sum(pi) if g == G - 1 else pi[g]
        prob = prob + mvn(Z, mu[g], sigma_diag[g]) * pi_g
    return prob


def log_prob_gm(Z, mu, log_sigma_diag, logit_pi):     # distributions.py:72-83
    pi = sigmoid(logit_pi)
    assert mu.shape[0] == pi.shape[0] + 1, 'Number of means does not match number of components'
    assert Z.shape[1] == mu.shape[1], 'Dimensions of random variable and mean vector not aligned'
    return np.log(prob_gm(Z, mu, np.exp(log_sigma_diag), pi))


def log_bern_mult(X, p):                              # distributions.py:86-89
    eps = 1e-7
    if _is_t(X, p):                                   # fused kernel (nf_bernoulli_ll), one scalar for the batch
        from nfcore import tape
        p = p if isinstance(p, torch.Tensor) else torch.as_tensor(p)
        X = torch.as_tensor(X, dtype=p.dtype, device=p.device)
        return tape.BernoulliLL.apply(X, p, eps)
    return np.sum(X * np.log(eps + p) + (1 - X) * np.log(eps + 1 - p))


def sample_from_pz(mu, log_sigma_diag, W, U, b, K):   # distributions.py:92-102
    """randn(N, D) * sd + mu from the global NumPy stream (:96-98, the draws the reference makes), then the K planar
    steps with per-sample parameters (:100-101) -- on the GPU in ONE launch (nf_planar_amortised_fwd_*, whose
    reparameterisation sqrt(eps + exp(log_sigma)) and step are optimization.py:74-75,84 = this function's)."""
    from nfcore import ops
    mu = np.asarray(mu)
    N, D = mu.shape
    dt = np.float32 if mu.dtype == np.float32 else np.float64
    z0 = np.random.randn(N, D)                        # global NumPy RNG, as the reference (:97)
    if K == 0:
        return (z0 * np.sqrt(1e-7 + np.exp(np.asarray(log_sigma_diag))) + mu).astype(dt)
    ops.require_cuda()
    dev = torch.device('cuda', torch.cuda.current_device())
    def t(a, shape):
        return torch.as_tensor(np.ascontiguousarray(a, dtype=dt).reshape(shape)).to(dev)
    zK, _, _ = ops.amortised_forward(t(z0, (N, D)), t(mu, (N, D)), t(log_sigma_diag, (N, D)), t(np.asarray(W)[:K], (K, N, D)),
                                     t(np.asarray(U)[:K], (K, N, D)), t(np.asarray(b)[:K], (K, N)), 1e-7, save=False)
    return zK.cpu().numpy()


def make_samples_z(mu, log_sigma_diag, W, U, b, K):   # distributions.py:105-118
    return sample_from_pz(mu, log_sigma_diag, W, U, b, K)
