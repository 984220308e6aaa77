"""normflows/normflows/optimization.py on the GPU: the amortised (flow-VAE) free energy and its
Adam loop.

`optimize(logp, X, D, K, N, init_params, unpack_params, encode, decode, max_iter, batch_size,
step_size, verbose=True)` keeps the HEAD signature (optimization.py:40-42), the β schedule
(:71-72), the callback cadence and its extra RNG consumption (:97-116), the
`results/free_energy.txt` append (:113-116) and returns `unpack_params(final_params)`.
`F(phi, theta, t)` (:66-92) is evaluated on the device: the user's `encode` / `decode` /
`logp` closures run on CUDA tensors (their slicing / reshape / transpose is recorded by
torch's tape), and every arithmetic node is a libnfcore kernel (nfcore.tape).
Gradients therefore come from the kernels' reverse passes, not from a CPU tape.
"""
import os

import numpy as np
import numpy.random as npr
import torch
from tqdm import tqdm

from autograd.misc.optimizers import adam, rmsprop  # noqa: F401  (same names the reference imports)
from nfcore import ops, tape

from .flows import planar_flow  # noqa: F401
from .utils import (clear_figs, get_samples_from_params, compare_reconstruction, make_batch_iter,  # noqa: F401
                    make_batch_schedule)
from .config import results

rs = npr.RandomState(0)                     # optimization.py:13

_TORCH_DT = {'float32': torch.float32, 'float64': torch.float64, 'fp32': torch.float32, 'fp64': torch.float64}


def default_dtype():
    return _TORCH_DT[os.environ.get('NORMFLOWS_PRECISION', 'fp32').lower()]


def gradient_create(F, D, N, unpack_params, dtype=None):
    """optimization.py:17-37 -> (variational_objective, gradient); both take float64 NumPy params."""
    ops.require_cuda()
    dt = default_dtype() if dtype is None else dtype
    dev = torch.device('cuda', torch.cuda.current_device())

    def _eval(params, t, need_grad):
        p = torch.as_tensor(np.asarray(params, dtype=np.float64)).to(dev, dt).requires_grad_(need_grad)
        with torch.set_grad_enabled(need_grad):
            phi, theta = unpack_params(p)
            free_energy = F(phi, theta, t)
        return p, free_energy

    def variational_objective(params, t):
        _, fe = _eval(params, t, False)
        return float(fe)

    def gradient(params, t):
        p, fe = _eval(params, t, True)
        g, = torch.autograd.grad(fe, p)
        return g.double().cpu().numpy()

    variational_objective.gradient = gradient
    return variational_objective, gradient


def make_free_energy(logp, X, D, K, encode, decode, max_iter, batch_size, dtype=None, batch_iter=None, device_rng=None):
    """The F(phi, theta, t) of optimization.py:66-92 as a function of CUDA tensors.

    device_rng: None (default) draws z0 from the module's RandomState like the reference (:69) -- 164 k host draws per
    step at the BASELINE batch; an integer seed (or NORMFLOWS_DEVICE_RNG=<seed> in the environment) draws z0 on the
    device instead (opt-in: another stream than the reference's, nothing crosses PCIe)."""
    dt = default_dtype() if dtype is None else dtype
    dev = torch.device('cuda', torch.cuda.current_device())
    if device_rng is None and os.environ.get('NORMFLOWS_DEVICE_RNG'):
        device_rng = int(os.environ['NORMFLOWS_DEVICE_RNG'])
    gen = None if device_rng is None else torch.Generator(device=dev).manual_seed(int(device_rng))
    X_dev = None
    if batch_iter is None:
        # optimization.py:55: same schedule, same consumption of the global NumPy RNG as make_batch_iter,
        # but the data set is uploaded once and mini-batches are gathered on the device
        batch_indices = make_batch_schedule(X.shape[0], batch_size, max_iter)
        X_dev = torch.as_tensor(np.ascontiguousarray(X)).to(dev, dt)

    def F(phi, theta, t):
        eps = 1e-7
        if X_dev is not None:
            Xbatch = X_dev.index_select(0, torch.as_tensor(batch_indices(t), device=dev))  # :68
        else:
            Xbatch = torch.as_tensor(np.ascontiguousarray(batch_iter(t))).to(dev, dt)
        if gen is None:
            z0_np = rs.randn(Xbatch.shape[0], D)                                          # :69 (module RandomState)
            z0 = torch.as_tensor(z0_np).to(dev, dt)
            # first = mean over d of logq0 summed over SAMPLES (:56-59, :77); constant in the parameters
            first = float(np.mean(-D / 2 * np.log(2 * np.pi) - 0.5 * np.sum(z0_np ** 2, axis=0)))
        else:
            z0 = torch.randn(Xbatch.shape[0], D, dtype=dt, device=dev, generator=gen)
            first = torch.mean(-D / 2 * np.log(2 * np.pi) - 0.5 * torch.sum(z0.double() ** 2, dim=0)).to(dt)
        mu0, log_sigma_diag0, W, U, B = encode(phi, Xbatch)                               # :70
        cooling_max = min(max_iter / 4, 10000)                                            # :71
        beta_t = min(1.0, 0.001 + t / cooling_max)                                        # :72
        zk, logdet = tape.AmortisedFlow.apply(z0, mu0, log_sigma_diag0, W, U, B, eps)     # :74-84
        third = torch.mean(logdet)                                                        # :85
        logits = decode(theta, zk)                                                        # :87
        second = torch.mean(logp(Xbatch, zk, logits)) * beta_t                            # :89
        return first - second - third                                                     # :92
    return F


def _parse_optimize_args(args, kwargs):
    """HEAD signature plus the two stale ones still used by src/ (SURVEY.md §2.3)."""
    names = ['logp', 'X', 'D', 'K', 'N', 'init_params', 'unpack_params', 'encode', 'decode', 'max_iter', 'batch_size',
             'step_size']
    if len(args) == 12 or 'encode' in kwargs:
        a = dict(zip(names, args))
        a.update(kwargs)
        return a
    raise TypeError(
        'optimize(): this call uses a signature older than HEAD optimization.py:40-42 '
        '(src/learning_simple_gaussian.py:169-171 / learning_gaussian_mixture.py:181-183 pass no encode/decode); '
        'those drivers rely on autograd tracing arbitrary NumPy code in `logp`, which has no CUDA counterpart. '
        'Use normflows.objectives.gradient_create for shared-parameter flows.')


def optimize(*args, verbose=True, **kwargs):
    a = _parse_optimize_args(args, kwargs)
    logp, X, D, K, N = a['logp'], a['X'], a['D'], a['K'], a['N']
    init_params, unpack_params, encode, decode = a['init_params'], a['unpack_params'], a['encode'], a['decode']
    max_iter, batch_size, step_size = a['max_iter'], a['batch_size'], a['step_size']

    F = make_free_energy(logp, X, D, K, encode, decode, max_iter, batch_size)
    objective, gradient = gradient_create(F, D, N, unpack_params)      # optimization.py:94
    pbar = tqdm(total=max_iter, disable=not verbose)

    def callback(params, t, g):                                        # optimization.py:97-116
        pbar.update()
        if verbose:
            if t % 100 == 0:
                grad_mag = np.linalg.norm(gradient(params, t))
                tqdm.write(f"Iteration {t}; objective: {objective(params, t)} gradient mag: {grad_mag:.3f}")
            if t % 200 == 0:
                try:
                    Xtrue = X[101].reshape(1, -1)
                    phi, theta = unpack_params(params)
                    compare_reconstruction(phi, theta, Xtrue, encode, decode, K, t)
                except (ValueError, IndexError):
                    tqdm.write("nan gradient")
            if t == max_iter - 1:
                variational_lower_bound = objective(params, t)
                try:
                    with (results / 'free_energy.txt').open('a') as f:
                        f.write(f'\n{K} flows: {variational_lower_bound}')
                except OSError:
                    pass

    # adam(gradient, init_params, step_size=step_size, callback=callback, num_iters=max_iter) (optimization.py:118) with
    # the parameters, both moment vectors and the update on the device (nf_adam_step, float64 like the reference): per
    # iteration nothing crosses PCIe but the mini-batch indices; the callback sees host copies only when it prints.
    dev = torch.device('cuda', torch.cuda.current_device())
    dt = default_dtype()
    x = torch.as_tensor(np.asarray(init_params, dtype=np.float64)).to(dev)
    m, v = torch.zeros_like(x), torch.zeros_like(x)
    for t in range(max_iter):
        p = x.to(dt, copy=True).requires_grad_(True)
        phi, theta = unpack_params(p)
        fe = F(phi, theta, t)
        g, = torch.autograd.grad(fe, p)
        if verbose and (t % 100 == 0 or t == max_iter - 1):
            callback(x.cpu().numpy(), t, None)
        else:
            pbar.update()
        ops.adam_step(x, g, m, v, t, step_size)
    variational_params = x.cpu().numpy()
    pbar.close()
    return unpack_params(variational_params)
