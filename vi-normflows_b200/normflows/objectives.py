"""Shared-parameter free energy with the call signature the reference's scripts use.

The reference defines this function inline, four times over (get_data.py:72-117,
experimentation.py:46-74, ``Final (master).ipynb`` cell 16, ``updated_experimentation.ipynb``
cell 0).  This module is the single replacement: same arguments, same return tuple,
same RNG consumption (one ``np.random.multivariate_normal`` draw per objective /
gradient / other_metrics call, get_data.py:100-101) -- the sample loop runs on the GPU.
"""
import numpy as np

from nfcore.objective import SharedFlowObjective


def gradient_create(target, eps, dim_z, num_samples, K, dtype='float32', semantics='script',
                    process_group=None):
    """-> (variational_objective, gradient, unpack_params, other_metrics)

    ``target`` is one of the reference potentials (its name, or the reference function
    object itself -- recognised by ``__name__``) or a ('mixdiag'|'gaussdiag', ...) tuple;
    see ``nfcore.objective.resolve_target``.  ``objective(params, t)`` and
    ``gradient(params, t)`` take and return float64 NumPy arrays like the reference.
    """
    obj = SharedFlowObjective(target, K, dim_z, log_eps=eps, dtype=dtype, semantics=semantics,
                              process_group=process_group)

    def unpack_params(params):                       # get_data.py:74-78
        W = params[:K * dim_z].reshape(K, dim_z)
        U = params[K * dim_z:2 * K * dim_z].reshape(K, dim_z)
        B = params[-K:]
        return W, U, B

    def _draw():                                     # get_data.py:100-101 (global NumPy RNG)
        return np.random.multivariate_normal(np.zeros(dim_z), np.eye(dim_z), num_samples)

    def variational_objective(params, t):
        return float(obj.value(params, _draw()))

    def gradient(params, t):
        return obj.value_and_grad(params, _draw())[1]

    def other_metrics(params, t):
        mq, mp = obj.metrics(params, _draw())
        return float(mq), float(mp)

    variational_objective.engine = obj
    gradient.engine = obj
    return variational_objective, gradient, unpack_params, other_metrics


def optimise(func, num_samples, num_iter, lr, K, dim_z=2, optimizer='rmsprop', eps=1e-7, dtype='float32',
             verbose=True, init_params=None):
    """get_data.py:119-142 `optimise` with the loop on the device.

    Same initialisation (0.1 * ones, :122-126), same optimiser (rmsprop(gradient, init, callback, num_iter, lr),
    :140), same callback cadence and printed lines (:128-138), and the SAME consumption of the global NumPy RNG:
    per iteration one draw for the gradient, plus the callback's extra draws every 100 iterations and at the end.
    Between callbacks the iterations run back to back on the GPU (nf_script_train_*): no parameter or gradient
    crosses PCIe.  `optimizer` may also be 'adam' (updated_experimentation.ipynb cell 1) or 'sgd'
    (experimentation.py:109)."""
    obj = SharedFlowObjective(func, K, dim_z, log_eps=eps, dtype=dtype)
    if init_params is None:
        init_params = np.concatenate([0.1 * np.ones(K * dim_z), 0.1 * np.ones(K * dim_z), 0.1 * np.ones(K)])
    x = np.array(init_params, dtype=np.float64)

    def draw():
        return np.random.multivariate_normal(np.zeros(dim_z), np.eye(dim_z), num_samples)

    # Iterations run on the device in runs that end where the reference's callback has to print: the callback of
    # iteration t (get_data.py:128-138) sees the parameters BEFORE update t and draws after gradient(params, t)
    # has drawn, so a run covers iterations [t0, t1) where t1 is the next iteration whose callback prints.  One
    # optimiser state serves all runs (rmsprop(gradient, init_params, callback, num_iter, lr) keeps one, :140).
    def prints(t):
        return t % 100 == 0 or (t + 1) % num_iter == 0

    state = None
    pending = []                 # draws of iterations whose update has not been applied yet
    for t in range(num_iter):
        if prints(t) and pending:
            x, _ = obj.train(x, np.stack(pending), optimizer=optimizer, step_size=lr, state=state)
            state, pending = obj.train_state, []
        pending.append(draw())                               # gradient(params, t)
        if t % 100 == 0:                                     # callback :129-132 (objective, then other_metrics)
            z_obj, z_met = draw(), draw()
            if verbose:
                energy = obj.value(x, z_obj)
                joint, entropy = obj.metrics(x, z_met)
                print("Iteration {}; Energy: {}; Joint: {}; Entropy: {}".format(t, energy, joint, entropy))
        if (t + 1) % num_iter == 0:                          # callback :133-138 (other_metrics, then objective)
            z_met, z_obj = draw(), draw()
            if verbose:
                print("\nFINAL METRICS\n")
                joint, entropy = obj.metrics(x, z_met)
                print("Free energy: ", obj.value(x, z_obj))
                print("Joint: ", joint)
                print("Entropy: ", entropy)
    if pending:
        x, _ = obj.train(x, np.stack(pending), optimizer=optimizer, step_size=lr, state=state)
    return x


def target_density(target, Z, dim_z=None, dtype='float64'):
    """The target density at the points Z (N, dim_z) on the GPU: what visualise_flow (Final (master).ipynb cell 16)
    evaluates on its contour grid with `func(z)`.  `target` as for gradient_create.  NumPy in, NumPy out (N,)."""
    import torch
    from nfcore import ops
    from nfcore.objective import resolve_target
    Z = np.asarray(Z)
    D = Z.shape[1] if dim_z is None else dim_z
    name, tparams = resolve_target(target, D)
    ops.require_cuda()
    dev = torch.device('cuda', torch.cuda.current_device())
    dt = torch.float32 if np.dtype(dtype) == np.float32 else torch.float64
    tp = None if tparams is None else torch.as_tensor(tparams).to(dev, dt)
    return ops.target_density(torch.as_tensor(np.ascontiguousarray(Z)).to(dev, dt), name, tp).cpu().numpy()


def density_grid(target, xs, ys, dtype='float64'):
    """Density of a 2-D target on the grid xs x ys -> array (len(ys), len(xs)), as np.meshgrid(xs, ys) orders it."""
    X, Y = np.meshgrid(np.asarray(xs, dtype=np.float64), np.asarray(ys, dtype=np.float64))
    return target_density(target, np.stack([X.ravel(), Y.ravel()], axis=1), 2, dtype).reshape(Y.shape)


def flow_samples(params, z0, K, dim_z, dtype='float64', semantics='script', log_eps=1e-7):
    """Samples z0 (N, dim_z) pushed through the K shared-parameter layers described by `params` ([W | U | B],
    get_data.py:74-78): returns (z_K, log q_K) -- the scatter visualise_flow draws over the contours, and the
    variational log-density of those points (get_data.py:104-111).  NumPy in, NumPy out."""
    import ctypes
    import torch
    from nfcore import _lib, ops, semantics as sem
    ops.require_cuda()
    lib = _lib.load()
    p = np.ascontiguousarray(params, dtype=np.float64).ravel()
    lc = np.empty(2 * K * dim_z + 2 * K, dtype=np.float64)
    _lib.check(lib.nf_script_constants_host(ctypes.c_void_p(p.ctypes.data), K, dim_z, sem.SEMANTICS.index(semantics), ctypes.c_void_p(lc.ctypes.data)))
    dev = torch.device('cuda', torch.cuda.current_device())
    dt = torch.float32 if np.dtype(dtype) == np.float32 else torch.float64
    t = torch.as_tensor(lc).to(dev, dt)
    KD = K * dim_z
    zK, lq = ops.chain_forward(torch.as_tensor(np.ascontiguousarray(z0)).to(dev, dt), t[:KD].view(K, dim_z).contiguous(),
                               t[KD:2 * KD].view(K, dim_z).contiguous(), t[2 * KD:2 * KD + K].contiguous(), t[2 * KD + K:].contiguous(), log_eps)
    return zK.cpu().numpy(), lq.cpu().numpy()
