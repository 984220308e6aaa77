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

    t = 0
    while t < num_iter:
        # iterations [t, end) run on the device; iteration t is the only one of the chunk whose callback prints
        end = min(num_iter, (t // 100 + 1) * 100)
        z_first = draw()                                     # gradient(params, t)
        if t % 100 == 0:                                     # callback :129-132 (objective, then other_metrics)
            z_obj, z_met = draw(), draw()
            if verbose:
                energy = obj.value(x, z_obj)
                joint, entropy = obj.metrics(x, z_met)
                print("Iteration {}; Energy: {}; Joint: {}; Entropy: {}".format(t, energy, joint, entropy))
        last_in_chunk = end == num_iter
        n_mid = end - t - 1 - (1 if last_in_chunk and end - t > 1 else 0)
        zs = [z_first] + [draw() for _ in range(max(n_mid, 0))]
        x, _ = obj.train(x, np.stack(zs), optimizer=optimizer, step_size=lr)
        t += len(zs)
        if last_in_chunk:
            if t < num_iter:                                 # the final iteration: its draws come before the final metrics
                z_last = draw()
            else:
                z_last = None
            z_met, z_obj = draw(), draw()                    # callback :133-138 (other_metrics, then objective)
            if verbose:
                print("\nFINAL METRICS\n")
                joint, entropy = obj.metrics(x, z_met)
                print("Free energy: ", obj.value(x, z_obj))
                print("Joint: ", joint)
                print("Entropy: ", entropy)
            if z_last is not None:
                x, _ = obj.train(x, z_last[None], optimizer=optimizer, step_size=lr)
                t += 1
    return x
