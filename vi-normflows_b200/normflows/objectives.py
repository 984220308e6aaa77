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
