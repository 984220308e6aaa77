"""normflows/normflows/nn_models.py: the flat-weight-vector MLP, forward pass on the GPU.

`Feedforward(architecture, random=None, weights=None)`, `.forward(weights, x)`, `.D`, `.params`
keep the reference's signatures and layouts (weights (1, P) flat, x (D_in, N) feature-major,
result (1, D_out, N)).  The arithmetic is libnfcore's nf_mlp_forward / nf_mlp_backward.
`make_objective` / `fit` (the MSE trainer, nn_models.py:86-164) are outside the hot path and
are not provided.
"""
import numpy as np
import torch

from nfcore import ops, tape, NFError

_ACT_TYPES = {'relu': 'relu', 'rbf': 'rbf', 'identity': None, 'linear': None, 'none': None}


def _out_act_name(fn):
    if fn is None:
        return None
    name = getattr(fn, '__name__', '')
    if name == 'sigmoid':
        return 'sigmoid'
    if name == 'relu' or fn is _relu_marker:
        return 'relu'
    probe = np.array([-1.5, 0.0, 2.0])
    try:
        val = np.asarray(fn(probe), dtype=float)
    except Exception as e:      # pragma: no cover
        raise NFError('output_activation_fn could not be identified: %s' % e)
    if np.allclose(val, probe):
        return None
    if np.allclose(val, 1 / (1 + np.exp(-probe))):
        return 'sigmoid'
    if np.allclose(val, np.maximum(probe, 0)):
        return 'relu'
    raise NFError('output_activation_fn must be identity, sigmoid or relu to run on the GPU kernels')


_relu_marker = object()


class Feedforward:
    def __init__(self, architecture, random=None, weights=None):
        self.params = {'H': architecture['width'],                     # nn_models.py:9-14
                       'L': architecture['hidden_layers'],
                       'D_in': architecture['input_dim'],
                       'D_out': architecture['output_dim'],
                       'activation_type': architecture['activation_fn_type'],
                       'activation_params': architecture['activation_fn_params']}
        self.D = ((architecture['input_dim'] * architecture['width'] + architecture['width'])       # :16-19
                  + (architecture['output_dim'] * architecture['width'] + architecture['output_dim'])
                  + (architecture['hidden_layers'] - 1) * (architecture['width'] ** 2 + architecture['width']))
        self.random = random if random is not None else np.random.RandomState(101)
        self.h = architecture['activation_fn']
        atype = str(architecture['activation_fn_type']).lower()
        if atype not in _ACT_TYPES:
            raise NFError("activation_fn_type %r is not implemented by the GPU kernels (relu, rbf, identity)" % atype)
        self._hidden_act = _ACT_TYPES[atype]
        self.weights = self.random.normal(0, 1, size=(1, self.D)) if weights is None else weights
        self.output_activation_fn = architecture.get('output_activation_fn', lambda x: x)
        self._out_act = _out_act_name(architecture.get('output_activation_fn'))
        self.objective_trace = np.empty((1, 1))
        self.weight_trace = np.empty((1, self.D))

    @property
    def dims(self):
        p = self.params
        return (p['D_in'], p['H'], p['L'], p['D_out'])

    def forward(self, weights, x):
        """weights (1, P), x (D_in, N) [or (1, D_in, N)] -> (1, D_out, N)   -- nn_models.py:41-84"""
        D_in, H, L, D_out = self.dims
        assert weights.shape[1] == self.D, f'Incorrect input shape {tuple(weights.shape)}'
        if len(x.shape) == 2:
            assert x.shape[0] == D_in
        else:
            assert x.shape[1] == D_in
            x = x.reshape(D_in, -1)
        ops.require_cuda()
        if isinstance(weights, torch.Tensor) or isinstance(x, torch.Tensor):
            dev = weights.device if isinstance(weights, torch.Tensor) else x.device
            dtype = weights.dtype if isinstance(weights, torch.Tensor) else x.dtype
            wt = torch.as_tensor(weights, dtype=dtype, device=dev).reshape(-1)
            xt = torch.as_tensor(x, dtype=dtype, device=dev).t()              # [N, D_in] (a view of the caller's X)
            out = tape.MLP.apply(wt, xt, self.dims, self._hidden_act, self._out_act)
            return out.t().unsqueeze(0)                                        # (1, D_out, N), feature-major view
        np_dtype = np.float32 if np.asarray(x).dtype == np.float32 else np.float64
        dev = torch.device('cuda', torch.cuda.current_device())
        wt = torch.as_tensor(np.ascontiguousarray(weights, dtype=np_dtype).reshape(-1)).to(dev)
        xt = torch.as_tensor(np.ascontiguousarray(np.asarray(x).T, dtype=np_dtype)).to(dev)
        out, _ = ops.mlp_forward(wt, xt, self.dims, self._hidden_act, self._out_act)
        return out.t().cpu().numpy().reshape(1, D_out, -1)

    def make_objective(self, *a, **k):
        raise NotImplementedError('Feedforward.make_objective/fit (nn_models.py:86-164) is the unrelated MSE trainer; '
                                  'it is outside the flow hot path and not part of this package')

    fit = make_objective


K = 3                                              # nn_models.py:167-175
D = 1
default_architecture = {'width': 8,
                        'hidden_layers': 3,
                        'input_dim': 1,
                        'output_dim': 2 * D + 2 * D * K + 1 * K,
                        'activation_fn_type': 'rbf',
                        'activation_fn_params': 'c=0, alpha=1',
                        'activation_fn': lambda x: np.exp(-1 * (x - 0) ** 2)}
nn = Feedforward(default_architecture)             # expected by src/learning_gaussian_mixture.py:12
