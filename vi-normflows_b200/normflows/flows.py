"""normflows/normflows/flows.py on the GPU.

``planar_flow(z, w, u, b, h=np.tanh)`` keeps the reference's two calling conventions:
per-sample parameters ``w,u:(N,D), b:(N,)`` (HEAD, flows.py:26-27) and shared parameters
``w,u:(D,)`` or ``(1,D)``, scalar ``b`` (the stale callers, src/learning_basic_flow.py:33,97,
src/learning_simple_gaussian.py:75).  The value is what flows.py:29-32 computes, including the
sum-over-d-then-broadcast of line 32; pass ``semantics='paper'`` for
z + u_hat tanh(w.z+b).  NumPy arrays in -> NumPy array out (float64 unless the input is
float32); CUDA torch tensors in -> CUDA tensor out.
"""
import numpy as np
import torch

from nfcore import ops, NFError


def _is_tanh(h):
    return h is np.tanh or h is torch.tanh or getattr(h, '__name__', '') == 'tanh'


def planar_flow(z, w, u, b, h=np.tanh, semantics='package'):
    if not _is_tanh(h):
        # the reference ignores `h` and hard-codes tanh (flows.py:32); refuse anything else loudly
        raise NFError('planar_flow: only h=tanh is implemented (the reference ignores h, flows.py:32)')
    ops.require_cuda()
    as_torch = isinstance(z, torch.Tensor)
    if as_torch:
        dev, dtype = z.device, z.dtype
        zt, wt, ut = z, torch.as_tensor(w, dtype=dtype, device=dev), torch.as_tensor(u, dtype=dtype, device=dev)
        bt = torch.as_tensor(b, dtype=dtype, device=dev)
    else:
        z = np.asarray(z)
        np_dtype = np.float32 if z.dtype == np.float32 else np.float64
        dev = torch.device('cuda', torch.cuda.current_device())
        conv = lambda a: torch.as_tensor(np.ascontiguousarray(a, dtype=np_dtype)).to(dev)  # noqa: E731
        zt, wt, ut, bt = conv(z), conv(w), conv(u), conv(np.asarray(b))
    assert zt.ndim == 2, 'z must be (n_samples, n_dim)'
    N, D = zt.shape
    per_sample = wt.ndim == 2 and wt.shape[0] == N and not (N == 1 and bt.ndim == 0)
    if per_sample:
        # flows.py:26-27
        assert wt.shape[0] == N and ut.shape[0] == N and bt.reshape(-1).shape[0] == N, 'Incorrect first dimension'
        assert wt.shape[1] == D and ut.shape[1] == D, 'Incorrect second dimension'
        bt = bt.reshape(N)
    else:
        wt, ut, bt = wt.reshape(-1), ut.reshape(-1), bt.reshape(-1)[:1]
        assert wt.shape[0] == D and ut.shape[0] == D, 'Incorrect second dimension'
    out = ops.planar_flow(zt.contiguous(), wt.contiguous(), ut.contiguous(), bt.contiguous(),
                          per_sample=per_sample, semantics=semantics)
    assert out.shape == zt.shape, f'Incorrect output shape: {tuple(out.shape)}'      # flows.py:33
    return out if as_torch else out.cpu().numpy()


def m(x):                                     # flows.py:42-43
    return -1 + np.log(1 + np.exp(x))


def _get_uhat(u, w):                          # flows.py:37-39 (host-side helper; K*D numbers)
    N, D = u.shape
    wu = np.sum(w * u, axis=1).reshape(-1, 1)
    return u + (m(wu) - wu) * (w / (np.linalg.norm(w, axis=1).reshape(-1, 1) ** 2)).reshape(N, D)
