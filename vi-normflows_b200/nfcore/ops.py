"""Thin torch-tensor wrappers over the C ABI (device pointers in, device tensors out).

torch is used for device memory and streams only; the arithmetic is libnfcore's.
"""
import ctypes

import torch

from . import _lib
from ._lib import NFError

TARGET_IDS = {'p1': 0, 'U1': 0, 'p2': 1, 'U2': 1, 'p3': 2, 'U3': 2, 'p4': 3, 'U4': 3, 'trial1': 4,
              'mixdiag': 5, 'gaussdiag': 6}

_DT = {torch.float32: ('f32', 4, ctypes.c_float), torch.float64: ('f64', 8, ctypes.c_double)}


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def require_cuda():
    if not torch.cuda.is_available():
        raise NFError('no CUDA device: libnfcore has no CPU fallback')
    _lib.check(_lib.load().nf_check_device())


_workspaces = {}


def _workspace(K, D, dtype, device):
    lib = _lib.load()
    nbytes = lib.nf_planar_elbo_workspace_bytes(K, D, _DT[dtype][1])
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


def planar_elbo(eps, w, s, b, c, target, target_params=None, log_eps=1e-7, backward=True, out=None):
    """Launch nf_planar_elbo_fwd_bwd_{f32,f64} on the current stream.

    eps [N,D]; w,s [K,D]; b,c [K]; all CUDA tensors of one dtype.  Returns
    (out_F[3], g_w, g_s, g_b, g_c) as device tensors (gradients None when
    backward=False).  Nothing is synchronised.
    """
    lib = _lib.load()
    if not eps.is_cuda:
        raise NFError('planar_elbo needs CUDA tensors (no CPU fallback)')
    dtype = eps.dtype
    if dtype not in _DT:
        raise NFError('unsupported dtype %s' % dtype)
    sfx, _, creal = _DT[dtype]
    N, D = eps.shape
    K = w.shape[0]
    for t in (eps, w, s, b, c):
        if t.dtype != dtype or not t.is_contiguous() or t.device != eps.device:
            raise NFError('all inputs must be contiguous, same dtype, same device')
    tid = TARGET_IDS[target] if isinstance(target, str) else int(target)
    ntp = 0
    if target_params is not None:
        if target_params.dtype != dtype or not target_params.is_contiguous():
            raise NFError('target_params must be contiguous and of the same dtype')
        ntp = target_params.numel()
    dev = eps.device
    if out is None:
        out_F = torch.empty(3, dtype=dtype, device=dev)
        if backward:
            g_w = torch.empty(K, D, dtype=dtype, device=dev)
            g_s = torch.empty(K, D, dtype=dtype, device=dev)
            g_b = torch.empty(K, dtype=dtype, device=dev)
            g_c = torch.empty(K, dtype=dtype, device=dev)
        else:
            g_w = g_s = g_b = g_c = None
    else:
        out_F, g_w, g_s, g_b, g_c = out
    ws = _workspace(K, D, dtype, dev)
    fn = getattr(lib, 'nf_planar_elbo_fwd_bwd_' + sfx)
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev).cuda_stream
        rc = fn(_ptr(eps), N, D, K, _ptr(w), _ptr(s), _ptr(b), _ptr(c), tid, _ptr(target_params), ntp,
                creal(log_eps), _ptr(out_F), _ptr(g_w), _ptr(g_s), _ptr(g_b), _ptr(g_c),
                _ptr(ws), ws.numel(), ctypes.c_void_p(stream))
    _lib.check(rc)
    return out_F, g_w, g_s, g_b, g_c
