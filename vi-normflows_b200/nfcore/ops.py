"""Thin torch-tensor wrappers over the C ABI (device pointers in, device tensors out).

torch is used for device memory and streams only; the arithmetic is libnfcore's.
"""
import ctypes

import torch

from . import _lib
from ._lib import NFError

TARGET_IDS = {'p1': 0, 'U1': 0, 'p2': 1, 'U2': 1, 'p3': 2, 'U3': 2, 'p4': 3, 'U4': 3, 'trial1': 4,
              'mixdiag': 5, 'gaussdiag': 6, 'mixfull': 7}

_DT = {torch.float32: ('f32', 4, ctypes.c_float), torch.float64: ('f64', 8, ctypes.c_double)}


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


_checked_devices = set()


def require_cuda():
    if not torch.cuda.is_available():
        raise NFError('no CUDA device: libnfcore has no CPU fallback')
    dev = torch.cuda.current_device()
    if dev not in _checked_devices:
        _lib.check(_lib.load().nf_check_device())
        _checked_devices.add(dev)


_workspaces = {}


def _workspace(K, D, dtype, device, N=0):
    lib = _lib.load()
    nbytes = lib.nf_planar_elbo_workspace_bytes_n(N, K, D, _DT[dtype][1])
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
    ws = _workspace(K, D, dtype, dev, N)
    fn = getattr(lib, 'nf_planar_elbo_fwd_bwd_' + sfx)
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev).cuda_stream
        rc = fn(_ptr(eps), N, D, K, _ptr(w), _ptr(s), _ptr(b), _ptr(c), tid, _ptr(target_params), ntp,
                creal(log_eps), _ptr(out_F), _ptr(g_w), _ptr(g_s), _ptr(g_b), _ptr(g_c),
                _ptr(ws), ws.numel(), ctypes.c_void_p(stream))
    _lib.check(rc)
    return out_F, g_w, g_s, g_b, g_c


def planar_elbo_exchange(eps, w, s, b, c, target, target_params, log_eps, out_vec, ar):
    """nf_planar_elbo_fwd_bwd_exchange_f32 on the current stream: this rank's sums are exchanged with the peers of
    `ar` (nfcore.peer.SmallAllReduce) before they land in out_vec = [F parts | g_w | g_s | g_b | g_c | n_hi | n_lo | status]."""
    lib = _lib.load()
    if eps.dtype != torch.float32 or not eps.is_cuda:
        raise NFError('planar_elbo_exchange: float32 CUDA tensors')
    N, D = eps.shape
    K = w.shape[0]
    for t in (eps, w, s, b, c, out_vec):
        if t.dtype != torch.float32 or not t.is_contiguous() or t.device != eps.device:
            raise NFError('all inputs must be contiguous float32 on one device')
    if out_vec.numel() < 3 + 2 * K * D + 2 * K + 3:
        raise NFError('out_vec needs 3 + 2KD + 2K + 3 elements')
    tid = TARGET_IDS[target] if isinstance(target, str) else int(target)
    ntp = 0 if target_params is None else target_params.numel()
    dev = eps.device
    ws = _workspace(K, D, torch.float32, dev, N)
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev).cuda_stream
        rc = lib.nf_planar_elbo_fwd_bwd_exchange_f32(_ptr(eps), N, D, K, _ptr(w), _ptr(s), _ptr(b), _ptr(c), tid, _ptr(target_params), ntp,
                                                     ctypes.c_float(log_eps), _ptr(out_vec), _ptr(ws), ws.numel(), ar._ptrs, ar.world, ar.rank,
                                                     ctypes.c_uint64(ar.next_seq()), ctypes.c_void_p(stream))
    _lib.check(rc)
    return out_vec


def target_density(z, target, target_params=None):
    """nf_target_density_*: out[n] = target(z[n]) for CUDA z [N, D] (float32 or float64)."""
    lib = _lib.load()
    if not z.is_cuda or z.dtype not in _DT:
        raise NFError('target_density needs a float32/float64 CUDA tensor (no CPU fallback)')
    z = z.contiguous()
    N, D = z.shape
    tid = TARGET_IDS[target] if isinstance(target, str) else int(target)
    ntp = 0 if target_params is None else target_params.numel()
    if target_params is not None:
        _same(z.dtype, z.device, target_params)
    out = torch.empty(N, dtype=z.dtype, device=z.device)
    fn = getattr(lib, 'nf_target_density_' + _DT[z.dtype][0])
    with torch.cuda.device(z.device):
        st = torch.cuda.current_stream(z.device).cuda_stream
        _lib.check(fn(_ptr(z), N, D, tid, _ptr(target_params), ntp, _ptr(out), ctypes.c_void_p(st)))
    return out


def chain_forward(z0, w, s, b, c, log_eps=1e-7, want_logq=True):
    """nf_planar_chain_forward_*: (z_K [N, D], log q_K [N] or None) for CUDA samples z0 and layer constants."""
    lib = _lib.load()
    if not z0.is_cuda or z0.dtype not in _DT:
        raise NFError('chain_forward needs float32/float64 CUDA tensors (no CPU fallback)')
    z0 = z0.contiguous()
    _same(z0.dtype, z0.device, w, s, b, c)
    N, D = z0.shape
    K = w.shape[0]
    zK = torch.empty_like(z0)
    lq = torch.empty(N, dtype=z0.dtype, device=z0.device) if want_logq else None
    fn = getattr(lib, 'nf_planar_chain_forward_' + _DT[z0.dtype][0])
    with torch.cuda.device(z0.device):
        st = torch.cuda.current_stream(z0.device).cuda_stream
        _lib.check(fn(_ptr(z0), N, D, K, _ptr(w), _ptr(s), _ptr(b), _ptr(c), _DT[z0.dtype][2](log_eps), _ptr(zK), _ptr(lq), ctypes.c_void_p(st)))
    return zK, lq


def philox_normal(seed, step, N, D, device=None, out=None):
    """nf_philox_normal_f32: the [N, D] float32 samples the free-energy kernel draws in place for (seed, step)."""
    lib = _lib.load()
    dev = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
    if out is None:
        out = torch.empty(N, D, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        st = torch.cuda.current_stream(dev).cuda_stream
        _lib.check(lib.nf_philox_normal_f32(ctypes.c_uint64(seed), ctypes.c_uint64(step), N, D, _ptr(out), ctypes.c_void_p(st)))
    return out


def planar_elbo_rng(seed, step, N, w, s, b, c, target, target_params=None, log_eps=1e-7, out=None):
    """nf_planar_elbo_fwd_bwd_rng_f32 on the current stream: free energy and gradient sums over N samples drawn inside
    the kernel (float32, D <= 2).  Returns (out_F[3], g_w, g_s, g_b, g_c); nothing is synchronised."""
    lib = _lib.load()
    K, D = w.shape
    dev = w.device
    for t in (w, s, b, c):
        if t.dtype != torch.float32 or not t.is_contiguous() or t.device != dev or not t.is_cuda:
            raise NFError('planar_elbo_rng: contiguous float32 CUDA tensors on one device')
    tid = TARGET_IDS[target] if isinstance(target, str) else int(target)
    ntp = 0 if target_params is None else target_params.numel()
    if out is None:
        out = (torch.empty(3, dtype=torch.float32, device=dev), torch.empty(K, D, dtype=torch.float32, device=dev),
               torch.empty(K, D, dtype=torch.float32, device=dev), torch.empty(K, dtype=torch.float32, device=dev),
               torch.empty(K, dtype=torch.float32, device=dev))
    out_F, g_w, g_s, g_b, g_c = out
    ws = _workspace(K, D, torch.float32, dev, N)
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev).cuda_stream
        rc = lib.nf_planar_elbo_fwd_bwd_rng_f32(ctypes.c_uint64(seed), ctypes.c_uint64(step), N, D, K, _ptr(w), _ptr(s), _ptr(b), _ptr(c), tid,
                                                _ptr(target_params), ntp, ctypes.c_float(log_eps), _ptr(out_F), _ptr(g_w), _ptr(g_s),
                                                _ptr(g_b), _ptr(g_c), _ptr(ws), ws.numel(), ctypes.c_void_p(stream))
    _lib.check(rc)
    return out


def _same(dtype, dev, *ts):
    for t in ts:
        if t is not None and (t.dtype != dtype or not t.is_contiguous() or t.device != dev):
            raise NFError('all tensors must be contiguous, of one dtype and on one device')


def planar_flow(z, w, u, b, per_sample=True, semantics='package', out=None):
    """nf_planar_flow_*: one planar step on CUDA tensors (flows.py:8-34)."""
    lib = _lib.load()
    if not z.is_cuda:
        raise NFError('planar_flow needs CUDA tensors (no CPU fallback)')
    dtype, dev = z.dtype, z.device
    _same(dtype, dev, z, w, u, b)
    N, D = z.shape
    if out is None:
        out = torch.empty_like(z)
    fn = getattr(lib, 'nf_planar_flow_' + _DT[dtype][0])
    with torch.cuda.device(dev):
        st = torch.cuda.current_stream(dev).cuda_stream
        rc = fn(_ptr(z), _ptr(w), _ptr(u), _ptr(b), N, D, 1 if per_sample else 0,
                {'package': 0, 'paper': 1}[semantics], _ptr(out), ctypes.c_void_p(st))
    _lib.check(rc)
    return out


def amortised_forward(z0, mu0, ls0, W, U, Bb, log_eps=1e-7, save=True):
    """nf_planar_amortised_fwd_*: (zK [N,D], logdet [N], t_save [K,N] or None)."""
    lib = _lib.load()
    dtype, dev = z0.dtype, z0.device
    if not z0.is_cuda:
        raise NFError('amortised_forward needs CUDA tensors (no CPU fallback)')
    _same(dtype, dev, z0, mu0, ls0, W, U, Bb)
    N, D = z0.shape
    K = W.shape[0]
    zK = torch.empty_like(z0)
    logdet = torch.empty(N, dtype=dtype, device=dev)
    tsave = torch.empty(K, N, dtype=dtype, device=dev) if save else None
    fn = getattr(lib, 'nf_planar_amortised_fwd_' + _DT[dtype][0])
    with torch.cuda.device(dev):
        st = torch.cuda.current_stream(dev).cuda_stream
        rc = fn(_ptr(z0), _ptr(mu0), _ptr(ls0), _ptr(W), _ptr(U), _ptr(Bb), N, D, K, _DT[dtype][2](log_eps),
                _ptr(zK), _ptr(logdet), _ptr(tsave), ctypes.c_void_p(st))
    _lib.check(rc)
    return zK, logdet, tsave


def amortised_backward(z0, ls0, W, U, zK, tsave, g_zK, g_logdet, log_eps=1e-7):
    """nf_planar_amortised_bwd_*: gradients wrt (mu0, ls0, W, U, Bb)."""
    lib = _lib.load()
    dtype, dev = z0.dtype, z0.device
    _same(dtype, dev, z0, ls0, W, U, zK, tsave, g_zK)
    N, D = z0.shape
    K = W.shape[0]
    g_mu0, g_ls0 = torch.empty_like(z0), torch.empty_like(z0)
    g_W, g_U = torch.empty_like(W), torch.empty_like(U)
    g_B = torch.empty(K, N, dtype=dtype, device=dev)
    creal = _DT[dtype][2]
    vec = isinstance(g_logdet, torch.Tensor)         # per-sample dF/dlogdet in device memory: no host round trip
    if vec:
        g_logdet = g_logdet.detach().to(dtype).reshape(N).contiguous()
    fn = getattr(lib, 'nf_planar_amortised_bwd_' + ('v_' if vec else '') + _DT[dtype][0])
    with torch.cuda.device(dev):
        st = torch.cuda.current_stream(dev).cuda_stream
        rc = fn(_ptr(z0), _ptr(ls0), _ptr(W), _ptr(U), _ptr(zK), _ptr(tsave), _ptr(g_zK), _ptr(g_logdet) if vec else creal(g_logdet),
                N, D, K, creal(log_eps), _ptr(g_mu0), _ptr(g_ls0), _ptr(g_W), _ptr(g_U), _ptr(g_B),
                ctypes.c_void_p(st))
    _lib.check(rc)
    return g_mu0, g_ls0, g_W, g_U, g_B


def adam_step(params64, grad, m, v, t, step_size=1e-3, b1=0.9, b2=0.999, eps=1e-8):
    """nf_adam_step: one Adam update of the float64 master parameters on the device (autograd.misc.optimizers.adam)."""
    lib = _lib.load()
    for x in (params64, m, v):
        if x.dtype != torch.float64 or not x.is_cuda or not x.is_contiguous():
            raise NFError('adam_step: params, m, v must be contiguous float64 CUDA tensors')
    grad = grad.contiguous()
    if grad.dtype not in _DT or grad.numel() != params64.numel():
        raise NFError('adam_step: gradient of the wrong type or size')
    with torch.cuda.device(params64.device):
        st = torch.cuda.current_stream(params64.device).cuda_stream
        _lib.check(lib.nf_adam_step(_ptr(params64), _ptr(grad), _DT[grad.dtype][1], _ptr(m), _ptr(v), params64.numel(), float(step_size),
                                    float(b1), float(b2), float(eps), int(t), ctypes.c_void_p(st)))


ACT_IDS = {None: 0, 'none': 0, 'identity': 0, 'relu': 1, 'sigmoid': 2, 'rbf': 3}


def mlp_forward(weights, x, dims, hidden_act='relu', out_act=None):
    """nf_mlp_forward_*: weights flat [P], x [N, D_in] -> (out [N, D_out], fwd_ws)."""
    lib = _lib.load()
    if not x.is_cuda:
        raise NFError('mlp_forward needs CUDA tensors (no CPU fallback)')
    dtype, dev = x.dtype, x.device
    _same(dtype, dev, weights, x)
    D_in, H, L, D_out = dims
    N = x.shape[0]
    if x.shape[1] != D_in or weights.numel() != lib.nf_mlp_param_count(D_in, H, L, D_out):
        raise NFError('mlp_forward: shape mismatch (x %r, weights %d)' % (tuple(x.shape), weights.numel()))
    es = _DT[dtype][1]
    ha, oa = ACT_IDS[hidden_act], ACT_IDS[out_act]
    ws = torch.empty(max(1, lib.nf_mlp_forward_workspace_bytes(N, D_in, H, L, D_out, ha, es) // es), dtype=dtype, device=dev)
    out = torch.empty(N, D_out, dtype=dtype, device=dev)
    fn = getattr(lib, 'nf_mlp_forward_' + _DT[dtype][0])
    with torch.cuda.device(dev):
        st = torch.cuda.current_stream(dev).cuda_stream
        _lib.check(fn(_ptr(weights), _ptr(x), N, D_in, H, L, D_out, ha, oa, _ptr(ws), _ptr(out), ctypes.c_void_p(st)))
    return out, ws


def mlp_backward(weights, x, dims, hidden_act, out_act, fwd_ws, out, g_out, need_gx=False):
    """nf_mlp_backward_*: -> (g_weights [P], g_x [N, D_in] or None)."""
    lib = _lib.load()
    dtype, dev = x.dtype, x.device
    _same(dtype, dev, weights, x, fwd_ws, out, g_out)
    D_in, H, L, D_out = dims
    N = x.shape[0]
    es = _DT[dtype][1]
    bws = torch.empty(max(1, lib.nf_mlp_backward_workspace_bytes(N, D_in, H, L, D_out, es) // es), dtype=dtype, device=dev)
    g_w = torch.empty_like(weights)
    g_x = torch.empty_like(x) if need_gx else None
    fn = getattr(lib, 'nf_mlp_backward_' + _DT[dtype][0])
    with torch.cuda.device(dev):
        st = torch.cuda.current_stream(dev).cuda_stream
        _lib.check(fn(_ptr(weights), _ptr(x), N, D_in, H, L, D_out, ACT_IDS[hidden_act], ACT_IDS[out_act], _ptr(fwd_ws),
                      _ptr(out), _ptr(g_out), _ptr(g_w), _ptr(g_x), _ptr(bws), ctypes.c_void_p(st)))
    return g_w, g_x


def bernoulli_ll(x, p, log_eps=1e-7, g_scale=None):
    """nf_bernoulli_ll_*: (sum [1], g_p or None) with g_p = g_scale * d sum / d p.  g_scale may be a Python float or a
    one-element CUDA tensor (then it is read on the device: nothing synchronises)."""
    lib = _lib.load()
    if isinstance(g_scale, torch.Tensor):
        dtype, dev = p.dtype, p.device
        _same(dtype, dev, x, p)
        es = _DT[dtype][1]
        gs = g_scale.detach().to(dtype).reshape(1).contiguous()
        ws = torch.empty(lib.nf_bernoulli_workspace_bytes(es) // es, dtype=dtype, device=dev)
        out = torch.empty(1, dtype=dtype, device=dev)
        g_p = torch.empty_like(p)
        fn = getattr(lib, 'nf_bernoulli_ll_dev_' + _DT[dtype][0])
        with torch.cuda.device(dev):
            st = torch.cuda.current_stream(dev).cuda_stream
            _lib.check(fn(_ptr(x), _ptr(p), p.numel(), _DT[dtype][2](log_eps), _ptr(gs), _ptr(out), _ptr(g_p), _ptr(ws), ctypes.c_void_p(st)))
        return out, g_p
    if not p.is_cuda:
        raise NFError('bernoulli_ll needs CUDA tensors (no CPU fallback)')
    dtype, dev = p.dtype, p.device
    _same(dtype, dev, x, p)
    es, creal = _DT[dtype][1], _DT[dtype][2]
    ws = torch.empty(lib.nf_bernoulli_workspace_bytes(es) // es, dtype=dtype, device=dev)
    out = torch.empty(1, dtype=dtype, device=dev)
    g_p = torch.empty_like(p) if g_scale is not None else None
    fn = getattr(lib, 'nf_bernoulli_ll_' + _DT[dtype][0])
    with torch.cuda.device(dev):
        st = torch.cuda.current_stream(dev).cuda_stream
        _lib.check(fn(_ptr(x), _ptr(p), p.numel(), creal(log_eps), creal(0.0 if g_scale is None else g_scale),
                      _ptr(out), _ptr(g_p), _ptr(ws), ctypes.c_void_p(st)))
    return out, g_p
