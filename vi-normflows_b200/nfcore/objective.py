"""Host-side driver of the shared-parameter free energy (the object behind
``gradient_create`` of get_data.py:72-117).

NumPy in / NumPy out, exactly like the reference's ``objective(params, t)`` and
``gradient(params, t)``; the arithmetic over the samples happens in libnfcore on
the GPU.  No CPU fallback: constructing the object without a CUDA device raises.
"""
import ctypes

import numpy as np
import torch

from . import _lib, ops, semantics as sem, sharding
from ._lib import NFError

_NP2T = {np.dtype('float32'): torch.float32, np.dtype('float64'): torch.float64}


def resolve_target(target, D):
    """Map what the reference's callers pass as ``target`` to (kernel id name, raw params).

    Accepted: the names 'p1'..'p4', 'trial1', 'gmm' (get_data.py:61-66 constants),
    'gmm_exp' (experimentation.py:18-23 constants); a callable whose __name__ is one of
    those (the reference's own functions are recognised by name); or a tuple
    ('mixdiag', pi[G], mu[G,D], var[G,D] or covariance matrices [G,D,D] -- full matrices for D <= 4, as
    distributions.py:20-25 takes them) / ('gaussdiag', mu[D], log_sigma_diag[D]).
    Arbitrary Python callables cannot be fused into the kernel and are rejected.
    """
    if callable(target) and not isinstance(target, (str, tuple)):
        name = getattr(target, '__name__', None)
        if name not in ('p1', 'p2', 'p3', 'p4', 'trial1', 'gmm'):
            raise NFError('target callable %r is not one of the reference potentials; pass a name or a '
                          "('mixdiag'|'gaussdiag', ...) tuple (arbitrary callables cannot be fused)" % (name,))
        target = name
    if isinstance(target, str):
        if target in ('p1', 'p2', 'p3', 'p4', 'trial1', 'U1', 'U2', 'U3', 'U4'):
            if D != 2:
                raise NFError('%s is a 2-D potential (get_data.py:29), dim_z=%d' % (target, D))
            return target, None
        if target in ('gmm', 'gmm_exp'):
            if D != 1:
                raise NFError('gmm is the 1-D mixture of get_data.py:61-66, dim_z=%d' % D)
            if target == 'gmm':
                pi, mu, sg = [0.3, 0.7], [-1.0, 3.0], [1.0, 1.0]
            else:
                pi, mu, sg = [0.5, 0.5], [-1.0, 1.0], [0.5, 0.5]
            return 'mixdiag', np.concatenate([pi, mu, np.square(sg)]).astype(np.float64)
        raise NFError('unknown target %r' % target)
    kind = target[0]
    if kind == 'mixdiag':
        pi, mu, var = (np.asarray(a, dtype=np.float64) for a in target[1:])
        mu = mu.reshape(-1, D)
        if var.ndim == 3:      # distributions.py:23-25 takes (G, D, D) covariance matrices
            off = var - np.array([np.diag(np.diag(v)) for v in var])
            if np.abs(off).max() > 0:
                # full covariances: their own target (the kernel inverts the matrices, D <= 4)
                if D > 4:
                    raise NFError('mixture with full covariance matrices: dim_z <= 4 (diagonal covariances have no limit)')
                G = mu.shape[0]
                if pi.shape[0] == G - 1:
                    pi = np.concatenate([pi, [1 - pi.sum()]])
                if pi.shape[0] != G or var.shape != (G, D, D):
                    raise NFError('mixture: inconsistent component count')
                return 'mixfull', np.concatenate([pi, mu.ravel(), var.ravel()])
            var = np.array([np.diag(v) for v in var])
        var = var.reshape(-1, D)
        G = mu.shape[0]
        if pi.shape[0] == G - 1:            # distributions.py:63-66: last weight is 1 - sum(pi)
            pi = np.concatenate([pi, [1 - pi.sum()]])
        if pi.shape[0] != G or var.shape[0] != G:
            raise NFError('mixdiag: inconsistent component count')
        return 'mixdiag', np.concatenate([pi, mu.ravel(), var.ravel()])
    if kind == 'gaussdiag':
        mu, ls = (np.asarray(a, dtype=np.float64).reshape(-1) for a in target[1:])
        if mu.shape[0] != D or ls.shape[0] != D:
            raise NFError('gaussdiag: mu/log_sigma_diag must have dim_z entries')
        return 'gaussdiag', np.concatenate([mu, ls])
    raise NFError('unknown target %r' % (target,))


class SharedFlowObjective:
    """Free energy of K shared-parameter planar layers and its gradient on the GPU.

    value_and_grad(params, z0) returns what get_data.py:113/115 return on the same z0:
    (mean_n(log q_K - log p), d/dparams of it), params laid out [W(K*D) | U(K*D) | B(K)].
    """

    def __init__(self, target, K, D, log_eps=1e-7, dtype='float32', semantics='script', device=None,
                 process_group=None):
        ops.require_cuda()
        self.K, self.D = int(K), int(D)
        self.log_eps = float(log_eps)
        self.np_dtype = np.dtype(dtype)
        self.dtype = _NP2T[self.np_dtype]
        self.semantics = semantics
        self.device = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
        self.target_name, tparams = resolve_target(target, self.D)
        self.tparams = None
        if tparams is not None:
            self.tparams = torch.as_tensor(tparams.astype(self.np_dtype)).to(self.device)
        K_, D_ = self.K, self.D
        self.n_lc = 2 * K_ * D_ + 2 * K_
        # one device buffer for the per-layer constants, one for every result: one H2D, one D2H per call
        self._lc_host = torch.empty(self.n_lc, dtype=self.dtype).pin_memory()
        self._lc = torch.empty(self.n_lc, dtype=self.dtype, device=self.device)
        # results [F parts (3) | g_w | g_s | g_b | g_c] followed by three slots the sharded exchange fills
        # (sample count in two parts, status): everything a step returns crosses PCIe in one copy
        self.n_res = 3 + self.n_lc
        self._res = torch.zeros(self.n_res + 3, dtype=self.dtype, device=self.device)
        self._res_host = torch.empty(self.n_res + 3, dtype=self.dtype).pin_memory()
        self._eps_dev = None
        self._eps_pinned = None
        self.group = process_group
        self.launches = 0
        self._small_ar = None
        if process_group is not None:
            import torch.distributed as dist
            ok = 0
            if self.dtype == torch.float32:
                try:    # one-shot exchange over NVLink peer memory; NCCL when the peers cannot be mapped
                    from .peer import SmallAllReduce
                    ar = SmallAllReduce(process_group, self.device)
                    ok = 1 if self.n_res <= ar.max_count else 0
                except Exception:
                    ar, ok = None, 0
            # every rank must take the same path: one rank in the peer kernel and another in NCCL would hang
            flag = torch.tensor([ok], dtype=torch.int32, device=self.device)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=process_group)
            if int(flag.item()) == 1:
                self._small_ar = ar
        self._session = None
        self._session_cap = 0
        self._tparams_host = None if tparams is None else np.ascontiguousarray(tparams.astype(self.np_dtype))

    def __del__(self):
        try:
            if self._session is not None:
                _lib.load().nf_session_destroy(self._session)
                self._session = None
        except Exception:
            pass

    # -- host-side parameter algebra (libnfcore, plain C: the NumPy version costs ~85 us per call) ----
    def _constants(self, params):
        """[W | U | B] float64 -> ([w | s | b | c] float64, the contiguous float64 parameter vector)."""
        lib = _lib.load()
        p = np.ascontiguousarray(params, dtype=np.float64).ravel()
        if p.size != 2 * self.K * self.D + self.K:
            raise NFError('params must have 2*K*D + K = %d entries, got %d' % (2 * self.K * self.D + self.K, p.size))
        lc = np.empty(self.n_lc, dtype=np.float64)
        _lib.check(lib.nf_script_constants_host(ctypes.c_void_p(p.ctypes.data), self.K, self.D,
                                                sem.SEMANTICS.index(self.semantics), ctypes.c_void_p(lc.ctypes.data)))
        return lc, p

    def _chain(self, p, res, n):
        """res = [F parts (3) | g_w | g_s | g_b | g_c] float64 -> d(sum F)/d[W | U | B] / n."""
        lib = _lib.load()
        g = np.ascontiguousarray(res[3:], dtype=np.float64)
        out = np.empty(p.size, dtype=np.float64)
        _lib.check(lib.nf_script_chain_host(ctypes.c_void_p(p.ctypes.data), ctypes.c_void_p(g.ctypes.data), self.K, self.D,
                                            sem.SEMANTICS.index(self.semantics), ctypes.c_double(1.0 / n),
                                            ctypes.c_void_p(out.ctypes.data)))
        return out

    # -- host-buffer path: nf_session_elbo_host_* (chunked H2D overlapped with the kernels) ----
    def _host_eval(self, params, z0, backward):
        """z0 in host memory (NumPy array or CPU torch tensor, pinned or not) -> float64 result vector
        [F parts (3) | g_w | g_s | g_b | g_c] (sums), N and the float64 parameter vector."""
        lib = _lib.load()
        if isinstance(z0, torch.Tensor):
            if z0.dtype != self.dtype or not z0.is_contiguous():
                z0 = z0.to(self.dtype).contiguous()
            ptr, N = z0.data_ptr(), z0.shape[0]
            shape = tuple(z0.shape)
        else:
            z0 = np.ascontiguousarray(z0, dtype=self.np_dtype)
            ptr, N = z0.ctypes.data, z0.shape[0]
            shape = z0.shape
        if len(shape) != 2 or shape[1] != self.D:
            raise NFError('z0 must have shape (N, %d), got %r' % (self.D, shape))
        K, D = self.K, self.D
        if self._session is None or N > self._session_cap:
            if self._session is not None:
                lib.nf_session_destroy(self._session)
                self._session = None
            cap = max(int(N), 1)
            h = ctypes.c_void_p()
            with torch.cuda.device(self.device):
                _lib.check(lib.nf_session_create(ctypes.byref(h), cap, D, K, self.np_dtype.itemsize))
            self._session, self._session_cap = h, cap
        lc64, p64 = self._constants(params)
        lc = lc64.astype(self.np_dtype)
        out = np.empty(3 + self.n_lc, dtype=self.np_dtype)
        tp = self._tparams_host
        fn = lib.nf_session_elbo_host_f32 if self.np_dtype == np.float32 else lib.nf_session_elbo_host_f64
        creal = ctypes.c_float if self.np_dtype == np.float32 else ctypes.c_double
        rc = fn(self._session, ctypes.c_void_p(ptr), N, ctypes.c_void_p(lc.ctypes.data), ops.TARGET_IDS[self.target_name],
                ctypes.c_void_p(tp.ctypes.data) if tp is not None else None, 0 if tp is None else tp.size,
                creal(self.log_eps), 1 if backward else 0, ctypes.c_void_p(out.ctypes.data))
        _lib.check(rc)
        self.launches += 2
        return out.astype(np.float64), N, p64

    def _host_objective(self, params, z0, backward, exchange=False):
        """nf_session_objective_host_*: parameters in, (F parts, gradient) out -- one library call per objective call.
        exchange: this rank's shard; the sums go over the ranks of the box (nf_session_objective_host_exchange_f32) and the
        returned count is the global one."""
        lib = _lib.load()
        if isinstance(z0, torch.Tensor):
            if z0.dtype != self.dtype or not z0.is_contiguous():
                z0 = z0.to(self.dtype).contiguous()
            ptr, shape = z0.data_ptr(), tuple(z0.shape)
        else:
            z0 = np.ascontiguousarray(z0, dtype=self.np_dtype)
            ptr, shape = z0.ctypes.data, z0.shape
        if len(shape) != 2 or shape[1] != self.D:
            raise NFError('z0 must have shape (N, %d), got %r' % (self.D, shape))
        N = shape[0]
        p = np.ascontiguousarray(params, dtype=np.float64).ravel()
        if p.size != 2 * self.K * self.D + self.K:
            raise NFError('params must have 2*K*D + K = %d entries, got %d' % (2 * self.K * self.D + self.K, p.size))
        if self._session is None or N > self._session_cap:
            if self._session is not None:
                lib.nf_session_destroy(self._session)
                self._session = None
            h = ctypes.c_void_p()
            with torch.cuda.device(self.device):
                _lib.check(lib.nf_session_create(ctypes.byref(h), max(int(N), 1), self.D, self.K, self.np_dtype.itemsize))
            self._session, self._session_cap = h, max(int(N), 1)
        parts = np.empty(3, dtype=np.float64)
        grad = np.empty(p.size, dtype=np.float64) if backward else None
        tp = self._tparams_host
        f32 = self.np_dtype == np.float32
        if exchange:
            ar = self._small_ar
            ng = ctypes.c_int64(0)
            _lib.check(lib.nf_session_objective_host_exchange_f32(
                self._session, ptr, N, p.ctypes.data, sem.SEMANTICS.index(self.semantics), ops.TARGET_IDS[self.target_name],
                tp.ctypes.data if tp is not None else None, 0 if tp is None else tp.size, ctypes.c_float(self.log_eps),
                ar._ptrs, ar.world, ar.rank, ctypes.c_uint64(ar.next_seq()), parts.ctypes.data, grad.ctypes.data, ctypes.byref(ng)))
            self.launches += 3
            return parts, int(ng.value), grad
        fn = lib.nf_session_objective_host_f32 if f32 else lib.nf_session_objective_host_f64
        _lib.check(fn(self._session, ptr, N, p.ctypes.data, sem.SEMANTICS.index(self.semantics), ops.TARGET_IDS[self.target_name],
                      tp.ctypes.data if tp is not None else None, 0 if tp is None else tp.size,
                      (ctypes.c_float if f32 else ctypes.c_double)(self.log_eps), 1 if backward else 0, parts.ctypes.data,
                      grad.ctypes.data if backward else None))
        self.launches += 2
        return parts, N, grad

    def _is_host(self, z0):
        return self.group is None and not (isinstance(z0, torch.Tensor) and z0.is_cuda)

    # -- sample staging ------------------------------------------------------------------
    def stage_samples(self, z0):
        """Copy host samples (N,D) into device memory (pinned staging buffer, async H2D)."""
        if isinstance(z0, torch.Tensor) and z0.is_cuda:
            if z0.dtype != self.dtype:
                z0 = z0.to(self.dtype)
            return z0.contiguous()
        if isinstance(z0, torch.Tensor):
            if z0.dtype == self.dtype and z0.is_pinned() and z0.is_contiguous():
                # caller-owned pinned buffer: straight async H2D, no staging copy
                if z0.ndim != 2 or z0.shape[1] != self.D:
                    raise NFError('z0 must have shape (N, %d), got %r' % (self.D, tuple(z0.shape)))
                n = z0.numel()
                if self._eps_dev is None or self._eps_dev.numel() < n:
                    self._eps_dev = torch.empty(max(n, 1), dtype=self.dtype, device=self.device)
                dev = self._eps_dev[:n]
                dev.copy_(z0.view(-1), non_blocking=True)
                return dev.view(z0.shape[0], self.D)
            z0 = z0.numpy()
        z0 = np.ascontiguousarray(z0, dtype=self.np_dtype)
        if z0.ndim != 2 or z0.shape[1] != self.D:
            raise NFError('z0 must have shape (N, %d), got %r' % (self.D, z0.shape))
        n = z0.shape[0] * self.D
        if self._eps_pinned is None or self._eps_pinned.numel() < n:
            self._eps_pinned = torch.empty(max(n, 1), dtype=self.dtype).pin_memory()
        if self._eps_dev is None or self._eps_dev.numel() < n:
            self._eps_dev = torch.empty(max(n, 1), dtype=self.dtype, device=self.device)
        self._eps_pinned[:n].numpy()[:] = z0.reshape(-1)
        dev = self._eps_dev[:n]
        dev.copy_(self._eps_pinned[:n], non_blocking=True)
        return dev.view(z0.shape[0], self.D)

    # -- evaluation ----------------------------------------------------------------------
    def _launch(self, params, eps_dev, backward):
        K, D = self.K, self.D
        lc64, p64 = self._constants(params)
        self._lc_host.numpy()[:] = lc64
        self._lc.copy_(self._lc_host, non_blocking=True)
        lc = self._lc
        w = lc[:K * D].view(K, D)
        s = lc[K * D:2 * K * D].view(K, D)
        b = lc[2 * K * D:2 * K * D + K]
        cc = lc[2 * K * D + K:]
        r = self._res
        if backward and self._small_ar is not None:
            # objective + exchange in one call (one launch when the packed-pair kernel covers the shape)
            ops.planar_elbo_exchange(eps_dev, w, s, b, cc, self.target_name, self.tparams, self.log_eps, r, self._small_ar)
            self.launches += 1
            self._exchanged = True
            return p64
        out = (r[:3], r[3:3 + K * D].view(K, D), r[3 + K * D:3 + 2 * K * D].view(K, D),
               r[3 + 2 * K * D:3 + 2 * K * D + K], r[3 + 2 * K * D + K:self.n_res]) if backward else (r[:3], None, None, None, None)
        ops.planar_elbo(eps_dev, w, s, b, cc, self.target_name, self.tparams, self.log_eps, backward, out)
        self.launches += 2
        self._exchanged = False
        return p64

    def _exchange_only(self, n_local):
        """The exchange of `_fetch` without the read-back (enqueued on the current stream): for device-side loops."""
        if self._small_ar is not None:
            self._small_ar.counted(self._res, self.n_res, n_local)
        else:
            sharding.reduce_with_count(self._res[:self.n_res], n_local, self.group)

    def _fetch(self, n_local):
        """exchange (when sharded) + D2H of [F parts | grads | count, status]; returns float64 array and global N."""
        n_global = n_local
        counted = False
        if self.group is not None:
            if self._small_ar is not None:
                if not self._exchanged:
                    self._small_ar.counted(self._res, self.n_res, n_local)
                counted = True
            else:
                # one NCCL all-reduce carries the sums and this rank's sample count (nothing cached between calls)
                n_global = sharding.reduce_with_count(self._res[:self.n_res], n_local, self.group)
        self._res_host.copy_(self._res, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        res = self._res_host.numpy().astype(np.float64)
        if counted:
            from .peer import SmallAllReduce
            n_global = SmallAllReduce.global_count(res[self.n_res:])
        return res[:self.n_res], n_global

    def value_and_grad(self, params, z0):
        if (self.group is not None and self._small_ar is not None and self.np_dtype == np.float32
                and not (isinstance(z0, torch.Tensor) and z0.is_cuda) and np.shape(z0)[0] > 0
                and 3 + self.n_lc <= self._small_ar.max_count):
            # sharded, host samples: chunked upload + kernels + peer exchange + chain rule in one library call
            parts, n, grad = self._host_objective(params, z0, True, exchange=True)
            return parts[0] / n, grad
        if self._is_host(z0):
            if np.shape(z0)[0] > 0:
                parts, n, grad = self._host_objective(params, z0, True)
                return parts[0] / n, grad
            res, n, p64 = self._host_eval(params, z0, True)
        else:
            eps_dev = self.stage_samples(z0)
            p64 = self._launch(params, eps_dev, True)
            res, n = self._fetch(eps_dev.shape[0])
        return res[0] / n, self._chain(p64, res, n)

    def value_and_grad_rng(self, params, num_samples, seed, step):
        """value_and_grad with the samples drawn inside the kernel (nf_planar_elbo_fwd_bwd_rng_f32): the library's own
        counter-based stream for (seed, step) instead of the NumPy draw of get_data.py:100-101 -- opt-in, for runs that
        do not have to reproduce the reference's draws.  Nothing but the K x D constants goes up, nothing but F and the
        gradient comes back.  float32, D <= 2, single process."""
        if self.group is not None or self.dtype != torch.float32:
            raise NFError('value_and_grad_rng: float32, single process')
        lib = _lib.load()
        N = int(num_samples)
        p = np.ascontiguousarray(params, dtype=np.float64).ravel()
        if p.size != 2 * self.K * self.D + self.K:
            raise NFError('params must have 2*K*D + K = %d entries, got %d' % (2 * self.K * self.D + self.K, p.size))
        if self._session is None:
            h = ctypes.c_void_p()
            with torch.cuda.device(self.device):
                _lib.check(lib.nf_session_create(ctypes.byref(h), 1, self.D, self.K, 4))
            self._session, self._session_cap = h, 1
        parts = np.empty(3, dtype=np.float64)
        grad = np.empty(p.size, dtype=np.float64)
        tp = self._tparams_host
        # nf_session_objective_rng_f32: constants up, one launch, F and the gradient back -- one library call
        _lib.check(lib.nf_session_objective_rng_f32(self._session, int(seed), int(step), N, p.ctypes.data,
                                                    sem.SEMANTICS.index(self.semantics), ops.TARGET_IDS[self.target_name],
                                                    tp.ctypes.data if tp is not None else None, 0 if tp is None else tp.size,
                                                    ctypes.c_float(self.log_eps), parts.ctypes.data, grad.ctypes.data))
        self.launches += 1
        return parts[0] / N, grad

    def value(self, params, z0):
        if self._is_host(z0):
            res, n, _ = self._host_eval(params, z0, False)
            return res[0] / max(n, 1)
        eps_dev = self.stage_samples(z0)
        self._launch(params, eps_dev, False)
        res, n = self._fetch(eps_dev.shape[0])
        return res[0] / n

    def metrics(self, params, z0):
        """(mean log q_K, mean log p) -- get_data.py:80-95 ``other_metrics``."""
        if self._is_host(z0):
            res, n, _ = self._host_eval(params, z0, False)
            return res[1] / max(n, 1), res[2] / max(n, 1)
        eps_dev = self.stage_samples(z0)
        self._launch(params, eps_dev, False)
        res, n = self._fetch(eps_dev.shape[0])
        return res[1] / n, res[2] / n


    # -- whole optimisation loop on the device (nf_script_train_*) ---------------------------------------
    _OPT = {'sgd': (0, lambda h: (h.get('mass', 0.9), 0.0, 0.0)),
            'rmsprop': (1, lambda h: (h.get('gamma', 0.9), h.get('eps', 1e-8), 0.0)),
            'adam': (2, lambda h: (h.get('b1', 0.9), h.get('b2', 0.999), h.get('eps', 1e-8)))}

    def train(self, params, eps_all, optimizer='adam', step_size=1e-3, state=None, **hyper):
        """Run len(eps_all) optimiser iterations without leaving the GPU.

        eps_all: (n_iters, N, D) samples, slice t being the draw of iteration t.  Update rules and defaults
        are autograd 1.3's (autograd.misc.optimizers).  Returns (final params float64, trace of F per iteration).
        After the call ``self.train_state`` describes the optimiser (its running averages and the number of
        iterations done); passing it back as ``state`` continues the same run -- a run cut into several calls
        follows the trajectory of one call (the reference keeps one optimiser for all iterations, get_data.py:140).
        Script semantics only.  With a process group (float32, NVLink peer exchange available) every rank passes ITS
        shard of the samples, (n_iters, N_local, D): each iteration's sums are exchanged on the device and every rank
        applies the same update (nf_script_train_sharded_f32); all ranks return the same parameters."""
        if self.semantics != 'script':
            raise NFError('train(): script semantics only')
        if self.group is not None and self._small_ar is None:
            raise NFError('train() with a process group needs the NVLink peer exchange (float32, <= 8 ranks of one box)')
        lib = _lib.load()
        eps_all = np.ascontiguousarray(eps_all, dtype=self.np_dtype)
        if eps_all.ndim != 3 or eps_all.shape[2] != self.D:
            raise NFError('eps_all must have shape (n_iters, N, %d)' % self.D)
        n_iters, N, D = eps_all.shape
        K = self.K
        P = 2 * K * D + K
        dev = self.device
        code, hp = self._OPT[optimizer]
        h1, h2, h3 = hp(hyper)
        # the kernels read every iteration's slice with 16-byte vector loads: pad the slices to a multiple of 16 bytes
        per16 = 16 // self.np_dtype.itemsize
        stride = -(-(N * D) // per16) * per16
        if stride != N * D and n_iters > 1:
            padded = np.zeros((n_iters, stride), dtype=self.np_dtype)
            padded[:, :N * D] = eps_all.reshape(n_iters, N * D)
            eps_all = padded
        else:
            stride = N * D
        eps_dev = torch.as_tensor(eps_all).to(dev)
        p_dev = torch.as_tensor(np.asarray(params, dtype=np.float64).astype(self.np_dtype)).to(dev)
        if state is not None:
            if state['optimizer'] != optimizer or state['opt'].numel() != 2 * P or state['opt'].dtype != self.dtype:
                raise NFError('train(): `state` belongs to a different optimiser or problem')
            opt, first_iter, resume = state['opt'], int(state['iters']), 1
        else:
            opt, first_iter, resume = torch.empty(2 * P, dtype=self.dtype, device=dev), 0, 0
        trace = torch.empty(max(n_iters, 1), dtype=self.dtype, device=dev)
        nws = lib.nf_script_train_workspace_bytes_n(N, K, D, self.np_dtype.itemsize)
        ws = torch.empty(nws, dtype=torch.uint8, device=dev)
        fn = lib.nf_script_train_f32 if self.np_dtype == np.float32 else lib.nf_script_train_f64
        creal = ctypes.c_float if self.np_dtype == np.float32 else ctypes.c_double
        vp = lambda t: ctypes.c_void_p(t.data_ptr()) if t is not None else None
        with torch.cuda.device(dev):
            st = torch.cuda.current_stream(dev).cuda_stream
            if self.group is None:
                rc = fn(vp(eps_dev), n_iters, N, D, K, vp(p_dev), ops.TARGET_IDS[self.target_name], vp(self.tparams),
                        0 if self.tparams is None else self.tparams.numel(), creal(self.log_eps), code, creal(step_size),
                        creal(h1), creal(h2), creal(h3), first_iter, resume, stride, vp(opt), vp(trace), vp(ws), nws,
                        ctypes.c_void_p(st))
            else:
                import torch.distributed as dist
                cnt = torch.tensor([N, n_iters], dtype=torch.int64, device=dev)
                mx = cnt.clone()
                dist.all_reduce(cnt, op=dist.ReduceOp.SUM, group=self.group)
                dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=self.group)
                if int(mx[1]) * self._small_ar.world != int(cnt[1]):
                    raise NFError('train(): every rank must run the same number of iterations')
                ar = self._small_ar
                status = torch.zeros(1, dtype=torch.float32, device=dev)
                rc = lib.nf_script_train_sharded_f32(vp(eps_dev), n_iters, N, int(cnt[0]), D, K, vp(p_dev), ops.TARGET_IDS[self.target_name],
                                                     vp(self.tparams), 0 if self.tparams is None else self.tparams.numel(),
                                                     ctypes.c_float(self.log_eps), code, ctypes.c_float(step_size), ctypes.c_float(h1),
                                                     ctypes.c_float(h2), ctypes.c_float(h3), first_iter, resume, stride, vp(opt), vp(trace),
                                                     vp(ws), nws, ar._ptrs, ar.world, ar.rank, ctypes.c_uint64(ar.seq + 1), vp(status),
                                                     ctypes.c_void_p(st))
                ar.seq += max(n_iters, 0)
                if rc == 0 and float(status.item()) != 0.0:
                    raise NFError('train(): a rank did not arrive at an exchange within the time limit; the parameters are not valid')
        _lib.check(rc)
        self.launches += 3 * n_iters + 1
        self.train_state = {'opt': opt, 'iters': first_iter + n_iters, 'optimizer': optimizer}
        return p_dev.double().cpu().numpy(), trace[:n_iters].double().cpu().numpy()
