"""GPU tests of the in-kernel sample stream (csrc/nf_philox.cuh, nf_planar_elbo_fwd_bwd_rng_f32, nf_philox_normal_f32).

The stream is regenerated on the host (oracle/philox_np.py: the same Philox-4x32-10 counters, Box-Muller in float64)
and fed to the oracle, so F and the gradient the kernel computes on its own draws are checked against the reference
math on the very same samples: float32, 1e-4 (north_star).
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


def _rel(a, ref):
    return float(np.abs(np.asarray(a) - ref).max() / max(np.abs(ref).max(), 1e-300))


@pytest.mark.parametrize('D', [1, 2])
def test_stream_matches_host_regeneration_and_is_grid_independent(D):
    from nfcore import ops
    from oracle import philox_np as ph
    seed, step, N = 0x1234567890ABCDEF, 41, 100003
    dev = ops.philox_normal(seed, step, N, D, DEV).cpu().numpy()
    host = ph.normal(seed, step, N, D)
    assert dev.shape == (N, D)
    # MUFU lg2 / sin / cos against float64 libm on identical uniforms: absolute agreement to a few 1e-6
    assert np.abs(dev - host).max() <= 1e-5, np.abs(dev - host).max()
    # sample n depends on (seed, step, n) only: a shorter request is a prefix, another step or seed is another stream
    pre = ops.philox_normal(seed, step, 4097, D, DEV).cpu().numpy()
    assert np.array_equal(pre, dev[:4097])
    assert not np.array_equal(ops.philox_normal(seed, step + 1, 4097, D, DEV).cpu().numpy(), pre)
    assert not np.array_equal(ops.philox_normal(seed + 1, step, 4097, D, DEV).cpu().numpy(), pre)
    # moments of 100 003 x D standard normals
    assert abs(dev.mean()) < 0.01 and abs(dev.std() - 1) < 0.01
    assert abs(np.mean(dev ** 3)) < 0.03 and abs(np.mean(dev ** 4) - 3) < 0.08
    if D == 2:
        assert abs(np.corrcoef(dev.T)[0, 1]) < 0.01
        assert abs(np.corrcoef(dev[:-1, 0], dev[1:, 0])[0, 1]) < 0.01


@pytest.mark.parametrize('target,D,K,N', [('p1', 2, 32, 1 << 18), ('p3', 2, 16, 200001), ('p2', 2, 64, 160001), ('p4', 2, 5, 4097),
                                          ('gauss', 1, 8, 1 << 18), ('gauss', 2, 32, 70001)])
def test_free_energy_on_in_kernel_draws_matches_oracle(target, D, K, N):
    from nfcore.objective import SharedFlowObjective
    from oracle import philox_np as ph, planar_torch as pt
    seed, step = 20240607, 3
    rs = np.random.RandomState(K + D)
    params = np.concatenate([0.1 + 0.01 * rs.randn(2 * K * D), 0.1 + 0.01 * rs.randn(K)])
    if target == 'gauss':
        mu, ls = 0.5 * rs.randn(D), 0.4 * rs.randn(D)
        arg, fn = ('gaussdiag', mu, ls), (lambda z: torch.exp(pt.log_mvn(z, mu, ls)))
    else:
        arg, fn = target, getattr(pt, target)
    z0 = ph.normal(seed, step, N, D)
    F_ref, g_ref = pt.script_value_and_grad(params, z0, fn, 1e-7, K)
    obj = SharedFlowObjective(arg, K, D, dtype='float32')
    F, grad = obj.value_and_grad_rng(params, N, seed, step)
    eF, eg = abs(F - F_ref) / max(1.0, abs(F_ref)), _rel(grad, g_ref)
    print('rng %s D=%d K=%d N=%d: F err %.2e, grad err %.2e' % (target, D, K, N, eF, eg))
    assert eF <= 1e-4 and eg <= 1e-4, (eF, eg)


def test_in_kernel_draws_equal_generated_samples_bitwise():
    """The kernel that draws in place and the kernel that reads the same stream from memory
    (nf_philox_normal_f32 -> nf_planar_elbo_fwd_bwd_f32) run the same tiles on the same grid: identical bits.
    Two calls with the same (seed, step) agree bit for bit; another step does not."""
    from nfcore import ops, semantics as sem
    K, D, N = 32, 2, (1 << 18) + 77
    rs = np.random.RandomState(1)
    W = 0.1 + 0.01 * rs.randn(K, D); U = 0.1 + 0.01 * rs.randn(K, D); B = 0.1 * np.ones(K)
    S, c = sem.forward(W, U)
    a = [torch.as_tensor(x, dtype=torch.float32).to(DEV).contiguous() for x in (W, S, B, c)]
    seed, step = 99, 12
    r1 = [t.clone() for t in ops.planar_elbo_rng(seed, step, N, *a, 'p1')]
    r2 = [t.clone() for t in ops.planar_elbo_rng(seed, step, N, *a, 'p1')]
    r3 = [t.clone() for t in ops.planar_elbo_rng(seed, step + 1, N, *a, 'p1')]
    eps = ops.philox_normal(seed, step, N, D, DEV)
    mem = ops.planar_elbo(eps, *a, 'p1')
    torch.cuda.synchronize()
    for x, y, z, m in zip(r1, r2, r3, mem):
        assert torch.equal(x, y)
        assert torch.equal(x, m), float((x - m).abs().max())
        assert not torch.equal(x, z)


def test_rng_entry_argument_checks():
    from nfcore import _lib
    lib = _lib.load()
    dev = torch.device(DEV)
    K, D = 4, 3
    z = torch.zeros(64, device=dev)
    ws = torch.empty(lib.nf_planar_elbo_workspace_bytes(K, D, 4), dtype=torch.uint8, device=dev)
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    rc = lib.nf_planar_elbo_fwd_bwd_rng_f32(ctypes.c_uint64(1), ctypes.c_uint64(0), 1000, D, K, p(z), p(z), p(z), p(z), 6, p(z), 2 * D,
                                            ctypes.c_float(1e-7), p(z), p(z), p(z), p(z), p(z), p(ws), ws.numel(), None)
    assert rc == -2                                          # in-kernel draws exist for D <= 2
    assert lib.nf_philox_normal_f32(ctypes.c_uint64(1), ctypes.c_uint64(0), 10, 3, p(z), None) == -2
