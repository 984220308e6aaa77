"""GPU parity for the per-sample-parameter flow kernels (flows.py:8-43, optimization.py:74-85)
and for the wide-D free energy kernel, against golden vectors produced by the reference and
against the oracle.  float64 1e-10, float32 1e-4 (relative to the largest reference entry)."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


def _rel(a, ref):
    return float(np.abs(np.asarray(a) - ref).max() / max(np.abs(ref).max(), 1e-300))


@pytest.mark.parametrize('D', [1, 2, 40])
@pytest.mark.parametrize('dtype,tol', [(np.float64, 1e-10), (np.float32, 1e-4)])
def test_planar_flow_matches_reference(D, dtype, tol):
    from normflows import flows
    g = np.load(os.path.join(GOLDEN, f'planar_flow_D{D}.npz'))
    out = flows.planar_flow(g['z'].astype(dtype), g['w'].astype(dtype), g['u'].astype(dtype), g['b'].astype(dtype))
    assert out.shape == g['out'].shape and out.dtype == dtype
    assert _rel(out, g['out']) <= tol


def test_planar_flow_shared_params_and_errors():
    import nfcore
    from normflows import flows
    from oracle import planar_np as pn
    rs = np.random.RandomState(3)
    z = rs.randn(300, 2)
    w, u, b = np.array([-5., 1.]), np.array([-2., 1.]), 0.0      # src/learning_basic_flow.py:19-21
    out = flows.planar_flow(z, w, u, b)
    assert _rel(out, pn.planar_flow_shared(z, w, u, b)) <= 1e-10
    out2 = flows.planar_flow(z, w.reshape(1, 2), u.reshape(1, 2), np.array([0.0]))
    np.testing.assert_array_equal(out, out2)
    # paper semantics differs from the package's sum-and-broadcast for D > 1
    outp = flows.planar_flow(z, w, u, b, semantics='paper')
    uhat = pn.get_uhat(u.reshape(1, 2), w.reshape(1, 2))[0]
    ref = z + np.tanh(z @ w + b)[:, None] * uhat[None, :]
    assert _rel(outp, ref) <= 1e-10
    with pytest.raises(AssertionError):
        flows.planar_flow(z, rs.randn(300, 3), rs.randn(300, 3), rs.randn(300))
    with pytest.raises(nfcore.NFError):
        flows.planar_flow(z, w, u, b, h=np.sin)
    # torch tensors stay on the device
    zt = torch.as_tensor(z, device=DEV)
    assert flows.planar_flow(zt, torch.as_tensor(w, device=DEV), torch.as_tensor(u, device=DEV), 0.0).is_cuda


@pytest.mark.parametrize('name', ['amortised_K4_D8', 'amortised_K16_D40', 'amortised_K3_D1', 'amortised_K8_D2'])
@pytest.mark.parametrize('dtype,tol', [(torch.float64, 1e-10), (torch.float32, 1e-4)])
def test_amortised_flow_forward_backward(name, dtype, tol):
    from nfcore import ops
    g = np.load(os.path.join(GOLDEN, name + '.npz'))
    t = lambda k: torch.as_tensor(g[k]).to(DEV, dtype).contiguous()
    z0, mu0, ls0, W, U, B = (t(k) for k in ('z0', 'mu0', 'ls0', 'W', 'U', 'B'))
    zK, logdet, tsave = ops.amortised_forward(z0, mu0, ls0, W, U, B)
    assert _rel(zK.cpu().numpy(), g['zK_ref']) <= tol
    assert _rel(logdet.cpu().numpy(), g['running_sum']) <= tol
    N, D = z0.shape
    # F = first - beta * mean(log N(zK; 0, I)) - mean(logdet)  (oracle/make_golden.py gen_amortised)
    beta = float(g['beta_t'])
    F = np.mean(-D / 2 * np.log(2 * np.pi) - 0.5 * np.sum(g['z0'] ** 2, axis=0)) \
        - beta * float(torch.mean(-D / 2 * np.log(2 * np.pi) - 0.5 * (zK.double() ** 2).sum(1))) \
        - float(logdet.double().mean())
    assert abs(F - float(g['F_ref'])) <= tol * abs(float(g['F_ref']))
    g_zK = (beta / N) * zK
    grads = ops.amortised_backward(z0, ls0, W, U, zK, tsave, g_zK.contiguous(), -1.0 / N)
    for got, key in zip(grads, ('g_mu0', 'g_ls0', 'g_W', 'g_U', 'g_B')):
        assert _rel(got.cpu().numpy(), g[key]) <= tol, key


@pytest.mark.parametrize('D,K,N', [(5, 6, 300), (16, 8, 513), (33, 5, 200), (64, 7, 130), (100, 4, 77), (256, 6, 40)])
def test_wide_free_energy_matches_oracle(D, K, N):
    from nfcore.objective import SharedFlowObjective
    from oracle import planar_torch as pt
    rs = np.random.RandomState(D)
    z0 = rs.randn(N, D)
    sc = 1.0 / np.sqrt(D / 2)
    params = np.concatenate([sc * 0.5 * rs.randn(2 * K * D), 0.3 * rs.randn(K)])
    mu, ls = 0.2 * rs.randn(D), 0.3 * rs.randn(D) - 4.0 / D      # keeps exp(log_mvn) away from total underflow for small D
    tgt = lambda z: torch.exp(pt.log_mvn(z, mu, ls))
    F_ref, g_ref = pt.script_value_and_grad(params, z0, tgt, 1e-7, K)
    for dtype, tol in (('float64', 1e-10), ('float32', 1e-4)):
        obj = SharedFlowObjective(('gaussdiag', mu, ls), K, D, dtype=dtype)
        F, grad = obj.value_and_grad(params, z0)
        assert abs(F - F_ref) <= tol * max(1.0, abs(F_ref)), (dtype, F, F_ref)
        assert _rel(grad, g_ref) <= tol, (dtype, _rel(grad, g_ref))
        assert abs(obj.value(params, z0) - F_ref) <= tol * max(1.0, abs(F_ref))
    if D <= 33:
        G = 2
        pi = np.array([0.4]); mmu = 0.3 * rs.randn(G, D); var = 0.8 + 0.4 * rs.rand(G, D)
        sig = torch.tensor(np.array([np.diag(v) for v in var]))
        tgt = lambda z: pt.prob_gm(z, torch.tensor(mmu), sig, torch.tensor(pi))
        F_ref, g_ref = pt.script_value_and_grad(params, z0, tgt, 1e-7, K)
        obj = SharedFlowObjective(('mixdiag', pi, mmu, var), K, D, dtype='float64')
        F, grad = obj.value_and_grad(params, z0)
        assert abs(F - F_ref) <= 1e-10 * max(1.0, abs(F_ref))
        assert _rel(grad, g_ref) <= 1e-9


def test_amortised_flow_full_size_float32_vs_float64():
    """BASELINE config 4 shapes (B=4096, D=40, K=16): float32 kernels against the float64 kernels (which the
    small golden cases pin to the reference), forward and backward."""
    from nfcore import ops
    torch.manual_seed(7)
    K, N, D = 16, 4096, 40
    g = lambda *shape: torch.randn(*shape, dtype=torch.float64, device=DEV)
    z0, mu0, ls0 = g(N, D), 0.3 * g(N, D), 0.3 * g(N, D)
    # flow parameters of the size an encoder at its initialisation emits: keeps every sample away from the
    # log|1 + h'c| singularity, where float32 inputs alone (not the kernel) decide the last digits
    W, U, B = 0.3 * g(K, N, D) / D ** 0.5, 0.3 * g(K, N, D) / D ** 0.5, 0.3 * g(K, N)
    g_zK = g(N, D) / N
    out = {}
    for dt in (torch.float64, torch.float32):
        a = [t.to(dt).contiguous() for t in (z0, mu0, ls0, W, U, B)]
        zK, logdet, tsave = ops.amortised_forward(*a)
        grads = ops.amortised_backward(a[0], a[2], a[3], a[4], zK, tsave, g_zK.to(dt).contiguous(), -1.0 / N)
        out[dt] = [t.double().cpu().numpy() for t in (zK, logdet) + tuple(grads)]
    for x32, x64 in zip(out[torch.float32], out[torch.float64]):
        assert _rel(x32, x64) <= 1e-4


# ---- matrix-product ("Gram") formulation of the wide-D free energy: csrc/nf_elbo_gram.cu ----
def _gram_case(D, K, N, seed):
    rs = np.random.RandomState(seed)
    z0 = rs.randn(N, D)
    sc = 1.0 / np.sqrt(D / 2)
    params = np.concatenate([sc * 0.5 * rs.randn(2 * K * D), 0.3 * rs.randn(K)])
    mu, ls = 0.2 * rs.randn(D), 0.3 * rs.randn(D) - 4.0 / D
    return rs, z0, params, mu, ls


@pytest.mark.parametrize('D,K,N,chunk', [(32, 5, 1500, None), (64, 7, 2000, None), (100, 16, 1100, None),
                                         (256, 64, 1200, None), (64, 32, 3000, 512), (128, 9, 2500, 1024), (512, 33, 1050, None), (32, 1, 700, None), (64, 2, 513, None)])
def test_gram_free_energy_matches_oracle(D, K, N, chunk, monkeypatch):
    """float32, D % 4 == 0, K <= 64 can take the tensor-core path (forced here); ragged N, K that is not a multiple
    of 4, and (through the chunk knob) several chunks with a short last one.  Same oracle and tolerance as the
    direct wide-D kernel above; the direct kernel is run on the same inputs as a cross-check."""
    from nfcore.objective import SharedFlowObjective
    from oracle import planar_torch as pt
    monkeypatch.setenv('NF_GRAM', '1')          # small shapes would otherwise be routed to the direct kernel
    if chunk:
        monkeypatch.setenv('NF_GRAM_CHUNK', str(chunk))
    rs, z0, params, mu, ls = _gram_case(D, K, N, 11 * D + K)
    tgt = lambda z: torch.exp(pt.log_mvn(z, mu, ls))
    F_ref, g_ref = pt.script_value_and_grad(params, z0, tgt, 1e-7, K)
    tol = 1e-4
    obj = SharedFlowObjective(('gaussdiag', mu, ls), K, D, dtype='float32')
    z0_dev = torch.tensor(z0, device=DEV)       # device samples go through nf_planar_elbo_fwd_bwd_f32
    F, grad = obj.value_and_grad(params, z0_dev)
    assert abs(F - F_ref) <= tol * max(1.0, abs(F_ref)), (F, F_ref)
    assert _rel(grad, g_ref) <= tol, _rel(grad, g_ref)
    assert abs(obj.value(params, z0_dev) - F_ref) <= tol * max(1.0, abs(F_ref))
    monkeypatch.delenv('NF_GRAM')
    monkeypatch.setenv('NF_NO_GRAM', '1')
    F_d, grad_d = obj.value_and_grad(params, z0_dev)
    assert abs(F - F_d) <= tol * max(1.0, abs(F_ref))
    assert _rel(grad, grad_d) <= tol


def test_gram_mixture_target_matches_oracle(monkeypatch):
    from nfcore.objective import SharedFlowObjective
    from oracle import planar_torch as pt
    monkeypatch.setenv('NF_GRAM', '1')
    D, K, N, G = 32, 6, 1300, 3
    rs, z0, params, _, _ = _gram_case(D, K, N, 5)
    pi = np.array([0.3, 0.45]); mmu = 0.3 * rs.randn(G, D); var = 0.8 + 0.4 * rs.rand(G, D)
    sig = torch.tensor(np.array([np.diag(v) for v in var]))
    tgt = lambda z: pt.prob_gm(z, torch.tensor(mmu), sig, torch.tensor(pi))
    F_ref, g_ref = pt.script_value_and_grad(params, z0, tgt, 1e-7, K)
    obj = SharedFlowObjective(('mixdiag', pi, mmu, var), K, D, dtype='float32')
    F, grad = obj.value_and_grad(params, torch.tensor(z0, device=DEV))
    assert abs(F - F_ref) <= 1e-4 * max(1.0, abs(F_ref)), (F, F_ref)
    assert _rel(grad, g_ref) <= 1e-4, _rel(grad, g_ref)


@pytest.mark.parametrize('D,N', [(64, 1 << 19), (256, 1 << 17)])
def test_gram_full_size_agrees_with_direct_kernel(D, N, monkeypatch):
    """BASELINE config 5 sizes (K = 64): the two formulations of the same sums, float32, on the same samples."""
    from nfcore import ops
    K = 64
    torch.manual_seed(D)
    eps = torch.randn(N, D, device=DEV)
    sc = 0.5 / (D / 2) ** 0.5
    w, s = sc * torch.randn(K, D, device=DEV), sc * torch.randn(K, D, device=DEV)
    b, c = 0.3 * torch.randn(K, device=DEV), 0.3 * torch.randn(K, device=DEV)
    tp = torch.cat([0.2 * torch.randn(D, device=DEV), 0.3 * torch.randn(D, device=DEV) - 4.0 / D])
    got = [t.double().cpu().numpy() for t in ops.planar_elbo(eps, w, s, b, c, 'gaussdiag', tp)]
    monkeypatch.setenv('NF_NO_GRAM', '1')
    ref = [t.double().cpu().numpy() for t in ops.planar_elbo(eps, w, s, b, c, 'gaussdiag', tp)]
    assert np.all(np.isfinite(got[0]))
    assert np.abs(got[0] - ref[0]).max() <= 1e-4 * np.abs(ref[0]).max()
    for a, r in zip(got[1:], ref[1:]):
        assert _rel(a, r) <= 1e-4


def test_device_training_loop_wide_matrix_product_vs_direct(monkeypatch):
    """nf_script_train_f32 at D=64: thirty Adam steps with the per-iteration objective evaluated by the
    matrix-product path and by the direct kernel, same draws; and the first gradient step against the oracle."""
    from nfcore.objective import SharedFlowObjective
    from oracle import planar_torch as pt
    D, K, N, T = 64, 8, 1536, 30
    rs, _, params, mu, ls = _gram_case(D, K, N, 3)
    eps_all = rs.randn(T, N, D)
    obj = SharedFlowObjective(('gaussdiag', mu, ls), K, D, dtype='float32')
    monkeypatch.setenv('NF_GRAM', '1')
    x_g, tr_g = obj.train(params, eps_all, optimizer='adam', step_size=1e-3)
    monkeypatch.delenv('NF_GRAM')
    monkeypatch.setenv('NF_NO_GRAM', '1')
    x_d, tr_d = obj.train(params, eps_all, optimizer='adam', step_size=1e-3)
    assert np.abs(x_g - params).max() > 1e-3
    assert np.abs(x_g - x_d).max() <= 1e-4 * max(1.0, np.abs(x_d).max())
    assert np.abs(tr_g - tr_d).max() <= 1e-4 * max(1.0, np.abs(tr_d).max())
    tgt = lambda z: torch.exp(pt.log_mvn(z, mu, ls))
    F0, _ = pt.script_value_and_grad(params, eps_all[0], tgt, 1e-7, K)
    assert abs(tr_g[0] - F0) <= 1e-4 * max(1.0, abs(F0))


@pytest.mark.parametrize('D,K,N,chunk', [(64, 7, 3000, None), (128, 16, 5200, 512)])
def test_gram_host_samples_session_matches_oracle(D, K, N, chunk, monkeypatch):
    """NumPy samples go through the host-buffer session (nf_session_elbo_host_f32), which for wide D feeds the
    matrix-product path piece by piece while the next piece uploads; with the chunk knob: eleven pieces, more than
    the session has events.  Against the oracle, and against the same session forced onto the direct kernel."""
    from nfcore.objective import SharedFlowObjective
    from oracle import planar_torch as pt
    monkeypatch.setenv('NF_GRAM', '1')
    if chunk:
        monkeypatch.setenv('NF_GRAM_CHUNK', str(chunk))
    rs, z0, params, mu, ls = _gram_case(D, K, N, 17 * D + K)
    tgt = lambda z: torch.exp(pt.log_mvn(z, mu, ls))
    F_ref, g_ref = pt.script_value_and_grad(params, z0, tgt, 1e-7, K)
    obj = SharedFlowObjective(('gaussdiag', mu, ls), K, D, dtype='float32')
    F, grad = obj.value_and_grad(params, z0)
    assert abs(F - F_ref) <= 1e-4 * max(1.0, abs(F_ref)), (F, F_ref)
    assert _rel(grad, g_ref) <= 1e-4, _rel(grad, g_ref)
    assert abs(obj.value(params, z0) - F_ref) <= 1e-4 * max(1.0, abs(F_ref))
    monkeypatch.delenv('NF_GRAM')
    monkeypatch.setenv('NF_NO_GRAM', '1')
    F_d, grad_d = obj.value_and_grad(params, z0)
    assert abs(F - F_d) <= 1e-4 * max(1.0, abs(F_ref))
    assert _rel(grad, grad_d) <= 1e-4


# ---- BASELINE config 5 shapes against the oracle (K = 64, D in {2, 16, 64, 256}), every path that serves them ----
@pytest.mark.parametrize('D,N,paths', [(2, 1 << 15, ('default',)), (2, 1 << 18, ('default',)), (16, 1 << 15, ('default',)),
                                       (64, 1 << 15, ('gram', 'direct')), (256, 1 << 14, ('gram', 'direct'))])
def test_c5_shapes_match_oracle(D, N, paths, monkeypatch):
    """SURVEY.md section 8d C5 inputs: W, U = (0.1 + 0.01 N(0,1)) / sqrt(D/2), b = 0.1; target U1 for D = 2, a diagonal
    Gaussian through log_mvn math for D > 2.  D = 2 runs the scalar kernel (2^15 samples) and the packed-pair kernel
    (2^18); D = 16 the thread-per-sample kernel; D = 64 and 256 the matrix-product path and the direct wide kernel.
    float32 against the torch float64 oracle, 1e-4."""
    from nfcore.objective import SharedFlowObjective
    from oracle import planar_torch as pt
    K = 64
    rs = np.random.RandomState(500 + D)
    sc = 1.0 / np.sqrt(D / 2)
    params = np.concatenate([sc * (0.1 + 0.01 * rs.randn(2 * K * D)), 0.1 * np.ones(K)])
    z0 = rs.randn(N, D)
    if D == 2:
        target, fn = 'p1', pt.p1
    else:
        mu, ls = rs.uniform(-1, 1, D), rs.uniform(-1, 1, D)
        target, fn = ('gaussdiag', mu, ls), (lambda z: torch.exp(pt.log_mvn(z, mu, ls)))
    F_ref, g_ref = pt.script_value_and_grad(params, z0, fn, 1e-7, K)
    z0_dev = torch.tensor(z0, device=DEV)
    for path in paths:
        monkeypatch.delenv('NF_GRAM', raising=False)
        monkeypatch.delenv('NF_NO_GRAM', raising=False)
        if path == 'gram':
            monkeypatch.setenv('NF_GRAM', '1')
        elif path == 'direct':
            monkeypatch.setenv('NF_NO_GRAM', '1')
        obj = SharedFlowObjective(target, K, D, dtype='float32')
        F, grad = obj.value_and_grad(params, z0_dev)
        eF, eg = abs(F - F_ref) / max(1.0, abs(F_ref)), _rel(grad, g_ref)
        print('C5 D=%d N=%d %s: F err %.2e, grad err %.2e' % (D, N, path, eF, eg))
        assert eF <= 1e-4, (path, eF)
        assert eg <= 1e-4, (path, eg)


def test_sample_from_pz_matches_oracle_on_pinned_draws(monkeypatch):
    """distributions.sample_from_pz (distributions.py:92-102): randn(N, D) * sd + mu from the GLOBAL NumPy stream, then
    K planar_flow steps with per-sample parameters.  Same seed -> same draws as the reference makes; the flow steps
    against the oracle's restatement of flows.py:8-34."""
    from normflows import distributions
    from oracle import planar_np as pn
    rs = np.random.RandomState(8)
    N, D, K = 257, 5, 3
    mu, lsd = rs.randn(N, D), 0.3 * rs.randn(N, D)
    W, U, B = rs.randn(K, N, D), rs.randn(K, N, D), rs.randn(K, N)
    np.random.seed(123)
    got = distributions.sample_from_pz(mu, lsd, W, U, B, K)
    assert np.random.randn() == np.random.RandomState(123).randn(N * D + 1)[-1]      # exactly N*D draws were consumed
    np.random.seed(123)
    z = np.random.randn(N, D) * np.sqrt(1e-7 + np.exp(lsd)) + mu         # distributions.py:94-98
    for k in range(K):
        z = pn.planar_flow(z, W[k], U[k], B[k])                          # :100-101
    assert got.shape == (N, D) and got.dtype == np.float64
    assert _rel(got, z) <= 1e-10
    assert _rel(distributions.make_samples_z(mu, lsd, W, U, B, 0), mu) < 10      # K = 0: no flow step, still samples
    got32 = distributions.sample_from_pz(mu.astype(np.float32), lsd.astype(np.float32), W.astype(np.float32), U.astype(np.float32),
                                         B.astype(np.float32), K)
    assert got32.dtype == np.float32
