"""torch float64 restatement of ``planar_np`` (TEST INFRASTRUCTURE).

Same op sequence as the reference, written with torch so that
``torch.autograd`` plays the role HIPS autograd 1.3 plays in the reference
(get_data.py:115, optimization.py:35).  Forward values are checked against
``planar_np`` / the reference to ~1e-15; gradients against central differences
of the reference forward (tests/test_oracle_*.py).  It is also the CPU baseline
``bench.py`` times (all host cores), never a product fallback.
"""
import math
import torch

DT = torch.float64
LOG2PI = math.log(2 * math.pi)


def _t(x, dtype=DT):
    return x if isinstance(x, torch.Tensor) else torch.as_tensor(x, dtype=dtype)


def sigmoid(x):
    return 1 / (1 + torch.exp(-x))


def m(x):                            # flows.py:42-43 / get_data.py:68
    return -1 + torch.log(1 + torch.exp(x))


def h_prime(x):
    return 1 - torch.tanh(x) ** 2


def w1(x):
    return torch.sin(2 * math.pi * x / 4)


def w2(x):
    return 3 * torch.exp(-0.5 * ((x - 1) / 0.6) ** 2)


def w3(x):
    return 3 * sigmoid((x - 1) / 0.3)


def trial1(z):                       # get_data.py:20-26
    z1 = z[:, 0]
    norm = torch.sqrt(z[:, 0] ** 2 + z[:, 1] ** 2)
    exp1 = torch.exp(-0.5 * ((z1 - 2) / 0.8) ** 2)
    exp2 = torch.exp(-0.5 * ((z1 + 2) / 0.8) ** 2)
    u = 0.5 * ((norm - 4) / 0.4) ** 2 - torch.log(exp1 + exp2)
    return torch.exp(-u)


def p1(z):                           # get_data.py:28-35
    z1 = z[:, 0]
    first = (torch.linalg.norm(z, 2, 1) - 2) / 0.4
    exp1 = torch.exp(-0.5 * ((z1 - 2) / 0.6) ** 2)
    exp2 = torch.exp(-0.5 * ((z1 + 2) / 0.6) ** 2)
    u = 0.5 * first ** 2 - torch.log(exp1 + exp2)
    return torch.exp(-u)


def p2(z):                           # get_data.py:37-43
    z1, z2 = z[:, 0], z[:, 1]
    u = 0.5 * ((z2 - w1(z1)) / 0.4) ** 2
    u = torch.where(torch.abs(z1) <= 4, u, torch.ones_like(u) * 1e7)
    return torch.exp(-u)


def p3(z):                           # get_data.py:45-51
    z1, z2 = z[:, 0], z[:, 1]
    exp1 = torch.exp(-0.5 * ((z2 - w1(z1)) / 0.35) ** 2)
    exp2 = torch.exp(-0.5 * ((z2 - w1(z1) + w2(z1)) / 0.35) ** 2)
    return torch.exp(torch.log(exp1 + exp2))


def p4(z):                           # get_data.py:53-59
    z1, z2 = z[:, 0], z[:, 1]
    exp1 = torch.exp(-0.5 * ((z2 - w1(z1)) / 0.4) ** 2)
    exp2 = torch.exp(-0.5 * ((z2 - w1(z1) + w3(z1)) / 0.35) ** 2)
    return torch.exp(torch.log(exp1 + exp2))


def make_gmm1d(pi=(0.3, 0.7), mu=(-1.0, 3.0), sigma=(1.0, 1.0)):   # get_data.py:61-66
    def gmm(z):
        def g(x, mu_, s_):
            return 1 / math.sqrt(2 * math.pi * s_ ** 2) * torch.exp(-0.5 * ((x - mu_) / s_) ** 2)
        return pi[0] * g(z, mu[0], sigma[0]) + pi[1] * g(z, mu[1], sigma[1])
    return gmm


def mvn(Z, mu, sigma_diag):          # distributions.py:10-32
    N, D = Z.shape
    mu = _t(mu, Z.dtype).reshape(-1, D)
    sigma_diag = _t(sigma_diag, Z.dtype)
    if sigma_diag.shape[0] == 1:
        siginv = (1 / sigma_diag).reshape(1, 1)
        det_sigma = sigma_diag.reshape(())
    else:
        siginv = torch.linalg.inv(sigma_diag)
        det_sigma = torch.linalg.det(sigma_diag)
    zdiff = Z - mu
    const = (2 * math.pi) ** (-D / 2) * det_sigma ** (-1 / 2)
    prod = (zdiff @ siginv).reshape(N, D)
    return const * torch.exp(-0.5 * torch.sum(prod * zdiff, dim=1)).reshape(N, 1)


def log_mvn(Z, mu, log_sigma_diag):  # distributions.py:35-54
    N, D = Z.shape
    mu = _t(mu, Z.dtype).reshape(-1, D)
    log_sigma_diag = _t(log_sigma_diag, Z.dtype)
    logdet_sigma = torch.sum(log_sigma_diag)
    if log_sigma_diag.shape[0] == 1:
        siginv = torch.exp(1 - log_sigma_diag).reshape(1, 1)
    else:
        siginv = torch.linalg.inv(torch.diag(torch.exp(log_sigma_diag)))
    const = -D / 2 * LOG2PI - 0.5 * logdet_sigma
    zdiff = Z - mu
    return const - 0.5 * torch.sum((zdiff @ siginv) * zdiff, dim=1)


def log_std_norm(x):                 # distributions.py:57
    return -x.shape[1] / 2 * LOG2PI - 0.5 * torch.sum(x * x, dim=1)


def prob_gm(Z, mu, sigma_diag, pi):  # distributions.py:60-69
    pi = _t(pi, Z.dtype)
    G = pi.shape[0] + 1
    prob = 0
    for g in range(G):
        pi_g = 1 - torch.sum(pi) if g == G - 1 else pi[g]
        prob = prob + mvn(Z, mu[g], sigma_diag[g]) * pi_g
    return prob


def log_bern_mult(X, p):             # distributions.py:86-89
    eps = 1e-7
    return torch.sum(X * torch.log(eps + p) + (1 - X) * torch.log(eps + 1 - p))


def get_uhat(u, w):                  # flows.py:37-39
    wu = torch.sum(w * u, dim=1).reshape(-1, 1)
    return u + (m(wu) - wu) * (w / (torch.linalg.norm(w, dim=1).reshape(-1, 1) ** 2))


def planar_flow(z, w, u, b):         # flows.py:8-34
    uhat = get_uhat(u, w)
    a = torch.sum(z * w, dim=1) + b
    return z + torch.sum(uhat * torch.tanh(a).reshape(-1, 1), dim=1).reshape(z.shape[0], -1)


def unpack_script(params, K, D):     # get_data.py:74-78
    W = params[:K * D].reshape(K, D)
    U = params[K * D:2 * K * D].reshape(K, D)
    B = params[-K:]
    return W, U, B


def script_terms(params, z0, target, eps, K):      # get_data.py:80-95 / 97-113
    N, D = z0.shape
    W, U, B = unpack_script(params, K, D)
    z_prev = z0
    sum_log_det_jacob = 0.
    for k in range(K):
        w, u, b = W[k], U[k], B[k]
        wu = torch.dot(w, u)
        u_hat = (m(wu) - wu) * (w / torch.linalg.norm(w)) + u                       # :106
        affine = torch.outer(h_prime(z_prev @ w + b), w)                            # :107
        sum_log_det_jacob = sum_log_det_jacob + torch.log(eps + torch.abs(1 + affine @ u))  # :108
        z_prev = z_prev + torch.outer(torch.tanh(z_prev @ w + b), u_hat)            # :109
    log_q_K = -0.5 * torch.sum(LOG2PI + z0 ** 2, 1) - sum_log_det_jacob             # :111
    log_p = torch.log(eps + target(z_prev))                                         # :112
    return log_q_K, log_p, z_prev


def chain_terms(z0, w, s, b, c, target, eps):
    """The layer chain in the form the C ABI takes it (include/nfcore.h, nf_planar_elbo_fwd_bwd_*): per layer the
    direction w_k, the step vector s_k (= u_hat of get_data.py:106), the offset b_k and the log-det coefficient c_k
    (= w_k.u_k, :107-108) are given; returns the per-sample (log q_K, log p) of :111-112."""
    K = w.shape[0]
    z = z0
    sld = 0.
    for k in range(K):
        a = z @ w[k] + b[k]
        sld = sld + torch.log(eps + torch.abs(1 + h_prime(a) * c[k]))              # :107-108 with affine @ u = h'(a) (w.u)
        z = z + torch.outer(torch.tanh(a), s[k])                                   # :109
    log_q_K = -0.5 * torch.sum(LOG2PI + z0 ** 2, 1) - sld                          # :111
    lp = torch.log(eps + target(z))                                                # :112
    return log_q_K, lp.reshape(-1)


def script_min_det_factor(params, z0, K):
    """min over samples and layers of |1 + h'(w.z+b) w.u| (get_data.py:107-108): how close the parameters bring the
    chain to the log|.| singularity, where gradients grow like 1/|.| and float32 loses digits accordingly."""
    z0 = _t(z0)
    params = _t(params)
    N, D = z0.shape
    W, U, B = unpack_script(params, K, D)
    z, mn = z0, float('inf')
    for k in range(K):
        w, u, b = W[k], U[k], B[k]
        wu = torch.dot(w, u)
        a = z @ w + b
        mn = min(mn, float(torch.abs(1 + h_prime(a) * wu).min()))
        z = z + torch.outer(torch.tanh(a), (m(wu) - wu) * (w / torch.linalg.norm(w)) + u)
    return mn


def script_objective(params, z0, target, eps, K):  # get_data.py:97-113
    log_q_K, log_p, _ = script_terms(params, z0, target, eps, K)
    # (N,) - (N,1) broadcasts to (N,N) in the reference for D=1 targets; the mean is
    # identical to mean(log_q_K) - mean(log_p), which is what is formed here.
    return torch.mean(log_q_K) - torch.mean(log_p)


def script_value_and_grad(params, z0, target, eps, K):
    """(F, dF/dparams) in float64 -- stands in for autograd.grad (get_data.py:115)."""
    p = torch.tensor(params, dtype=DT, requires_grad=True)
    F = script_objective(p, _t(z0), target, eps, K)
    g, = torch.autograd.grad(F, p)
    return float(F.detach()), g.numpy()


def amortised_flow_terms(z0, mu0, log_sigma_diag0, W, U, B, eps=1e-7):   # optimization.py:74-85
    K = W.shape[0]
    sd = torch.sqrt(eps + torch.exp(log_sigma_diag0))
    zk = z0 * sd + mu0
    running_sum = 0.
    for k in range(K):
        w, u, b = W[k], U[k], B[k]
        ldj = torch.sum(w * h_prime(torch.sum(w * zk, dim=1) + b).reshape(-1, 1), dim=1)
        running_sum = running_sum + torch.log(eps + torch.abs(1. + torch.sum(u * ldj.reshape(-1, 1), dim=1)))
        zk = planar_flow(zk, w, u, b)
    return zk, running_sum


def amortised_min_det_factor(z0, mu0, log_sigma_diag0, W, U, B, eps=1e-7):
    """min over samples and layers of |1 + (sum u)(sum w) h'(w.z+b)| (optimization.py:64,83): distance of the batch from
    the log|.| singularity, where float32 gradients lose their digits."""
    K = W.shape[0]
    zk = z0 * torch.sqrt(eps + torch.exp(log_sigma_diag0)) + mu0
    mn = float('inf')
    for k in range(K):
        w, u, b = W[k], U[k], B[k]
        ldj = torch.sum(w * h_prime(torch.sum(w * zk, dim=1) + b).reshape(-1, 1), dim=1)
        mn = min(mn, float(torch.abs(1. + torch.sum(u * ldj.reshape(-1, 1), dim=1)).min()))
        zk = planar_flow(zk, w, u, b)
    return mn


def beta_schedule(t, max_iter):      # optimization.py:71-72
    return min(1.0, 0.001 + t / min(max_iter / 4, 10000))


def logq0_axis0(z, D):               # optimization.py:56-59
    return -D / 2 * LOG2PI - 0.5 * torch.sum(z ** 2, dim=0)


def mlp_forward(weights, x, D_in, H, L, D_out, out_sigmoid=False):      # nn_models.py:41-84 (ReLU hidden)
    w = weights.reshape(-1)
    x = x.reshape(D_in, -1)
    i = 0
    W = w[i:i + H * D_in].reshape(H, D_in); i += H * D_in
    b = w[i:i + H].reshape(H, 1); i += H
    a = torch.relu(W @ x + b)
    for _ in range(L - 1):
        W = w[i:i + H * H].reshape(H, H); i += H * H
        b = w[i:i + H].reshape(H, 1); i += H
        a = torch.relu(W @ a + b)
    W = w[i:i + H * D_out].reshape(D_out, H); i += H * D_out
    b = w[i:].reshape(D_out, 1)
    out = W @ a + b
    return sigmoid(out) if out_sigmoid else out


def mnist_encode(phi, X, enc_dims, D, K):        # src/learning_mnist.py:60-69
    D_in, H, L, D_out = enc_dims
    N = X.shape[0]
    out = mlp_forward(phi, X.T, D_in, H, L, D_out)
    mu0 = out[:D].reshape(N, D)
    ls0 = out[D:2 * D].reshape(N, D)
    W = out[2 * D:2 * D + K * D].reshape(K, N, D)
    U = out[2 * D + K * D:2 * D + 2 * K * D].reshape(K, N, D)
    b = out[-K:].reshape(K, N)
    return mu0, ls0, W, U, b


def mnist_decode(theta, Z, dec_dims):            # src/learning_mnist.py:71-73
    D_in, H, L, D_out = dec_dims
    return mlp_forward(theta, Z.T, D_in, H, L, D_out, out_sigmoid=True).T


def mnist_F(params, z0, Xbatch, t, max_iter, enc_dims, dec_dims, D, K, eps=1e-7):
    """optimization.py:66-92 with learning_mnist.py's encode/decode/logp."""
    n_phi = (enc_dims[0] * enc_dims[1] + enc_dims[1]) + (enc_dims[3] * enc_dims[1] + enc_dims[3]) \
        + (enc_dims[2] - 1) * (enc_dims[1] ** 2 + enc_dims[1])
    phi, theta = params[:n_phi], params[n_phi:]
    mu0, ls0, W, U, B = mnist_encode(phi, Xbatch, enc_dims, D, K)
    beta_t = beta_schedule(t, max_iter)
    first = torch.mean(logq0_axis0(z0, D))
    zk, running_sum = amortised_flow_terms(z0, mu0, ls0, W, U, B, eps)
    third = torch.mean(running_sum)
    logits = mnist_decode(theta, zk, dec_dims)
    second = torch.mean(log_bern_mult(Xbatch, logits) + log_std_norm(zk)) * beta_t
    return first - second - third


TARGETS = {'p1': p1, 'p2': p2, 'p3': p3, 'p4': p4, 'trial1': trial1}
