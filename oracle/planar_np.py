"""NumPy float64 restatement of the reference forward math (TEST INFRASTRUCTURE).

Every function names the reference lines it restates (paths relative to the
reference checkout).  Quirks are reproduced on purpose -- see SURVEY.md §8-Q.
Nothing here is imported by the product path.
"""
import numpy as np

LOG2PI = np.log(2 * np.pi)


# ----------------------------------------------------------------------------
# helpers -- get_data.py:15-18, 68-70 ; flows.py:42-43 ; transformations.py:4-21
# ----------------------------------------------------------------------------
def sigmoid(x):                      # get_data.py:15, transformations.py:7-8
    return 1 / (1 + np.exp(-x))


def relu(x):                         # transformations.py:21
    return np.maximum(x, 0)


def m(x):                            # get_data.py:68, flows.py:42-43
    return -1 + np.log(1 + np.exp(x))


def h_prime(x):                      # get_data.py:70, optimization.py:61-62
    return 1 - np.tanh(x) ** 2


def w1(x):                           # get_data.py:16
    return np.sin(2 * np.pi * x / 4)


def w2(x):                           # get_data.py:17
    return 3 * np.exp(-0.5 * ((x - 1) / 0.6) ** 2)


def w3(x):                           # get_data.py:18 (no 4th power, unlike theano_implement.py:72)
    return 3 * sigmoid((x - 1) / 0.3)


# ----------------------------------------------------------------------------
# targets: un-normalised densities exp(-U(z)) -- get_data.py:20-66
# ----------------------------------------------------------------------------
def trial1(z):                       # get_data.py:20-26
    z1 = z[:, 0]
    norm = np.sqrt(z[:, 0] ** 2 + z[:, 1] ** 2)
    exp1 = np.exp(-0.5 * ((z1 - 2) / 0.8) ** 2)
    exp2 = np.exp(-0.5 * ((z1 + 2) / 0.8) ** 2)
    u = 0.5 * ((norm - 4) / 0.4) ** 2 - np.log(exp1 + exp2)
    return np.exp(-u)


def p1(z):                           # get_data.py:28-35
    z1 = z[:, 0]
    first = (np.linalg.norm(z, 2, 1) - 2) / 0.4
    exp1 = np.exp(-0.5 * ((z1 - 2) / 0.6) ** 2)
    exp2 = np.exp(-0.5 * ((z1 + 2) / 0.6) ** 2)
    u = 0.5 * first ** 2 - np.log(exp1 + exp2)
    return np.exp(-u)


def p2(z):                           # get_data.py:37-43 (wall at |z1|>4, U=1e7)
    z1, z2 = z[:, 0], z[:, 1]
    u = 0.5 * ((z2 - w1(z1)) / 0.4) ** 2
    u = np.where(np.abs(z1) <= 4, u, np.ones(u.shape) * 1e7)
    return np.exp(-u)


def p3(z):                           # get_data.py:45-51
    z1, z2 = z[:, 0], z[:, 1]
    exp1 = np.exp(-0.5 * ((z2 - w1(z1)) / 0.35) ** 2)
    exp2 = np.exp(-0.5 * ((z2 - w1(z1) + w2(z1)) / 0.35) ** 2)
    u = -np.log(exp1 + exp2)
    return np.exp(-u)


def p4(z):                           # get_data.py:53-59
    z1, z2 = z[:, 0], z[:, 1]
    exp1 = np.exp(-0.5 * ((z2 - w1(z1)) / 0.4) ** 2)
    exp2 = np.exp(-0.5 * ((z2 - w1(z1) + w3(z1)) / 0.35) ** 2)
    u = -np.log(exp1 + exp2)
    return np.exp(-u)


def gmm1d(z, pi=(0.3, 0.7), mu=(-1.0, 3.0), sigma=(1.0, 1.0)):
    """get_data.py:61-66 (defaults) / experimentation.py:18-23 (pi=.5,.5 mu=-1,1 sigma=.5,.5).

    ``z`` is (N,1); the result keeps that shape, so ``log_q_K - log_p`` in the
    objective broadcasts (N,) against (N,1) -> (N,N) and ``np.mean`` of it equals
    mean(log_q_K) - mean(log_p).  The restated objective reproduces that value.
    """
    def gauss_uni(x, mu_, s_):
        return 1 / np.sqrt(2 * np.pi * s_ ** 2) * np.exp(-0.5 * ((x - mu_) / s_) ** 2)
    return pi[0] * gauss_uni(z, mu[0], sigma[0]) + pi[1] * gauss_uni(z, mu[1], sigma[1])


# ----------------------------------------------------------------------------
# distributions.py
# ----------------------------------------------------------------------------
def mvn(Z, mu, sigma_diag):          # distributions.py:10-32
    N, D = Z.shape
    mu = np.asarray(mu).reshape(-1, D)
    sigma_diag = np.asarray(sigma_diag)
    if len(sigma_diag) == 1:         # :20-22 scalar variance
        siginv = 1 / sigma_diag
        det_sigma = sigma_diag
    else:                            # :23-25 treated as a full matrix (Q9)
        siginv = np.linalg.inv(sigma_diag)
        det_sigma = np.linalg.det(sigma_diag)
    zdiff = Z - mu
    const = (2 * np.pi) ** (-D / 2) * det_sigma ** (-1 / 2)
    prod = np.dot(zdiff, siginv).reshape(N, D)
    return const * np.exp(-0.5 * np.sum(prod * zdiff, axis=1)).reshape(N, 1)


def log_mvn(Z, mu, log_sigma_diag):  # distributions.py:35-54
    N, D = Z.shape
    mu = np.asarray(mu).reshape(-1, D)
    log_sigma_diag = np.asarray(log_sigma_diag)
    logdet_sigma = np.sum(log_sigma_diag)
    if len(log_sigma_diag) == 1:     # :46-47 exp(1 - logsig) quirk (Q8)
        siginv = np.exp(1 - log_sigma_diag)
    else:
        siginv = np.linalg.inv(np.diag(np.exp(log_sigma_diag)))
    const = -D / 2 * np.log(2 * np.pi) - 0.5 * logdet_sigma
    zdiff = Z - mu
    return const - 0.5 * np.sum(np.matmul(zdiff, siginv) * zdiff, axis=1)


def log_std_norm(x):                 # distributions.py:57
    return -x.shape[1] / 2 * np.log(2 * np.pi) - 0.5 * np.sum(x * x, axis=1)


def prob_gm(Z, mu, sigma_diag, pi):  # distributions.py:60-69
    pi = np.asarray(pi)
    G = pi.shape[0] + 1
    prob = 0
    for g in range(G):
        pi_g = 1 - np.sum(pi) if g == G - 1 else pi[g]
        prob = prob + mvn(Z, mu[g], sigma_diag[g]) * pi_g
    return prob


def log_prob_gm(Z, mu, log_sigma_diag, logit_pi):   # distributions.py:72-83
    pi = sigmoid(np.asarray(logit_pi))
    return np.log(prob_gm(Z, mu, np.exp(log_sigma_diag), pi))


def log_bern_mult(X, p):             # distributions.py:86-89 (scalar over the whole batch, Q5)
    eps = 1e-7
    return np.sum(X * np.log(eps + p) + (1 - X) * np.log(eps + 1 - p))


# ----------------------------------------------------------------------------
# flows.py -- per-sample-parameter planar step ("package" semantics)
# ----------------------------------------------------------------------------
def get_uhat(u, w):                  # flows.py:37-39 (||w||^2)
    wu = np.sum(w * u, axis=1).reshape(-1, 1)
    return u + (m(wu) - wu) * (w / (np.linalg.norm(w, axis=1).reshape(-1, 1) ** 2))


def planar_flow(z, w, u, b):         # flows.py:8-34 (sum over d then broadcast, Q1)
    uhat = get_uhat(u, w)
    a = np.sum(z * w, axis=1) + b
    return z + np.sum(uhat * np.tanh(a).reshape(-1, 1), axis=1).reshape(z.shape[0], -1)


def planar_flow_shared(z, w, u, b):
    """The call the stale drivers make (learning_basic_flow.py:33,97;
    learning_simple_gaussian.py:75): w,u of shape (D,), scalar b.  HEAD
    flows.py:26 raises IndexError on these; the value the callers expect is the
    HEAD formula with the single parameter row broadcast over samples."""
    N = z.shape[0]
    w2_ = np.broadcast_to(np.asarray(w, dtype=float).reshape(1, -1), z.shape)
    u2_ = np.broadcast_to(np.asarray(u, dtype=float).reshape(1, -1), z.shape)
    return planar_flow(z, w2_, u2_, np.full(N, float(np.asarray(b).reshape(-1)[0])))


# ----------------------------------------------------------------------------
# shared-parameter free energy ("script" semantics) -- get_data.py:72-117
# ----------------------------------------------------------------------------
def unpack_script(params, K, D):     # get_data.py:74-78
    W = params[:K * D].reshape(K, D)
    U = params[K * D:2 * K * D].reshape(K, D)
    B = params[-K:]
    return W, U, B


def script_layer_constants(W, U):
    """Per-layer step vector s = u_hat (get_data.py:106, ||w|| un-squared) and
    log-det coefficient c = w.u with the RAW u (get_data.py:107-108)."""
    wu = np.sum(W * U, axis=1)
    S = (m(wu) - wu)[:, None] * (W / np.linalg.norm(W, axis=1)[:, None]) + U
    return S, wu


def script_terms(params, z0, target, eps, K):
    """get_data.py:80-95 ``other_metrics`` without the RNG draw: returns
    (log_q_K (N,), log_p (N,) or (N,1))."""
    N, D = z0.shape
    W, U, B = unpack_script(params, K, D)
    z_prev = z0
    sum_log_det_jacob = 0.
    for k in range(K):
        w, u, b = W[k], U[k], B[k]
        u_hat = (m(np.dot(w, u)) - np.dot(w, u)) * (w / np.linalg.norm(w)) + u      # :106
        affine = np.outer(h_prime(np.matmul(z_prev, w) + b), w)                     # :107
        sum_log_det_jacob = sum_log_det_jacob + np.log(eps + np.abs(1 + np.matmul(affine, u)))  # :108
        z_prev = z_prev + np.outer(np.tanh(np.matmul(z_prev, w) + b), u_hat)        # :109
    log_q_K = -0.5 * np.sum(np.log(2 * np.pi) + z0 ** 2, 1) - sum_log_det_jacob     # :111
    log_p = np.log(eps + target(z_prev))                                            # :112
    return log_q_K, log_p, z_prev


def script_objective(params, z0, target, eps, K):
    """get_data.py:97-113 with z0 supplied instead of drawn at :100-101."""
    log_q_K, log_p, _ = script_terms(params, z0, target, eps, K)
    return np.mean(log_q_K - log_p)                                                 # :113


# ----------------------------------------------------------------------------
# amortised free energy ("package" semantics) -- optimization.py:56-92
# ----------------------------------------------------------------------------
def amortised_flow_terms(z0, mu0, log_sigma_diag0, W, U, B, eps=1e-7):
    """optimization.py:74-85: reparameterise, K per-sample planar layers,
    running log-det with raw u and (sum w)(sum u) (Q2).  Returns (zK, running_sum (N,))."""
    K = W.shape[0]
    sd = np.sqrt(eps + np.exp(log_sigma_diag0))                                     # :74
    zk = z0 * sd + mu0                                                              # :75
    running_sum = 0.
    for k in range(K):
        w, u, b = W[k], U[k], B[k]
        ldj = np.sum(w * h_prime(np.sum(w * zk, axis=1) + b).reshape(-1, 1), axis=1)   # :64
        running_sum = running_sum + np.log(eps + np.abs(1. + np.sum(u * ldj.reshape(-1, 1), axis=1)))  # :83
        zk = planar_flow(zk, w, u, b)                                               # :84
    return zk, running_sum


def beta_schedule(t, max_iter):      # optimization.py:71-72
    cooling_max = np.min(np.array([max_iter / 4, 10000]))
    return np.min(np.array([1, 0.001 + t / cooling_max]))


def logq0_axis0(z, D):               # optimization.py:56-59 (sum over SAMPLES, Q4)
    return -D / 2 * np.log(2 * np.pi) - 0.5 * np.sum(z ** 2, axis=0)


def amortised_F(z0, enc_out, decode, theta, Xbatch, logp, t, max_iter, eps=1e-7):
    """optimization.py:66-92 with the batch and z0 supplied."""
    mu0, log_sigma_diag0, W, U, B = enc_out
    D = z0.shape[1]
    beta_t = beta_schedule(t, max_iter)
    first = np.mean(logq0_axis0(z0, D))                                             # :77
    zk, running_sum = amortised_flow_terms(z0, mu0, log_sigma_diag0, W, U, B, eps)
    third = np.mean(running_sum)                                                    # :85
    logits = decode(theta, zk)                                                      # :87
    second = np.mean(logp(Xbatch, zk, logits)) * beta_t                             # :89
    return first - second - third                                                   # :92


def mnist_logp(X, Z, logits):        # src/learning_mnist.py:89-99
    return log_bern_mult(X, logits) + log_std_norm(Z)


# ----------------------------------------------------------------------------
# nn_models.Feedforward.forward -- nn_models.py:41-84 (flat weights, feature-major)
# ----------------------------------------------------------------------------
def mlp_param_count(D_in, H, L, D_out):      # nn_models.py:16-19
    return (D_in * H + H) + (D_out * H + D_out) + (L - 1) * (H * H + H)


def mlp_forward(weights, x, D_in, H, L, D_out, act=relu, out_act=None):
    """weights (1,P) flat; x (D_in, N) -> (1, D_out, N)."""
    weights = np.asarray(weights).reshape(1, -1).T
    x = x.reshape((1, D_in, -1))
    W = weights[:H * D_in].T.reshape((-1, H, D_in))                                 # :59
    b = weights[H * D_in:H * D_in + H].T.reshape((-1, H, 1))                        # :60
    inp = act(np.matmul(W, x) + b)                                                  # :61
    index = H * D_in + H
    for _ in range(L - 1):                                                          # :67-76
        W = weights[index:index + H * H].T.reshape((-1, H, H)); index += H * H
        b = weights[index:index + H].T.reshape((-1, H, 1)); index += H
        inp = act(np.matmul(W, inp) + b)
    W = weights[index:index + H * D_out].T.reshape((-1, D_out, H))                  # :79
    b = weights[index + H * D_out:].T.reshape((-1, D_out, 1))                       # :80
    out = np.matmul(W, inp) + b
    return out if out_act is None else out_act(out)


def mnist_encode(phi, X, enc_dims, D, K):
    """src/learning_mnist.py:60-69 -- reshape-not-transpose scramble kept (Q7)."""
    D_in, H, L, D_out = enc_dims
    N = X.shape[0]
    out = mlp_forward(phi.reshape(1, -1), X.T, D_in, H, L, D_out)[0]
    mu0 = out[:D].reshape(N, D)
    ls0 = out[D:2 * D].reshape(N, D)
    W = out[2 * D:2 * D + K * D].reshape(K, N, D)
    U = out[2 * D + K * D:2 * D + 2 * K * D].reshape(K, N, D)
    b = out[-K:].reshape(K, N)
    return mu0, ls0, W, U, b


def mnist_decode(theta, Z, dec_dims):        # src/learning_mnist.py:71-73
    D_in, H, L, D_out = dec_dims
    return mlp_forward(theta.reshape(1, -1), Z.T, D_in, H, L, D_out, out_act=sigmoid)[0].T


TARGETS = {'p1': p1, 'p2': p2, 'p3': p3, 'p4': p4, 'trial1': trial1}
