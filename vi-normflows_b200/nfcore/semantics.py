"""Host-side (K x D, float64) parameter algebra around the kernels.

The kernels see per-layer (w, s, b, c): direction, step vector, offset and
log-det coefficient.  Which (s, c) a raw (w, u) pair maps to is
the "semantics" of the caller:

  script   get_data.py:106-108   s = u + (m(w.u) - w.u) w/||w||     c = w.u   (raw u, norm not squared)
  paper    Rezende & Mohamed eq. 10-12 / app/js/flows.js:42-53
                                 s = u + (m(w.u) - w.u) w/||w||^2   c = w.s

`forward` maps raw parameters to kernel constants, `backward` pulls the kernel's
gradients (g_w direct, g_s, g_b, g_c) back to (dW, dU, dB).  K*D numbers: this
is bookkeeping, not the hot path.
"""
import numpy as np

SEMANTICS = ('script', 'paper')


def _m(x):          # get_data.py:68 / flows.py:42-43
    return -1 + np.log(1 + np.exp(x))


def _sigmoid(x):
    return 1 / (1 + np.exp(-x))


def forward(W, U, semantics='script'):
    W = np.asarray(W, dtype=np.float64)
    U = np.asarray(U, dtype=np.float64)
    wu = np.sum(W * U, axis=1)
    alpha = _m(wu) - wu
    n2 = np.sum(W * W, axis=1)
    if semantics == 'script':
        S = U + alpha[:, None] * W / np.sqrt(n2)[:, None]
        c = wu
    elif semantics == 'paper':
        S = U + alpha[:, None] * W / n2[:, None]
        c = np.sum(W * S, axis=1)
    else:
        raise ValueError('unknown semantics %r (expected one of %s)' % (semantics, SEMANTICS))
    return S, c


def backward(W, U, g_w, g_s, g_b, g_c, semantics='script'):
    """Chain rule from (g_w direct, g_s, g_b, g_c) to (dW, dU, dB)."""
    W = np.asarray(W, dtype=np.float64)
    U = np.asarray(U, dtype=np.float64)
    g_w = np.asarray(g_w, dtype=np.float64)
    g_s = np.asarray(g_s, dtype=np.float64)
    g_c = np.asarray(g_c, dtype=np.float64)
    wu = np.sum(W * U, axis=1)
    alpha = _m(wu) - wu
    dalpha = _sigmoid(wu) - 1.0
    n2 = np.sum(W * W, axis=1)
    if semantics == 'paper':
        # c = w.s : fold its gradient into g_s and the direct g_w first
        S, _ = forward(W, U, 'paper')
        g_w = g_w + g_c[:, None] * S
        g_s = g_s + g_c[:, None] * W
        g_c = np.zeros_like(g_c)
        p = 2.0      # s = u + alpha w / n^p, n = ||w||
    else:
        p = 1.0
    n = np.sqrt(n2)
    npow = n ** p
    gs_w = np.sum(g_s * W, axis=1)                     # g_s . w
    # s_j = u_j + alpha(wu) w_j / n^p
    dU = g_s + (dalpha * gs_w / npow)[:, None] * W + g_c[:, None] * W
    dW = (g_w + (dalpha * gs_w / npow)[:, None] * U
          + (alpha / npow)[:, None] * g_s
          - (alpha * p * gs_w / (npow * n2))[:, None] * W
          + g_c[:, None] * U)
    return dW, dU, np.asarray(g_b, dtype=np.float64)


def pack(W, U, B):
    return np.concatenate([np.ravel(W), np.ravel(U), np.ravel(B)])


def unpack(params, K, D):       # get_data.py:74-78
    params = np.asarray(params)
    return params[:K * D].reshape(K, D), params[K * D:2 * K * D].reshape(K, D), params[-K:]
