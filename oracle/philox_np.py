"""Host restatement of the library's in-kernel sample stream (csrc/nf_philox.cuh) -- TEST INFRASTRUCTURE ONLY.

Philox-4x32-10 (Salmon, Moraes, Dror, Shaw, "Parallel random numbers: as easy as 1, 2, 3", SC'11; Random123 constants),
key = the 64-bit seed, counter = (pair index q, step); 24-bit uniforms; Box-Muller in float64.  The reference has no
counterpart: it draws z0 with NumPy's Mersenne Twister inside the objective (get_data.py:100-101).  This file lets a
test feed the oracle the very samples the kernel drew.
"""
import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = 0x9E3779B9, 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised over the counters (uint64 arrays holding 32-bit values); returns four uint64 arrays of 32-bit values."""
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint64) & MASK for c in (c0, c1, c2, c3))
    k0, k1 = int(k0) & 0xFFFFFFFF, int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0, p1 = M0 * c0, M1 * c2                      # 32 x 32 -> 64 bits, exact in uint64
        n0 = (p1 >> np.uint64(32)) ^ c1 ^ np.uint64(k0)
        n1 = p1 & MASK
        n2 = (p0 >> np.uint64(32)) ^ c3 ^ np.uint64(k1)
        n3 = p0 & MASK
        c0, c1, c2, c3 = n0, n1, n2, n3
        k0, k1 = (k0 + W0) & 0xFFFFFFFF, (k1 + W1) & 0xFFFFFFFF
    return c0, c1, c2, c3


def _box_muller(xa, xb):
    ua = ((xa >> np.uint64(8)).astype(np.float64) + 0.5) * 2.0 ** -24
    ub = ((xb >> np.uint64(8)).astype(np.float64) + 0.5) * 2.0 ** -24
    r = np.sqrt(-2.0 * np.log(ua))
    return r * np.cos(2 * np.pi * ub), r * np.sin(2 * np.pi * ub)


def normal(seed, step, N, D):
    """The [N][D] float64 samples of step `step` (D = 1 or 2), as nf_philox_normal_f32 lays them out."""
    assert D in (1, 2)
    npairs = (N + 1) // 2
    q = np.arange(npairs, dtype=np.uint64)
    x = philox4x32_10(q & MASK, q >> np.uint64(32), np.uint64(step) & MASK, np.uint64(step) >> np.uint64(32),
                      seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    out = np.empty((2 * npairs, D))
    if D == 2:
        out[0::2, 0], out[0::2, 1] = _box_muller(x[0], x[1])
        out[1::2, 0], out[1::2, 1] = _box_muller(x[2], x[3])
    else:
        out[0::2, 0], out[1::2, 0] = _box_muller(x[0], x[1])
    return out[:N]
