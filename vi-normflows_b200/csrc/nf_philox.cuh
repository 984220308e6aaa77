// Counter-based standard-normal samples for the free-energy kernels (float32).
//
// The reference draws z0 = np.random.multivariate_normal(0, I, N) inside every objective call
// (get_data.py:100-101); on the GPU the draw can be made where it is consumed, which removes the only stream
// the kernel reads.  Philox-4x32-10 (Salmon et al., SC'11; the Random123 constants) keyed by the caller's seed, with
// the counter (pair index, step): sample n of step t depends on (seed, t, n) only -- not on the grid, the tile size or
// the kernel that asks for it -- so the stream can be regenerated anywhere (tests/test_gpu_rng.py does it in NumPy).
//   one Philox call -> 4 x 32 bits -> 4 uniforms u = ((x >> 8) + 0.5) 2^-24 in (0, 1) -> two Box-Muller pairs
//   D = 2: call number q gives samples 2q and 2q+1: (z[2q][0], z[2q][1]) = BM(u0, u1), (z[2q+1][0], z[2q+1][1]) = BM(u2, u3)
//   D = 1: call number q gives samples 2q and 2q+1: (z[2q], z[2q+1]) = BM(u0, u1)
//   BM(a, b) = sqrt(-2 ln a) (cos 2 pi b, sin 2 pi b)
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

namespace nf {

struct RngDesc { unsigned long long seed, step; };

__host__ __device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1,
                                                      uint32_t (&out)[4]) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1, n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

__device__ __forceinline__ void box_muller(uint32_t xa, uint32_t xb, float& za, float& zb) {
    const float ua = ((float)(xa >> 8) + 0.5f) * 5.9604644775390625e-8f;       // 2^-24
    const float ub = ((float)(xb >> 8) + 0.5f) * 5.9604644775390625e-8f;
    float l;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l) : "f"(ua));
    const float r = sqrtf(-1.3862943611198906f * l);                          // sqrt(-2 ln u) = sqrt(-2 ln2 lg2 u)
    float sn, cs;
    __sincosf(6.283185307179586f * ub, &sn, &cs);
    za = r * cs;
    zb = r * sn;
}

// the normals of samples 2q and 2q+1 (D coordinates each) of step `rng.step`
template <int D>
__device__ __forceinline__ void rng_pair(const RngDesc& rng, unsigned long long q, float (&za)[D], float (&zb)[D]) {
    static_assert(D == 1 || D == 2, "in-kernel draws: D <= 2");
    uint32_t x[4];
    philox4x32_10((uint32_t)q, (uint32_t)(q >> 32), (uint32_t)rng.step, (uint32_t)(rng.step >> 32), (uint32_t)rng.seed,
                  (uint32_t)(rng.seed >> 32), x);
    if (D == 2) {
        box_muller(x[0], x[1], za[0], za[D - 1]);
        box_muller(x[2], x[3], zb[0], zb[D - 1]);
    } else {
        box_muller(x[0], x[1], za[0], zb[0]);
    }
}

}  // namespace nf
