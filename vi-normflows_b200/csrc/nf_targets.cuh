// Target densities for the free energy: value log(eps + target(z)) and, when BWD,
// gz[d] = -d/dz_d log(eps + target(z))  (the seed of the reverse sweep).
// Reference: get_data.py:15-66 (potentials), distributions.py:10-69 (mvn / prob_gm / log_mvn).
#pragma once
#include "nf_common.cuh"

namespace nf {

// number of preprocessed target parameters held in shared memory
__host__ __device__ inline int target_smem_count(int tgt, int D, int ntp) {
    if (tgt == NF_TARGET_MIXDIAG) return ntp;          // coef(G) | mu(G*D) | inv_var(G*D)
    if (tgt == NF_TARGET_GAUSSDIAG) return 1 + 2 * D;  // const | mu(D) | inv_var(D)
    if (tgt == NF_TARGET_MIXFULL) return ntp;          // coef(G) | mu(G*D) | precision(G*D*D)
    return 0;
}

// One block-wide pass that turns the caller's raw target parameters into the
// form the per-sample code wants.  Call before __syncthreads().
template <typename T>
__device__ inline void target_prepare(int tgt, int D, const T* __restrict__ raw, int ntp, T* tp) {
    const double LOG2PI = 1.8378770664093453;
    if (tgt == NF_TARGET_MIXDIAG) {
        const int G = ntp / (1 + 2 * D);
        for (int i = threadIdx.x; i < G * D; i += blockDim.x) {
            tp[G + i] = raw[G + i];                                    // mu
            tp[G + G * D + i] = (T)(1.0 / (double)raw[G + G * D + i]);  // 1/var
        }
        for (int g = threadIdx.x; g < G; g += blockDim.x) {
            double logdet = 0.0;
            for (int d = 0; d < D; ++d) logdet += ::log((double)raw[G + G * D + g * D + d]);
            // pi_g (2 pi)^{-D/2} det^{-1/2}   -- distributions.py:29
            tp[g] = (T)((double)raw[g] * ::exp(-0.5 * D * LOG2PI - 0.5 * logdet));
        }
    } else if (tgt == NF_TARGET_MIXFULL) {
        // distributions.py:23-25: siginv = inv(Sigma), det(Sigma); one thread per component, D <= 4, float64
        const int G = ntp / (1 + D + D * D);
        for (int i = threadIdx.x; i < G * D; i += blockDim.x) tp[G + i] = raw[G + i];
        for (int g = threadIdx.x; g < G; g += blockDim.x) {
            double a[4][8];
            const T* S = raw + G + G * D + g * D * D;
            for (int i = 0; i < D && i < 4; ++i)
                for (int j = 0; j < D && j < 4; ++j) { a[i][j] = (double)S[i * D + j]; a[i][4 + j] = i == j ? 1.0 : 0.0; }
            double det = 1.0;
            for (int c0 = 0; c0 < D && c0 < 4; ++c0) {
                int piv = c0;
                for (int r = c0 + 1; r < D && r < 4; ++r) if (::fabs(a[r][c0]) > ::fabs(a[piv][c0])) piv = r;
                if (piv != c0) {
                    for (int j = 0; j < 8; ++j) { const double t = a[c0][j]; a[c0][j] = a[piv][j]; a[piv][j] = t; }
                    det = -det;
                }
                const double pv = a[c0][c0];
                det *= pv;
                const double ip = 1.0 / pv;
                for (int j = 0; j < 8; ++j) a[c0][j] *= ip;
                for (int r = 0; r < D && r < 4; ++r) {
                    if (r == c0) continue;
                    const double f = a[r][c0];
                    for (int j = 0; j < 8; ++j) a[r][j] -= f * a[c0][j];
                }
            }
            for (int i = 0; i < D && i < 4; ++i)
                for (int j = 0; j < D && j < 4; ++j) tp[G + G * D + g * D * D + i * D + j] = (T)a[i][4 + j];
            // pi_g (2 pi)^{-D/2} det^{-1/2}   -- distributions.py:29
            tp[g] = (T)((double)raw[g] * ::exp(-0.5 * D * LOG2PI) / ::sqrt(det));
        }
    } else if (tgt == NF_TARGET_GAUSSDIAG) {
        for (int d = threadIdx.x; d < D; d += blockDim.x) {
            tp[1 + d] = raw[d];
            // distributions.py:46-49: exp(1 - logsig) when len == 1, else inv(diag(exp(logsig)))
            tp[1 + D + d] = (T)(D == 1 ? ::exp(1.0 - (double)raw[D + d]) : ::exp(-(double)raw[D + d]));
        }
        if (threadIdx.x == 0) {
            double s = 0.0;
            for (int d = 0; d < D; ++d) s += (double)raw[D + d];
            tp[0] = (T)(-0.5 * D * LOG2PI - 0.5 * s);                  // distributions.py:50
        }
    }
}

// ring-shaped potentials p1 (get_data.py:28-35) and trial1 (:20-26)
template <typename T, bool BWD>
__device__ __forceinline__ void target_ring(const T* z, T R, T inv_sr, T inv_sx, T eps, T& lp, T* gz) {
    using M = Mth<T>;
    const T z1 = z[0], z2 = z[1];
    const T n = M::fsqrt(z1 * z1 + z2 * z2);
    const T f = (n - R) * inv_sr;
    const T q1 = (z1 - (T)2) * inv_sx, q2 = (z1 + (T)2) * inv_sx;
    const T a1 = (T)-0.5 * q1 * q1, a2 = (T)-0.5 * q2 * q2;
    // log(exp(a1) + exp(a2)) in max + log1p form (same value, no underflow to log(0))
    const T mx = M::fmax(a1, a2);
    const T ed = M::fexp(-M::fabs(a1 - a2));
    const T lse = mx + M::flog1p(ed);
    const T u = (T)0.5 * f * f - lse;
    const T dens = M::fexp(-u);
    const T den = eps + dens;
    lp = M::flog(den);
    if (BWD) {
        const T rho = M::div(dens, den);
        const T wbig = M::div((T)1, (T)1 + ed);
        const T pw1 = (a1 >= a2) ? wbig : (T)1 - wbig;       // e^{a1}/(e^{a1}+e^{a2})
        const T dlse = -(pw1 * q1 + ((T)1 - pw1) * q2) * inv_sx;
        const T fr = M::div(f * inv_sr, n);
        gz[0] = rho * (fr * z1 - dlse);
        gz[1] = rho * (fr * z2);
    }
}

template <typename T, int D, int TGT, bool BWD>
struct Target {
    static __device__ __forceinline__ void eval(const T* z, const T* tp, int ntp, T eps, T& lp, T* gz);
};

template <typename T, bool BWD> struct Target<T, 2, NF_TARGET_U1, BWD> {
    static __device__ __forceinline__ void eval(const T* z, const T*, int, T eps, T& lp, T* gz) {
        target_ring<T, BWD>(z, (T)2, (T)2.5, (T)(1.0 / 0.6), eps, lp, gz);
    }
};
template <typename T, bool BWD> struct Target<T, 2, NF_TARGET_TRIAL1, BWD> {
    static __device__ __forceinline__ void eval(const T* z, const T*, int, T eps, T& lp, T* gz) {
        target_ring<T, BWD>(z, (T)4, (T)2.5, (T)1.25, eps, lp, gz);
    }
};

// p2 -- get_data.py:37-43
template <typename T, bool BWD> struct Target<T, 2, NF_TARGET_U2, BWD> {
    static __device__ __forceinline__ void eval(const T* z, const T*, int, T eps, T& lp, T* gz) {
        using M = Mth<T>;
        const T HPI = (T)1.5707963267948966;
        const T z1 = z[0], z2 = z[1];
        const T w1 = M::sin(HPI * z1);
        const T q = (z2 - w1) * (T)2.5;
        const bool inside = M::fabs(z1) <= (T)4;
        const T dens = inside ? M::exp((T)-0.5 * q * q) : (T)0;   // exp(-1e7) == 0
        const T den = eps + dens;
        lp = M::logp(den);
        if (BWD) {
            const T rho = M::div(dens, den);
            const T du2 = q * (T)2.5;
            gz[0] = -rho * du2 * (M::cos(HPI * z1) * HPI);
            gz[1] = rho * du2;
        }
    }
};

// p3 / p4 -- get_data.py:45-59: two ridges, second one displaced by w2 (p3) or w3 (p4)
template <typename T, bool BWD, bool IS_P4>
__device__ __forceinline__ void target_ridges(const T* z, T eps, T& lp, T* gz) {
    using M = Mth<T>;
    const T HPI = (T)1.5707963267948966;
    const T z1 = z[0], z2 = z[1];
    const T w1 = M::sin(HPI * z1);
    T wb, dwb;
    if (IS_P4) {                                  // w3 = 3 sigmoid((z1-1)/0.3)      get_data.py:18
        const T sg = M::div((T)1, (T)1 + M::exp(-(z1 - (T)1) * (T)(1.0 / 0.3)));
        wb = (T)3 * sg;
        dwb = (T)3 * sg * ((T)1 - sg) * (T)(1.0 / 0.3);
    } else {                                      // w2 = 3 exp(-.5((z1-1)/0.6)^2)   get_data.py:17
        const T g = (z1 - (T)1) * (T)(1.0 / 0.6);
        wb = (T)3 * M::exp((T)-0.5 * g * g);
        dwb = -wb * g * (T)(1.0 / 0.6);
    }
    const T i1 = IS_P4 ? (T)2.5 : (T)(1.0 / 0.35);
    const T i2 = (T)(1.0 / 0.35);
    const T r1 = (z2 - w1) * i1;
    const T r2 = (z2 - w1 + wb) * i2;
    const T e1 = M::exp((T)-0.5 * r1 * r1);
    const T e2 = M::exp((T)-0.5 * r2 * r2);
    const T dens = e1 + e2;
    const T den = eps + dens;
    lp = M::logp(den);
    if (BWD) {
        const T dw1 = M::cos(HPI * z1) * HPI;
        const T k1 = e1 * r1 * i1, k2 = e2 * r2 * i2;       // -d dens/d z2 = k1 + k2
        const T inv = M::div((T)1, den);
        gz[1] = (k1 + k2) * inv;
        gz[0] = (k1 * (-dw1) + k2 * (dwb - dw1)) * inv;
    }
}
template <typename T, bool BWD> struct Target<T, 2, NF_TARGET_U3, BWD> {
    static __device__ __forceinline__ void eval(const T* z, const T*, int, T eps, T& lp, T* gz) {
        target_ridges<T, BWD, false>(z, eps, lp, gz);
    }
};
template <typename T, bool BWD> struct Target<T, 2, NF_TARGET_U4, BWD> {
    static __device__ __forceinline__ void eval(const T* z, const T*, int, T eps, T& lp, T* gz) {
        target_ridges<T, BWD, true>(z, eps, lp, gz);
    }
};

// mixture of diagonal Gaussians -- distributions.py:60-69 (prob_gm) over :10-32 (mvn)
template <typename T, int D, bool BWD> struct Target<T, D, NF_TARGET_MIXDIAG, BWD> {
    static __device__ __forceinline__ void eval(const T* z, const T* tp, int ntp, T eps, T& lp, T* gz) {
        using M = Mth<T>;
        const int G = ntp / (1 + 2 * D);
        T dens = (T)0;
        T dd[D];
#pragma unroll
        for (int d = 0; d < D; ++d) dd[d] = (T)0;
        for (int g = 0; g < G; ++g) {
            T q = (T)0;
            T sc[D];
#pragma unroll
            for (int d = 0; d < D; ++d) {
                const T df = z[d] - tp[G + g * D + d];
                sc[d] = df * tp[G + G * D + g * D + d];
                q = M::fma(df, sc[d], q);
            }
            const T e = tp[g] * M::fexp((T)-0.5 * q);
            dens += e;
            if (BWD) {
#pragma unroll
                for (int d = 0; d < D; ++d) dd[d] = M::fma(e, sc[d], dd[d]);   // -d dens/d z_d
            }
        }
        const T den = eps + dens;
        lp = M::flog(den);
        if (BWD) {
            const T inv = M::div((T)1, den);
#pragma unroll
            for (int d = 0; d < D; ++d) gz[d] = dd[d] * inv;
        }
    }
};

// mixture of Gaussians with full covariance matrices -- distributions.py:60-69 (prob_gm) over :10-32 (mvn, matrix branch)
template <typename T, int D, bool BWD> struct Target<T, D, NF_TARGET_MIXFULL, BWD> {
    static __device__ __forceinline__ void eval(const T* z, const T* tp, int ntp, T eps, T& lp, T* gz) {
        using M = Mth<T>;
        const int G = ntp / (1 + D + D * D);
        T dens = (T)0;
        T dd[D];
#pragma unroll
        for (int d = 0; d < D; ++d) dd[d] = (T)0;
        for (int g = 0; g < G; ++g) {
            T df[D], sc[D];
#pragma unroll
            for (int d = 0; d < D; ++d) df[d] = z[d] - tp[G + g * D + d];
            T q = (T)0;
            const T* P = tp + G + G * D + g * D * D;
#pragma unroll
            for (int i = 0; i < D; ++i) {
                T a = (T)0;
#pragma unroll
                for (int j = 0; j < D; ++j) a = M::fma(df[j], P[j * D + i], a);      // (zdiff @ siginv)_i, :30
                sc[i] = a;
                q = M::fma(a, df[i], q);
            }
            const T e = tp[g] * M::fexp((T)-0.5 * q);
            dens += e;
            if (BWD) {
                // d q / d z_i = sum_j (P_ij + P_ji) df_j; the reference's inverse of a symmetric matrix is symmetric up to
                // rounding, the gradient follows the formula as written
#pragma unroll
                for (int i = 0; i < D; ++i) {
                    T b = (T)0;
#pragma unroll
                    for (int j = 0; j < D; ++j) b = M::fma(P[i * D + j], df[j], b);
                    dd[i] = M::fma(e, (T)0.5 * (sc[i] + b), dd[i]);                    // -d dens/d z_i
                }
            }
        }
        const T den = eps + dens;
        lp = M::flog(den);
        if (BWD) {
            const T inv = M::div((T)1, den);
#pragma unroll
            for (int d = 0; d < D; ++d) gz[d] = dd[d] * inv;
        }
    }
};

// exp(log_mvn) -- distributions.py:35-54
template <typename T, int D, bool BWD> struct Target<T, D, NF_TARGET_GAUSSDIAG, BWD> {
    static __device__ __forceinline__ void eval(const T* z, const T* tp, int, T eps, T& lp, T* gz) {
        using M = Mth<T>;
        T q = (T)0;
        T sc[D];
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const T df = z[d] - tp[1 + d];
            sc[d] = df * tp[1 + D + d];
            q = M::fma(df, sc[d], q);
        }
        const T dens = M::fexp(tp[0] - (T)0.5 * q);
        const T den = eps + dens;
        lp = M::flog(den);
        if (BWD) {
            const T rho = M::div(dens, den);
#pragma unroll
            for (int d = 0; d < D; ++d) gz[d] = rho * sc[d];
        }
    }
};

}  // namespace nf
