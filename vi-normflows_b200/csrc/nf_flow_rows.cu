// Per-sample-parameter ("amortised") planar flows: flows.py:8-43 and the K-layer
// loop of optimization.py:74-85, forward and backward (sm_100a).
//
// Every sample row carries its own (w, u, b) for every layer, so there is no
// cross-sample reduction and the arithmetic intensity is a few flop per byte: these
// kernels are HBM/L2-streaming kernels.  One warp owns one row; lanes stride over the
// D coordinates, which makes every global access a contiguous, coalesced row segment;
// the five per-row dot products a layer needs are reduced with one halving butterfly.
// The backward pass walks z back with z_{k-1} = z_k - A_k t_k from the tanh values the
// forward pass parked in the workspace ([K][N]), so nothing is recomputed through tanh.
#include "nf_common.cuh"

namespace nf {

constexpr int kRowThreads = 256;
constexpr int kRowWarps = kRowThreads / kWarp;

template <typename T> __device__ __forceinline__ T softplus_m1(T x) {   // m(x) = -1 + log(1+e^x), flows.py:42-43
    using M = Mth<T>;
    // log(1+e^x) = max(x,0) + log1p(e^{-|x|}) : same value, no overflow for large x
    return (T)-1 + M::fmax(x, (T)0) + M::log1p(M::exp(-M::fabs(x)));
}

template <typename T, int DPL>
__device__ __forceinline__ void load_row(const T* __restrict__ p, int D, int lane, T (&r)[DPL]) {
#pragma unroll
    for (int i = 0; i < DPL; ++i) {
        const int d = lane + i * kWarp;
        r[i] = d < D ? p[d] : (T)0;
    }
}
template <typename T, int DPL>
__device__ __forceinline__ void store_row(T* __restrict__ p, int D, int lane, const T (&r)[DPL]) {
#pragma unroll
    for (int i = 0; i < DPL; ++i) {
        const int d = lane + i * kWarp;
        if (d < D) p[d] = r[i];
    }
}

// all-lanes totals of 5 per-lane partials (halving butterfly + broadcast)
template <typename T>
__device__ __forceinline__ void warp_sum5(T (&v)[5], int lane) {
    // plain xor butterflies keep the totals on every lane, which is what the row math needs
#pragma unroll
    for (int i = 0; i < 5; ++i) v[i] = warp_sum(v[i]);
}

// ---- flows.planar_flow (flows.py:8-34) -------------------------------------------------
// semantics 0 = package (sum over d of u_hat_d tanh(.) broadcast to every coordinate, Q1)
//           1 = paper   (z + u_hat tanh(.))
template <typename T, int DPL>
__global__ void __launch_bounds__(kRowThreads)
planar_flow_rows_kernel(const T* __restrict__ z, const T* __restrict__ w, const T* __restrict__ u,
                        const T* __restrict__ b, long long N, int D, int per_sample, int semantics,
                        T* __restrict__ out) {
    using M = Mth<T>;
    const int lane = threadIdx.x & 31;
    const long long warp0 = (long long)blockIdx.x * kRowWarps + (threadIdx.x >> 5);
    const long long nwarps = (long long)gridDim.x * kRowWarps;
    for (long long n = warp0; n < N; n += nwarps) {
        T zr[DPL], wr[DPL], ur[DPL];
        load_row<T, DPL>(z + n * D, D, lane, zr);
        load_row<T, DPL>(w + (per_sample ? n * D : 0), D, lane, wr);
        load_row<T, DPL>(u + (per_sample ? n * D : 0), D, lane, ur);
        T v[5] = {(T)0, (T)0, (T)0, (T)0, (T)0};     // w.u, w.w, w.z, sum u, sum w
#pragma unroll
        for (int i = 0; i < DPL; ++i) {
            v[0] = M::fma(wr[i], ur[i], v[0]);
            v[1] = M::fma(wr[i], wr[i], v[1]);
            v[2] = M::fma(wr[i], zr[i], v[2]);
            v[3] += ur[i];
            v[4] += wr[i];
        }
        warp_sum5(v, lane);
        const T alpha = softplus_m1(v[0]) - v[0];
        const T t = M::tanh(v[2] + b[per_sample ? n : 0]);
        T o[DPL];
        if (semantics == 0) {
            const T A = v[3] + alpha * v[4] / v[1];          // sum_d u_hat_d
#pragma unroll
            for (int i = 0; i < DPL; ++i) o[i] = M::fma(A, t, zr[i]);
        } else {
            const T q = alpha / v[1];
#pragma unroll
            for (int i = 0; i < DPL; ++i) o[i] = M::fma(M::fma(q, wr[i], ur[i]), t, zr[i]);
        }
        store_row<T, DPL>(out + n * D, D, lane, o);
    }
}

// ---- amortised K-layer forward (optimization.py:74-85) ----------------------------------
template <typename T, int DPL>
__global__ void __launch_bounds__(kRowThreads)
amortised_fwd_kernel(const T* __restrict__ z0, const T* __restrict__ mu0, const T* __restrict__ ls0,
                     const T* __restrict__ W, const T* __restrict__ U, const T* __restrict__ Bb,
                     long long N, int D, int K, T epsv,
                     T* __restrict__ zK, T* __restrict__ logdet /*[N]*/, T* __restrict__ tsave /*[K][N] or null*/) {
    using M = Mth<T>;
    const int lane = threadIdx.x & 31;
    const long long warp0 = (long long)blockIdx.x * kRowWarps + (threadIdx.x >> 5);
    const long long nwarps = (long long)gridDim.x * kRowWarps;
    for (long long n = warp0; n < N; n += nwarps) {
        T zr[DPL], a1[DPL], a2[DPL];
        load_row<T, DPL>(z0 + n * D, D, lane, zr);
        load_row<T, DPL>(mu0 + n * D, D, lane, a1);
        load_row<T, DPL>(ls0 + n * D, D, lane, a2);
#pragma unroll
        for (int i = 0; i < DPL; ++i) {
            const bool ok = lane + i * kWarp < D;
            const T sd = M::sqrt(epsv + M::exp(a2[i]));                  // optimization.py:74
            zr[i] = ok ? M::fma(zr[i], sd, a1[i]) : (T)0;                // :75
        }
        T rs = (T)0;
        for (int k = 0; k < K; ++k) {
            const size_t off = ((size_t)k * N + n) * D;
            load_row<T, DPL>(W + off, D, lane, a1);
            load_row<T, DPL>(U + off, D, lane, a2);
            T v[5] = {(T)0, (T)0, (T)0, (T)0, (T)0};
#pragma unroll
            for (int i = 0; i < DPL; ++i) {
                v[0] = M::fma(a1[i], a2[i], v[0]);
                v[1] = M::fma(a1[i], a1[i], v[1]);
                v[2] = M::fma(a1[i], zr[i], v[2]);
                v[3] += a2[i];
                v[4] += a1[i];
            }
            warp_sum5(v, lane);
            const T t = M::tanh(v[2] + Bb[(size_t)k * N + n]);
            const T hp = M::fma(-t, t, (T)1);
            const T x = M::fma(v[3] * v[4], hp, (T)1);                   // :64,83  1 + (sum u)(sum w) h'
            rs += M::logp(epsv + M::fabs(x));
            const T alpha = softplus_m1(v[0]) - v[0];
            const T A = v[3] + alpha * v[4] / v[1];                      // flows.py:32,39
#pragma unroll
            for (int i = 0; i < DPL; ++i) {
                const bool ok = lane + i * kWarp < D;
                zr[i] = ok ? M::fma(A, t, zr[i]) : (T)0;
            }
            if (tsave && lane == 0) tsave[(size_t)k * N + n] = t;
        }
        store_row<T, DPL>(zK + n * D, D, lane, zr);
        if (lane == 0) logdet[n] = rs;
    }
}

// ---- amortised K-layer backward ------------------------------------------------------------
// Given g_zK = dF/dz_K (from the decoder + prior) and the scalar g_ld = dF/d logdet[n]
// (-1/N for optimization.py:85,92), produces dF/d(mu0, ls0, W, U, Bb).
template <typename T, int DPL>
__global__ void __launch_bounds__(kRowThreads)
amortised_bwd_kernel(const T* __restrict__ z0, const T* __restrict__ ls0,
                     const T* __restrict__ W, const T* __restrict__ U,
                     const T* __restrict__ zK, const T* __restrict__ tsave, const T* __restrict__ g_zK, T g_ld_all,
                     const T* __restrict__ g_ld_vec /* dF/dlogdet[n] per sample, or NULL: g_ld_all for every n */,
                     long long N, int D, int K, T epsv,
                     T* __restrict__ g_mu0, T* __restrict__ g_ls0, T* __restrict__ g_W, T* __restrict__ g_U,
                     T* __restrict__ g_Bb) {
    using M = Mth<T>;
    const int lane = threadIdx.x & 31;
    const long long warp0 = (long long)blockIdx.x * kRowWarps + (threadIdx.x >> 5);
    const long long nwarps = (long long)gridDim.x * kRowWarps;
    for (long long n = warp0; n < N; n += nwarps) {
        T zr[DPL], gz[DPL], wr[DPL], ur[DPL];
        load_row<T, DPL>(zK + n * D, D, lane, zr);
        load_row<T, DPL>(g_zK + n * D, D, lane, gz);
        const T g_ld = g_ld_vec ? g_ld_vec[n] : g_ld_all;
        for (int k = K - 1; k >= 0; --k) {
            const size_t off = ((size_t)k * N + n) * D;
            load_row<T, DPL>(W + off, D, lane, wr);
            load_row<T, DPL>(U + off, D, lane, ur);
            const T t = tsave[(size_t)k * N + n];
            T v[5] = {(T)0, (T)0, (T)0, (T)0, (T)0};      // w.u, w.w, sum gz, sum u, sum w
#pragma unroll
            for (int i = 0; i < DPL; ++i) {
                v[0] = M::fma(wr[i], ur[i], v[0]);
                v[1] = M::fma(wr[i], wr[i], v[1]);
                v[2] += gz[i];
                v[3] += ur[i];
                v[4] += wr[i];
            }
            warp_sum5(v, lane);
            const T wu = v[0], ww = v[1], G = v[2], su = v[3], sw = v[4];
            const T alpha = softplus_m1(wu) - wu;
            const T dalpha = M::div((T)1, (T)1 + M::exp(-wu)) - (T)1;      // sigmoid(wu) - 1
            const T iww = M::div((T)1, ww);
            const T A = su + alpha * sw * iww;
            const T hp = M::fma(-t, t, (T)1);
            const T x = M::fma(su * sw, hp, (T)1);
            const T gx = g_ld * M::copysign(M::div((T)1, epsv + M::fabs(x)), x);
            const T gt = A * G - (T)2 * t * (gx * su * sw);
            const T ga = gt * hp;
            const T gA = t * G;
            const T g_su = gA + gx * sw * hp;
            const T g_sw = gA * alpha * iww + gx * su * hp;
            const T g_wu = gA * dalpha * sw * iww;
            const T g_ww = -gA * alpha * sw * iww * iww;
            T gw[DPL], gu[DPL];
#pragma unroll
            for (int i = 0; i < DPL; ++i) {
                const bool ok = lane + i * kWarp < D;
                zr[i] = ok ? M::fma(-A, t, zr[i]) : (T)0;                   // z_{k-1}
                gu[i] = M::fma(g_wu, wr[i], g_su);
                gw[i] = M::fma(ga, zr[i], M::fma(g_wu, ur[i], M::fma((T)2 * g_ww, wr[i], g_sw)));
                gz[i] = ok ? M::fma(ga, wr[i], gz[i]) : (T)0;
            }
            store_row<T, DPL>(g_W + off, D, lane, gw);
            store_row<T, DPL>(g_U + off, D, lane, gu);
            if (lane == 0) g_Bb[(size_t)k * N + n] = ga;
        }
        // z_0 = z0 * sqrt(eps + exp(ls0)) + mu0
        T e0[DPL], l0[DPL];
        load_row<T, DPL>(z0 + n * D, D, lane, e0);
        load_row<T, DPL>(ls0 + n * D, D, lane, l0);
        T gl[DPL];
#pragma unroll
        for (int i = 0; i < DPL; ++i) {
            const T ex = M::exp(l0[i]);
            const T sd = M::sqrt(epsv + ex);
            gl[i] = gz[i] * e0[i] * M::div(ex, (T)2 * sd);
        }
        store_row<T, DPL>(g_mu0 + n * D, D, lane, gz);
        store_row<T, DPL>(g_ls0 + n * D, D, lane, gl);
    }
}

static int rows_grid(long long N) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    long long blocks = (N + kRowWarps - 1) / kRowWarps;
    const long long cap = (long long)sms * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

#define NF_DPL_DISPATCH(D, CALL)                                      \
    if ((D) <= 32) { constexpr int DPL = 1; CALL; }                   \
    else if ((D) <= 64) { constexpr int DPL = 2; CALL; }              \
    else if ((D) <= 128) { constexpr int DPL = 4; CALL; }             \
    else if ((D) <= 256) { constexpr int DPL = 8; CALL; }             \
    else return NF_ERR_UNSUPPORTED;

template <typename T>
static int planar_flow_entry(const T* z, const T* w, const T* u, const T* b, int64_t N, int D, int per_sample,
                             int semantics, T* out, void* stream_) {
    if (N < 0 || D < 1 || (N > 0 && (!z || !out)) || !w || !u || !b) return NF_ERR_INVALID;
    if (semantics != 0 && semantics != 1) return NF_ERR_INVALID;
    if (N == 0) return NF_OK;
    cudaStream_t st = (cudaStream_t)stream_;
    NF_DPL_DISPATCH(D, (planar_flow_rows_kernel<T, DPL><<<rows_grid(N), kRowThreads, 0, st>>>(
                           z, w, u, b, N, D, per_sample, semantics, out)));
    NF_CUDA_OK(cudaGetLastError());
    return NF_OK;
}

template <typename T>
static int amortised_fwd_entry(const T* z0, const T* mu0, const T* ls0, const T* W, const T* U, const T* Bb,
                               int64_t N, int D, int K, T epsv, T* zK, T* logdet, T* tsave, void* stream_) {
    if (N < 0 || D < 1 || K < 0 || !zK || !logdet) return NF_ERR_INVALID;
    if (N > 0 && (!z0 || !mu0 || !ls0 || (K > 0 && (!W || !U || !Bb)))) return NF_ERR_INVALID;
    if (N == 0) return NF_OK;
    cudaStream_t st = (cudaStream_t)stream_;
    NF_DPL_DISPATCH(D, (amortised_fwd_kernel<T, DPL><<<rows_grid(N), kRowThreads, 0, st>>>(
                           z0, mu0, ls0, W, U, Bb, N, D, K, epsv, zK, logdet, tsave)));
    NF_CUDA_OK(cudaGetLastError());
    return NF_OK;
}

template <typename T>
static int amortised_bwd_entry(const T* z0, const T* ls0, const T* W, const T* U, const T* zK, const T* tsave,
                               const T* g_zK, T g_ld, const T* g_ld_vec, int64_t N, int D, int K, T epsv, T* g_mu0, T* g_ls0,
                               T* g_W, T* g_U, T* g_Bb, void* stream_) {
    if (N < 0 || D < 1 || K < 0) return NF_ERR_INVALID;
    if (N > 0 && (!z0 || !ls0 || !zK || !g_zK || !g_mu0 || !g_ls0)) return NF_ERR_INVALID;
    if (N > 0 && K > 0 && (!W || !U || !tsave || !g_W || !g_U || !g_Bb)) return NF_ERR_INVALID;
    if (N == 0) return NF_OK;
    cudaStream_t st = (cudaStream_t)stream_;
    NF_DPL_DISPATCH(D, (amortised_bwd_kernel<T, DPL><<<rows_grid(N), kRowThreads, 0, st>>>(
                           z0, ls0, W, U, zK, tsave, g_zK, g_ld, g_ld_vec, N, D, K, epsv, g_mu0, g_ls0, g_W, g_U, g_Bb)));
    NF_CUDA_OK(cudaGetLastError());
    return NF_OK;
}

}  // namespace nf

extern "C" {

int nf_planar_flow_f32(const float* z, const float* w, const float* u, const float* b, int64_t N, int D,
                       int per_sample, int semantics, float* out, void* stream) {
    return nf::planar_flow_entry<float>(z, w, u, b, N, D, per_sample, semantics, out, stream);
}
int nf_planar_flow_f64(const double* z, const double* w, const double* u, const double* b, int64_t N, int D,
                       int per_sample, int semantics, double* out, void* stream) {
    return nf::planar_flow_entry<double>(z, w, u, b, N, D, per_sample, semantics, out, stream);
}

int nf_planar_amortised_fwd_f32(const float* z0, const float* mu0, const float* log_sigma0, const float* W,
                                const float* U, const float* Bb, int64_t N, int D, int K, float log_eps,
                                float* zK, float* logdet, float* t_save, void* stream) {
    return nf::amortised_fwd_entry<float>(z0, mu0, log_sigma0, W, U, Bb, N, D, K, log_eps, zK, logdet, t_save, stream);
}
int nf_planar_amortised_fwd_f64(const double* z0, const double* mu0, const double* log_sigma0, const double* W,
                                const double* U, const double* Bb, int64_t N, int D, int K, double log_eps,
                                double* zK, double* logdet, double* t_save, void* stream) {
    return nf::amortised_fwd_entry<double>(z0, mu0, log_sigma0, W, U, Bb, N, D, K, log_eps, zK, logdet, t_save, stream);
}

int nf_planar_amortised_bwd_f32(const float* z0, const float* log_sigma0, const float* W, const float* U,
                                const float* zK, const float* t_save, const float* g_zK, float g_logdet,
                                int64_t N, int D, int K, float log_eps, float* g_mu0, float* g_log_sigma0,
                                float* g_W, float* g_U, float* g_Bb, void* stream) {
    return nf::amortised_bwd_entry<float>(z0, log_sigma0, W, U, zK, t_save, g_zK, g_logdet, nullptr, N, D, K, log_eps, g_mu0,
                                          g_log_sigma0, g_W, g_U, g_Bb, stream);
}
int nf_planar_amortised_bwd_v_f32(const float* z0, const float* log_sigma0, const float* W, const float* U,
                                  const float* zK, const float* t_save, const float* g_zK, const float* g_logdet_vec,
                                  int64_t N, int D, int K, float log_eps, float* g_mu0, float* g_log_sigma0,
                                  float* g_W, float* g_U, float* g_Bb, void* stream) {
    if (N > 0 && !g_logdet_vec) return NF_ERR_INVALID;
    return nf::amortised_bwd_entry<float>(z0, log_sigma0, W, U, zK, t_save, g_zK, 0.f, g_logdet_vec, N, D, K, log_eps, g_mu0,
                                          g_log_sigma0, g_W, g_U, g_Bb, stream);
}
int nf_planar_amortised_bwd_f64(const double* z0, const double* log_sigma0, const double* W, const double* U,
                                const double* zK, const double* t_save, const double* g_zK, double g_logdet,
                                int64_t N, int D, int K, double log_eps, double* g_mu0, double* g_log_sigma0,
                                double* g_W, double* g_U, double* g_Bb, void* stream) {
    return nf::amortised_bwd_entry<double>(z0, log_sigma0, W, U, zK, t_save, g_zK, g_logdet, nullptr, N, D, K, log_eps, g_mu0,
                                           g_log_sigma0, g_W, g_U, g_Bb, stream);
}
int nf_planar_amortised_bwd_v_f64(const double* z0, const double* log_sigma0, const double* W, const double* U,
                                  const double* zK, const double* t_save, const double* g_zK, const double* g_logdet_vec,
                                  int64_t N, int D, int K, double log_eps, double* g_mu0, double* g_log_sigma0,
                                  double* g_W, double* g_U, double* g_Bb, void* stream) {
    if (N > 0 && !g_logdet_vec) return NF_ERR_INVALID;
    return nf::amortised_bwd_entry<double>(z0, log_sigma0, W, U, zK, t_save, g_zK, 0.0, g_logdet_vec, N, D, K, log_eps, g_mu0,
                                           g_log_sigma0, g_W, g_U, g_Bb, stream);
}

}  // extern "C"
