// Evaluation surface of the shared-parameter path: forward-only kernels behind
// visualise_flow (Final (master).ipynb cell 16) -- the target density on a grid of points, and samples pushed through
// the learned flow together with their log q_K -- and behind the density helpers of distributions.py for (N, D) inputs.
//   nf_target_density_*      out[n] = target(z[n])                      get_data.py:20-66, distributions.py:10-69
//   nf_planar_chain_forward_* z_K[n] and log q_K[n] of get_data.py:100-111 for given samples z0
// One thread per sample for D <= 4 (coordinates in registers); one warp per sample for larger D (lanes over the
// coordinates, up to 512).  HBM-streaming: D floats in, D+1 out per sample.
#include "nf_common.cuh"
#include "nf_launch.cuh"
#include "nf_targets.cuh"

namespace nf {

constexpr int kEvalThreads = 256;
constexpr int kEvalMaxPerLane = 16;       // 32 lanes x 16 coordinates = 512

template <typename T, int D, int TGT>
__global__ void __launch_bounds__(kEvalThreads)
density_small_kernel(const T* __restrict__ z, long long N, const T* __restrict__ raw, int ntp, T* __restrict__ out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* tp = reinterpret_cast<T*>(smem_raw);
    asm volatile("griddepcontrol.wait;" ::: "memory");
    target_prepare<T>(TGT, D, raw, ntp, tp);
    __syncthreads();
    for (long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x; n < N; n += (long long)gridDim.x * blockDim.x) {
        T zz[D], lp, g[D];
#pragma unroll
        for (int d = 0; d < D; ++d) zz[d] = z[n * D + d];
        // log(0 + target) evaluated by the same code the objective uses; the density is its exponential
        Target<T, D, TGT, false>::eval(zz, tp, ntp, (T)0, lp, g);
        out[n] = Mth<T>::exp(lp);
    }
}

// diagonal mixtures / Gaussians in any dimension: lane l holds coordinates l, l+32, ...
template <typename T, int TGT>
__global__ void __launch_bounds__(kEvalThreads)
density_wide_kernel(const T* __restrict__ z, long long N, int D, const T* __restrict__ raw, int ntp, T* __restrict__ out) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* tp = reinterpret_cast<T*>(smem_raw);
    asm volatile("griddepcontrol.wait;" ::: "memory");
    target_prepare<T>(TGT, D, raw, ntp, tp);
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long n = warp; n < N; n += nwarps) {
        T dens = (T)0;
        if (TGT == NF_TARGET_GAUSSDIAG) {
            T q = (T)0;
            for (int d = lane; d < D; d += 32) { const T df = z[n * D + d] - tp[1 + d]; q += df * df * tp[1 + D + d]; }
            q = warp_sum(q);
            dens = Mth<T>::exp(tp[0] - (T)0.5 * q);
        } else {
            const int G = ntp / (1 + 2 * D);
            for (int g = 0; g < G; ++g) {
                T q = (T)0;
                for (int d = lane; d < D; d += 32) { const T df = z[n * D + d] - tp[G + g * D + d]; q += df * df * tp[G + G * D + g * D + d]; }
                q = warp_sum(q);
                dens += tp[g] * Mth<T>::exp((T)-0.5 * q);
            }
        }
        if (lane == 0) out[n] = dens;
    }
}

template <typename T, int D>
__global__ void __launch_bounds__(kEvalThreads)
chain_small_kernel(const T* __restrict__ z0, long long N, int K, const T* __restrict__ w, const T* __restrict__ s,
                   const T* __restrict__ b, const T* __restrict__ c, T epsv, T* __restrict__ zK, T* __restrict__ logq) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* lc = reinterpret_cast<T*>(smem_raw);          // [K][2D+2]: w | s | b | c
    asm volatile("griddepcontrol.wait;" ::: "memory");
    constexpr int R = 2 * D + 2;
    for (int i = threadIdx.x; i < K; i += blockDim.x) {
        for (int d = 0; d < D; ++d) { lc[i * R + d] = w[i * D + d]; lc[i * R + D + d] = s[i * D + d]; }
        lc[i * R + 2 * D] = b[i];
        lc[i * R + 2 * D + 1] = c[i];
    }
    __syncthreads();
    const T LOG2PI = (T)1.8378770664093453;
    for (long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x; n < N; n += (long long)gridDim.x * blockDim.x) {
        T zz[D], q = (T)0;
#pragma unroll
        for (int d = 0; d < D; ++d) { zz[d] = z0[n * D + d]; q += zz[d] * zz[d]; }
        T lq = (T)-0.5 * (D * LOG2PI + q);                                 // get_data.py:111
        for (int k = 0; k < K; ++k) {
            T a = lc[k * R + 2 * D];
#pragma unroll
            for (int d = 0; d < D; ++d) a += lc[k * R + d] * zz[d];
            const T t = Mth<T>::tanh(a);
            lq -= Mth<T>::logp(epsv + Mth<T>::fabs((T)1 + ((T)1 - t * t) * lc[k * R + 2 * D + 1]));     // :107-108
#pragma unroll
            for (int d = 0; d < D; ++d) zz[d] += lc[k * R + D + d] * t;     // :109
        }
#pragma unroll
        for (int d = 0; d < D; ++d) zK[n * D + d] = zz[d];
        if (logq) logq[n] = lq;
    }
}

template <typename T>
__global__ void __launch_bounds__(kEvalThreads)
chain_wide_kernel(const T* __restrict__ z0, long long N, int D, int K, const T* __restrict__ w, const T* __restrict__ s,
                  const T* __restrict__ b, const T* __restrict__ c, T epsv, T* __restrict__ zK, T* __restrict__ logq) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    const T LOG2PI = (T)1.8378770664093453;
    for (long long n = warp; n < N; n += nwarps) {
        T zz[kEvalMaxPerLane], q = (T)0;
#pragma unroll
        for (int i = 0; i < kEvalMaxPerLane; ++i) {
            const int d = lane + 32 * i;
            zz[i] = d < D ? z0[n * D + d] : (T)0;
            q += zz[i] * zz[i];
        }
        T lq = (T)-0.5 * (D * LOG2PI + warp_sum(q));
        for (int k = 0; k < K; ++k) {
            T a = (T)0;
#pragma unroll
            for (int i = 0; i < kEvalMaxPerLane; ++i) { const int d = lane + 32 * i; if (d < D) a += __ldg(w + (size_t)k * D + d) * zz[i]; }
            a = warp_sum(a) + b[k];
            const T t = Mth<T>::tanh(a);
            lq -= Mth<T>::logp(epsv + Mth<T>::fabs((T)1 + ((T)1 - t * t) * c[k]));
#pragma unroll
            for (int i = 0; i < kEvalMaxPerLane; ++i) { const int d = lane + 32 * i; if (d < D) zz[i] += __ldg(s + (size_t)k * D + d) * t; }
        }
#pragma unroll
        for (int i = 0; i < kEvalMaxPerLane; ++i) { const int d = lane + 32 * i; if (d < D) zK[n * D + d] = zz[i]; }
        if (logq && lane == 0) logq[n] = lq;
    }
}

static inline unsigned eval_grid(long long work_items, int sms) {
    long long blocks = (work_items + kEvalThreads - 1) / kEvalThreads;
    const long long cap = (long long)(sms > 0 ? sms : 148) * 8;
    if (blocks > cap) blocks = cap;
    return (unsigned)(blocks < 1 ? 1 : blocks);
}

template <typename T, int D>
static int density_small(const T* z, long long N, int tgt, const T* raw, int ntp, T* out, cudaStream_t st, int sms) {
    const size_t smem = (size_t)target_smem_count(tgt, D, ntp) * sizeof(T);
    if (smem > 48 * 1024) return NF_ERR_UNSUPPORTED;
    const dim3 grid(eval_grid(N, sms)), block(kEvalThreads);
#define NF_DS(TG) case TG: NF_CUDA_OK(launch_pdl(density_small_kernel<T, D, TG>, grid, block, smem, st, z, N, raw, ntp, out)); return NF_OK
    switch (tgt) {
        NF_DS(NF_TARGET_MIXDIAG);
        NF_DS(NF_TARGET_GAUSSDIAG);
        NF_DS(NF_TARGET_MIXFULL);
        default: break;
    }
    if constexpr (D == 2) {
        switch (tgt) {
            NF_DS(NF_TARGET_U1);
            NF_DS(NF_TARGET_U2);
            NF_DS(NF_TARGET_U3);
            NF_DS(NF_TARGET_U4);
            NF_DS(NF_TARGET_TRIAL1);
            default: break;
        }
    }
#undef NF_DS
    return NF_ERR_UNSUPPORTED;
}

template <typename T>
static int target_check(int D, int tgt, const T* raw, int ntp) {
    if (tgt == NF_TARGET_MIXDIAG) return (raw && ntp >= 1 + 2 * D && ntp % (1 + 2 * D) == 0) ? NF_OK : NF_ERR_INVALID;
    if (tgt == NF_TARGET_GAUSSDIAG) return (raw && ntp == 2 * D) ? NF_OK : NF_ERR_INVALID;
    if (tgt == NF_TARGET_MIXFULL) {
        if (D > 4) return NF_ERR_UNSUPPORTED;
        return (raw && ntp >= 1 + D + D * D && ntp % (1 + D + D * D) == 0) ? NF_OK : NF_ERR_INVALID;
    }
    if (tgt < 0 || tgt > NF_TARGET_MIXFULL) return NF_ERR_INVALID;
    return D == 2 ? NF_OK : NF_ERR_INVALID;
}

template <typename T>
static int target_density(const T* z, int64_t N, int D, int tgt, const T* raw, int ntp, T* out, void* stream_) {
    cudaStream_t st = (cudaStream_t)stream_;
    if (N < 0 || D < 1 || (N > 0 && (!z || !out))) return NF_ERR_INVALID;
    int rc = target_check<T>(D, tgt, raw, ntp);
    if (rc != NF_OK) return rc;
    const DevInfo di = dev_info();
    if (!di.ok) return NF_ERR_NO_DEVICE;
    if (N == 0) return NF_OK;
    switch (D) {
        case 1: return density_small<T, 1>(z, N, tgt, raw, ntp, out, st, di.sms);
        case 2: return density_small<T, 2>(z, N, tgt, raw, ntp, out, st, di.sms);
        case 3: return density_small<T, 3>(z, N, tgt, raw, ntp, out, st, di.sms);
        case 4: return density_small<T, 4>(z, N, tgt, raw, ntp, out, st, di.sms);
        default: break;
    }
    if (D > 32 * kEvalMaxPerLane) return NF_ERR_UNSUPPORTED;
    const size_t smem = (size_t)target_smem_count(tgt, D, ntp) * sizeof(T);
    if (smem > 48 * 1024) return NF_ERR_UNSUPPORTED;
    const dim3 grid(eval_grid(N * 32, di.sms)), block(kEvalThreads);
    if (tgt == NF_TARGET_GAUSSDIAG) NF_CUDA_OK(launch_pdl(density_wide_kernel<T, NF_TARGET_GAUSSDIAG>, grid, block, smem, st, z, (long long)N, D, raw, ntp, out));
    else NF_CUDA_OK(launch_pdl(density_wide_kernel<T, NF_TARGET_MIXDIAG>, grid, block, smem, st, z, (long long)N, D, raw, ntp, out));
    return NF_OK;
}

template <typename T>
static int chain_forward(const T* z0, int64_t N, int D, int K, const T* w, const T* s, const T* b, const T* c, T epsv, T* zK, T* logq,
                         void* stream_) {
    cudaStream_t st = (cudaStream_t)stream_;
    if (N < 0 || D < 1 || K < 0 || (N > 0 && (!z0 || !zK)) || (K > 0 && (!w || !s || !b || !c))) return NF_ERR_INVALID;
    if (D > 32 * kEvalMaxPerLane) return NF_ERR_UNSUPPORTED;
    const DevInfo di = dev_info();
    if (!di.ok) return NF_ERR_NO_DEVICE;
    if (N == 0) return NF_OK;
    const size_t smem = (size_t)K * (2 * D + 2) * sizeof(T);
    const dim3 block(kEvalThreads);
    if (D <= 4 && smem <= 48 * 1024) {
        const dim3 grid(eval_grid(N, di.sms));
        switch (D) {
            case 1: NF_CUDA_OK(launch_pdl(chain_small_kernel<T, 1>, grid, block, smem, st, z0, (long long)N, K, w, s, b, c, epsv, zK, logq)); break;
            case 2: NF_CUDA_OK(launch_pdl(chain_small_kernel<T, 2>, grid, block, smem, st, z0, (long long)N, K, w, s, b, c, epsv, zK, logq)); break;
            case 3: NF_CUDA_OK(launch_pdl(chain_small_kernel<T, 3>, grid, block, smem, st, z0, (long long)N, K, w, s, b, c, epsv, zK, logq)); break;
            default: NF_CUDA_OK(launch_pdl(chain_small_kernel<T, 4>, grid, block, smem, st, z0, (long long)N, K, w, s, b, c, epsv, zK, logq)); break;
        }
        return NF_OK;
    }
    NF_CUDA_OK(launch_pdl(chain_wide_kernel<T>, dim3(eval_grid(N * 32, di.sms)), block, 0, st, z0, (long long)N, D, K, w, s, b, c, epsv, zK, logq));
    return NF_OK;
}

}  // namespace nf

extern "C" {
int nf_target_density_f32(const float* z, int64_t N, int D, int target_id, const float* target_params, int n_target_params, float* out, void* stream) {
    return nf::target_density<float>(z, N, D, target_id, target_params, n_target_params, out, stream);
}
int nf_target_density_f64(const double* z, int64_t N, int D, int target_id, const double* target_params, int n_target_params, double* out, void* stream) {
    return nf::target_density<double>(z, N, D, target_id, target_params, n_target_params, out, stream);
}
int nf_planar_chain_forward_f32(const float* z0, int64_t N, int D, int K, const float* w, const float* s, const float* b, const float* c,
                                float log_eps, float* zK, float* log_qK, void* stream) {
    return nf::chain_forward<float>(z0, N, D, K, w, s, b, c, log_eps, zK, log_qK, stream);
}
int nf_planar_chain_forward_f64(const double* z0, int64_t N, int D, int K, const double* w, const double* s, const double* b, const double* c,
                                double log_eps, double* zK, double* log_qK, void* stream) {
    return nf::chain_forward<double>(z0, N, D, K, w, s, b, c, log_eps, zK, log_qK, stream);
}
}
