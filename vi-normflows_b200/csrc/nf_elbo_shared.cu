// Shared-parameter planar-flow free energy: fused forward + backward (sm_100a).
//
// Replaces get_data.py:97-113 (forward) and the autograd tape replay of
// get_data.py:115 (backward).  One launch per step:
//   * every thread owns S samples; z stays in registers across all K layers,
//     tanh activations t_k are parked in shared memory for the reverse sweep
//     (z is walked back with z_{k-1} = z_k - s_k t_k, nothing is recomputed);
//   * per layer the 2D+2 gradient partials of the thread's samples are reduced
//     across the warp with a halving butterfly and accumulated in a per-warp
//     shared-memory table, so cross-sample reduction costs O(1) per layer per warp;
//   * blocks are persistent (grid = SMs x occupancy) and stride over sample tiles;
//     each block writes one row of partial sums, a second tiny kernel adds the
//     rows in a fixed order (bitwise reproducible for a fixed launch shape).
#include "nf_common.cuh"
#include "nf_targets.cuh"

namespace nf {

constexpr int kElboThreads = 128;
constexpr int kMaxBlocksPerSM = 8;

template <int D> __host__ __device__ constexpr int grad_vals() { return 2 * D + 2; }   // w[D] s[D] b c

// dynamic shared memory carve-up (in elements of T)
template <typename T>
struct ElboSmem {
    int lc, tp, acc, hist, total;
    __host__ __device__ ElboSmem(int K, int D, int S, int ntp_s, bool bwd, int nthreads) {
        const int V = 2 * D + 2;
        lc = 0;
        tp = lc + K * V;
        acc = tp + ntp_s;
        hist = acc + (bwd ? (nthreads / kWarp) * K * V : 0);
        total = hist + (bwd ? K * S * nthreads : 0);
    }
};

// A thread owns S consecutive samples = S*D consecutive elements of eps.  Full
// tiles are fetched with the widest aligned vector load that divides S*D elements
// (LDG.128 for float D=2 S=2: lanes are 16 B apart, perfectly coalesced); the
// ragged tail falls back to guarded scalar loads and a 0/1 mask.
template <typename T, int D, int S>
__device__ __forceinline__ void load_samples(const T* __restrict__ eps, long long n0, long long N,
                                             T (&z)[S][D], T (&msk)[S]) {
    constexpr int E = S * D;
    constexpr int BYTES = E * (int)sizeof(T);
    if (n0 + S <= N) {
        T buf[E];
        const T* src = eps + n0 * D;
        if (BYTES % 16 == 0) {
            const int4* p = reinterpret_cast<const int4*>(src);
            int4* q = reinterpret_cast<int4*>(buf);
#pragma unroll
            for (int i = 0; i < BYTES / 16; ++i) q[i] = __ldg(p + i);
        } else if (BYTES % 8 == 0) {
            const int2* p = reinterpret_cast<const int2*>(src);
            int2* q = reinterpret_cast<int2*>(buf);
#pragma unroll
            for (int i = 0; i < BYTES / 8; ++i) q[i] = __ldg(p + i);
        } else {
#pragma unroll
            for (int i = 0; i < E; ++i) buf[i] = __ldg(src + i);
        }
#pragma unroll
        for (int j = 0; j < S; ++j) {
            msk[j] = (T)1;
#pragma unroll
            for (int d = 0; d < D; ++d) z[j][d] = buf[j * D + d];
        }
    } else {
#pragma unroll
        for (int j = 0; j < S; ++j) {
            const bool ok = n0 + j < N;
            msk[j] = ok ? (T)1 : (T)0;
#pragma unroll
            for (int d = 0; d < D; ++d) z[j][d] = ok ? eps[(n0 + j) * D + d] : (T)0.5;
        }
    }
}

template <typename T, int D, int S, int TGT, bool BWD>
__global__ void __launch_bounds__(kElboThreads)
elbo_shared_kernel(const T* __restrict__ eps, long long N, int K,
                   const T* __restrict__ w, const T* __restrict__ s,
                   const T* __restrict__ b, const T* __restrict__ c,
                   const T* __restrict__ tparams, int ntp, T epsv,
                   T* __restrict__ partials /* [gridDim.x][K*V + 2] */) {
    using M = Mth<T>;
    constexpr int V = grad_vals<D>();
    constexpr int NT = kElboThreads;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* sm = reinterpret_cast<T*>(smem_raw);
    const int ntp_s = target_smem_count(TGT, D, ntp);
    const ElboSmem<T> L(K, D, S, ntp_s, BWD, NT);
    T* lc = sm + L.lc;        // [K][V] : w[D] s[D] b c
    T* tp = sm + L.tp;
    T* acc = sm + L.acc;      // [NW][K][V]
    T* hist = sm + L.hist;    // [K][S][NT]

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    for (int i = tid; i < K * D; i += NT) {
        const int k = i / D, d = i - k * D;
        lc[k * V + d] = w[i];
        lc[k * V + D + d] = s[i];
    }
    for (int k = tid; k < K; k += NT) {
        lc[k * V + 2 * D] = b[k];
        lc[k * V + 2 * D + 1] = c[k];
    }
    target_prepare<T>(TGT, D, tparams, ntp, tp);
    if (BWD)
        for (int i = tid; i < (NT / kWarp) * K * V; i += NT) acc[i] = (T)0;
    __syncthreads();

    const T LOG2PI = (T)1.8378770664093453;
    const int my_slot = warp_slot<V>(lane);
    const bool my_own = warp_slot_owner<V>(lane);
    T* acc_w = acc + warp * K * V;

    T sum_q = (T)0, sum_p = (T)0;
    const long long tile_samples = (long long)NT * S;
    const long long ntiles = (N + tile_samples - 1) / tile_samples;

    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const long long base = tile * tile_samples;
        T z[S][D], msk[S], ld[S], lq0[S];
        load_samples<T, D, S>(eps, base + (long long)tid * S, N, z, msk);
#pragma unroll
        for (int j = 0; j < S; ++j) {
            T zz = (T)0;
#pragma unroll
            for (int d = 0; d < D; ++d) zz = M::fma(z[j][d], z[j][d], zz);
            lq0[j] = (T)-0.5 * (D * LOG2PI + zz);          // get_data.py:111, first term
            ld[j] = (T)0;
        }
        // ---- forward over the K layers (get_data.py:104-109) ----
        for (int k = 0; k < K; ++k) {
            T wk[D], sk[D];
#pragma unroll
            for (int d = 0; d < D; ++d) { wk[d] = lc[k * V + d]; sk[d] = lc[k * V + D + d]; }
            const T bk = lc[k * V + 2 * D], ck = lc[k * V + 2 * D + 1];
#pragma unroll
            for (int j = 0; j < S; ++j) {
                T a = bk;
#pragma unroll
                for (int d = 0; d < D; ++d) a = M::fma(wk[d], z[j][d], a);
                const T t = M::tanh(a);
                const T hp = M::fma(-t, t, (T)1);
                const T x = M::fma(hp, ck, (T)1);
                ld[j] += M::log(epsv + M::fabs(x));
#pragma unroll
                for (int d = 0; d < D; ++d) z[j][d] = M::fma(sk[d], t, z[j][d]);
                if (BWD) hist[(k * S + j) * NT + tid] = t;
            }
        }
        // ---- target: log(eps + target(z_K)) and the reverse seed ----
        T gz[S][D];
#pragma unroll
        for (int j = 0; j < S; ++j) {
            T lp;
            Target<T, D, TGT, BWD>::eval(z[j], tp, ntp, epsv, lp, gz[j]);
            // select, not multiply: a masked tail sample must not leak NaN/Inf
            const bool ok = msk[j] != (T)0;
            sum_q += ok ? (lq0[j] - ld[j]) : (T)0;
            sum_p += ok ? lp : (T)0;
            if (BWD) {
#pragma unroll
                for (int d = 0; d < D; ++d) gz[j][d] = ok ? gz[j][d] : (T)0;
            }
        }
        if (BWD) {
            // ---- reverse sweep ----
            for (int k = K - 1; k >= 0; --k) {
                T wk[D], sk[D];
#pragma unroll
                for (int d = 0; d < D; ++d) { wk[d] = lc[k * V + d]; sk[d] = lc[k * V + D + d]; }
                const T ck = lc[k * V + 2 * D + 1];
                T v[V];
#pragma unroll
                for (int i = 0; i < V; ++i) v[i] = (T)0;
#pragma unroll
                for (int j = 0; j < S; ++j) {
                    const T t = hist[(k * S + j) * NT + tid];
#pragma unroll
                    for (int d = 0; d < D; ++d) z[j][d] = M::fma(-sk[d], t, z[j][d]);   // z_{k-1}
                    const T hp = M::fma(-t, t, (T)1);
                    const T x = M::fma(hp, ck, (T)1);
                    const T r = M::copysign(M::rcp(epsv + M::fabs(x)), x);
                    const T gx = -r * msk[j];                    // d(-ld)/dx
                    v[2 * D + 1] = M::fma(gx, hp, v[2 * D + 1]);  // g_c
                    T gt = (T)-2 * t * (gx * ck);
#pragma unroll
                    for (int d = 0; d < D; ++d) {
                        v[D + d] = M::fma(gz[j][d], t, v[D + d]);  // g_s
                        gt = M::fma(gz[j][d], sk[d], gt);
                    }
                    const T ga = gt * hp;
                    v[2 * D] += ga;                                // g_b
#pragma unroll
                    for (int d = 0; d < D; ++d) {
                        v[d] = M::fma(ga, z[j][d], v[d]);          // g_w (direct)
                        gz[j][d] = M::fma(ga, wk[d], gz[j][d]);
                    }
                }
                warp_reduce_multi<T, V>(v, lane);
                if (my_own) acc_w[k * V + my_slot] += v[0];
            }
        }
    }

    // ---- block partials ----
    sum_q = warp_sum(sum_q);
    sum_p = warp_sum(sum_p);
    __shared__ T red[2 * (NT / kWarp)];
    if (lane == 0) { red[warp] = sum_q; red[NT / kWarp + warp] = sum_p; }
    __syncthreads();
    const int row = K * V + 2;
    T* out = partials + (size_t)blockIdx.x * row;
    if (BWD) {
        for (int i = tid; i < K * V; i += NT) {
            T a = (T)0;
#pragma unroll
            for (int wi = 0; wi < NT / kWarp; ++wi) a += acc[wi * K * V + i];
            out[i] = a;
        }
    }
    if (tid == 0) {
        T q = (T)0, p = (T)0;
#pragma unroll
        for (int wi = 0; wi < NT / kWarp; ++wi) { q += red[wi]; p += red[NT / kWarp + wi]; }
        out[K * V] = q;
        out[K * V + 1] = p;
    }
}

// Adds the per-block rows in a fixed order: one warp per output value, lanes
// stride over blocks, then a shuffle tree.  Scatters into the caller's arrays.
template <typename T>
__global__ void elbo_finalize_kernel(const T* __restrict__ partials, int nblocks, int K, int D, int bwd,
                                     T* __restrict__ out_F, T* __restrict__ g_w, T* __restrict__ g_s,
                                     T* __restrict__ g_b, T* __restrict__ g_c) {
    const int V = 2 * D + 2;
    const int row = K * V + 2;
    const int lane = threadIdx.x & 31;
    const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nw = (gridDim.x * blockDim.x) >> 5;
    for (int i = gw; i < row; i += nw) {
        if (!bwd && i < K * V) continue;
        T a = (T)0;
        for (int bl = lane; bl < nblocks; bl += 32) a += partials[(size_t)bl * row + i];
        a = warp_sum(a);
        if (lane == 0) {
            if (i >= K * V) {
                out_F[1 + (i - K * V)] = a;
            } else {
                const int k = i / V, j = i - k * V;
                if (j < D) g_w[k * D + j] = a;
                else if (j < 2 * D) g_s[k * D + (j - D)] = a;
                else if (j == 2 * D) g_b[k] = a;
                else g_c[k] = a;
            }
        }
    }
    // F = sum log q_K - sum log p : recomputed by one warp from the same rows (same order => same bits)
    if (gw == 0) {
        T q = (T)0, p = (T)0;
        for (int bl = lane; bl < nblocks; bl += 32) {
            q += partials[(size_t)bl * row + K * V];
            p += partials[(size_t)bl * row + K * V + 1];
        }
        q = warp_sum(q);
        p = warp_sum(p);
        if (lane == 0) out_F[0] = q - p;
    }
}

struct DevInfo { int sms; int ok; };
static DevInfo dev_info() {
    static thread_local int cached_dev = -1;
    static thread_local DevInfo info = {0, 0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return {0, 0};
    if (dev != cached_dev) {
        cudaDeviceProp p;
        if (cudaGetDeviceProperties(&p, dev) != cudaSuccess) return {0, 0};
        info.sms = p.multiProcessorCount;
        info.ok = (p.major == 10);
        cached_dev = dev;
    }
    return info;
}

template <typename T, int D, int S, int TGT, bool BWD>
static int launch_elbo(const T* eps, long long N, int K, const T* w, const T* s, const T* b, const T* c,
                       const T* tparams, int ntp, T epsv, T* partials, int max_blocks, int* nblocks_out,
                       cudaStream_t stream) {
    auto kern = elbo_shared_kernel<T, D, S, TGT, BWD>;
    const int ntp_s = target_smem_count(TGT, D, ntp);
    const ElboSmem<T> L(K, D, S, ntp_s, BWD, kElboThreads);
    const size_t smem = (size_t)L.total * sizeof(T);
    if (smem > 227 * 1024) return NF_ERR_UNSUPPORTED;
    NF_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    NF_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kElboThreads, smem));
    if (per_sm < 1) return NF_ERR_UNSUPPORTED;
    if (per_sm > kMaxBlocksPerSM) per_sm = kMaxBlocksPerSM;
    const DevInfo di = dev_info();
    const long long tile = (long long)kElboThreads * S;
    long long ntiles = (N + tile - 1) / tile;
    long long grid = (long long)di.sms * per_sm;
    if (grid > ntiles) grid = ntiles;
    if (grid < 1) grid = 1;
    if (grid > max_blocks) grid = max_blocks;
    kern<<<(int)grid, kElboThreads, smem, stream>>>(eps, N, K, w, s, b, c, tparams, ntp, epsv, partials);
    NF_CUDA_OK(cudaGetLastError());
    *nblocks_out = (int)grid;
    return NF_OK;
}

template <typename T, int D, int TGT>
static int dispatch_bwd(bool bwd, const T* eps, long long N, int K, const T* w, const T* s, const T* b, const T* c,
                        const T* tp, int ntp, T epsv, T* partials, int max_blocks, int* nb, cudaStream_t st) {
    constexpr int S = 2;
    // large K: the activation history of two samples per thread no longer fits next to a second block
    if (bwd) {
        const ElboSmem<T> L2(K, D, S, target_smem_count(TGT, D, ntp), true, kElboThreads);
        if ((size_t)L2.total * sizeof(T) <= 113 * 1024)
            return launch_elbo<T, D, S, TGT, true>(eps, N, K, w, s, b, c, tp, ntp, epsv, partials, max_blocks, nb, st);
        return launch_elbo<T, D, 1, TGT, true>(eps, N, K, w, s, b, c, tp, ntp, epsv, partials, max_blocks, nb, st);
    }
    return launch_elbo<T, D, S, TGT, false>(eps, N, K, w, s, b, c, tp, ntp, epsv, partials, max_blocks, nb, st);
}

template <typename T, int D>
static int dispatch_target(int tgt, bool bwd, const T* eps, long long N, int K, const T* w, const T* s, const T* b,
                           const T* c, const T* tp, int ntp, T epsv, T* partials, int max_blocks, int* nb,
                           cudaStream_t st) {
#define NF_CASE(TG) case TG: return dispatch_bwd<T, D, TG>(bwd, eps, N, K, w, s, b, c, tp, ntp, epsv, partials, max_blocks, nb, st)
    switch (tgt) {
        NF_CASE(NF_TARGET_MIXDIAG);
        NF_CASE(NF_TARGET_GAUSSDIAG);
        default: break;
    }
    if (D == 2) {
        switch (tgt) {
            case NF_TARGET_U1: return dispatch_bwd<T, 2, NF_TARGET_U1>(bwd, eps, N, K, w, s, b, c, tp, ntp, epsv, partials, max_blocks, nb, st);
            case NF_TARGET_U2: return dispatch_bwd<T, 2, NF_TARGET_U2>(bwd, eps, N, K, w, s, b, c, tp, ntp, epsv, partials, max_blocks, nb, st);
            case NF_TARGET_U3: return dispatch_bwd<T, 2, NF_TARGET_U3>(bwd, eps, N, K, w, s, b, c, tp, ntp, epsv, partials, max_blocks, nb, st);
            case NF_TARGET_U4: return dispatch_bwd<T, 2, NF_TARGET_U4>(bwd, eps, N, K, w, s, b, c, tp, ntp, epsv, partials, max_blocks, nb, st);
            case NF_TARGET_TRIAL1: return dispatch_bwd<T, 2, NF_TARGET_TRIAL1>(bwd, eps, N, K, w, s, b, c, tp, ntp, epsv, partials, max_blocks, nb, st);
            default: break;
        }
    }
#undef NF_CASE
    return NF_ERR_UNSUPPORTED;
}

static int max_grid_blocks() {
    const DevInfo di = dev_info();
    return (di.sms > 0 ? di.sms : 148) * kMaxBlocksPerSM;
}

template <typename T>
static int elbo_entry(const T* eps, int64_t N, int D, int K, const T* w, const T* s, const T* b, const T* c,
                      int target_id, const T* tparams, int ntp, T epsv, T* out_F, T* g_w, T* g_s, T* g_b, T* g_c,
                      void* workspace, size_t workspace_bytes, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!eps && N > 0) return NF_ERR_INVALID;
    if ((reinterpret_cast<uintptr_t>(eps) & 15) != 0) return NF_ERR_INVALID;   // vector loads
    if (N < 0 || D < 1 || K < 1 || !w || !s || !b || !c || !out_F || !workspace) return NF_ERR_INVALID;
    const bool any_g = g_w || g_s || g_b || g_c;
    const bool all_g = g_w && g_s && g_b && g_c;
    if (any_g && !all_g) return NF_ERR_INVALID;
    if (target_id == NF_TARGET_MIXDIAG) {
        if (!tparams || ntp < 1 + 2 * D || ntp % (1 + 2 * D) != 0) return NF_ERR_INVALID;
    } else if (target_id == NF_TARGET_GAUSSDIAG) {
        if (!tparams || ntp != 2 * D) return NF_ERR_INVALID;
    } else if (target_id < 0 || target_id > NF_TARGET_GAUSSDIAG) {
        return NF_ERR_INVALID;
    } else if (D != 2) {
        return NF_ERR_INVALID;      // the 2-D potentials index z[:,0], z[:,1] (get_data.py:29)
    }
    const DevInfo di = dev_info();
    if (!di.ok) return NF_ERR_NO_DEVICE;
    if (workspace_bytes < nf_planar_elbo_workspace_bytes(K, D, (int)sizeof(T))) return NF_ERR_WORKSPACE;
    T* partials = reinterpret_cast<T*>(workspace);
    const int max_blocks = max_grid_blocks();
    int nb = 0, rc;
    switch (D) {
        case 1: rc = dispatch_target<T, 1>(target_id, all_g, eps, N, K, w, s, b, c, tparams, ntp, epsv, partials, max_blocks, &nb, stream); break;
        case 2: rc = dispatch_target<T, 2>(target_id, all_g, eps, N, K, w, s, b, c, tparams, ntp, epsv, partials, max_blocks, &nb, stream); break;
        case 3: rc = dispatch_target<T, 3>(target_id, all_g, eps, N, K, w, s, b, c, tparams, ntp, epsv, partials, max_blocks, &nb, stream); break;
        case 4: rc = dispatch_target<T, 4>(target_id, all_g, eps, N, K, w, s, b, c, tparams, ntp, epsv, partials, max_blocks, &nb, stream); break;
        default: rc = NF_ERR_UNSUPPORTED; break;
    }
    if (rc != NF_OK) return rc;
    const int row = K * (2 * D + 2) + 2;
    int fblocks = (row + 7) / 8;
    if (fblocks > 64) fblocks = 64;
    elbo_finalize_kernel<T><<<fblocks, 256, 0, stream>>>(partials, nb, K, D, all_g ? 1 : 0, out_F, g_w, g_s, g_b, g_c);
    NF_CUDA_OK(cudaGetLastError());
    return NF_OK;
}

}  // namespace nf

extern "C" {

size_t nf_planar_elbo_workspace_bytes(int K, int D, int elem_size) {
    if (K < 1 || D < 1 || elem_size < 1) return 0;
    const size_t row = (size_t)K * (2 * D + 2) + 2;
    return row * (size_t)nf::max_grid_blocks() * (size_t)elem_size;
}

int nf_planar_elbo_fwd_bwd_f32(const float* eps, int64_t N, int D, int K, const float* w, const float* s,
                               const float* b, const float* c, int target_id, const float* target_params,
                               int n_target_params, float log_eps, float* out_F, float* g_w, float* g_s,
                               float* g_b, float* g_c, void* workspace, size_t workspace_bytes, void* stream) {
    return nf::elbo_entry<float>(eps, N, D, K, w, s, b, c, target_id, target_params, n_target_params, log_eps,
                                 out_F, g_w, g_s, g_b, g_c, workspace, workspace_bytes, stream);
}

int nf_planar_elbo_fwd_bwd_f64(const double* eps, int64_t N, int D, int K, const double* w, const double* s,
                               const double* b, const double* c, int target_id, const double* target_params,
                               int n_target_params, double log_eps, double* out_F, double* g_w, double* g_s,
                               double* g_b, double* g_c, void* workspace, size_t workspace_bytes, void* stream) {
    return nf::elbo_entry<double>(eps, N, D, K, w, s, b, c, target_id, target_params, n_target_params, log_eps,
                                  out_F, g_w, g_s, g_b, g_c, workspace, workspace_bytes, stream);
}

}  // extern "C"
