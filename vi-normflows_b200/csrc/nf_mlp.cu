// Encoder / decoder MLP of the flow-VAE (nn_models.py:41-84 `Feedforward.forward`) and the
// Bernoulli log-likelihood (distributions.py:86-89), forward and backward, sm_100a.
//
// Two GEMM back ends: float32 problems whose dimensions TMA can describe (multiples of 4, batch >= 128)
// run on the tensor cores (nf_gemm_tc.cu: TMA + tcgen05.mma kind::tf32 with a hi/lo operand split that
// keeps float32 accuracy); everything else (float64, rbf activations, odd shapes) runs on the
// register-blocked shared-memory CUDA-core GEMM below.  Bias + activation are fused in both.  Activations are
// kept sample-major [N][H] so that every layer is C[m][n] = sum_k A[m][k] W[n][k] with both
// operands contiguous along k; the flat weight vector has exactly the reference's layout
// (nn_models.py:59-80): W1 (H x D_in) | b1 | {W_l (H x H) | b_l} x (L-1) | W_out (D_out x H) | b_out.
#include <cstdlib>
#include "nf_common.cuh"

namespace nf {

// tensor-core GEMM (nf_gemm_tc.cu): C = epi(A B^T), both operands K-contiguous
int tc_gemm_f32(const float* A, long long lda, const float* B, long long ldb, int M, int N, int K, float* C, long long ldc,
                float* CT, long long ldct, const float* bias, int act, const float* mask_src, long long ldm, int mask,
                cudaStream_t st, int splits = 1, float* split_ws = nullptr, bool accumulate = false);

enum { ACT_NONE = 0, ACT_RELU = 1, ACT_SIGMOID = 2, ACT_RBF = 3 };   // rbf: exp(-x^2) (nn_models.py:175)

template <typename T> __device__ __forceinline__ T act_fwd(T x, int act) {
    using M = Mth<T>;
    switch (act) {
        case ACT_RELU: return x > (T)0 ? x : (T)0;                       // transformations.py:21
        case ACT_SIGMOID: return M::div((T)1, (T)1 + M::exp(-x));        // transformations.py:7-8
        case ACT_RBF: return M::exp(-x * x);
        default: return x;
    }
}
// derivative expressed with the saved OUTPUT y (and, for rbf, the pre-activation is recovered
// from y = exp(-x^2) only up to sign, so rbf layers save the pre-activation instead: see below)
template <typename T> __device__ __forceinline__ T act_bwd_from_out(T y, int act) {
    switch (act) {
        case ACT_RELU: return y > (T)0 ? (T)1 : (T)0;
        case ACT_SIGMOID: return y * ((T)1 - y);
        default: return (T)1;
    }
}

constexpr int TM = 64, TN = 64, TK = 16, GT = 256;   // block tile and threads (16x16, 4x4 outputs each)

// C[m][n] (ldc) = epi( sum_k A(m,k) * B(k,n) ),  A(m,k) = a[m*sam + k*sak],  B(k,n) = b[k*sbk + n*sbn].
// Split-K: blockIdx.z handles k in [z*kchunk, (z+1)*kchunk) and writes to c + z*M*ldc (partials,
// epilogue skipped) when gridDim.z > 1.
template <typename T>
__global__ void __launch_bounds__(GT)
gemm_kernel(const T* __restrict__ a, long long sam, long long sak, const T* __restrict__ b, long long sbk,
            long long sbn, T* __restrict__ c, long long ldc, int M, int N, int K, int kchunk,
            const T* __restrict__ bias, int act, T* __restrict__ preact /* optional copy of the pre-activation */) {
    __shared__ T As[TK][TM + 4];
    __shared__ T Bs[TK][TN + 4];
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int m0 = blockIdx.y * TM, n0 = blockIdx.x * TN;
    const int kbeg = blockIdx.z * kchunk;
    const int kend = min(K, kbeg + kchunk);
    T acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = (T)0;
    // loader mapping: make the contiguous global index the fastest-varying thread index
    const bool a_kfast = sak == 1;       // A contiguous along k
    const bool b_kfast = sbk == 1;       // B contiguous along k
    for (int k0 = kbeg; k0 < kend; k0 += TK) {
#pragma unroll
        for (int r = 0; r < (TM * TK) / GT; ++r) {
            const int e = tid + r * GT;
            const int kk = a_kfast ? (e % TK) : (e / TM);
            const int mm = a_kfast ? (e / TK) : (e % TM);
            const int gm = m0 + mm, gk = k0 + kk;
            As[kk][mm] = (gm < M && gk < kend) ? a[(long long)gm * sam + (long long)gk * sak] : (T)0;
        }
#pragma unroll
        for (int r = 0; r < (TN * TK) / GT; ++r) {
            const int e = tid + r * GT;
            const int kk = b_kfast ? (e % TK) : (e / TN);
            const int nn = b_kfast ? (e / TK) : (e % TN);
            const int gn = n0 + nn, gk = k0 + kk;
            Bs[kk][nn] = (gn < N && gk < kend) ? b[(long long)gk * sbk + (long long)gn * sbn] : (T)0;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < TK; ++kk) {
            T av[4], bv[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) av[i] = As[kk][ty * 4 + i];
#pragma unroll
            for (int j = 0; j < 4; ++j) bv[j] = Bs[kk][tx * 4 + j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = Mth<T>::fma(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }
    const bool partial = gridDim.z > 1;
    T* cz = c + (partial ? (long long)blockIdx.z * M * ldc : 0);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int gm = m0 + ty * 4 + i;
        if (gm >= M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int gn = n0 + tx * 4 + j;
            if (gn >= N) continue;
            T v = acc[i][j];
            if (!partial) {
                if (bias) v += bias[gn];
                if (preact) preact[(long long)gm * ldc + gn] = v;
                v = act_fwd(v, act);
            }
            cz[(long long)gm * ldc + gn] = v;
        }
    }
}

// out[i] = sum_z partial[z][i] (fixed order), i < count
template <typename T>
__global__ void splitk_reduce_kernel(const T* __restrict__ partial, int nz, long long count, T* __restrict__ out) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < count; i += (long long)gridDim.x * blockDim.x) {
        T s = (T)0;
        for (int z = 0; z < nz; ++z) s += partial[(long long)z * count + i];
        out[i] = s;
    }
}

// dz = dy * act'(.)   (relu / sigmoid from the saved output, rbf from the saved pre-activation)
template <typename T>
__global__ void act_backward_kernel(const T* __restrict__ dy, const T* __restrict__ saved, long long count, int act,
                                    T* __restrict__ dz) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < count; i += (long long)gridDim.x * blockDim.x) {
        const T sv = saved[i];
        T d;
        if (act == ACT_RBF) d = (T)-2 * sv * Mth<T>::exp(-sv * sv);
        else d = act_bwd_from_out(sv, act);
        dz[i] = dy[i] * d;
    }
}

// column sums out[n] = sum_m x[m][n], two fixed-order stages: blockIdx.y sums a slab of rows for 32
// columns into part[slab][n]; with gridDim.y == 1 the result goes straight to out.
constexpr int kColsumSlabs = 64;
template <typename T>
__global__ void colsum_kernel(const T* __restrict__ x, int M, int N, T* __restrict__ out, int rows_per_slab) {
    __shared__ T red[8][33];
    const int n = blockIdx.x * 32 + (threadIdx.x & 31);
    const int r = threadIdx.x >> 5;
    const int m_lo = blockIdx.y * rows_per_slab;
    const int m_hi = min(M, m_lo + rows_per_slab);
    T s = (T)0;
    if (n < N)
        for (int m = m_lo + r; m < m_hi; m += 8) s += x[(long long)m * N + n];
    red[r][threadIdx.x & 31] = s;
    __syncthreads();
    if (r == 0 && n < N) {
        T t = (T)0;
#pragma unroll
        for (int i = 0; i < 8; ++i) t += red[i][threadIdx.x & 31];
        out[(long long)blockIdx.y * N + n] = t;
    }
}
// bias gradient: column sums of dz [M][N] via `scratch` (kColsumSlabs * N elements)
template <typename T>
static int launch_colsum(const T* x, int M, int N, T* out, T* scratch, cudaStream_t st) {
    if (M < 4 * kColsumSlabs || !scratch) {
        colsum_kernel<T><<<dim3((N + 31) / 32, 1), 256, 0, st>>>(x, M, N, out, M);
    } else {
        const int rps = (M + kColsumSlabs - 1) / kColsumSlabs;
        colsum_kernel<T><<<dim3((N + 31) / 32, kColsumSlabs), 256, 0, st>>>(x, M, N, scratch, rps);
        colsum_kernel<T><<<dim3((N + 31) / 32, 1), 256, 0, st>>>(scratch, kColsumSlabs, N, out, kColsumSlabs);
    }
    NF_CUDA_OK(cudaGetLastError());
    return NF_OK;
}

template <typename T>
static int launch_gemm(const T* a, long long sam, long long sak, const T* b, long long sbk, long long sbn, T* c,
                       long long ldc, int M, int N, int K, const T* bias, int act, T* preact, int splits,
                       T* split_ws, cudaStream_t st) {
    dim3 grid((N + TN - 1) / TN, (M + TM - 1) / TM, splits);
    int kchunk = (K + splits - 1) / splits;
    kchunk = ((kchunk + TK - 1) / TK) * TK;
    if (splits > 1) {
        gemm_kernel<T><<<grid, GT, 0, st>>>(a, sam, sak, b, sbk, sbn, split_ws, ldc, M, N, K, kchunk, nullptr, ACT_NONE, nullptr);
        NF_CUDA_OK(cudaGetLastError());
        const long long count = (long long)M * ldc;
        splitk_reduce_kernel<T><<<(int)min((count + 255) / 256, (long long)1184), 256, 0, st>>>(split_ws, splits, count, c);
    } else {
        gemm_kernel<T><<<grid, GT, 0, st>>>(a, sam, sak, b, sbk, sbn, c, ldc, M, N, K, kchunk, bias, act, preact);
    }
    NF_CUDA_OK(cudaGetLastError());
    return NF_OK;
}


// out[c][r] = in[r][c]   (32x32 tiles through shared memory; both sides coalesced)
template <typename T>
__global__ void transpose_kernel(const T* __restrict__ in, int R, int Ccols, T* __restrict__ out) {
    __shared__ T tile[32][33];
    const int r0 = blockIdx.y * 32, c0 = blockIdx.x * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int r = r0 + i, c = c0 + threadIdx.x;
        tile[i][threadIdx.x] = (r < R && c < Ccols) ? in[(long long)r * Ccols + c] : (T)0;
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int c = c0 + i, r = r0 + threadIdx.x;
        if (c < Ccols && r < R) out[(long long)c * R + r] = tile[threadIdx.x][i];
    }
}
template <typename T>
static int launch_transpose(const T* in, int R, int Ccols, T* out, cudaStream_t st) {
    dim3 grid((Ccols + 31) / 32, (R + 31) / 32), block(32, 8);
    transpose_kernel<T><<<grid, block, 0, st>>>(in, R, Ccols, out);
    NF_CUDA_OK(cudaGetLastError());
    return NF_OK;
}

struct MlpDims { int N, D_in, H, L, D_out; };
static inline long long mlp_param_count(const MlpDims& d) {       // nn_models.py:16-19
    return ((long long)d.D_in * d.H + d.H) + ((long long)d.D_out * d.H + d.D_out) + (long long)(d.L - 1) * ((long long)d.H * d.H + d.H);
}
constexpr int kSplits = 32;      // split-K factor of the weight-gradient GEMMs (contraction over the batch)

// forward workspace: L hidden activation buffers [N][H] (+ L pre-activation buffers for rbf)
// (the tensor-core path also keeps the transposed activations [L][H][N] and x^T [D_in][N])
static inline size_t mlp_fwd_ws_elems(const MlpDims& d, int hidden_act) {
    return (size_t)d.L * d.N * d.H * 2 + (size_t)d.D_in * d.N;
}
// backward workspace: two ping-pong dz buffers [N][max(H, D_out)] + split-K partials of the largest dW
static inline size_t mlp_bwd_ws_elems(const MlpDims& d) {
    const size_t wide = (size_t)max(d.H, d.D_out);
    const size_t maxw = (size_t)max((long long)d.H * d.D_in, max((long long)d.D_out * d.H, (long long)d.H * d.H));
    return 4 * (size_t)d.N * wide + (size_t)(kSplits + 1) * maxw;      // dz x2, dz^T x2, W^T, split-K partials
}

static inline bool mlp_use_tc(const MlpDims& d, int hidden_act, int out_act, size_t elem) {
    if (elem != 4 || getenv("NF_MLP_SIMT")) return false;
    if (hidden_act != ACT_RELU && hidden_act != ACT_NONE) return false;
    if (out_act == ACT_RBF) return false;
    return d.N % 4 == 0 && d.D_in % 4 == 0 && d.H % 4 == 0 && d.D_out % 4 == 0 && d.N >= 128;
}

// ---- tensor-core path (float32): every GEMM is C = A B^T with K-contiguous operands; the epilogues
// also write the transposed activations / gradients the weight-gradient GEMMs contract over ----
static int mlp_forward_tc(const float* wts, const float* x, MlpDims d, int hidden_act, int out_act, float* ws, float* out,
                          cudaStream_t st) {
    const size_t layer = (size_t)d.N * d.H;
    float* acts = ws;
    float* actsT = ws + (size_t)d.L * layer;
    float* xT = ws + 2 * (size_t)d.L * layer;
    int rc = launch_transpose<float>(x, d.N, d.D_in, xT, st);
    if (rc != NF_OK) return rc;
    const float* in = x;
    int kin = d.D_in;
    const float* p = wts;
    for (int l = 0; l < d.L; ++l) {
        const float* W = p;
        const float* bvec = p + (size_t)d.H * kin;
        rc = tc_gemm_f32(in, kin, W, kin, d.N, d.H, kin, acts + (size_t)l * layer, d.H, actsT + (size_t)l * layer, d.N, bvec,
                         hidden_act, nullptr, 0, 0, st);
        if (rc != NF_OK) return rc;
        p = bvec + d.H;
        in = acts + (size_t)l * layer;
        kin = d.H;
    }
    const float* W = p;
    const float* bvec = p + (size_t)d.D_out * d.H;
    return tc_gemm_f32(in, d.H, W, d.H, d.N, d.D_out, d.H, out, d.D_out, nullptr, 0, bvec, out_act, nullptr, 0, 0, st);
}

static int mlp_backward_tc(const float* wts, const float* x, MlpDims d, int hidden_act, int out_act, const float* fws,
                           const float* out, const float* g_out, float* g_wts, float* g_x, float* bws, cudaStream_t st);


template <typename T>
static int mlp_forward(const T* wts, const T* x, MlpDims d, int hidden_act, int out_act, T* ws, T* out, cudaStream_t st) {
    if (d.N < 0 || d.D_in < 1 || d.H < 1 || d.L < 1 || d.D_out < 1 || !wts || !ws || (d.N > 0 && (!x || !out))) return NF_ERR_INVALID;
    if (hidden_act < 0 || hidden_act > ACT_RBF || out_act < 0 || out_act > ACT_RBF) return NF_ERR_INVALID;
    if (d.N == 0) return NF_OK;
    if constexpr (sizeof(T) == 4) {
        if (mlp_use_tc(d, hidden_act, out_act, sizeof(T))) return mlp_forward_tc(wts, x, d, hidden_act, out_act, ws, out, st);
    }
    const bool rbf = hidden_act == ACT_RBF;
    const size_t layer = (size_t)d.N * d.H;
    T* pre = rbf ? ws + (size_t)d.L * layer : nullptr;
    const T* in = x;
    int kin = d.D_in;
    const T* p = wts;
    for (int l = 0; l < d.L; ++l) {
        const T* W = p;
        const T* bvec = p + (size_t)d.H * kin;
        T* a = ws + (size_t)l * layer;
        int rc = launch_gemm<T>(in, kin, 1, W, 1, kin, a, d.H, d.N, d.H, kin, bvec, hidden_act,
                                rbf ? pre + (size_t)l * layer : nullptr, 1, nullptr, st);
        if (rc != NF_OK) return rc;
        p = bvec + d.H;
        in = a;
        kin = d.H;
    }
    const T* W = p;
    const T* bvec = p + (size_t)d.D_out * d.H;
    return launch_gemm<T>(in, d.H, 1, W, 1, d.H, out, d.D_out, d.N, d.D_out, d.H, bvec, out_act, nullptr, 1, nullptr, st);
}

template <typename T>
static int mlp_backward(const T* wts, const T* x, MlpDims d, int hidden_act, int out_act, const T* fws, const T* out,
                        const T* g_out, T* g_wts, T* g_x, T* bws, cudaStream_t st) {
    if (d.N < 1 || !wts || !x || !fws || !out || !g_out || !g_wts || !bws) return NF_ERR_INVALID;
    if (out_act == ACT_RBF) return NF_ERR_UNSUPPORTED;
    if constexpr (sizeof(T) == 4) {
        if (mlp_use_tc(d, hidden_act, out_act, sizeof(T)))
            return mlp_backward_tc(wts, x, d, hidden_act, out_act, fws, out, g_out, g_wts, g_x, bws, st);
    }
    const bool rbf = hidden_act == ACT_RBF;
    const size_t layer = (size_t)d.N * d.H;
    const size_t wide = (size_t)max(d.H, d.D_out);
    T* dz_a = bws;
    T* dz_b = bws + (size_t)d.N * wide;
    T* split_ws = bws + 4 * (size_t)d.N * wide + (size_t)max((long long)d.H * d.D_in, max((long long)d.D_out * d.H, (long long)d.H * d.H));
    const int eb = 1184;
    // parameter offsets
    const long long P = mlp_param_count(d);
    long long off_out = P - ((long long)d.D_out * d.H + d.D_out);
    // ---- output layer ----
    const long long cnt_out = (long long)d.N * d.D_out;
    act_backward_kernel<T><<<eb, 256, 0, st>>>(g_out, out, cnt_out, out_act, dz_a);     // dz_a: [N][D_out]
    NF_CUDA_OK(cudaGetLastError());
    const T* a_prev = fws + (size_t)(d.L - 1) * layer;                                   // last hidden activation
    // dW_out[j][h] = sum_n dz[n][j] a_prev[n][h] : A(m=j,k=n) = dz[n*D_out + j], B(k=n, n'=h) = a_prev[n*H + h]
    int rc = launch_gemm<T>(dz_a, 1, d.D_out, a_prev, d.H, 1, g_wts + off_out, d.H, d.D_out, d.H, d.N, nullptr, ACT_NONE,
                            nullptr, kSplits, split_ws, st);
    if (rc != NF_OK) return rc;
    if ((rc = launch_colsum<T>(dz_a, d.N, d.D_out, g_wts + off_out + (long long)d.D_out * d.H, split_ws, st)) != NF_OK) return rc;
    // d a_prev[n][h] = sum_j dz[n][j] W_out[j][h] : B(k=j, n'=h) = W_out[j*H + h]
    rc = launch_gemm<T>(dz_a, d.D_out, 1, wts + off_out, d.H, 1, dz_b, d.H, d.N, d.H, d.D_out, nullptr, ACT_NONE, nullptr, 1,
                        nullptr, st);
    if (rc != NF_OK) return rc;
    // ---- hidden layers, last to first ----
    long long off = off_out;
    T* g_act = dz_b;      // gradient wrt the activation of layer l
    T* other = dz_a;
    for (int l = d.L - 1; l >= 0; --l) {
        const int kin = l == 0 ? d.D_in : d.H;
        off -= (long long)d.H * kin + d.H;
        const T* saved = rbf ? fws + (size_t)d.L * layer + (size_t)l * layer : fws + (size_t)l * layer;
        act_backward_kernel<T><<<eb, 256, 0, st>>>(g_act, saved, (long long)layer, hidden_act, other);   // other: dz [N][H]
        NF_CUDA_OK(cudaGetLastError());
        const T* in = l == 0 ? x : fws + (size_t)(l - 1) * layer;
        rc = launch_gemm<T>(other, 1, d.H, in, kin, 1, g_wts + off, kin, d.H, kin, d.N, nullptr, ACT_NONE, nullptr, kSplits,
                            split_ws, st);
        if (rc != NF_OK) return rc;
        if ((rc = launch_colsum<T>(other, d.N, d.H, g_wts + off + (long long)d.H * kin, split_ws, st)) != NF_OK) return rc;
        if (l > 0) {
            rc = launch_gemm<T>(other, d.H, 1, wts + off, kin, 1, g_act, d.H, d.N, d.H, d.H, nullptr, ACT_NONE, nullptr, 1, nullptr, st);
            if (rc != NF_OK) return rc;
        } else if (g_x) {
            rc = launch_gemm<T>(other, d.H, 1, wts + off, kin, 1, g_x, d.D_in, d.N, d.D_in, d.H, nullptr, ACT_NONE, nullptr, 1, nullptr, st);
            if (rc != NF_OK) return rc;
        }
    }
    return NF_OK;
}

static int mlp_backward_tc(const float* wts, const float* x, MlpDims d, int hidden_act, int out_act, const float* fws,
                           const float* out, const float* g_out, float* g_wts, float* g_x, float* bws, cudaStream_t st) {
    const size_t layer = (size_t)d.N * d.H;
    const size_t wide = (size_t)max(d.H, d.D_out);
    const float* acts = fws;
    const float* actsT = fws + (size_t)d.L * layer;
    const float* xT = fws + 2 * (size_t)d.L * layer;
    float* dz_cur = bws;                                  // [N][<=wide]
    float* dz_oth = bws + (size_t)d.N * wide;
    float* dzT_cur = bws + 2 * (size_t)d.N * wide;        // [<=wide][N]
    float* dzT_oth = bws + 3 * (size_t)d.N * wide;
    float* WT = bws + 4 * (size_t)d.N * wide;             // transposed weight scratch
    float* split_ws = WT + (size_t)max((long long)d.H * d.D_in, max((long long)d.D_out * d.H, (long long)d.H * d.H));
    const int ksplit = d.N >= 2048 ? 16 : (d.N >= 512 ? 4 : 1);       // weight gradients contract over the batch
    const int mask = hidden_act == ACT_RELU ? 1 : 0;
    const long long P = mlp_param_count(d);
    long long off = P - ((long long)d.D_out * d.H + d.D_out);
    int rc;
    // ---- output layer: dz = g_out * act'(out) ----
    act_backward_kernel<float><<<1184, 256, 0, st>>>(g_out, out, (long long)d.N * d.D_out, out_act, dz_cur);
    NF_CUDA_OK(cudaGetLastError());
    if ((rc = launch_transpose<float>(dz_cur, d.N, d.D_out, dzT_cur, st)) != NF_OK) return rc;
    // dW_out[j][h] = sum_n dz^T[j][n] a^T[h][n]
    if ((rc = tc_gemm_f32(dzT_cur, d.N, actsT + (size_t)(d.L - 1) * layer, d.N, d.D_out, d.H, d.N, g_wts + off, d.H, nullptr, 0,
                          nullptr, 0, nullptr, 0, 0, st, ksplit, split_ws)) != NF_OK) return rc;
    if ((rc = launch_colsum<float>(dz_cur, d.N, d.D_out, g_wts + off + (long long)d.D_out * d.H, split_ws, st)) != NF_OK) return rc;
    // dz_{L-1}[n][h] = (sum_j dz[n][j] W_out[j][h]) * act'(a_{L-1}[n][h])
    if ((rc = launch_transpose<float>(wts + off, d.D_out, d.H, WT, st)) != NF_OK) return rc;
    if ((rc = tc_gemm_f32(dz_cur, d.D_out, WT, d.D_out, d.N, d.H, d.D_out, dz_oth, d.H, dzT_oth, d.N, nullptr, 0,
                          acts + (size_t)(d.L - 1) * layer, d.H, mask, st)) != NF_OK) return rc;
    { float* t = dz_cur; dz_cur = dz_oth; dz_oth = t; t = dzT_cur; dzT_cur = dzT_oth; dzT_oth = t; }
    for (int l = d.L - 1; l >= 0; --l) {
        const int kin = l == 0 ? d.D_in : d.H;
        off -= (long long)d.H * kin + d.H;
        const float* inT = l == 0 ? xT : actsT + (size_t)(l - 1) * layer;
        // dW_l[j][k] = sum_n dz^T[j][n] in^T[k][n]
        if ((rc = tc_gemm_f32(dzT_cur, d.N, inT, d.N, d.H, kin, d.N, g_wts + off, kin, nullptr, 0, nullptr, 0, nullptr, 0, 0, st, ksplit, split_ws)) != NF_OK) return rc;
        if ((rc = launch_colsum<float>(dz_cur, d.N, d.H, g_wts + off + (long long)d.H * kin, split_ws, st)) != NF_OK) return rc;
        if (l > 0 || g_x) {
            if ((rc = launch_transpose<float>(wts + off, d.H, kin, WT, st)) != NF_OK) return rc;      // [kin][H]
            if (l > 0) {
                if ((rc = tc_gemm_f32(dz_cur, d.H, WT, d.H, d.N, kin, d.H, dz_oth, kin, dzT_oth, d.N, nullptr, 0,
                                      acts + (size_t)(l - 1) * layer, d.H, mask, st)) != NF_OK) return rc;
                float* t = dz_cur; dz_cur = dz_oth; dz_oth = t; t = dzT_cur; dzT_cur = dzT_oth; dzT_oth = t;
            } else {
                if ((rc = tc_gemm_f32(dz_cur, d.H, WT, d.H, d.N, kin, d.H, g_x, kin, nullptr, 0, nullptr, 0, nullptr, 0, 0, st)) != NF_OK) return rc;
            }
        }
    }
    return NF_OK;
}

// ---- Bernoulli log-likelihood: sum x log(eps+p) + (1-x) log(eps+1-p), and d/dp ------------------
template <typename T>
__global__ void __launch_bounds__(256)
bernoulli_kernel(const T* __restrict__ x, const T* __restrict__ p, long long count, T epsv, T g_scale_host,
                 const T* __restrict__ g_scale_dev /* upstream scalar in device memory, or NULL */,
                 T* __restrict__ partial, T* __restrict__ g_p) {
    using M = Mth<T>;
    const T g_scale = g_scale_dev ? g_scale_dev[0] * g_scale_host : g_scale_host;
    T s = (T)0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < count; i += (long long)gridDim.x * blockDim.x) {
        const T xv = x[i], pv = p[i];
        const T a = epsv + pv, b = epsv + (T)1 - pv;
        s += xv * M::logp(a) + ((T)1 - xv) * M::logp(b);
        if (g_p) g_p[i] = g_scale * (M::div(xv, a) - M::div((T)1 - xv, b));
    }
    s = warp_sum(s);
    __shared__ T red[8];
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        T t = (T)0;
#pragma unroll
        for (int i = 0; i < 8; ++i) t += red[i];
        partial[blockIdx.x] = t;
    }
}
template <typename T>
__global__ void bernoulli_final_kernel(const T* __restrict__ partial, int n, T* __restrict__ out) {
    T s = (T)0;
    for (int i = threadIdx.x; i < n; i += 32) s += partial[i];
    s = warp_sum(s);
    if (threadIdx.x == 0) out[0] = s;
}
constexpr int kBernBlocks = 592;

template <typename T>
static int bernoulli_entry(const T* x, const T* p, long long count, T epsv, T g_scale, T* out, T* g_p, T* ws, cudaStream_t st,
                           const T* g_scale_dev = nullptr) {
    if (count < 0 || !out || !ws || (count > 0 && (!x || !p))) return NF_ERR_INVALID;
    const int blocks = (int)max(1LL, min((long long)kBernBlocks, (count + 255) / 256));
    bernoulli_kernel<T><<<blocks, 256, 0, st>>>(x, p, count, epsv, g_scale, g_scale_dev, ws, g_p);
    NF_CUDA_OK(cudaGetLastError());
    bernoulli_final_kernel<T><<<1, 32, 0, st>>>(ws, blocks, out);
    NF_CUDA_OK(cudaGetLastError());
    return NF_OK;
}

}  // namespace nf

extern "C" {

int64_t nf_mlp_param_count(int D_in, int H, int L, int D_out) {
    nf::MlpDims d{0, D_in, H, L, D_out};
    return nf::mlp_param_count(d);
}
size_t nf_mlp_forward_workspace_bytes(int64_t N, int D_in, int H, int L, int D_out, int hidden_act, int elem_size) {
    nf::MlpDims d{(int)N, D_in, H, L, D_out};
    return nf::mlp_fwd_ws_elems(d, hidden_act) * (size_t)elem_size;
}
size_t nf_mlp_backward_workspace_bytes(int64_t N, int D_in, int H, int L, int D_out, int elem_size) {
    nf::MlpDims d{(int)N, D_in, H, L, D_out};
    return nf::mlp_bwd_ws_elems(d) * (size_t)elem_size;
}
int nf_mlp_forward_f32(const float* weights, const float* x, int64_t N, int D_in, int H, int L, int D_out, int hidden_act,
                       int out_act, float* fwd_ws, float* out, void* stream) {
    return nf::mlp_forward<float>(weights, x, nf::MlpDims{(int)N, D_in, H, L, D_out}, hidden_act, out_act, fwd_ws, out, (cudaStream_t)stream);
}
int nf_mlp_forward_f64(const double* weights, const double* x, int64_t N, int D_in, int H, int L, int D_out, int hidden_act,
                       int out_act, double* fwd_ws, double* out, void* stream) {
    return nf::mlp_forward<double>(weights, x, nf::MlpDims{(int)N, D_in, H, L, D_out}, hidden_act, out_act, fwd_ws, out, (cudaStream_t)stream);
}
int nf_mlp_backward_f32(const float* weights, const float* x, int64_t N, int D_in, int H, int L, int D_out, int hidden_act,
                        int out_act, const float* fwd_ws, const float* out, const float* g_out, float* g_weights,
                        float* g_x, float* bwd_ws, void* stream) {
    return nf::mlp_backward<float>(weights, x, nf::MlpDims{(int)N, D_in, H, L, D_out}, hidden_act, out_act, fwd_ws, out, g_out,
                                   g_weights, g_x, bwd_ws, (cudaStream_t)stream);
}
int nf_mlp_backward_f64(const double* weights, const double* x, int64_t N, int D_in, int H, int L, int D_out, int hidden_act,
                        int out_act, const double* fwd_ws, const double* out, const double* g_out, double* g_weights,
                        double* g_x, double* bwd_ws, void* stream) {
    return nf::mlp_backward<double>(weights, x, nf::MlpDims{(int)N, D_in, H, L, D_out}, hidden_act, out_act, fwd_ws, out, g_out,
                                    g_weights, g_x, bwd_ws, (cudaStream_t)stream);
}
size_t nf_bernoulli_workspace_bytes(int elem_size) { return (size_t)nf::kBernBlocks * (size_t)elem_size; }
int nf_bernoulli_ll_f32(const float* x, const float* p, int64_t count, float log_eps, float g_scale, float* out_sum,
                        float* g_p, float* workspace, void* stream) {
    return nf::bernoulli_entry<float>(x, p, count, log_eps, g_scale, out_sum, g_p, workspace, (cudaStream_t)stream);
}
int nf_bernoulli_ll_dev_f32(const float* x, const float* p, int64_t count, float log_eps, const float* g_scale_dev, float* out_sum,
                            float* g_p, float* workspace, void* stream) {
    if (!g_scale_dev) return NF_ERR_INVALID;
    return nf::bernoulli_entry<float>(x, p, count, log_eps, 1.f, out_sum, g_p, workspace, (cudaStream_t)stream, g_scale_dev);
}
int nf_bernoulli_ll_dev_f64(const double* x, const double* p, int64_t count, double log_eps, const double* g_scale_dev, double* out_sum,
                            double* g_p, double* workspace, void* stream) {
    if (!g_scale_dev) return NF_ERR_INVALID;
    return nf::bernoulli_entry<double>(x, p, count, log_eps, 1.0, out_sum, g_p, workspace, (cudaStream_t)stream, g_scale_dev);
}
int nf_bernoulli_ll_f64(const double* x, const double* p, int64_t count, double log_eps, double g_scale, double* out_sum,
                        double* g_p, double* workspace, void* stream) {
    return nf::bernoulli_entry<double>(x, p, count, log_eps, g_scale, out_sum, g_p, workspace, (cudaStream_t)stream);
}

}  // extern "C"
