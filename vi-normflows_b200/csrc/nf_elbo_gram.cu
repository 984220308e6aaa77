// Shared-parameter free energy for wide latent spaces through matrix products ("Gram" path), sm_100a.
//
// get_data.py:104-109 moves z only inside span{s_j}:  z_k = z_0 + sum_{j<=k} s_j t_j.  With
// M[k][j] = w_k.s_j every size-D operation of the layer chain and of its reverse pass becomes a
// product with a K-column matrix, and what is sequential shrinks to K^2/2 scalar multiply-adds
// per sample and direction (tests/test_gram_identity.py restates the algebra in NumPy):
//
//   A0 = z_0 W^T + b                       product 1   [N x D] x [D x K]
//   a_k = A0[k] + sum_{j<k} M[k][j] t_j    forward recurrence, one thread per sample
//   z_K = z_0 + T S                        product 2   [N x K] x [K x D]
//   gz_K = dF/dz_K, log p, log q_0         target kernel (distributions.py:10-54)
//   P = gz_K S^T                           product 3   [N x D] x [D x K]
//   gz_k.s_k = P[k] + sum_{j>k} delta_j M[j][k]        reverse recurrence
//   g_s = T^T gz_K + triu(T^T Delta,1) W   products 4, 6   (contraction over the samples)
//   g_w = Delta^T z_0 + triu(T^T Delta,1)^T S          product 5
//
// The products run on the tensor cores (nf_gemm_tc.cu: tcgen05.mma kind::tf32, 3 passes for
// float32 accuracy).  The direct kernel (nf_elbo_wide.cuh) spends 7 D multiply-adds per layer and
// sample on the CUDA cores and, worse, has to add a [K][2D] gradient record per warp and layer
// to global memory; here the gradient accumulators are output tiles of products 4-6.
// Samples are processed in chunks of up to 32M/D (the workspace holds one chunk's intermediates).
// float32, D % 4 == 0, K <= 64, targets GAUSSDIAG / MIXDIAG (<= 16 components).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include "nf_common.cuh"
#include "nf_elbo_pair.cuh"

namespace nf {

int tc_gemm_f32(const float* A, long long lda, const float* B, long long ldb, int M, int N, int K, float* C, long long ldc,
                float* CT, long long ldct, const float* bias, int act, const float* mask_src, long long ldm, int mask,
                cudaStream_t st, int splits = 1, float* split_ws = nullptr, bool accumulate = false);

int tc_gemm_stacked_f32(const float* A, long long lda, int M1, const float* A2, long long lda2, int M2, const float* B, long long ldb,
                        int N, int K, float* C, long long ldc, float* CT, long long ldct, const float* bias, int act,
                        const float* mask_src, long long ldm, int mask, cudaStream_t st, int splits, float* split_ws, bool accumulate);

namespace gram {

#ifndef NF_REC_THREADS
#define NF_REC_THREADS 128
#endif
// Recurrence kernels: 128-thread blocks, four per SM.  Their bodies are ~4000 instructions of straight-line code (the
// running values sit in registers, so nothing can be rolled) and instruction fetch is their largest stall reason; one
// 512-thread block per SM whose warps re-align every sixteenth layer (to share fetched lines) was measured slower
// (D=64, N=2^19: forward 127 vs 118 us, reverse 165 vs 146 us).  -DNF_REC_THREADS=512 rebuilds that variant.
constexpr int kRecThreads = NF_REC_THREADS;
constexpr int kRecBlocksPerSM = 512 / NF_REC_THREADS;
constexpr int kMaxK = 64;
constexpr int kMaxMix = 16;
constexpr int kTgtRows = 32;         // rows per block of the target kernel
constexpr int kTgtThreads = 256;

static inline int pad_k(int K) { return K <= 16 ? 16 : (K <= 32 ? 32 : 64); }
static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// chunk pitch (samples): up to 32M elements of [Nc][D] (small chunks leave the GPU under-filled: every kernel of the
// chain is a separate launch), a multiple of 512, no more than the job needs
static inline long long chunk_samples(long long N, int D) {
    long long nc = ((long long)32 << 20) / D;
    if (const char* e = getenv("NF_GRAM_CHUNK")) { const long long v = atoll(e); if (v >= 128) nc = v; }
    constexpr long long kBlk = 512;       // whole blocks of the recurrence kernels
    nc = (nc + kBlk - 1) / kBlk * kBlk;
    const long long need = (N + kBlk - 1) / kBlk * kBlk;
    return nc < need ? nc : need;
}
static inline int grad_splits(long long nc, int D) {
    const int tiles_m = (D + 127) / 128;
    long long s = (296 + tiles_m - 1) / tiles_m;          // about two CTAs per SM
    const long long nkb = (nc + 31) / 32;
    if (s > nkb / 4) s = nkb / 4;                          // at least four k-blocks per CTA
    return (int)(s < 1 ? 1 : s);
}

struct Layout {
    int KP; long long Nc; int splits;
    size_t acc, M, Mt, ST, G4, G56, part_ld, part_g, part_t, split4, split56;   // offsets in floats
    size_t AT, DT, TT, Y, total;
};
constexpr int kG5Row = 64;       // row of the stacked product 5/6 result where the z_0^T Delta block starts (KP <= 64 rounded up to 64)
static Layout make_layout(long long N, int D, int K) {
    Layout L{};
    L.KP = pad_k(K);
    L.Nc = chunk_samples(N, D);
    L.splits = grad_splits(L.Nc, D);
    const size_t KP = (size_t)L.KP, Nc = (size_t)L.Nc;
    size_t o = 0;
    auto take = [&](size_t n) { const size_t at = o; o = align_up(o + n, 64); return at; };   // 256-byte aligned pieces
    L.acc = take(3 + 2 * kMaxK);
    L.M = take(KP * KP);
    L.Mt = take(KP * KP);
    L.ST = take((size_t)D * KP);
    L.G4 = take((size_t)D * K);                 // gz_K^T T             [D][K]
    L.G56 = take((size_t)(kG5Row + D) * K);     // [T^T ; 0 ; z_0^T] Delta: rows [0, KP) and [64, 64 + D)
    L.part_ld = take(Nc / kRecThreads);
    L.part_g = take(Nc / kRecThreads * 2 * KP);
    L.part_t = take(2 * (Nc / kTgtRows));
    L.split4 = take((size_t)L.splits * D * K);
    L.split56 = take((size_t)L.splits * (kG5Row + D) * K);
    L.AT = take(KP * Nc);                       // A0^T, later P^T (rows K.. stay zero)
    L.DT = take(KP * Nc);                       // Delta^T
    L.TT = take(KP * Nc);                       // T^T
    L.Y = take(Nc * D);                         // T S, later gz_K
    L.total = o;
    return L;
}

// ------------------------------------------------------------------------------------------------
// set-up: M = W S^T (both orientations, scaled for the exp2-based tanh, strictly triangular parts only),
// S^T, and cleared running totals.  Grid (KP, KP / 8), 256 threads: block (k, y) owns the eight products
// w_k . s_j, j = 8 y .. 8 y + 7, one per warp.
__global__ void prep_kernel(const float* __restrict__ w, const float* __restrict__ s, int K, int D, int KP,
                            float* __restrict__ M, float* __restrict__ Mt, float* __restrict__ ST, float* __restrict__ acc) {
    const int k = blockIdx.x;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (k == 0 && blockIdx.y == 0)
        for (int i = threadIdx.x; i < 3 + 2 * kMaxK; i += blockDim.x) acc[i] = 0.f;
    const int j = blockIdx.y * 8 + warp;
    if (j < KP) {
        float v = 0.f;
        if (k < K && j < K && j != k) {
            for (int d = lane; d < D; d += 32) v = fmaf(w[(size_t)k * D + d], s[(size_t)j * D + d], v);
            v = warp_sum(v);
        }
        if (lane == 0) {
            // forward uses column j of M below the diagonal: Mt[j][k], k > j, in units of the scaled tanh argument
            Mt[(size_t)j * KP + k] = (k > j) ? v * Mth<float>::kArgScale : 0.f;
            // reverse uses row k of M left of the diagonal: M[k][j], j < k
            M[(size_t)k * KP + j] = (j < k) ? v : 0.f;
        }
    }
    if (blockIdx.y == 0)
        for (int d = threadIdx.x; d < D; d += blockDim.x) ST[(size_t)d * KP + k] = k < K ? s[(size_t)k * D + d] : 0.f;
}

// block-wide sum in a fixed order (warp trees, then warp 0 over the warp results)
template <int NT>
__device__ __forceinline__ float block_sum(float v, float* red /* [NT/32] */) {
    v = warp_sum(v);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float t = 0.f;
    if (warp == 0) {
        t = lane < NT / 32 ? red[lane] : 0.f;
        t = warp_sum(t);
    }
    return t;      // valid on warp 0
}

// ------------------------------------------------------------------------------------------------
// Forward recurrence.  One thread per sample; the K pre-activations live in registers and are turned
// into t_j = tanh(a_j) in place, column j of M being added to the later ones as soon as t_j is known.
// Mt rows are zero up to and including the diagonal, so whole float4 groups can be applied blindly.
// The loop body has no control flow that depends on K or on the sample index (padded layers have zero
// rows everywhere and a unit log-det factor; samples past the end of the chunk compute on whatever the
// buffers hold and are masked out of the sums), so the whole recurrence is one basic block.
// The product-1 values of a thread (one per layer, needed at step k) are brought into a thread-private
// column of shared memory by cp.async, sixteen layers per group, and waited for group by group: the
// fetches cost no registers and overlap the arithmetic.
// Outputs: T^T ([KP][Nc], coalesced; operand of products 2, 4 and 6), per-block sum of the log-det terms.
__device__ __forceinline__ void cp_async4(float* smem_dst, const float* gsrc) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_le(int pending) {      // `pending` folds to a constant where this is used
    if (pending <= 0) asm volatile("cp.async.wait_group 0;" ::: "memory");
    else if (pending == 1) asm volatile("cp.async.wait_group 1;" ::: "memory");
    else if (pending == 2) asm volatile("cp.async.wait_group 2;" ::: "memory");
    else if (pending == 3) asm volatile("cp.async.wait_group 3;" ::: "memory");
    else asm volatile("cp.async.wait_group 4;" ::: "memory");
}

// S samples per thread (columns tid, tid + 128, ... of the block) share every M value fetched from shared memory.
// S = 2 halves the broadcast LDS.128 traffic but was measured slower on B200 (D=256: 539 vs 504 us per call, D=64:
// 830 vs 765): at 219-255 registers only 8 warps fit an SM, and the recurrences are latency-bound, not LSU-bound.
constexpr int kRecS = 1;
template <int KP, int S> constexpr size_t rec_fwd_smem() { return (size_t)(KP * KP + KP * S * kRecThreads) * sizeof(float); }
template <int KP, int S> constexpr size_t rec_bwd_smem() { return (size_t)(KP * KP + (KP < 32 ? KP : 32) * 2 * S * kRecThreads) * sizeof(float); }

template <int KP, int S>
struct FwdCtx {
    const float* stage;    // [KP][S * 128], already offset by the thread index
    const float* sM; const float* sc; const float* se;
    float* dst; long long ldn;
};

// one layer of the forward recurrence; the layers are peeled by template recursion so that a[][] stays in registers
template <int KP, int S, int j>
struct FwdLayer {
    static __device__ __forceinline__ void run(const FwdCtx<KP, S>& x_, float (&a)[S][KP], float (&prod)[S], float (&ld)[S]) {
        using M_ = Mth<float>;
        constexpr int NG = KP / 16, W = S * kRecThreads;
        if (j > 0 && (j & 15) == 0) { cp_async_wait_le(NG - 1 - j / 16); if (kRecThreads > 128) __syncthreads(); }
        float t[S];
#pragma unroll
        for (int u = 0; u < S; ++u) {
            t[u] = M_::tanh_pre(fmaf(x_.stage[j * W + u * kRecThreads], M_::kArgScale, a[u][j]));
            const float hp = fmaf(-t[u], t[u], 1.f);
            prod[u] *= x_.se[j] + fabsf(fmaf(hp, x_.sc[j], 1.f));
            if ((j & 3) == 3) { ld[u] += M_::flog(prod[u]); prod[u] = 1.f; }      // MUFU.LG2, as in the direct kernels
            x_.dst[(size_t)j * x_.ldn + u * kRecThreads] = t[u];
        }
#pragma unroll
        for (int k4 = (j + 1) / 4; k4 < KP / 4; ++k4) {
            const float4 m = *reinterpret_cast<const float4*>(&x_.sM[j * KP + 4 * k4]);
#pragma unroll
            for (int u = 0; u < S; ++u) {
                if (4 * k4 + 0 > j) a[u][4 * k4 + 0] = fmaf(m.x, t[u], a[u][4 * k4 + 0]);
                if (4 * k4 + 1 > j) a[u][4 * k4 + 1] = fmaf(m.y, t[u], a[u][4 * k4 + 1]);
                if (4 * k4 + 2 > j) a[u][4 * k4 + 2] = fmaf(m.z, t[u], a[u][4 * k4 + 2]);
                if (4 * k4 + 3 > j) a[u][4 * k4 + 3] = fmaf(m.w, t[u], a[u][4 * k4 + 3]);
            }
        }
        if constexpr (j + 1 < KP) FwdLayer<KP, S, j + 1>::run(x_, a, prod, ld);
    }
};

template <int KP, int S>
__global__ void __launch_bounds__(kRecThreads, S == 1 ? kRecBlocksPerSM : 1)
rec_fwd_kernel(const float* __restrict__ AT, long long ldn, int n_act, int K, const float* __restrict__ Mt,
               const float* __restrict__ c, float epsv, float* __restrict__ TT, float* __restrict__ part_ld) {
    constexpr int W = S * kRecThreads;            // samples per block
    extern __shared__ __align__(16) float dyn[];
    float* sM = dyn;                              // [KP][KP]
    float* stage = dyn + KP * KP;                 // [KP][W]: column = sample inside the block
    __shared__ float sc[KP], se[KP];
    __shared__ float red[kRecThreads / 32];
    constexpr int NG = KP / 16;
    const int n = blockIdx.x * W + threadIdx.x;   // first sample of this thread; all < Nc: the buffers are padded to whole blocks
    const float* src = AT + n;
    for (int i = threadIdx.x; i < KP * KP / 4; i += kRecThreads) cp_async16(sM + 4 * i, Mt + 4 * i);
#pragma unroll
    for (int g = 0; g < NG; ++g) {
#pragma unroll
        for (int k = 16 * g; k < 16 * g + 16; ++k)
#pragma unroll
            for (int u = 0; u < S; ++u)
                cp_async4(stage + k * W + u * kRecThreads + threadIdx.x, src + (size_t)k * ldn + u * kRecThreads);
        cp_async_commit();                        // group g (group 0 also carries M)
    }
    for (int i = threadIdx.x; i < KP; i += kRecThreads) { sc[i] = i < K ? c[i] : 0.f; se[i] = i < K ? epsv : 0.f; }
    cp_async_wait_le(NG - 1);
    __syncthreads();                              // M, c and eps visible to all
    FwdCtx<KP, S> x_;
    x_.stage = stage + threadIdx.x; x_.sM = sM; x_.sc = sc; x_.se = se; x_.dst = TT + n; x_.ldn = ldn;
    float a[S][KP];
    float prod[S], ld[S];
#pragma unroll
    for (int u = 0; u < S; ++u) {
        prod[u] = 1.f; ld[u] = 0.f;
#pragma unroll
        for (int k = 0; k < KP; ++k) a[u][k] = 0.f;
    }
    FwdLayer<KP, S, 0>::run(x_, a, prod, ld);
    float mine = 0.f;
#pragma unroll
    for (int u = 0; u < S; ++u) mine += (n + u * kRecThreads < n_act) ? ld[u] : 0.f;
    const float tot = block_sum<kRecThreads>(mine, red);
    if (threadIdx.x == 0) part_ld[blockIdx.x] = tot;
}

// Reverse recurrence.  p_k starts as (gz_K.s_k) and receives delta_j M[j][k] from the later layers.
// Outputs: Delta^T ([KP][Nc]) and per-block sums over the samples of delta_k (= d F / d b_k) and of the
// d F / d c_k terms: sixteen layers at a time through the multi-value warp reduction, then across the four
// warps in order.  Branch-free body as in rec_fwd_kernel; the layers are peeled by template recursion
// (a 64-trip `#pragma unroll` of this body is declined by the compiler, which would put p[] in local memory).
// P^T and T^T arrive through cp.async in quarters of sixteen layers, two quarters resident (thread-private
// columns again): when a quarter has been consumed its slot is refilled with the quarter after next.
template <int KP, int S>
struct BwdCtx {
    const float* srcp; const float* srct; float* dst; long long ldn;
    const float* sM; const float* sc; const float* se; const float* smask; float* sg;   // sg: this warp's [2 KP] record
    float* stage;          // [2 slots][16 layers][2 (P^T, T^T)][S * 128], already offset by the thread index
    bool ok[S]; int lane;
};

template <int KP, int S>
__device__ __forceinline__ void bwd_issue_quarter(const BwdCtx<KP, S>& x_, int q) {
    constexpr int W = S * kRecThreads;
    float* slot = x_.stage + (q & 1) * 16 * 2 * W;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const int k = 16 * q + i;
#pragma unroll
        for (int u = 0; u < S; ++u) {
            cp_async4(slot + (2 * i) * W + u * kRecThreads, x_.srcp + (size_t)k * x_.ldn + u * kRecThreads);
            cp_async4(slot + (2 * i + 1) * W + u * kRecThreads, x_.srct + (size_t)k * x_.ldn + u * kRecThreads);
        }
    }
    cp_async_commit();
}

template <int KP, int S, int k>
struct BwdLayer {
    static __device__ __forceinline__ void run(const BwdCtx<KP, S>& x_, float (&p)[S][KP], float (&gbv)[16], float (&gcv)[16]) {
        using M_ = Mth<float>;
        constexpr int NQ = KP / 16, q = k / 16, W = S * kRecThreads;
        if ((k & 15) == 15) {                    // entering quarter q
            if (q < NQ - 1 && q >= 1) bwd_issue_quarter<KP, S>(x_, q - 1);     // quarter q+1 is consumed: refill its slot
            cp_async_wait_le(q >= 1 ? 1 : 0);
            if (kRecThreads > 128 && q < NQ - 1) __syncthreads();     // large-block variant: keep the warps in step
        }
        const float* slot = x_.stage + (q & 1) * 16 * 2 * W + 2 * (k & 15) * W;
        const float ck = x_.sc[k];
        float ga[S];
        float sum_b = 0.f, sum_c = 0.f;
#pragma unroll
        for (int u = 0; u < S; ++u) {
            const float pin = slot[u * kRecThreads];
            const float t = slot[W + u * kRecThreads];
            const float hp = fmaf(-t, t, 1.f);
            const float x = fmaf(hp, ck, 1.f);
            const float y = x_.se[k] + fabsf(x);
            const float gx = M_::neg_copysign(M_::rcp(y), x) * x_.smask[k];
            const float gt = fmaf(-2.f * ck * gx, t, p[u][k] + pin);
            ga[u] = gt * hp;
            sum_c += x_.ok[u] ? gx * hp : 0.f;
            sum_b += x_.ok[u] ? ga[u] : 0.f;
            x_.dst[(size_t)k * x_.ldn + u * kRecThreads] = ga[u];
        }
        gcv[k & 15] = sum_c;
        gbv[k & 15] = sum_b;
#pragma unroll
        for (int i4 = 0; i4 < (k + 3) / 4; ++i4) {
            const float4 m = *reinterpret_cast<const float4*>(&x_.sM[k * KP + 4 * i4]);
#pragma unroll
            for (int u = 0; u < S; ++u) {
                if (4 * i4 + 0 < k) p[u][4 * i4 + 0] = fmaf(m.x, ga[u], p[u][4 * i4 + 0]);
                if (4 * i4 + 1 < k) p[u][4 * i4 + 1] = fmaf(m.y, ga[u], p[u][4 * i4 + 1]);
                if (4 * i4 + 2 < k) p[u][4 * i4 + 2] = fmaf(m.z, ga[u], p[u][4 * i4 + 2]);
                if (4 * i4 + 3 < k) p[u][4 * i4 + 3] = fmaf(m.w, ga[u], p[u][4 * i4 + 3]);
            }
        }
        if ((k & 15) == 0) {        // layers k .. k+15 are complete: reduce them over the warp
            warp_reduce_multi<float, 16>(gbv, x_.lane);
            warp_reduce_multi<float, 16>(gcv, x_.lane);
            if (warp_slot_owner<16>(x_.lane)) {
                x_.sg[k + warp_slot<16>(x_.lane)] = gbv[0];
                x_.sg[KP + k + warp_slot<16>(x_.lane)] = gcv[0];
            }
        }
        if constexpr (k > 0) BwdLayer<KP, S, k - 1>::run(x_, p, gbv, gcv);
    }
};

template <int KP, int S>
__global__ void __launch_bounds__(kRecThreads, S == 1 ? kRecBlocksPerSM : 1)
rec_bwd_kernel(const float* __restrict__ PT, const float* __restrict__ TT, long long ldn, int n_act, int K,
               const float* __restrict__ Mrow, const float* __restrict__ c, float epsv, float* __restrict__ DT,
               float* __restrict__ part_g /* [gridDim][2][KP] : g_b | g_c */) {
    extern __shared__ __align__(16) float dyn[];
    float* sM = dyn;                                // [KP][KP]
    __shared__ float sc[KP], se[KP], smask[KP];
    __shared__ float sg[kRecThreads / 32][2 * KP];
    constexpr int NQ = KP / 16;
    const int n = blockIdx.x * S * kRecThreads + threadIdx.x;
    BwdCtx<KP, S> x_;
    x_.srcp = PT + n; x_.srct = TT + n; x_.dst = DT + n; x_.ldn = ldn;
    x_.sM = sM; x_.sc = sc; x_.se = se; x_.smask = smask; x_.sg = sg[threadIdx.x >> 5];
    x_.stage = dyn + KP * KP + threadIdx.x;
#pragma unroll
    for (int u = 0; u < S; ++u) x_.ok[u] = n + u * kRecThreads < n_act;
    x_.lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < KP * KP / 4; i += kRecThreads) cp_async16(sM + 4 * i, Mrow + 4 * i);
    bwd_issue_quarter<KP, S>(x_, NQ - 1);           // this group also carries M
    if (NQ >= 2) bwd_issue_quarter<KP, S>(x_, NQ - 2);
    for (int i = threadIdx.x; i < KP; i += kRecThreads) {
        sc[i] = i < K ? c[i] : 0.f; se[i] = i < K ? epsv : 0.f; smask[i] = i < K ? 1.f : 0.f;
    }
    cp_async_wait_le(NQ >= 2 ? 1 : 0);
    __syncthreads();
    float p[S][KP];
#pragma unroll
    for (int u = 0; u < S; ++u)
#pragma unroll
        for (int k = 0; k < KP; ++k) p[u][k] = 0.f;
    float gcv[16], gbv[16];
    BwdLayer<KP, S, KP - 1>::run(x_, p, gbv, gcv);
    __syncthreads();
    if (threadIdx.x < 2 * KP) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < kRecThreads / 32; ++w) t += sg[w][threadIdx.x];
        part_g[(size_t)blockIdx.x * 2 * KP + threadIdx.x] = t;
    }
}

// ------------------------------------------------------------------------------------------------
// Packed-pair recurrences: a thread owns TWO samples (columns tid and tid + 128 of a 256-sample block) and keeps their
// running values as one 64-bit register pair per layer, so that the K^2/2 updates per sample are FFMA2 instructions
// (fma.rn.f32x2: both samples per issue slot, the M value as the scalar-broadcast operand).  The same values as the
// scalar kernels above, operation for operation (fma.rn.f32x2 rounds each half like fmaf); what differs is the block
// size of the partial sums.  These have half the instructions per sample of the one-sample kernels, at 8 warps per SM
// instead of 16: the updates of one layer are independent of each other, which is what the FMA pipe needs.
using pr::u64; using pr::pk; using pr::upk; using pr::fma2; using pr::mul2; using pr::add2;
constexpr int kRecPkW = 2 * kRecThreads;         // samples per block of the packed kernels

template <int KP, int j>
struct FwdLayerP {
    static __device__ __forceinline__ void run(const FwdCtx<KP, 2>& x_, u64 (&a)[KP], u64& prod, u64& ld) {
        using M_ = Mth<float>;
        constexpr int NG = KP / 16, W = kRecPkW;
        if (j > 0 && (j & 15) == 0) { cp_async_wait_le(NG - 1 - j / 16); if (kRecThreads > 128) __syncthreads(); }
        const u64 one2 = pk(1.f, 1.f);
        const u64 arg = fma2(pk(x_.stage[j * W], x_.stage[j * W + kRecThreads]), pk(M_::kArgScale, M_::kArgScale), a[j]);
        float a0, a1;
        upk(arg, a0, a1);
        const u64 den = add2(pk(M_::ex2(a0), M_::ex2(a1)), one2);
        float d0, d1;
        upk(den, d0, d1);
        const u64 r = pk(M_::rcp(d0), M_::rcp(d1));
        const u64 t = fma2(pk(-2.f, -2.f), r, one2);                 // tanh_pre
        const u64 nt = fma2(pk(2.f, 2.f), r, pk(-1.f, -1.f));        // -t, exactly
        const u64 hp = fma2(nt, t, one2);
        const float cj = x_.sc[j], ej = x_.se[j];
        float x0, x1;
        upk(fma2(hp, pk(cj, cj), one2), x0, x1);
        prod = mul2(prod, add2(pk(fabsf(x0), fabsf(x1)), pk(ej, ej)));
        if ((j & 3) == 3) {
            float p0, p1;
            upk(prod, p0, p1);
            ld = fma2(pk(M_::lg2(p0), M_::lg2(p1)), pk(0.6931471805599453f, 0.6931471805599453f), ld);
            prod = one2;
        }
        float t0, t1;
        upk(t, t0, t1);
        x_.dst[(size_t)j * x_.ldn] = t0;
        x_.dst[(size_t)j * x_.ldn + kRecThreads] = t1;
#pragma unroll
        for (int k4 = (j + 1) / 4; k4 < KP / 4; ++k4) {
            const float4 m = *reinterpret_cast<const float4*>(&x_.sM[j * KP + 4 * k4]);
            if (4 * k4 + 0 > j) a[4 * k4 + 0] = fma2(pk(m.x, m.x), t, a[4 * k4 + 0]);
            if (4 * k4 + 1 > j) a[4 * k4 + 1] = fma2(pk(m.y, m.y), t, a[4 * k4 + 1]);
            if (4 * k4 + 2 > j) a[4 * k4 + 2] = fma2(pk(m.z, m.z), t, a[4 * k4 + 2]);
            if (4 * k4 + 3 > j) a[4 * k4 + 3] = fma2(pk(m.w, m.w), t, a[4 * k4 + 3]);
        }
        if constexpr (j + 1 < KP) FwdLayerP<KP, j + 1>::run(x_, a, prod, ld);
    }
};

template <int KP>
__global__ void __launch_bounds__(kRecThreads, kRecThreads <= 128 ? 2 : 1)
rec_fwd_pk_kernel(const float* __restrict__ AT, long long ldn, int n_act, int K, const float* __restrict__ Mt,
                  const float* __restrict__ c, float epsv, float* __restrict__ TT, float* __restrict__ part_ld) {
    constexpr int W = kRecPkW;
    extern __shared__ __align__(16) float dyn[];
    float* sM = dyn;                              // [KP][KP]
    float* stage = dyn + KP * KP;                 // [KP][W]
    __shared__ float sc[KP], se[KP];
    __shared__ float red[kRecThreads / 32];
    constexpr int NG = KP / 16;
    const int n = blockIdx.x * W + threadIdx.x;
    const float* src = AT + n;
    for (int i = threadIdx.x; i < KP * KP / 4; i += kRecThreads) cp_async16(sM + 4 * i, Mt + 4 * i);
#pragma unroll
    for (int g = 0; g < NG; ++g) {
#pragma unroll
        for (int k = 16 * g; k < 16 * g + 16; ++k) {
            cp_async4(stage + k * W + threadIdx.x, src + (size_t)k * ldn);
            cp_async4(stage + k * W + kRecThreads + threadIdx.x, src + (size_t)k * ldn + kRecThreads);
        }
        cp_async_commit();
    }
    for (int i = threadIdx.x; i < KP; i += kRecThreads) { sc[i] = i < K ? c[i] : 0.f; se[i] = i < K ? epsv : 0.f; }
    cp_async_wait_le(NG - 1);
    __syncthreads();
    FwdCtx<KP, 2> x_;
    x_.stage = stage + threadIdx.x; x_.sM = sM; x_.sc = sc; x_.se = se; x_.dst = TT + n; x_.ldn = ldn;
    u64 a[KP];
#pragma unroll
    for (int k = 0; k < KP; ++k) a[k] = pk(0.f, 0.f);
    u64 prod = pk(1.f, 1.f), ld = pk(0.f, 0.f);
    FwdLayerP<KP, 0>::run(x_, a, prod, ld);
    float l0, l1;
    upk(ld, l0, l1);
    const float mine = (n < n_act ? l0 : 0.f) + (n + kRecThreads < n_act ? l1 : 0.f);
    const float tot = block_sum<kRecThreads>(mine, red);
    if (threadIdx.x == 0) part_ld[blockIdx.x] = tot;
}

template <int KP, int k>
struct BwdLayerP {
    static __device__ __forceinline__ void run(const BwdCtx<KP, 2>& x_, u64 (&p)[KP], float (&gbv)[16], float (&gcv)[16]) {
        using M_ = Mth<float>;
        constexpr int NQ = KP / 16, q = k / 16, W = kRecPkW;
        if ((k & 15) == 15) {
            if (q < NQ - 1 && q >= 1) bwd_issue_quarter<KP, 2>(x_, q - 1);
            cp_async_wait_le(q >= 1 ? 1 : 0);
            if (kRecThreads > 128 && q < NQ - 1) __syncthreads();
        }
        const float* slot = x_.stage + (q & 1) * 16 * 2 * W + 2 * (k & 15) * W;
        const float ck = x_.sc[k], ek = x_.se[k], mk = x_.smask[k];
        const u64 one2 = pk(1.f, 1.f);
        const u64 pin = pk(slot[0], slot[kRecThreads]);
        const u64 t = pk(slot[W], slot[W + kRecThreads]);
        const u64 hp = fma2(mul2(t, pk(-1.f, -1.f)), t, one2);
        float x0, x1;
        upk(fma2(hp, pk(ck, ck), one2), x0, x1);
        float y0, y1;
        upk(add2(pk(fabsf(x0), fabsf(x1)), pk(ek, ek)), y0, y1);
        const u64 gx = mul2(pk(M_::neg_copysign(M_::rcp(y0), x0), M_::neg_copysign(M_::rcp(y1), x1)), pk(mk, mk));
        const float m2ck = -2.f * ck;
        const u64 gt = fma2(mul2(pk(m2ck, m2ck), gx), t, add2(p[k], pin));
        const u64 ga = mul2(gt, hp);
        float c0, c1, b0, b1;
        upk(mul2(gx, hp), c0, c1);
        upk(ga, b0, b1);
        gcv[k & 15] = (x_.ok[0] ? c0 : 0.f) + (x_.ok[1] ? c1 : 0.f);
        gbv[k & 15] = (x_.ok[0] ? b0 : 0.f) + (x_.ok[1] ? b1 : 0.f);
        x_.dst[(size_t)k * x_.ldn] = b0;
        x_.dst[(size_t)k * x_.ldn + kRecThreads] = b1;
#pragma unroll
        for (int i4 = 0; i4 < (k + 3) / 4; ++i4) {
            const float4 m = *reinterpret_cast<const float4*>(&x_.sM[k * KP + 4 * i4]);
            if (4 * i4 + 0 < k) p[4 * i4 + 0] = fma2(pk(m.x, m.x), ga, p[4 * i4 + 0]);
            if (4 * i4 + 1 < k) p[4 * i4 + 1] = fma2(pk(m.y, m.y), ga, p[4 * i4 + 1]);
            if (4 * i4 + 2 < k) p[4 * i4 + 2] = fma2(pk(m.z, m.z), ga, p[4 * i4 + 2]);
            if (4 * i4 + 3 < k) p[4 * i4 + 3] = fma2(pk(m.w, m.w), ga, p[4 * i4 + 3]);
        }
        if ((k & 15) == 0) {
            warp_reduce_multi<float, 16>(gbv, x_.lane);
            warp_reduce_multi<float, 16>(gcv, x_.lane);
            if (warp_slot_owner<16>(x_.lane)) {
                x_.sg[k + warp_slot<16>(x_.lane)] = gbv[0];
                x_.sg[KP + k + warp_slot<16>(x_.lane)] = gcv[0];
            }
        }
        if constexpr (k > 0) BwdLayerP<KP, k - 1>::run(x_, p, gbv, gcv);
    }
};

template <int KP>
__global__ void __launch_bounds__(kRecThreads, kRecThreads <= 128 ? 2 : 1)
rec_bwd_pk_kernel(const float* __restrict__ PT, const float* __restrict__ TT, long long ldn, int n_act, int K,
                  const float* __restrict__ Mrow, const float* __restrict__ c, float epsv, float* __restrict__ DT,
                  float* __restrict__ part_g /* [gridDim][2][KP] : g_b | g_c */) {
    extern __shared__ __align__(16) float dyn[];
    float* sM = dyn;
    __shared__ float sc[KP], se[KP], smask[KP];
    __shared__ float sg[kRecThreads / 32][2 * KP];
    constexpr int NQ = KP / 16;
    const int n = blockIdx.x * kRecPkW + threadIdx.x;
    BwdCtx<KP, 2> x_;
    x_.srcp = PT + n; x_.srct = TT + n; x_.dst = DT + n; x_.ldn = ldn;
    x_.sM = sM; x_.sc = sc; x_.se = se; x_.smask = smask; x_.sg = sg[threadIdx.x >> 5];
    x_.stage = dyn + KP * KP + threadIdx.x;
    x_.ok[0] = n < n_act; x_.ok[1] = n + kRecThreads < n_act;
    x_.lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < KP * KP / 4; i += kRecThreads) cp_async16(sM + 4 * i, Mrow + 4 * i);
    bwd_issue_quarter<KP, 2>(x_, NQ - 1);
    if (NQ >= 2) bwd_issue_quarter<KP, 2>(x_, NQ - 2);
    for (int i = threadIdx.x; i < KP; i += kRecThreads) {
        sc[i] = i < K ? c[i] : 0.f; se[i] = i < K ? epsv : 0.f; smask[i] = i < K ? 1.f : 0.f;
    }
    cp_async_wait_le(NQ >= 2 ? 1 : 0);
    __syncthreads();
    u64 p[KP];
#pragma unroll
    for (int k = 0; k < KP; ++k) p[k] = pk(0.f, 0.f);
    float gcv[16], gbv[16];
    BwdLayerP<KP, KP - 1>::run(x_, p, gbv, gcv);
    __syncthreads();
    if (threadIdx.x < 2 * KP) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < kRecThreads / 32; ++w) t += sg[w][threadIdx.x];
        part_g[(size_t)blockIdx.x * 2 * KP + threadIdx.x] = t;
    }
}

// ------------------------------------------------------------------------------------------------
// Target kernel: z_K = z_0 + Y, log p(z_K), log q_0(z_0) and gz_K = dF/dz_K = -d log p / dz_K, which replaces Y in
// place (rows: operand of product 3 as it is, and -- read sample-major by the product kernel -- of product 4).
// Eight lanes share a sample (four samples per warp, 32 per block); a lane keeps its float4 slices of z_K in
// registers between the row sums and the gradient, so every byte is read once and written once, in 128-byte runs.
// Same density formulas as elbo_wide_kernel (distributions.py:10-54).
__device__ __forceinline__ float sum8(float v) {
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    v += __shfl_xor_sync(0xffffffffu, v, 4);
    return v;
}
__device__ __forceinline__ float f4_get(const float4& v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : (i == 2 ? v.z : v.w)); }


static inline size_t tgt_smem_bytes(int D) { return 2 * (size_t)D * sizeof(float); }

template <int TGT, bool BWD, int kTgtMaxVec /* float4 slices per lane: D <= 32 kTgtMaxVec */>
__global__ void __launch_bounds__(kTgtThreads)
target_kernel(const float* __restrict__ z0, float* __restrict__ Y, int n_act, int D, const float* __restrict__ tparams, int ntp,
              float epsv, float* __restrict__ part /* [gridDim][2] : sum log q_0, sum log p */) {
    using M_ = Mth<float>;
    extern __shared__ __align__(16) float smem[];
    __shared__ float red[kTgtThreads / 32][2];
    __shared__ float mixe[kTgtRows][kMaxMix];
    const int NV = D / 4;
    float* s_mu = smem;
    float* s_iv = smem + D;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const float LOG2PI = 1.8378770664093453f;
    float g_const = 0.f;
    if (TGT == NF_TARGET_GAUSSDIAG) {
        float ls = 0.f;
        for (int d = tid; d < D; d += kTgtThreads) { s_mu[d] = tparams[d]; s_iv[d] = M_::fexp(-tparams[D + d]); }
        for (int d = lane; d < D; d += 32) ls += tparams[D + d];
        g_const = -0.5f * (D * LOG2PI + warp_sum(ls));
        __syncthreads();
    }
    const int rl = 4 * warp + (lane >> 3), l = lane & 7;       // row inside the block, lane inside the row
    const int n = blockIdx.x * kTgtRows + rl;
    const bool live = n < n_act;
    const float4* z4 = reinterpret_cast<const float4*>(z0 + (size_t)(live ? n : 0) * D);
    float4* y4 = reinterpret_cast<float4*>(Y + (size_t)(live ? n : 0) * D);
    float4 z[kTgtMaxVec];
    float zz = 0.f;
#pragma unroll
    for (int i = 0; i < kTgtMaxVec; ++i) {
        const int c = l + 8 * i;
        if (c < NV && live) {
            const float4 a = z4[c], b = y4[c];
            zz = fmaf(a.x, a.x, fmaf(a.y, a.y, fmaf(a.z, a.z, fmaf(a.w, a.w, zz))));
            z[i] = make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
        } else {
            z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    zz = sum8(zz);
    const float lq = live ? -0.5f * (D * LOG2PI + zz) : 0.f;
    float lp = 0.f;
    if (TGT == NF_TARGET_GAUSSDIAG) {
        const float4* mu4 = reinterpret_cast<const float4*>(s_mu);
        const float4* iv4 = reinterpret_cast<const float4*>(s_iv);
        float q = 0.f;
#pragma unroll
        for (int i = 0; i < kTgtMaxVec; ++i) {
            const int c = l + 8 * i;
            if (c < NV) {
                const float4 m = mu4[c], iv = iv4[c];
                const float dx = z[i].x - m.x, dy = z[i].y - m.y, dz = z[i].z - m.z, dw = z[i].w - m.w;
                q = fmaf(dx * iv.x, dx, fmaf(dy * iv.y, dy, fmaf(dz * iv.z, dz, fmaf(dw * iv.w, dw, q))));
                z[i] = make_float4(dx * iv.x, dy * iv.y, dz * iv.z, dw * iv.w);      // (z - mu) / var, scaled by rho below
            }
        }
        q = sum8(q);
        const float dens = M_::fexp(g_const - 0.5f * q);       // MUFU-based exp / log: once per sample, rel. error ~1e-6
        const float den = epsv + dens;
        lp = live ? M_::flog(den) : 0.f;
        if (BWD) {
            const float rho = M_::div(dens, den);
#pragma unroll
            for (int i = 0; i < kTgtMaxVec; ++i) {
                const int c = l + 8 * i;
                if (c < NV && live) y4[c] = make_float4(rho * z[i].x, rho * z[i].y, rho * z[i].z, rho * z[i].w);
            }
        }
    } else {   // MIXDIAG: [pi (G) | mu (G*D) | var (G*D)]
        const int G = ntp / (1 + 2 * D);
        float dens = 0.f;
        for (int gc = 0; gc < G; ++gc) {
            const float* mu = tparams + G + (size_t)gc * D;
            const float* var = tparams + G + (size_t)G * D + (size_t)gc * D;
            float q = 0.f, lsum = 0.f;
#pragma unroll
            for (int i = 0; i < kTgtMaxVec; ++i) {
                const int c = l + 8 * i;
                if (c < NV) {
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const int d = 4 * c + e;
                        const float df = f4_get(z[i], e) - mu[d];
                        q = fmaf(df, M_::div(df, var[d]), q);
                        lsum += M_::logp(var[d]);
                    }
                }
            }
            q = sum8(q);
            lsum = sum8(lsum);
            const float e = tparams[gc] * M_::exp(-0.5f * (D * LOG2PI + lsum + q));
            dens += e;
            if (l == 0) mixe[rl][gc] = e;
        }
        const float den = epsv + dens;
        lp = live ? M_::logp(den) : 0.f;
        if (BWD) {
            __syncwarp();
            const float inv = M_::div(1.f, den);
#pragma unroll
            for (int i = 0; i < kTgtMaxVec; ++i) {
                const int c = l + 8 * i;
                if (c < NV && live) {
                    float g[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const int d = 4 * c + e;
                        float acc = 0.f;
                        for (int gc = 0; gc < G; ++gc)
                            acc = fmaf(mixe[rl][gc], M_::div(f4_get(z[i], e) - tparams[G + (size_t)gc * D + d], tparams[G + (size_t)G * D + (size_t)gc * D + d]), acc);
                        g[e] = acc * inv;
                    }
                    y4[c] = make_float4(g[0], g[1], g[2], g[3]);
                }
            }
        }
    }
    // block totals in a fixed order: lanes with l == 0 carry one row each
    float vq = (l == 0) ? lq : 0.f, vp = (l == 0) ? lp : 0.f;
    vq = warp_sum(vq);
    vp = warp_sum(vp);
    if (lane == 0) { red[warp][0] = vq; red[warp][1] = vp; }
    __syncthreads();
    if (tid == 0) {
        float tq = 0.f, tp = 0.f;
        for (int i = 0; i < kTgtThreads / 32; ++i) { tq += red[i][0]; tp += red[i][1]; }
        part[2 * blockIdx.x] = tq;
        part[2 * blockIdx.x + 1] = tp;
    }
}

// Per chunk: running totals.  Blocks [0, 2K): column sums of the per-block (g_b | g_c) records; block 2K: log-det,
// log q_0 and log p partial sums.  Fixed order, and chunks arrive in stream order.
__global__ void __launch_bounds__(256)
chunk_totals_kernel(const float* __restrict__ part_g, int KP, int K, int bwd, const float* __restrict__ part_ld, int n_rec,
                    const float* __restrict__ part_t, int n_t, float* __restrict__ acc /* [ld, q0, p | g_b (kMaxK) | g_c (kMaxK)] */) {
    __shared__ float red[8];
    const int b = blockIdx.x;
    if (b < 2 * K) {
        if (!bwd) return;
        const int col = b < K ? b : KP + (b - K);
        float v = 0.f;
        for (int i = threadIdx.x; i < n_rec; i += 256) v += part_g[(size_t)i * 2 * KP + col];
        const float t = block_sum<256>(v, red);
        if (threadIdx.x == 0) acc[3 + (b < K ? b : kMaxK + (b - K))] += t;
    } else {
        float a = 0.f, q = 0.f, p = 0.f;
        for (int i = threadIdx.x; i < n_rec; i += 256) a += part_ld[i];
        for (int i = threadIdx.x; i < n_t; i += 256) { q += part_t[2 * i]; p += part_t[2 * i + 1]; }
        const float ta = block_sum<256>(a, red);
        const float tq = block_sum<256>(q, red);
        const float tp = block_sum<256>(p, red);
        if (threadIdx.x == 0) { acc[0] += ta; acc[1] += tq; acc[2] += tp; }
    }
}

// g_s[k][d] = G4[d][k] + sum_{j>k} G6[k][j] w[j][d] ;  g_w[k][d] = G5[d][k] + sum_{j<k} G6[j][k] s[j][d]
// with G6 = rows [0, KP) and G5 = rows [64, 64 + D) of the stacked product; grid (K, ceil(D / 128)).
__global__ void finish_kernel(const float* __restrict__ acc, const float* __restrict__ G4, const float* __restrict__ G56,
                              const float* __restrict__ w, const float* __restrict__ s, int K, int KP, int D, int bwd,
                              float* __restrict__ out_F, float* __restrict__ g_w, float* __restrict__ g_s,
                              float* __restrict__ g_b, float* __restrict__ g_c) {
    const int k = blockIdx.x;
    if (k == 0 && blockIdx.y == 0 && threadIdx.x == 0) {
        const float sum_q = acc[1] - acc[0];
        out_F[0] = sum_q - acc[2];
        out_F[1] = sum_q;
        out_F[2] = acc[2];
    }
    if (!bwd) return;
    if (blockIdx.y == 0 && threadIdx.x == 0) { g_b[k] = acc[3 + k]; g_c[k] = acc[3 + kMaxK + k]; }
    const float* G6 = G56;
    const float* G5 = G56 + (size_t)kG5Row * K;
    const int d = blockIdx.y * blockDim.x + threadIdx.x;
    if (d >= D) return;
    float gs = G4[(size_t)d * K + k], gw = G5[(size_t)d * K + k];
    for (int j = k + 1; j < K; ++j) gs = fmaf(G6[(size_t)k * K + j], w[(size_t)j * D + d], gs);
    for (int j = 0; j < k; ++j) gw = fmaf(G6[(size_t)j * K + k], s[(size_t)j * D + d], gw);
    g_s[(size_t)k * D + d] = gs;
    g_w[(size_t)k * D + d] = gw;
}

template <int TGT, int VEC>
static cudaError_t launch_target_v(bool bwd, const float* z0, float* Y, int n_act, int D, const float* tp, int ntp, float epsv,
                                   float* part, cudaStream_t st) {
    const size_t smem = tgt_smem_bytes(D);
    const int blocks = (n_act + kTgtRows - 1) / kTgtRows;
    if (bwd) target_kernel<TGT, true, VEC><<<blocks, kTgtThreads, smem, st>>>(z0, Y, n_act, D, tp, ntp, epsv, part);
    else target_kernel<TGT, false, VEC><<<blocks, kTgtThreads, smem, st>>>(z0, Y, n_act, D, tp, ntp, epsv, part);
    return cudaGetLastError();
}
template <int TGT>
static cudaError_t launch_target(bool bwd, const float* z0, float* Y, int n_act, int D, const float* tp, int ntp, float epsv,
                                 float* part, cudaStream_t st) {
    if (D <= 64) return launch_target_v<TGT, 2>(bwd, z0, Y, n_act, D, tp, ntp, epsv, part, st);
    if (D <= 128) return launch_target_v<TGT, 4>(bwd, z0, Y, n_act, D, tp, ntp, epsv, part, st);
    if (D <= 256) return launch_target_v<TGT, 8>(bwd, z0, Y, n_act, D, tp, ntp, epsv, part, st);
    return launch_target_v<TGT, 16>(bwd, z0, Y, n_act, D, tp, ntp, epsv, part, st);
}

// samples per block of the recurrence kernels for a chunk of n_act samples: the packed-pair kernels (256) once they
// fill the GPU twice over, the one-sample kernels (128) below that.  NF_REC_PACKED=0/1 forces one (tuning knob).
static inline int rec_block_samples(int n_act) {
    if (const char* e = getenv("NF_REC_PACKED")) return atoi(e) ? kRecPkW : kRecS * kRecThreads;
    return n_act >= 2 * 148 * 2 * kRecPkW ? kRecPkW : kRecS * kRecThreads;
}
template <int KP>
static cudaError_t launch_rec(bool fwd, const Layout& L, float* ws, int n_act, int K, const float* c, float epsv, cudaStream_t st) {
    constexpr int S = kRecS;
    const int per_block = rec_block_samples(n_act);
    const int blocks = (n_act + per_block - 1) / per_block;
    static bool attr_done_dev[64] = {};           // the attribute is per device
    int dev_id = 0;
    cudaGetDevice(&dev_id);
    bool& attr_done = attr_done_dev[dev_id & 63];
    if (!attr_done) {
        cudaFuncSetAttribute(rec_fwd_kernel<KP, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rec_fwd_smem<KP, S>());
        cudaFuncSetAttribute(rec_bwd_kernel<KP, S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rec_bwd_smem<KP, S>());
        cudaFuncSetAttribute(rec_fwd_pk_kernel<KP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rec_fwd_smem<KP, 2>());
        cudaFuncSetAttribute(rec_bwd_pk_kernel<KP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rec_bwd_smem<KP, 2>());
        attr_done = true;
    }
    if (per_block == kRecPkW) {
        if (fwd)
            rec_fwd_pk_kernel<KP><<<blocks, kRecThreads, rec_fwd_smem<KP, 2>(), st>>>(ws + L.AT, L.Nc, n_act, K, ws + L.Mt, c, epsv,
                                                                                     ws + L.TT, ws + L.part_ld);
        else
            rec_bwd_pk_kernel<KP><<<blocks, kRecThreads, rec_bwd_smem<KP, 2>(), st>>>(ws + L.AT, ws + L.TT, L.Nc, n_act, K, ws + L.M, c,
                                                                                     epsv, ws + L.DT, ws + L.part_g);
        return cudaGetLastError();
    }
    if (fwd)
        rec_fwd_kernel<KP, S><<<blocks, kRecThreads, rec_fwd_smem<KP, S>(), st>>>(ws + L.AT, L.Nc, n_act, K, ws + L.Mt, c, epsv, ws + L.TT,
                                                                                 ws + L.part_ld);
    else
        rec_bwd_kernel<KP, S><<<blocks, kRecThreads, rec_bwd_smem<KP, S>(), st>>>(ws + L.AT, ws + L.TT, L.Nc, n_act, K, ws + L.M, c, epsv,
                                                                                 ws + L.DT, ws + L.part_g);
    return cudaGetLastError();
}
static cudaError_t launch_rec_any(bool fwd, const Layout& L, float* ws, int n_act, int K, const float* c, float epsv, cudaStream_t st) {
    if (L.KP == 16) return launch_rec<16>(fwd, L, ws, n_act, K, c, epsv, st);
    if (L.KP == 32) return launch_rec<32>(fwd, L, ws, n_act, K, c, epsv, st);
    return launch_rec<64>(fwd, L, ws, n_act, K, c, epsv, st);
}

}  // namespace gram

// ---- interface used by nf_elbo_f32.cu --------------------------------------------------------------
// Shapes the path supports, and among those the ones where it is expected to win.  Both cost models (microseconds) were
// fitted on a B200 over D = 32..256, K = 16..64, N = 2^10..2^19 (profiles/r01h_gram_crossover.txt):
//   this path     a floor of ~(78 + 0.45 K) us for its chain of launches, then
//                 per sample max(0.67 ns, 11.5 ps D) in the products and the target kernel + 0.143 ps K^2 in the recurrences;
//   direct kernel one under-filled wave costs T0 = 10 + K (0.35 + 0.0078 D) us (at least 29), beyond it
//                 0.79 ps per (sample, layer, coordinate), growing as (D/64)^0.65 above D = 64 where the per-warp
//                 gradient records spill out of L2 (1.05 ps at D <= 32, where few lanes share a sample).
// NF_GRAM=1 forces this path where it is supported, NF_NO_GRAM=1 forbids it.
bool gram_applicable(long long N, int D, int K, int target_id, int ntp) {
    if (getenv("NF_NO_GRAM")) return false;
    if (D < 32 || D > 512 || (D % 4) != 0 || K < 1 || K > gram::kMaxK || N < 1) return false;
    if (target_id == NF_TARGET_MIXDIAG) { if (ntp / (1 + 2 * D) > gram::kMaxMix) return false; }
    else if (target_id >= 0 && target_id != NF_TARGET_GAUSSDIAG) return false;
    if (const char* e = getenv("NF_GRAM")) { if (atoi(e) != 0) return true; }
    if (N < 512) return false;
    const double n = (double)N;
    const double floor_g = 78.0 + 0.45 * K;
    const double per_d = 11.5e-6 * D > 0.67e-3 ? 11.5e-6 * D : 0.67e-3;
    const double lin_g = floor_g - 30.0 + n * (per_d + 0.143e-6 * K * K);
    const double t_gram = lin_g > floor_g ? lin_g : floor_g;
    double t0 = 10.0 + K * (0.35 + 0.0078 * D);
    if (t0 < 29.0) t0 = 29.0;
    const double cd = D <= 32 ? 1.05e-6 : 0.79e-6 * (D <= 64 ? 1.0 : pow(D / 64.0, 0.65));
    const double lin_d = 0.6 * t0 + n * K * D * cd;
    const double t_direct = lin_d > t0 ? lin_d : t0;
    return t_gram < t_direct;
}

size_t gram_workspace_bytes(long long N, int D, int K) { return gram::make_layout(N, D, K).total * sizeof(float); }

// The evaluation in three parts, so that a caller that receives its samples piecewise (the host-buffer session, which
// overlaps the upload of one piece with the arithmetic on the previous one) can feed pieces of at most gram_chunk_pitch()
// samples: begin (layer products, cleared totals), one call per piece (running totals and gradient products accumulate),
// end (results).  `layout_n` is the sample count the workspace was sized for; it fixes the layout.
long long gram_chunk_pitch(long long layout_n, int D, int K) { return gram::make_layout(layout_n, D, K).Nc; }

int gram_begin(long long layout_n, int D, int K, const float* w, const float* s, void* workspace, cudaStream_t st) {
    using namespace gram;
    const Layout L = make_layout(layout_n, D, K);
    float* ws = reinterpret_cast<float*>(workspace);
    const int KP = L.KP;
    prep_kernel<<<dim3(KP, KP / 8), 256, 0, st>>>(w, s, K, D, KP, ws + L.M, ws + L.Mt, ws + L.ST, ws + L.acc);
    NF_CUDA_OK(cudaGetLastError());
    if (KP > K) NF_CUDA_OK(cudaMemsetAsync(ws + L.AT + (size_t)K * L.Nc, 0, (size_t)(KP - K) * L.Nc * sizeof(float), st));
    return NF_OK;
}

int gram_piece(bool bwd, bool first, const float* z0, int n_act, long long layout_n, int D, int K, const float* w, const float* s,
               const float* b, const float* c, int target_id, const float* tp, int ntp, float epsv, void* workspace, cudaStream_t st) {
    using namespace gram;
    const Layout L = make_layout(layout_n, D, K);
    if (n_act < 1 || n_act > L.Nc) return NF_ERR_INVALID;
    float* ws = reinterpret_cast<float*>(workspace);
    const int KP = L.KP;
    int rc;
    // product 1: A0^T[k][n] = sum_d z0[n][d] w[k][d] + b[k]
    if ((rc = tc_gemm_f32(z0, D, w, D, n_act, K, D, nullptr, 0, ws + L.AT, L.Nc, b, 0, nullptr, 0, 0, st)) != NF_OK) return rc;
    NF_CUDA_OK(launch_rec_any(true, L, ws, n_act, K, c, epsv, st));
    // product 2: Y[n][d] = sum_k T^T[k][n] S^T[d][k]   (T is read layer-major, as the recurrence wrote it)
    if ((rc = tc_gemm_stacked_f32(nullptr, 0, 0, ws + L.TT, L.Nc, n_act, ws + L.ST, KP, D, K, ws + L.Y, D, nullptr, 0, nullptr, 0,
                                  nullptr, 0, 0, st, 1, nullptr, false)) != NF_OK) return rc;
    if (target_id == NF_TARGET_GAUSSDIAG)
        NF_CUDA_OK(launch_target<NF_TARGET_GAUSSDIAG>(bwd, z0, ws + L.Y, n_act, D, tp, ntp, epsv, ws + L.part_t, st));
    else
        NF_CUDA_OK(launch_target<NF_TARGET_MIXDIAG>(bwd, z0, ws + L.Y, n_act, D, tp, ntp, epsv, ws + L.part_t, st));
    if (bwd) {
        // product 3: P^T[k][n] = sum_d gz[n][d] s[k][d]
        if ((rc = tc_gemm_f32(ws + L.Y, D, s, D, n_act, K, D, nullptr, 0, ws + L.AT, L.Nc, nullptr, 0, nullptr, 0, 0, st)) != NF_OK) return rc;
        NF_CUDA_OK(launch_rec_any(false, L, ws, n_act, K, c, epsv, st));
        // products 4-6 contract over the samples of the piece (split across CTAs, fixed-order sums).  gz_K and z_0 are
        // read as they lie in memory, sample-major; 5 and 6 share Delta^T and run as one product whose row blocks
        // are T^T (rows [0, KP)) and z_0 (rows [64, 64 + D))
        if ((rc = tc_gemm_stacked_f32(nullptr, 0, 0, ws + L.Y, D, D, ws + L.TT, L.Nc, K, n_act, ws + L.G4, K, nullptr, 0, nullptr, 0,
                                      nullptr, 0, 0, st, L.splits, ws + L.split4, !first)) != NF_OK) return rc;
        if ((rc = tc_gemm_stacked_f32(ws + L.TT, L.Nc, KP, z0, D, D, ws + L.DT, L.Nc, K, n_act, ws + L.G56, K, nullptr, 0, nullptr, 0,
                                      nullptr, 0, 0, st, L.splits, ws + L.split56, !first)) != NF_OK) return rc;
    }
    chunk_totals_kernel<<<2 * K + 1, 256, 0, st>>>(ws + L.part_g, KP, K, bwd ? 1 : 0, ws + L.part_ld,
                                              (n_act + rec_block_samples(n_act) - 1) / rec_block_samples(n_act), ws + L.part_t,
                                              (n_act + kTgtRows - 1) / kTgtRows, ws + L.acc);
    NF_CUDA_OK(cudaGetLastError());
    return NF_OK;
}

int gram_end(bool bwd, long long layout_n, int D, int K, const float* w, const float* s, float* out_F, float* g_w, float* g_s,
             float* g_b, float* g_c, void* workspace, cudaStream_t st) {
    using namespace gram;
    const Layout L = make_layout(layout_n, D, K);
    float* ws = reinterpret_cast<float*>(workspace);
    finish_kernel<<<dim3(K, (D + 127) / 128), 128, 0, st>>>(ws + L.acc, ws + L.G4, ws + L.G56, w, s, K, L.KP, D, bwd ? 1 : 0, out_F, g_w,
                                                           g_s, g_b, g_c);
    NF_CUDA_OK(cudaGetLastError());
    return NF_OK;
}

int gram_elbo_f32(bool bwd, const float* eps, long long N, int D, int K, const float* w, const float* s, const float* b,
                  const float* c, int target_id, const float* tp, int ntp, float epsv, float* out_F, float* g_w, float* g_s,
                  float* g_b, float* g_c, void* workspace, cudaStream_t st) {
    const long long pitch = gram_chunk_pitch(N, D, K);
    int rc = gram_begin(N, D, K, w, s, workspace, st);
    if (rc != NF_OK) return rc;
    for (long long c0 = 0; c0 < N; c0 += pitch) {
        const int n_act = (int)((N - c0 < pitch) ? (N - c0) : pitch);
        rc = gram_piece(bwd, c0 == 0, eps + (size_t)c0 * D, n_act, N, D, K, w, s, b, c, target_id, tp, ntp, epsv, workspace, st);
        if (rc != NF_OK) return rc;
    }
    return gram_end(bwd, N, D, K, w, s, out_F, g_w, g_s, g_b, g_c, workspace, st);
}

}  // namespace nf
