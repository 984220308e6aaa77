// Shared-parameter planar-flow free energy: fused forward + backward (sm_100a).
// Template implementation; nf_elbo_f32.cu / nf_elbo_f64.cu instantiate it per element type
// (two translation units so that nvcc can build them in parallel).
//
// Replaces get_data.py:97-113 (forward) and the autograd tape replay of
// get_data.py:115 (backward).  One launch per step:
//   * every thread owns S samples (S=8 for float): z, the running gradient and the
//     log-det product stay in registers across all K layers; the only per-layer state
//     that leaves the register file is the tanh value t_k, parked in shared memory so
//     that the reverse sweep can walk z back (z_{k-1} = z_k - s_k t_k) without a second
//     tanh;
//   * the per-layer cost that does not scale with the number of samples (constant
//     fetch, loop control, the cross-lane reduction of the 2D+2 gradient partials) is
//     paid once per S samples per lane; the reduction is a halving butterfly whose
//     result lands in a per-warp shared-memory table;
//   * 2*log2(e) is folded into w and b when the block stages the layer constants, so
//     tanh is FADD + FFMA + MUFU.EX2 + MUFU.RCP; log|det| is accumulated as a running
//     product with one MUFU.LG2 per four layers; the forward pass of the fused kernel
//     does not touch the log-det at all (the reverse sweep needs h' and x anyway);
//   * blocks are persistent (grid = SMs x occupancy) and stride over sample tiles; each
//     block writes one row of partial sums and a second small kernel adds the rows in a
//     fixed order (bitwise reproducible for a fixed launch shape).
#pragma once
#include <cstdio>
#include <cstdlib>
#include <ctime>
#include <mutex>
#include <vector>
#include "nf_common.cuh"
#include "nf_targets.cuh"
#include "nf_elbo_wide.cuh"
#include "nf_launch.cuh"
#include "nf_elbo_pair.cuh"

namespace nf {

constexpr int kElboThreads = 64;
constexpr int kMaxBlocksPerSM = 16;

// ---- tensor memory as a per-thread activation stash ---------------------------------------
// TMEM (256 KB per SM, 128 lanes x 512 32-bit columns) is otherwise idle in this kernel.  A
// 128-thread block maps thread t to TMEM lane t (warp w owns lanes 32w..32w+31, the only ones
// tcgen05.ld/st let it touch), so every thread gets `columns` private 32-bit slots: the tanh
// history of ST samples per thread lives there instead of in shared memory, which roughly
// doubles the number of samples an SM can keep in flight.  One tcgen05.st / tcgen05.ld moves
// the ST activations of a layer.
template <int N> struct TmemVec;
template <> struct TmemVec<2> {
    static __device__ __forceinline__ void st(uint32_t a, const float* r) {
        asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1, %2};" ::"r"(a), "r"(__float_as_uint(r[0])), "r"(__float_as_uint(r[1])) : "memory");
    }
    static __device__ __forceinline__ void ld(uint32_t a, float* r) {
        uint32_t v0, v1;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(v0), "=r"(v1) : "r"(a) : "memory");
        r[0] = __uint_as_float(v0); r[1] = __uint_as_float(v1);
    }
};
template <> struct TmemVec<4> {
    static __device__ __forceinline__ void st(uint32_t a, const float* r) {
        asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(__float_as_uint(r[0])),
                     "r"(__float_as_uint(r[1])), "r"(__float_as_uint(r[2])), "r"(__float_as_uint(r[3])) : "memory");
    }
    static __device__ __forceinline__ void ld(uint32_t a, float* r) {
        uint32_t v0, v1, v2, v3;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v0), "=r"(v1), "=r"(v2), "=r"(v3) : "r"(a) : "memory");
        r[0] = __uint_as_float(v0); r[1] = __uint_as_float(v1); r[2] = __uint_as_float(v2); r[3] = __uint_as_float(v3);
    }
};
template <> struct TmemVec<8> {
    static __device__ __forceinline__ void st(uint32_t a, const float* r) {
        asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(a),
                     "r"(__float_as_uint(r[0])), "r"(__float_as_uint(r[1])), "r"(__float_as_uint(r[2])), "r"(__float_as_uint(r[3])),
                     "r"(__float_as_uint(r[4])), "r"(__float_as_uint(r[5])), "r"(__float_as_uint(r[6])), "r"(__float_as_uint(r[7])) : "memory");
    }
    static __device__ __forceinline__ void ld(uint32_t a, float* r) {
        uint32_t v[8];
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "r"(a) : "memory");
#pragma unroll
        for (int i = 0; i < 8; ++i) r[i] = __uint_as_float(v[i]);
    }
};
// the double instantiation never uses TMEM (ST == 0); these overloads only keep templates well-formed
template <int N> struct TmemVecT {
    static __device__ __forceinline__ void st(uint32_t a, const float* r) { TmemVec<N>::st(a, r); }
    static __device__ __forceinline__ void ld(uint32_t a, float* r) { TmemVec<N>::ld(a, r); }
    static __device__ __forceinline__ void st(uint32_t, const double*) {}
    static __device__ __forceinline__ void ld(uint32_t, double*) {}
};
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__host__ __device__ constexpr int tmem_columns(int n) {       // allocation granule: power of two >= 32
    int c = 32;
    while (c < n) c <<= 1;
    return c;
}

template <int D> __host__ __device__ constexpr int grad_vals() { return 2 * D + 2; }   // w[D] s[D] b c
__host__ __device__ constexpr int pad4(int n) { return (n + 3) & ~3; }
// forward record per layer : w*kappa [D] | b*kappa | s [D]           (padded to a multiple of 4)
// backward record per layer: w [D] | s [D] | c | -2c                  (padded to a multiple of 4)
template <int D> __host__ __device__ constexpr int rec_f() { return pad4(2 * D + 1); }
template <int D> __host__ __device__ constexpr int rec_b() { return pad4(2 * D + 2); }

// dynamic shared memory carve-up (in elements of T)
template <typename T>
struct ElboSmem {
    int lcf, lcb, tp, acc, hist, total;
    // S = samples per thread whose tanh history lives in shared memory (the rest, if any, is in TMEM)
    __host__ __device__ ElboSmem(int K, int D, int S, int ntp_s, bool bwd, int nthreads) {
        const int V = 2 * D + 2;
        lcf = 0;
        lcb = lcf + K * pad4(2 * D + 1);
        tp = lcb + K * pad4(2 * D + 2);
        acc = pad4(tp + ntp_s);
        hist = acc + (bwd ? (nthreads / kWarp) * K * V : 0);
        total = hist + (bwd ? K * S * nthreads : 0);
    }
};

// A thread owns S consecutive samples = S*D consecutive elements of eps.  Full tiles
// are fetched with the widest aligned vector load that divides S*D elements (LDG.128;
// for float D=2 S=8 each lane reads 64 contiguous bytes); the ragged last tile falls
// back to guarded scalar loads.
template <typename T, int D, int S, bool MASKED>
__device__ __forceinline__ void load_samples(const T* __restrict__ eps, long long n0, long long N,
                                             T (&z)[S][D], bool (&ok)[S]) {
    constexpr int E = S * D;
    constexpr int BYTES = E * (int)sizeof(T);
    if (!MASKED) {
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
            ok[j] = true;
#pragma unroll
            for (int d = 0; d < D; ++d) z[j][d] = buf[j * D + d];
        }
    } else {
#pragma unroll
        for (int j = 0; j < S; ++j) {
            ok[j] = n0 + j < N;
#pragma unroll
            for (int d = 0; d < D; ++d) z[j][d] = ok[j] ? eps[(n0 + j) * D + d] : (T)0.5;
        }
    }
}

// One tile = S samples per thread through all K layers (and back).  MASKED is the
// ragged last tile: samples past N get weight 0 (select, never multiply: a masked
// sample must not be able to leak NaN/Inf into the sums).
template <typename T, int D, int S, int ST, int NT, int TGT, bool BWD, bool MASKED>
__device__ __forceinline__ void elbo_tile(const T* __restrict__ eps, long long n0, long long N, int K,
                                          const T* __restrict__ lcf, const T* __restrict__ lcb,
                                          const T* __restrict__ tp, int ntp, T epsv,
                                          T* __restrict__ hist_t /* hist + tid */, uint32_t taddr,
                                          T* __restrict__ acc_w, int lane, int my_slot, bool my_own, T& sum_q,
                                          T& sum_p) {
    using M = Mth<T>;
    constexpr int V = grad_vals<D>();
    constexpr int RF = rec_f<D>(), RB = rec_b<D>();
    constexpr int SS = S - ST;          // samples whose history goes to shared memory
    const T LOG2PI = (T)1.8378770664093453;

    T z[S][D], lq[S], prod[S];
    bool ok[S];
    load_samples<T, D, S, MASKED>(eps, n0, N, z, ok);
#pragma unroll
    for (int j = 0; j < S; ++j) {
        T zz = (T)0;
#pragma unroll
        for (int d = 0; d < D; ++d) zz = M::fma(z[j][d], z[j][d], zz);
        lq[j] = (T)-0.5 * (D * LOG2PI + zz);          // get_data.py:111, first term; log-det subtracted below
        prod[j] = (T)1;
    }
    // ---- forward over the K layers (get_data.py:104-109) ----
    for (int k = 0; k < K; ++k) {
        T rf[RF];
#pragma unroll
        for (int i = 0; i < RF; ++i) rf[i] = lcf[k * RF + i];
        T ck = (T)0;
        if (!BWD) ck = lcb[k * RB + 2 * D];
        T tt[S];
#pragma unroll
        for (int j = 0; j < S; ++j) {
            T a = rf[D];
#pragma unroll
            for (int d = 0; d < D; ++d) a = M::fma(rf[d], z[j][d], a);
            const T t = M::tanh_pre(a);
            tt[j] = t;
#pragma unroll
            for (int d = 0; d < D; ++d) z[j][d] = M::fma(rf[D + 1 + d], t, z[j][d]);
            if (BWD) {
                if (j >= ST) hist_t[(k * SS + (j - ST)) * NT] = t;
            } else {
                const T hp = M::fma(-t, t, (T)1);
                const T x = M::fma(hp, ck, (T)1);
                prod[j] *= epsv + M::fabs(x);
            }
        }
        if (BWD && ST > 0) TmemVecT<ST == 0 ? 2 : ST>::st(taddr + k * ST, tt);
        if (!BWD && ((k % kLogEvery) == kLogEvery - 1 || k == K - 1)) {
#pragma unroll
            for (int j = 0; j < S; ++j) { lq[j] -= M::log(prod[j]); prod[j] = (T)1; }
        }
    }
    if (BWD && ST > 0) tmem_wait_st();
    // ---- target: log(eps + target(z_K)) and the reverse seed gz = -d log p / dz ----
    T gz[S][D];
#pragma unroll
    for (int j = 0; j < S; ++j) {
        T lp;
        Target<T, D, TGT, BWD>::eval(z[j], tp, ntp, epsv, lp, gz[j]);
        if (MASKED) {
            sum_p += ok[j] ? lp : (T)0;
            if (BWD) {
#pragma unroll
                for (int d = 0; d < D; ++d) gz[j][d] = ok[j] ? gz[j][d] : (T)0;
            }
        } else {
            sum_p += lp;
        }
    }
    if (BWD) {
        // ---- reverse sweep: restores z, accumulates log|det| and the gradient partials ----
        // The cross-lane reduction of layer k+1's partials (a chain of five dependent
        // shuffles) is issued inside iteration k, next to arithmetic it does not depend on,
        // so an in-order warp can overlap the shuffle latency with the layer math.
        // TMEM-resident history is fetched one layer ahead (the load issued while layer k+1 is processed is
        // awaited at the top of layer k).  The loop is unrolled by two so that the "previous layer" partials and
        // the prefetched activations alternate between two register sets instead of being copied.
        T va[V], vb[V], ta[ST > 0 ? ST : 1], tb[ST > 0 ? ST : 1];
#pragma unroll
        for (int i = 0; i < V; ++i) { va[i] = (T)0; vb[i] = (T)0; }
        if (ST > 0) TmemVecT<ST == 0 ? 2 : ST>::ld(taddr + (K - 1) * ST, ta);
        auto layer = [&](int k, T (&tc)[ST > 0 ? ST : 1], T (&tn)[ST > 0 ? ST : 1], T (&v)[V], T (&vp)[V]) {
            if (ST > 0) {
                tmem_wait_ld();
                if (k > 0) TmemVecT<ST == 0 ? 2 : ST>::ld(taddr + (k - 1) * ST, tn);
            }
            T rb[RB];
#pragma unroll
            for (int i = 0; i < RB; ++i) rb[i] = lcb[k * RB + i];
            const T ck = rb[2 * D], c2 = rb[2 * D + 1];
#pragma unroll
            for (int i = 0; i < V; ++i) v[i] = (T)0;
#pragma unroll
            for (int j = 0; j < S; ++j) {
                const T t = (j < ST) ? tc[j < ST ? j : 0] : hist_t[(k * SS + (j - ST)) * NT];
#pragma unroll
                for (int d = 0; d < D; ++d) z[j][d] = M::fma(-rb[D + d], t, z[j][d]);   // z_{k-1}
                const T hp = M::fma(-t, t, (T)1);
                const T x = M::fma(hp, ck, (T)1);
                const T y = epsv + M::fabs(x);
                prod[j] *= y;
                T gx = M::neg_copysign(M::rcp(y), x);          // d(-log y)/dx
                if (MASKED) gx = ok[j] ? gx : (T)0;
                v[2 * D + 1] = M::fma(gx, hp, v[2 * D + 1]);    // g_c
                T gt = (gx * c2) * t;                           // via h' = 1 - t^2
#pragma unroll
                for (int d = 0; d < D; ++d) {
                    v[D + d] = M::fma(gz[j][d], t, v[D + d]);   // g_s
                    gt = M::fma(gz[j][d], rb[D + d], gt);
                }
                const T ga = gt * hp;
                v[2 * D] += ga;                                 // g_b
#pragma unroll
                for (int d = 0; d < D; ++d) {
                    v[d] = M::fma(ga, z[j][d], v[d]);           // g_w (direct)
                    gz[j][d] = M::fma(ga, rb[d], gz[j][d]);
                }
            }
            if ((k % kLogEvery) == 0) {
#pragma unroll
                for (int j = 0; j < S; ++j) { lq[j] -= M::log(prod[j]); prod[j] = (T)1; }
            }
            // cross-lane reduction of layer k+1's partials, issued next to arithmetic it does not depend on
            if (k < K - 1) {
                warp_reduce_multi<T, V>(vp, lane);
                if (my_own) acc_w[(k + 1) * V + my_slot] += vp[0];
#pragma unroll
                for (int jx = 1; jx < WarpSlotTable<V>::KEEP; ++jx)      // V > 32 (D = 16): a second value per lane
                    if (warp_slot_owner<V>(lane, jx)) acc_w[(k + 1) * V + warp_slot<V>(lane, jx)] += vp[jx];
            }
        };
        int k = K - 1;
        for (; k >= 1; k -= 2) {
            layer(k, ta, tb, va, vb);
            layer(k - 1, tb, ta, vb, va);
        }
        T (&vlast)[V] = (k == 0) ? va : vb;      // where layer 0's partials end up (odd K: the tail call below)
        if (k == 0) layer(0, ta, tb, va, vb);
        T vp[V];
#pragma unroll
        for (int i = 0; i < V; ++i) vp[i] = vlast[i];
        warp_reduce_multi<T, V>(vp, lane);
        if (my_own) acc_w[my_slot] += vp[0];
#pragma unroll
        for (int jx = 1; jx < WarpSlotTable<V>::KEEP; ++jx)
            if (warp_slot_owner<V>(lane, jx)) acc_w[warp_slot<V>(lane, jx)] += vp[jx];
    }
#pragma unroll
    for (int j = 0; j < S; ++j) sum_q += (MASKED && !ok[j]) ? (T)0 : lq[j];
}

template <typename T, int D, int S, int ST, int NT, int TGT, bool BWD>
__global__ void __launch_bounds__(NT)
elbo_shared_kernel(const T* __restrict__ eps, long long N, int K,
                   const T* __restrict__ w, const T* __restrict__ s,
                   const T* __restrict__ b, const T* __restrict__ c,
                   const T* __restrict__ tparams, int ntp, T epsv,
                   T* __restrict__ partials /* [gridDim.x][K*V + 2] */) {
    using M = Mth<T>;
    constexpr int V = grad_vals<D>();
    constexpr int RF = rec_f<D>(), RB = rec_b<D>();
    static_assert(ST == 0 || (NT == 128 && BWD), "the TMEM stash needs one warp per TMEM quadrant");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* sm = reinterpret_cast<T*>(smem_raw);
    const int ntp_s = target_smem_count(TGT, D, ntp);
    const ElboSmem<T> L(K, D, S - ST, ntp_s, BWD, NT);
    T* lcf = sm + L.lcf;
    T* lcb = sm + L.lcb;
    T* tp = sm + L.tp;
    T* acc = sm + L.acc;      // [NW][K][V]
    T* hist = sm + L.hist;    // [K][S][NT]

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // programmatic dependent launch: this grid may start while the previous kernel on the stream drains;
    // nothing produced by earlier work (constants, the workspace rows a row kernel is still reading) is
    // touched before this returns
    asm volatile("griddepcontrol.wait;" ::: "memory");

    __shared__ uint32_t tmem_base_smem;
    const int tmem_cols = tmem_columns(ST * K);
    if (ST > 0 && warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                         (uint32_t)__cvta_generic_to_shared(&tmem_base_smem)), "r"(tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }

    for (int k = tid; k < K; k += NT) {
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const T wv = w[k * D + d], sv = s[k * D + d];
            lcf[k * RF + d] = wv * M::kArgScale;
            lcf[k * RF + D + 1 + d] = sv;
            lcb[k * RB + d] = wv;
            lcb[k * RB + D + d] = sv;
        }
        lcf[k * RF + D] = b[k] * M::kArgScale;
        lcb[k * RB + 2 * D] = c[k];
        lcb[k * RB + 2 * D + 1] = (T)-2 * c[k];
    }
    target_prepare<T>(TGT, D, tparams, ntp, tp);
    if (BWD)
        for (int i = tid; i < (NT / kWarp) * K * V; i += NT) acc[i] = (T)0;
    if (ST > 0) asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    uint32_t taddr = 0;
    if (ST > 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        taddr = tmem_base_smem + ((uint32_t)(warp * 32) << 16);     // lane field: this warp's quadrant
    }

    const int my_slot = warp_slot<V>(lane);
    const bool my_own = warp_slot_owner<V>(lane);
    T* acc_w = acc + warp * K * V;

    T sum_q = (T)0, sum_p = (T)0;
    const long long tile_samples = (long long)NT * S;
    const long long ntiles = (N + tile_samples - 1) / tile_samples;

    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const long long base = tile * tile_samples;
        const long long n0 = base + (long long)tid * S;
        if (base + tile_samples <= N)
            elbo_tile<T, D, S, ST, NT, TGT, BWD, false>(eps, n0, N, K, lcf, lcb, tp, ntp, epsv, hist + tid, taddr, acc_w,
                                                        lane, my_slot, my_own, sum_q, sum_p);
        else
            elbo_tile<T, D, S, ST, NT, TGT, BWD, true>(eps, n0, N, K, lcf, lcb, tp, ntp, epsv, hist + tid, taddr, acc_w,
                                                       lane, my_slot, my_own, sum_q, sum_p);
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
    if (ST > 0) {
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        if (warp == 0)
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base_smem), "r"(tmem_cols) : "memory");
    }
}

// Adds the per-block (or per-warp) rows in a fixed order: one warp per output value, lanes
// stride over rows, then a shuffle tree.  Scatters into the caller's arrays.  A row is
// [K][Vk] gradient partials followed by (sum log q_K, sum log p); inside a layer record
// g_w starts at 0, g_s at off_s, (g_b, g_c) at off_b.
template <typename T>
__global__ void elbo_finalize_kernel(const T* __restrict__ partials, int nrows, int K, int D, int Vk, int off_s,
                                     int off_b, int bwd, T* __restrict__ out_F, T* __restrict__ g_w,
                                     T* __restrict__ g_s, T* __restrict__ g_b, T* __restrict__ g_c) {
    asm volatile("griddepcontrol.wait;" ::: "memory");     // see elbo_shared_kernel
    const size_t row = (size_t)K * Vk + 2;
    const int lane = threadIdx.x & 31;
    const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nw = (gridDim.x * blockDim.x) >> 5;
    for (size_t i = gw; i < row; i += nw) {
        const bool tail = i >= (size_t)K * Vk;
        int k = 0, j = 0;
        if (!tail) {
            if (!bwd) continue;
            k = (int)(i / Vk);
            j = (int)(i - (size_t)k * Vk);
            const bool used = j < D || (j >= off_s && j < off_s + D) || j == off_b || j == off_b + 1;
            if (!used) continue;
        }
        T a = (T)0;
        for (int r = lane; r < nrows; r += 32) a += partials[(size_t)r * row + i];
        a = warp_sum(a);
        if (lane == 0) {
            if (tail) out_F[1 + (int)(i - (size_t)K * Vk)] = a;
            else if (j < D) g_w[k * D + j] = a;
            else if (j < off_s + D && j >= off_s) g_s[k * D + (j - off_s)] = a;
            else if (j == off_b) g_b[k] = a;
            else g_c[k] = a;
        }
    }
    // F = sum log q_K - sum log p : recomputed by one warp from the same rows (same order => same bits)
    if (gw == 0) {
        T q = (T)0, p = (T)0;
        for (int r = lane; r < nrows; r += 32) {
            q += partials[(size_t)r * row + (size_t)K * Vk];
            p += partials[(size_t)r * row + (size_t)K * Vk + 1];
        }
        q = warp_sum(q);
        p = warp_sum(p);
        if (lane == 0) out_F[0] = q - p;
    }
}

template <typename T, int D, int S, int ST, int NT, int TGT, bool BWD>
static int launch_elbo(const T* eps, long long N, int K, const T* w, const T* s, const T* b, const T* c,
                       const T* tparams, int ntp, T epsv, T* partials, int max_blocks, int* nblocks_out,
                       cudaStream_t stream) {
    auto kern = elbo_shared_kernel<T, D, S, ST, NT, TGT, BWD>;
    const int ntp_s = target_smem_count(TGT, D, ntp);
    const ElboSmem<T> L(K, D, S - ST, ntp_s, BWD, NT);
    size_t smem = (size_t)L.total * sizeof(T);
    if (smem > 227 * 1024) return NF_ERR_UNSUPPORTED;
    int tmem_limit = kMaxBlocksPerSM;
    if (ST > 0) {
        // at most 512 / columns blocks can hold their TMEM slice at once; pad the shared-memory request so
        // that the hardware never co-schedules more than that on one SM (an extra block would spin in alloc)
        if (ST * K > 512) return NF_ERR_UNSUPPORTED;
        tmem_limit = 512 / tmem_columns(ST * K);
        const size_t pad = (size_t)(227 * 1024) / (tmem_limit + 1) + 1024;
        if (smem < pad) smem = pad;
    }
    NF_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    if (ST > 0) {
        // the occupancy calculator reports 1 block/SM for any kernel that allocates TMEM; work the limits out here
        cudaFuncAttributes fa;
        NF_CUDA_OK(cudaFuncGetAttributes(&fa, kern));
        const int by_regs = 65536 / (((fa.numRegs + 7) & ~7) * NT);
        const int by_smem = (int)((size_t)(227 * 1024) / (smem + fa.sharedSizeBytes + 1024));
        per_sm = by_regs < by_smem ? by_regs : by_smem;
    } else {
        NF_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, NT, smem));
    }
    if (per_sm < 1) return NF_ERR_UNSUPPORTED;
    if (per_sm > kMaxBlocksPerSM) per_sm = kMaxBlocksPerSM;
    if (per_sm > tmem_limit) per_sm = tmem_limit;
    const DevInfo di = dev_info();
    const long long tile = (long long)NT * S;
    long long ntiles = (N + tile - 1) / tile;
    long long grid = (long long)di.sms * per_sm;
    if (grid > ntiles) grid = ntiles;
    if (grid < 1) grid = 1;
    if (grid > max_blocks) grid = max_blocks;
    if (getenv("NF_DEBUG"))
        fprintf(stderr, "[nfcore] elbo launch S=%d ST=%d NT=%d smem=%zu per_sm=%d tmem_limit=%d grid=%lld ntiles=%lld\n", S, ST, NT,
                smem, per_sm, tmem_limit, grid, ntiles);
    NF_CUDA_OK(launch_pdl(kern, dim3((unsigned)grid), dim3(NT), smem, stream, eps, N, K, w, s, b, c, tparams, ntp, epsv, partials));
    *nblocks_out = (int)grid;
    return NF_OK;
}

// ---- packed-pair kernel (nf_elbo_pair.cuh, built in its own translation units): float32, D <= 2, K <= 64 ----
bool pair_applicable(size_t elem, bool bwd, long long N, int D, int K);
// rows only (sync == NULL: `*nb` rows of partial sums for elbo_finish) or the whole call in one launch
int pair_elbo(int tgt, const float* eps, long long N, int D, int K, const float* w, const float* s, const float* b,
              const float* c, const float* tp, int ntp, float epsv, float* partials, pr::SyncHeader* sync,
              float* out_F, float* g_w, float* g_s, float* g_b, float* g_c, int* nb, cudaStream_t st, const pr::ArDesc& ar,
              const RngDesc& rng, void* stash, size_t stash_bytes);
size_t pair_stash_bytes_max(int K);        // room for the L2-resident part of the tanh history (nf_elbo_pair_launch.cuh)
static inline RngDesc no_rng() { RngDesc r = {0, 0}; return r; }
static inline pr::ArDesc no_exchange() { pr::ArDesc a = {}; return a; }
// nf_allreduce.cu
int launch_allreduce_small(const pr::ArDesc& ar, const float* src, float* dst, int n, int* status_i, cudaStream_t stream);
int make_ar_desc(pr::ArDesc* ar, void* const* peer_buffers, int world, int rank, uint64_t seq, int64_t n_local, float* xout);

template <typename T> struct SamplesPerThread { static constexpr int big = 8, mid = 4, small = 2; };
template <> struct SamplesPerThread<double> { static constexpr int big = 4, mid = 2, small = 1; };

template <typename T, int D, int TGT>
static int dispatch_bwd(bool bwd, const T* eps, long long N, int K, const T* w, const T* s, const T* b, const T* c,
                        const T* tp, int ntp, T epsv, T* partials, int max_blocks, int* nb, cudaStream_t st) {
    constexpr int NT = kElboThreads;
    // D = 16 keeps 2 x 16 coordinates of z and of the gradient per sample in registers: two samples per thread at most
    constexpr int SB = D <= 4 ? SamplesPerThread<T>::big : 2, SM_ = D <= 4 ? SamplesPerThread<T>::mid : (sizeof(T) == 4 ? 2 : 1),
                  SS = D <= 4 ? SamplesPerThread<T>::small : 1;
#define NF_GO(S_, ST_, NT_, BW_) \
    return launch_elbo<T, D, S_, ST_, NT_, TGT, BW_>(eps, N, K, w, s, b, c, tp, ntp, epsv, partials, max_blocks, nb, st)
    if (!bwd) {
        if (N >= (long long)148 * NT * SB) { NF_GO(SB, 0, NT, false); }
        NF_GO(SS, 0, NT, false);
    }
    // tuning knob (float, D <= 2): NF_ELBO_CFG="S,ST" picks samples per thread and how many of them keep
    // their history in TMEM
    int cs = 0, ct = 0;
    if (const char* ov = getenv("NF_ELBO_CFG")) sscanf(ov, "%d,%d", &cs, &ct);
    if constexpr (sizeof(T) == 4 && D <= 2) {
        if (cs == 8 && ct == 8) { NF_GO(8, 8, 128, true); }
        if (cs == 8 && ct == 4) { NF_GO(8, 4, 128, true); }
        if (cs == 7 && ct == 4) { NF_GO(7, 4, 128, true); }
        if (cs == 6 && ct == 4) { NF_GO(6, 4, 128, true); }
        if (cs == 5 && ct == 4) { NF_GO(5, 4, 128, true); }
        if (cs == 4 && ct == 4) { NF_GO(4, 4, 128, true); }
        if (cs == 4 && ct == 2) { NF_GO(4, 2, 128, true); }
        if (cs == 8 && ct == 0) { NF_GO(8, 0, NT, true); }
        if (cs == 4 && ct == 0) { NF_GO(4, 0, NT, true); }
        if (cs == 2 && ct == 0) { NF_GO(2, 0, NT, true); }
    }
    // float, D <= 2, enough samples: 128-thread blocks whose tanh history is split between tensor memory and
    // shared memory (measured best on B200: profiles/r01_sweep_configs.txt)
    if constexpr (sizeof(T) == 4 && D <= 2) {
        if (cs == 0 && N >= (long long)148 * 128 * 8) {
            if (K <= 16) { NF_GO(8, 8, 128, true); }
            if (K <= 32) { NF_GO(7, 4, 128, true); }
            if (K <= 64) { NF_GO(7, 4, 128, true); }     // 256 columns, 96 KB: two blocks per SM
        }
    }
    // otherwise: as many samples per thread as keep >= 3 blocks of activation history resident per SM;
    // short launches (latency-bound) and very deep chains fall back to smaller tiles
    const int ntp_s = target_smem_count(TGT, D, ntp);
    const bool enough = N >= (long long)148 * NT * SM_;
    const ElboSmem<T> Lm(K, D, SM_, ntp_s, true, NT);
    if (enough && (size_t)Lm.total * sizeof(T) <= 75 * 1024) { NF_GO(SM_, 0, NT, true); }
    const ElboSmem<T> Ls(K, D, SS, ntp_s, true, NT);
    if ((size_t)Ls.total * sizeof(T) <= 113 * 1024) { NF_GO(SS, 0, NT, true); }
    NF_GO(1, 0, NT, true);
#undef NF_GO
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
    if constexpr (D <= 4) {
        if (tgt == NF_TARGET_MIXFULL) return dispatch_bwd<T, D, NF_TARGET_MIXFULL>(bwd, eps, N, K, w, s, b, c, tp, ntp, epsv, partials, max_blocks, nb, st);
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

// Rows of the wide kernel live in global memory and are read-modified-written once per layer: keep all of
// them L2-resident (126 MB) -- at most ~96 MB of rows, between 1 and 8 blocks per SM.
static inline int wide_blocks_per_sm(int K, int D, size_t elem) {
    const DevInfo di = dev_info();
    const size_t rowbytes = ((size_t)K * wide_vk(D) + 2) * elem;
    const size_t per_block = rowbytes * kWideWarps;
    long long bps = (long long)((size_t)96000000 / ((size_t)(di.sms > 0 ? di.sms : 148) * per_block));
    if (bps < 1) bps = 1;
    if (bps > 8) bps = 8;
    return (int)bps;
}

template <typename T, int C, int L, int TGT, bool BWD>
static int launch_wide(const T* eps, long long N, int D, int K, const T* w, const T* s, const T* b, const T* c,
                       const T* tparams, int ntp, T epsv, T* partials, int* nrows_out, cudaStream_t stream) {
    constexpr int S = (L >= 32 && C == 8) ? 6 : ((L >= 8) ? 4 : 2);      // more interleaved sample groups where a warp holds few samples
    auto kern = elbo_wide_kernel<T, C, L, S, TGT, BWD>;
    const size_t smem = BWD ? (size_t)K * S * kWideWarps * (32 / L) * sizeof(T) : 0;
    if (smem > 200 * 1024) return NF_ERR_UNSUPPORTED;
    NF_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const DevInfo di = dev_info();
    const long long tile = (long long)kWideWarps * S * (32 / L);
    int bps = wide_blocks_per_sm(K, D, sizeof(T));
    int occ = 0;
    NF_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kWideThreads, smem));
    if (occ < 1) return NF_ERR_UNSUPPORTED;
    if (bps > occ) bps = occ;
    long long grid = (long long)di.sms * bps;
    const long long ntiles = (N + tile - 1) / tile;
    if (grid > ntiles) grid = ntiles;
    if (grid < 1) grid = 1;
    kern<<<(int)grid, kWideThreads, smem, stream>>>(eps, N, D, K, w, s, b, c, tparams, ntp, epsv, partials);
    NF_CUDA_OK(cudaGetLastError());
    *nrows_out = (int)grid * kWideWarps;
    return NF_OK;
}

template <typename T, int TGT>
static int dispatch_wide(bool bwd, const T* eps, long long N, int D, int K, const T* w, const T* s, const T* b,
                         const T* c, const T* tp, int ntp, T epsv, T* partials, int* nrows, cudaStream_t st) {
#define NF_WIDE(C, L)                                                                                          \
    return bwd ? launch_wide<T, C, L, TGT, true>(eps, N, D, K, w, s, b, c, tp, ntp, epsv, partials, nrows, st) \
               : launch_wide<T, C, L, TGT, false>(eps, N, D, K, w, s, b, c, tp, ntp, epsv, partials, nrows, st)
    if (D <= 16) { NF_WIDE(8, 2); }
    if (D <= 32) { NF_WIDE(8, 4); }
    if (D <= 64) { NF_WIDE(8, 8); }
    if (D <= 128) { NF_WIDE(8, 16); }
    if (D <= 256) { NF_WIDE(8, 32); }
    if (D <= 512) { NF_WIDE(16, 32); }
#undef NF_WIDE
    return NF_ERR_UNSUPPORTED;
}

static inline int max_grid_blocks() {
    const DevInfo di = dev_info();
    return (di.sms > 0 ? di.sms : 148) * kMaxBlocksPerSM;
}

struct ElboRowLayout { int Vk, off_s, off_b; };
// D <= 4 and D == 16 run one thread per sample (elbo_shared_kernel); other D share a sample among lanes
static inline bool elbo_thread_per_sample(int D) { return D <= 4 || D == 16; }

static inline ElboRowLayout elbo_row_layout(int D) {
    if (elbo_thread_per_sample(D)) return {2 * D + 2, D, 2 * D};
    const int Dp = (D + 3) & ~3;
    return {wide_vk(D), Dp, 2 * Dp};
}

template <typename T>
static int elbo_validate(const T* eps, int64_t N, int D, int K, const T* w, const T* s, const T* b, const T* c,
                         int target_id, const T* tparams, int ntp) {
    if (!eps && N > 0) return NF_ERR_INVALID;
    if ((reinterpret_cast<uintptr_t>(eps) & 15) != 0) return NF_ERR_INVALID;   // vector loads
    if (N < 0 || D < 1 || K < 1 || !w || !s || !b || !c) return NF_ERR_INVALID;
    if (target_id == NF_TARGET_MIXDIAG) {
        if (!tparams || ntp < 1 + 2 * D || ntp % (1 + 2 * D) != 0) return NF_ERR_INVALID;
    } else if (target_id == NF_TARGET_GAUSSDIAG) {
        if (!tparams || ntp != 2 * D) return NF_ERR_INVALID;
    } else if (target_id == NF_TARGET_MIXFULL) {
        if (!tparams || ntp < 1 + D + D * D || ntp % (1 + D + D * D) != 0) return NF_ERR_INVALID;
        if (D > 4) return NF_ERR_UNSUPPORTED;
    } else if (target_id < 0 || target_id > NF_TARGET_MIXFULL) {
        return NF_ERR_INVALID;
    } else if (D != 2) {
        return NF_ERR_INVALID;      // the 2-D potentials index z[:,0], z[:,1] (get_data.py:29)
    }
    const DevInfo di = dev_info();
    if (!di.ok) return NF_ERR_NO_DEVICE;
    return NF_OK;
}

// objective kernel only: fills `partials` rows [0, *nrows)
template <typename T>
static int elbo_main(bool bwd, const T* eps, int64_t N, int D, int K, const T* w, const T* s, const T* b, const T* c,
                     int target_id, const T* tparams, int ntp, T epsv, T* partials, int* nrows, cudaStream_t stream) {
    const int max_blocks = max_grid_blocks();
    if constexpr (sizeof(T) == 4) {
        if (pair_applicable(sizeof(T), bwd, N, D, K))
            return pair_elbo(target_id, eps, N, D, K, w, s, b, c, tparams, ntp, epsv, partials, nullptr, nullptr, nullptr, nullptr,
                             nullptr, nullptr, nrows, stream, no_exchange(), no_rng(), nullptr, 0);
    }
    switch (D) {
        case 1: return dispatch_target<T, 1>(target_id, bwd, eps, N, K, w, s, b, c, tparams, ntp, epsv, partials, max_blocks, nrows, stream);
        case 2: return dispatch_target<T, 2>(target_id, bwd, eps, N, K, w, s, b, c, tparams, ntp, epsv, partials, max_blocks, nrows, stream);
        case 3: return dispatch_target<T, 3>(target_id, bwd, eps, N, K, w, s, b, c, tparams, ntp, epsv, partials, max_blocks, nrows, stream);
        case 4: return dispatch_target<T, 4>(target_id, bwd, eps, N, K, w, s, b, c, tparams, ntp, epsv, partials, max_blocks, nrows, stream);
        case 16: return dispatch_target<T, 16>(target_id, bwd, eps, N, K, w, s, b, c, tparams, ntp, epsv, partials, max_blocks, nrows, stream);
        default:
            if (target_id == NF_TARGET_GAUSSDIAG)
                return dispatch_wide<T, NF_TARGET_GAUSSDIAG>(bwd, eps, N, D, K, w, s, b, c, tparams, ntp, epsv, partials, nrows, stream);
            return dispatch_wide<T, NF_TARGET_MIXDIAG>(bwd, eps, N, D, K, w, s, b, c, tparams, ntp, epsv, partials, nrows, stream);
    }
}

// fixed-order sum of `nrows` partial rows into the caller's outputs
template <typename T>
static int elbo_finish(bool bwd, const T* partials, int nrows, int D, int K, T* out_F, T* g_w, T* g_s, T* g_b, T* g_c,
                       cudaStream_t stream) {
    const ElboRowLayout L = elbo_row_layout(D);
    const size_t row = (size_t)K * L.Vk + 2;
    size_t fblocks = (row + 7) / 8;
    if (fblocks > 592) fblocks = 592;
    NF_CUDA_OK(launch_pdl(elbo_finalize_kernel<T>, dim3((unsigned)fblocks), dim3(256), 0, stream, partials, nrows, K, D, L.Vk,
                          L.off_s, L.off_b, bwd ? 1 : 0, out_F, g_w, g_s, g_b, g_c));
    return NF_OK;
}

// objective + fixed-order sum into the caller's outputs.  `partials` is a workspace of nf_planar_elbo_workspace_bytes():
// its last kSyncHeaderBytes hold the ticket of the single-launch kernel.
template <typename T>
static int elbo_fused(bool bwd, const T* eps, int64_t N, int D, int K, const T* w, const T* s, const T* b, const T* c,
                      int target_id, const T* tparams, int ntp, T epsv, T* out_F, T* g_w, T* g_s, T* g_b, T* g_c, T* partials,
                      cudaStream_t stream, const pr::ArDesc* ar = nullptr, bool* exchanged = nullptr, const RngDesc* rng = nullptr,
                      size_t ws_bytes = 0 /* size of the workspace `partials` points to, when known: what lies beyond
                                             nf_planar_elbo_workspace_bytes() may hold part of the tanh history */) {
    int nb = 0;
    if (exchanged) *exchanged = false;
    if constexpr (sizeof(T) == 4) {
        // in-kernel draws (eps == NULL, rng given) exist in the packed-pair kernel only: it then serves every N
        if (rng && !(bwd && D <= 2 && K <= 64)) return NF_ERR_UNSUPPORTED;
        if (rng || pair_applicable(sizeof(T), bwd, N, D, K)) {
            unsigned char* base = reinterpret_cast<unsigned char*>(partials);
            pr::SyncHeader* sync = reinterpret_cast<pr::SyncHeader*>(base + nf_planar_elbo_workspace_bytes(K, D, 4) - kSyncHeaderBytes);
            if (exchanged) *exchanged = ar != nullptr;        // the last block also does the exchange
            const size_t min_bytes = (nf_planar_elbo_workspace_bytes(K, D, 4) + 255) & ~(size_t)255;
            void* stash = ws_bytes > min_bytes ? base + min_bytes : nullptr;
            return pair_elbo(target_id, eps, N, D, K, w, s, b, c, tparams, ntp, epsv, partials, sync, out_F, g_w, g_s, g_b, g_c, &nb, stream,
                             ar ? *ar : no_exchange(), rng ? *rng : no_rng(), stash, stash ? ws_bytes - min_bytes : 0);
        }
    }
    int rc = elbo_main<T>(bwd, eps, N, D, K, w, s, b, c, target_id, tparams, ntp, epsv, partials, &nb, stream);
    if (rc != NF_OK) return rc;
    return elbo_finish<T>(bwd, partials, nb, D, K, out_F, g_w, g_s, g_b, g_c, stream);
}

template <typename T>
static int elbo_entry(const T* eps, int64_t N, int D, int K, const T* w, const T* s, const T* b, const T* c,
                      int target_id, const T* tparams, int ntp, T epsv, T* out_F, T* g_w, T* g_s, T* g_b, T* g_c,
                      void* workspace, size_t workspace_bytes, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!out_F || !workspace) return NF_ERR_INVALID;
    const bool any_g = g_w || g_s || g_b || g_c;
    const bool all_g = g_w && g_s && g_b && g_c;
    if (any_g && !all_g) return NF_ERR_INVALID;
    int rc = elbo_validate<T>(eps, N, D, K, w, s, b, c, target_id, tparams, ntp);
    if (rc != NF_OK) return rc;
    if (workspace_bytes < nf_planar_elbo_workspace_bytes(K, D, (int)sizeof(T))) return NF_ERR_WORKSPACE;
    T* partials = reinterpret_cast<T*>(workspace);
    return elbo_fused<T>(all_g, eps, N, D, K, w, s, b, c, target_id, tparams, ntp, epsv, out_F, g_w, g_s, g_b, g_c, partials, stream,
                         nullptr, nullptr, nullptr, workspace_bytes);
}

// ---- host-buffer session: H2D of the samples overlapped with the objective kernel -------------
// The reference's objective takes host arrays (get_data.py:97: params and the freshly drawn z0 live
// in NumPy).  A session owns the device-side staging for that call shape: the samples are sent in
// chunks on a copy stream while the compute stream runs the objective kernel on the chunks that
// have landed; all chunks write partial-sum rows into one workspace and a single row kernel sums them.
// nf_elbo_gram.cu: the matrix-product formulation for wide latent spaces (float32 only)
bool gram_applicable(long long N, int D, int K, int target_id, int ntp);      // target_id < 0: any supported target
size_t gram_workspace_bytes(long long N, int D, int K);
int gram_elbo_f32(bool bwd, const float* eps, long long N, int D, int K, const float* w, const float* s, const float* b,
                  const float* c, int target_id, const float* tp, int ntp, float epsv, float* out_F, float* g_w, float* g_s,
                  float* g_b, float* g_c, void* workspace, cudaStream_t st);
long long gram_chunk_pitch(long long layout_n, int D, int K);
int gram_begin(long long layout_n, int D, int K, const float* w, const float* s, void* workspace, cudaStream_t st);
int gram_piece(bool bwd, bool first, const float* z0, int n_act, long long layout_n, int D, int K, const float* w, const float* s,
               const float* b, const float* c, int target_id, const float* tp, int ntp, float epsv, void* workspace, cudaStream_t st);
int gram_end(bool bwd, long long layout_n, int D, int K, const float* w, const float* s, float* out_F, float* g_w, float* g_s,
             float* g_b, float* g_c, void* workspace, cudaStream_t st);

constexpr int kSessionMaxChunks = 8;
constexpr int64_t kSessionChunkSamples = 1 << 18;

struct Session {
    int dev = 0, D = 0, K = 0;
    size_t elem = 0;
    int64_t cap = 0;
    cudaStream_t s_copy = nullptr, s_comp = nullptr;
    cudaEvent_t ev[kSessionMaxChunks] = {};
    void *eps_dev = nullptr, *lc_dev = nullptr, *tp_dev = nullptr, *out_dev = nullptr, *ws_dev = nullptr;
    void *lc_pin = nullptr, *out_pin = nullptr;
    size_t ws_chunk_bytes = 0, tp_cap = 0;
    int n_out = 0;
    void* gram_ws = nullptr;          // float32, wide D: workspace of the matrix-product path, sized for gram_layout_n samples
    long long gram_layout_n = 0;
};

static inline void session_free(Session* S) {
    if (!S) return;
    cudaSetDevice(S->dev);
    if (S->s_comp) cudaStreamSynchronize(S->s_comp);
    for (auto& e : S->ev) if (e) cudaEventDestroy(e);
    if (S->s_copy) cudaStreamDestroy(S->s_copy);
    if (S->s_comp) cudaStreamDestroy(S->s_comp);
    cudaFree(S->eps_dev); cudaFree(S->lc_dev); cudaFree(S->tp_dev); cudaFree(S->out_dev); cudaFree(S->ws_dev); cudaFree(S->gram_ws);
    cudaFreeHost(S->lc_pin); cudaFreeHost(S->out_pin);
    delete S;
}

static inline int session_create(Session** out, int64_t max_samples, int D, int K, int elem_size) {
    if (!out || max_samples < 1 || D < 1 || K < 1 || (elem_size != 4 && elem_size != 8)) return NF_ERR_INVALID;
    const DevInfo di = dev_info();
    if (!di.ok) return NF_ERR_NO_DEVICE;
    Session* S = new Session();
    cudaGetDevice(&S->dev);
    S->D = D; S->K = K; S->elem = (size_t)elem_size; S->cap = max_samples;
    S->n_out = 3 + 2 * K * D + 2 * K;
    S->ws_chunk_bytes = nf_planar_elbo_workspace_bytes(K, D, elem_size);
    S->tp_cap = 4096;
    cudaError_t e = cudaSuccess;
    auto ok = [&](cudaError_t r) { if (e == cudaSuccess && r != cudaSuccess) e = r; };
    ok(cudaStreamCreateWithFlags(&S->s_copy, cudaStreamNonBlocking));
    ok(cudaStreamCreateWithFlags(&S->s_comp, cudaStreamNonBlocking));
    for (auto& ev : S->ev) ok(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    ok(cudaMalloc(&S->eps_dev, (size_t)max_samples * D * S->elem));
    ok(cudaMalloc(&S->lc_dev, (size_t)(2 * K * D + 2 * K) * S->elem));
    ok(cudaMalloc(&S->tp_dev, S->tp_cap * S->elem));
    ok(cudaMalloc(&S->out_dev, (size_t)(S->n_out + 4) * S->elem));      // + (n_hi, n_lo, status) of the multi-GPU exchange
    ok(cudaMalloc(&S->ws_dev, S->ws_chunk_bytes * kSessionMaxChunks));
    S->gram_layout_n = max_samples < kSessionChunkSamples ? max_samples : kSessionChunkSamples;
    if (elem_size == 4 && gram_applicable(1 << 30, D, K, -1, 0))        // shape supported (the per-call cost model decides later)
        ok(cudaMalloc(&S->gram_ws, gram_workspace_bytes(S->gram_layout_n, D, K)));
    ok(cudaMallocHost(&S->lc_pin, (size_t)(2 * K * D + 2 * K + S->tp_cap) * S->elem));
    ok(cudaMallocHost(&S->out_pin, (size_t)(S->n_out + 4) * S->elem));
    if (e != cudaSuccess) { nf_set_last_cuda_error(e); session_free(S); return NF_ERR_CUDA; }
    *out = S;
    return NF_OK;
}

template <typename T>
static int session_elbo_host(Session* S, const T* eps_host, int64_t N, const T* lc_host, int target_id,
                             const T* tp_host, int ntp, T epsv, int backward, T* out_host,
                             const pr::ArDesc* ar = nullptr /* float32: sum the result over the ranks of one box (nf_allreduce.cu) */,
                             T* tail_host = nullptr /* with ar: (n_hi, n_lo, status) as nf_allreduce_small_counted_f32 leaves them */) {
    if (!S || S->elem != sizeof(T) || N < 0 || N > S->cap || !lc_host || !out_host || (N > 0 && !eps_host)) return NF_ERR_INVALID;
    if (ntp < 0 || (size_t)ntp > S->tp_cap || (ntp > 0 && !tp_host)) return NF_ERR_INVALID;
    static const bool timing = getenv("NF_SESSION_TIMING") != nullptr;      // debug aid: host-side time stamps of this call
    double tq[8] = {0};
    auto now = []() { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3; };
    if (timing) tq[0] = now();
    NF_CUDA_OK(cudaSetDevice(S->dev));
    const int D = S->D, K = S->K;
    const int nlc = 2 * K * D + 2 * K;
    T* lc_pin = reinterpret_cast<T*>(S->lc_pin);
    for (int i = 0; i < nlc; ++i) lc_pin[i] = lc_host[i];
    for (int i = 0; i < ntp; ++i) lc_pin[nlc + i] = tp_host[i];
    T* lc = reinterpret_cast<T*>(S->lc_dev);
    T* tp = reinterpret_cast<T*>(S->tp_dev);
    NF_CUDA_OK(cudaMemcpyAsync(lc, lc_pin, (size_t)nlc * sizeof(T), cudaMemcpyHostToDevice, S->s_comp));
    if (ntp > 0) NF_CUDA_OK(cudaMemcpyAsync(tp, lc_pin + nlc, (size_t)ntp * sizeof(T), cudaMemcpyHostToDevice, S->s_comp));
    const T *w = lc, *s = lc + K * D, *b = lc + 2 * K * D, *c = lc + 2 * K * D + K;
    T* eps = reinterpret_cast<T*>(S->eps_dev);
    int rc = elbo_validate<T>(eps, N, D, K, w, s, b, c, target_id, ntp > 0 ? tp : nullptr, ntp);
    if (rc != NF_OK) return rc;
    if constexpr (sizeof(T) == 4) {
        if (S->gram_ws && N > 0 && gram_applicable(N, D, K, target_id, ntp)) {
            // wide D: pieces of at most one chunk pitch, each uploaded on the copy stream while the previous one is worked on
            const long long pitch = gram_chunk_pitch(S->gram_layout_n, D, K);
            int64_t pieces = (N + pitch - 1) / pitch;
            if (pieces < 4 && N >= 4 * 4096) pieces = 4;                       // a few pieces even when one would do, for the overlap
            const int64_t per = ((N + pieces - 1) / pieces + 511) & ~(int64_t)511;
            T* out = reinterpret_cast<T*>(S->out_dev);
            rc = gram_begin(S->gram_layout_n, D, K, w, s, S->gram_ws, S->s_comp);
            if (rc != NF_OK) return rc;
            int i = 0;
            for (int64_t lo = 0; lo < N; lo += per, ++i) {
                const int64_t n = (N - lo < per) ? N - lo : per;
                NF_CUDA_OK(cudaMemcpyAsync(eps + lo * D, eps_host + lo * D, (size_t)n * D * sizeof(T), cudaMemcpyHostToDevice, S->s_copy));
                NF_CUDA_OK(cudaEventRecord(S->ev[i % kSessionMaxChunks], S->s_copy));
                NF_CUDA_OK(cudaStreamWaitEvent(S->s_comp, S->ev[i % kSessionMaxChunks], 0));
                rc = gram_piece(backward != 0, i == 0, eps + lo * D, (int)n, S->gram_layout_n, D, K, w, s, b, c, target_id,
                                ntp > 0 ? tp : nullptr, ntp, epsv, S->gram_ws, S->s_comp);
                if (rc != NF_OK) return rc;
            }
            rc = gram_end(backward != 0, S->gram_layout_n, D, K, w, s, out, out + 3, out + 3 + K * D, out + 3 + 2 * K * D,
                          out + 3 + 2 * K * D + K, S->gram_ws, S->s_comp);
            if (rc != NF_OK) return rc;
            if (ar) {
                pr::ArDesc a2 = *ar;
                a2.xout = reinterpret_cast<float*>(out) + S->n_out;
                rc = launch_allreduce_small(a2, reinterpret_cast<float*>(out), reinterpret_cast<float*>(out), S->n_out, nullptr, S->s_comp);
                if (rc != NF_OK) return rc;
            }
            NF_CUDA_OK(cudaMemcpyAsync(S->out_pin, out, (size_t)(S->n_out + (ar ? 3 : 0)) * sizeof(T), cudaMemcpyDeviceToHost, S->s_comp));
            NF_CUDA_OK(cudaStreamSynchronize(S->s_comp));
            const T* op = reinterpret_cast<const T*>(S->out_pin);
            for (int j = 0; j < S->n_out; ++j) out_host[j] = op[j];
            if (ar && tail_host) for (int j = 0; j < 3; ++j) tail_host[j] = op[S->n_out + j];
            return NF_OK;
        }
    }
    // upload granularity: NF_SESSION_CHUNK=<samples> overrides (tuning knob, read once)
    static const int64_t chunk_samples = []() {
        const char* e = getenv("NF_SESSION_CHUNK");
        const long long v = e ? atoll(e) : 0;
        return (int64_t)(v >= 1024 ? v : kSessionChunkSamples);
    }();
    int nch = (int)((N + chunk_samples - 1) / chunk_samples);
    if (nch < 1) nch = 1;
    if (nch > kSessionMaxChunks) nch = kSessionMaxChunks;
    int64_t per = ((N + nch - 1) / nch + 1023) & ~(int64_t)1023;       // keeps every chunk 16-byte aligned
    const ElboRowLayout L = elbo_row_layout(D);
    const size_t row = (size_t)K * L.Vk + 2;
    T* partials = reinterpret_cast<T*>(S->ws_dev);
    int rows = 0;
    for (int i = 0; i < nch; ++i) {
        const int64_t lo = (int64_t)i * per;
        const int64_t n = (lo >= N) ? 0 : ((N - lo < per) ? N - lo : per);
        if (n <= 0 && i > 0) break;
        if (n > 0) NF_CUDA_OK(cudaMemcpyAsync(eps + lo * D, eps_host + lo * D, (size_t)n * D * sizeof(T), cudaMemcpyHostToDevice, S->s_copy));
        NF_CUDA_OK(cudaEventRecord(S->ev[i], S->s_copy));
        NF_CUDA_OK(cudaStreamWaitEvent(S->s_comp, S->ev[i], 0));
        int nb = 0;
        rc = elbo_main<T>(backward != 0, eps + lo * D, n, D, K, w, s, b, c, target_id, ntp > 0 ? tp : nullptr, ntp, epsv,
                          partials + (size_t)rows * row, &nb, S->s_comp);
        if (rc != NF_OK) return rc;
        rows += nb;
    }
    if (timing) tq[1] = now();
    T* out = reinterpret_cast<T*>(S->out_dev);
    rc = elbo_finish<T>(backward != 0, partials, rows, D, K, out, out + 3, out + 3 + K * D, out + 3 + 2 * K * D,
                        out + 3 + 2 * K * D + K, S->s_comp);
    if (rc != NF_OK) return rc;
    if constexpr (sizeof(T) == 4) {
        if (ar) {
            pr::ArDesc a2 = *ar;
            a2.xout = reinterpret_cast<float*>(out) + S->n_out;
            rc = launch_allreduce_small(a2, reinterpret_cast<float*>(out), reinterpret_cast<float*>(out), S->n_out, nullptr, S->s_comp);
            if (rc != NF_OK) return rc;
        }
    }
    NF_CUDA_OK(cudaMemcpyAsync(S->out_pin, out, (size_t)(S->n_out + (ar ? 3 : 0)) * sizeof(T), cudaMemcpyDeviceToHost, S->s_comp));
    if (timing) tq[2] = now();
    NF_CUDA_OK(cudaStreamSynchronize(S->s_comp));
    if (timing) tq[3] = now();
    const T* op = reinterpret_cast<const T*>(S->out_pin);
    for (int i = 0; i < S->n_out; ++i) out_host[i] = op[i];
    if (ar && tail_host) for (int i = 0; i < 3; ++i) tail_host[i] = op[S->n_out + i];
    if (timing) fprintf(stderr, "[nf session] N=%lld enqueue chunks %.1f us | finish+D2H enqueue %.1f | wait %.1f | total %.1f\n", (long long)N,
                        tq[1] - tq[0], tq[2] - tq[1], tq[3] - tq[2], now() - tq[0]);
    return NF_OK;
}

// The same with the samples drawn inside the kernel (nf_philox.cuh): nothing but the K x D constants goes up, nothing but
// F and the gradient comes back.  float32, D <= 2, and N large enough for the packed-pair kernel (the one that draws).
static int session_objective_rng(Session* S, uint64_t seed, uint64_t step, int64_t N, const double* params, int semantics,
                                 int target_id, const float* tp_host, int ntp, float epsv, double* F_parts, double* grad) {
    if (!S || S->elem != 4 || !params || !F_parts || !grad || N < 1) return NF_ERR_INVALID;
    if (ntp < 0 || (size_t)ntp > S->tp_cap || (ntp > 0 && !tp_host)) return NF_ERR_INVALID;
    const int D = S->D, K = S->K;
    if (D != 1 && D != 2) return NF_ERR_UNSUPPORTED;
    NF_CUDA_OK(cudaSetDevice(S->dev));
    const int nlc = 2 * K * D + 2 * K;
    static thread_local std::vector<double> lc64, g64;
    lc64.resize(nlc); g64.resize(nlc);
    int rc = nf_script_constants_host(params, K, D, semantics, lc64.data());
    if (rc != NF_OK) return rc;
    float* lc_pin = reinterpret_cast<float*>(S->lc_pin);
    for (int i = 0; i < nlc; ++i) lc_pin[i] = (float)lc64[i];
    for (int i = 0; i < ntp; ++i) lc_pin[nlc + i] = tp_host[i];
    float* lc = reinterpret_cast<float*>(S->lc_dev);
    float* tp = reinterpret_cast<float*>(S->tp_dev);
    NF_CUDA_OK(cudaMemcpyAsync(lc, lc_pin, (size_t)nlc * sizeof(float), cudaMemcpyHostToDevice, S->s_comp));
    if (ntp > 0) NF_CUDA_OK(cudaMemcpyAsync(tp, lc_pin + nlc, (size_t)ntp * sizeof(float), cudaMemcpyHostToDevice, S->s_comp));
    const float *w = lc, *s = lc + K * D, *b = lc + 2 * K * D, *c = lc + 2 * K * D + K;
    rc = elbo_validate<float>(lc /* stands in for the absent sample pointer */, N, D, K, w, s, b, c, target_id, ntp > 0 ? tp : nullptr, ntp);
    if (rc != NF_OK) return rc;
    float* out = reinterpret_cast<float*>(S->out_dev);
    const RngDesc rng = {seed, step};
    rc = elbo_fused<float>(true, nullptr, N, D, K, w, s, b, c, target_id, ntp > 0 ? tp : nullptr, ntp, epsv, out, out + 3, out + 3 + K * D,
                           out + 3 + 2 * K * D, out + 3 + 2 * K * D + K, reinterpret_cast<float*>(S->ws_dev), S->s_comp, nullptr, nullptr, &rng,
                           S->ws_chunk_bytes * kSessionMaxChunks);
    if (rc != NF_OK) return rc;
    NF_CUDA_OK(cudaMemcpyAsync(S->out_pin, out, (size_t)S->n_out * sizeof(float), cudaMemcpyDeviceToHost, S->s_comp));
    NF_CUDA_OK(cudaStreamSynchronize(S->s_comp));
    const float* op = reinterpret_cast<const float*>(S->out_pin);
    for (int i = 0; i < 3; ++i) F_parts[i] = (double)op[i];
    for (int i = 0; i < nlc; ++i) g64[i] = (double)op[3 + i];
    return nf_script_chain_host(params, g64.data(), K, D, semantics, 1.0 / (double)N, grad);
}

// objective(params, t) / gradient(params, t) of get_data.py:97-115 in ONE call with host buffers in and out: the reference's
// parameter vector [W | U | B] (float64) -> kernel constants (nf_script_constants_host), the session evaluation above,
// and the chain rule back to d(mean F)/d[W | U | B] (nf_script_chain_host).  What the Python face does in three calls.
template <typename T>
static int session_objective_host(Session* S, const T* eps_host, int64_t N, const double* params, int semantics, int target_id,
                                  const T* tp_host, int ntp, T epsv, int backward, double* F_parts, double* grad,
                                  const pr::ArDesc* ar = nullptr, int64_t* n_global = nullptr) {
    if (!S || !params || !F_parts || (backward && !grad) || N < 1) return NF_ERR_INVALID;
    if (ar && (sizeof(T) != 4 || S->n_out > pr::kArStride - 2)) return NF_ERR_UNSUPPORTED;
    const int D = S->D, K = S->K;
    const int nlc = 2 * K * D + 2 * K;
    static thread_local std::vector<double> lc64, g64;
    static thread_local std::vector<T> lc, out;
    lc64.resize(nlc); g64.resize(nlc); lc.resize(nlc); out.resize(3 + nlc);
    int rc = nf_script_constants_host(params, K, D, semantics, lc64.data());
    if (rc != NF_OK) return rc;
    for (int i = 0; i < nlc; ++i) lc[i] = (T)lc64[i];
    T tail[3] = {0, 0, 0};
    rc = session_elbo_host<T>(S, eps_host, N, lc.data(), target_id, tp_host, ntp, epsv, backward, out.data(), ar, tail);
    if (rc != NF_OK) return rc;
    int64_t ng = N;
    if (ar) {
        if ((double)tail[2] != 0.0) return NF_ERR_PEER;       // a rank did not arrive: the sums are not valid
        ng = (int64_t)llround((double)tail[0]) * (1 << 20) + (int64_t)llround((double)tail[1]);
    }
    if (n_global) *n_global = ng;
    for (int i = 0; i < 3; ++i) F_parts[i] = (double)out[i];
    if (backward) {
        for (int i = 0; i < nlc; ++i) g64[i] = (double)out[3 + i];
        rc = nf_script_chain_host(params, g64.data(), K, D, semantics, 1.0 / (double)ng, grad);
    }
    return rc;
}

// ---- whole optimisation loop on the device (get_data.py:119-142 `optimise`, experimentation.py:99-109) ----
// One tiny kernel per iteration applies the previous iteration's update (chain rule from the kernel's
// (g_w, g_s, g_b, g_c) back to (dW, dU, dB), then sgd / rmsprop / adam with autograd 1.3's update rules) and
// prepares the next iteration's layer constants (s = u_hat, c = w.u, get_data.py:106-108); the objective kernel
// and the row kernel follow.  The host only enqueues: no parameter round trip per step.
enum { OPT_SGD = 0, OPT_RMSPROP = 1, OPT_ADAM = 2 };

template <typename T>
__global__ void __launch_bounds__(256)
script_step_kernel(T* __restrict__ params /*[2KD+K]*/, T* __restrict__ lc /*w|s|b|c*/, const T* __restrict__ res /*[3+2KD+2K]*/,
                   T* __restrict__ state /*[2P]*/, int K, int D, long long n_samples, int apply_update, long long iter,
                   long long trace_idx, int optimizer, T step, T h1, T h2, T h3, T* __restrict__ trace_F, int paper) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int P = 2 * K * D + K;
    T* W = params;
    T* U = params + K * D;
    T* B = params + 2 * K * D;
    // one thread per layer does the K x D algebra of that layer (K <= 256 per block pass)
    for (int k = threadIdx.x; k < K; k += blockDim.x) {
        if (apply_update) {
            const T inv_n = (T)1 / (T)n_samples;
            const T* g_w = res + 3 + k * D;
            const T* g_s = res + 3 + K * D + k * D;
            const T g_b = res[3 + 2 * K * D + k];
            const T g_c = res[3 + 2 * K * D + K + k];
            T wu = 0, n2 = 0, gsw = 0;
            for (int d = 0; d < D; ++d) {
                wu += W[k * D + d] * U[k * D + d];
                n2 += W[k * D + d] * W[k * D + d];
                gsw += g_s[d] * W[k * D + d];
            }
            const T alpha = (T)-1 + ::log((T)1 + ::exp(wu)) - wu;        // m(wu) - wu
            const T dalpha = (T)1 / ((T)1 + ::exp(-wu)) - (T)1;
            const T nrm = ::sqrt(n2);
            const T npow = paper ? n2 : nrm;
            const T pexp = paper ? (T)2 : (T)1;
            // (paper semantics folds g_c through c = w.s; the script path has c = w.u)
            for (int d = 0; d < D; ++d) {
                const T w = W[k * D + d], u = U[k * D + d];
                T gs = g_s[d], gw = g_w[d], gc_u = g_c, gc_w = g_c;
                T dU = gs + (dalpha * gsw / npow) * w + gc_u * w;
                T dW = gw + (dalpha * gsw / npow) * u + (alpha / npow) * gs - (alpha * pexp * gsw / (npow * n2)) * w + gc_w * u;
                T grads[2] = {dW * inv_n, dU * inv_n};
                T* ps[2] = {&W[k * D + d], &U[k * D + d]};
                const int idx[2] = {k * D + d, K * D + k * D + d};
                for (int q = 0; q < 2; ++q) {
                    const T g = grads[q];
                    T x = *ps[q];
                    T* s0 = state + idx[q];
                    T* s1 = state + P + idx[q];
                    if (optimizer == OPT_SGD) { *s0 = h1 * *s0 - ((T)1 - h1) * g; x += step * *s0; }
                    else if (optimizer == OPT_RMSPROP) { *s0 = *s0 * h1 + g * g * ((T)1 - h1); x -= step * g / (::sqrt(*s0) + h2); }
                    else {
                        *s0 = ((T)1 - h1) * g + h1 * *s0;
                        *s1 = ((T)1 - h2) * g * g + h2 * *s1;
                        const T mh = *s0 / ((T)1 - ::pow(h1, (T)(iter))), vh = *s1 / ((T)1 - ::pow(h2, (T)(iter)));
                        x -= step * mh / (::sqrt(vh) + h3);
                    }
                    *ps[q] = x;
                }
            }
            {
                const T g = g_b * inv_n;
                const int ib = 2 * K * D + k;
                T x = B[k];
                T* s0 = state + ib;
                T* s1 = state + P + ib;
                if (optimizer == OPT_SGD) { *s0 = h1 * *s0 - ((T)1 - h1) * g; x += step * *s0; }
                else if (optimizer == OPT_RMSPROP) { *s0 = *s0 * h1 + g * g * ((T)1 - h1); x -= step * g / (::sqrt(*s0) + h2); }
                else {
                    *s0 = ((T)1 - h1) * g + h1 * *s0;
                    *s1 = ((T)1 - h2) * g * g + h2 * *s1;
                    const T mh = *s0 / ((T)1 - ::pow(h1, (T)(iter))), vh = *s1 / ((T)1 - ::pow(h2, (T)(iter)));
                    x -= step * mh / (::sqrt(vh) + h3);
                }
                B[k] = x;
            }
        }
    }
    __syncthreads();
    if (apply_update && trace_F && threadIdx.x == 0) trace_F[trace_idx] = res[0] / (T)n_samples;
    // constants for the next objective evaluation
    for (int k = threadIdx.x; k < K; k += blockDim.x) {
        T wu = 0, n2 = 0;
        for (int d = 0; d < D; ++d) { wu += W[k * D + d] * U[k * D + d]; n2 += W[k * D + d] * W[k * D + d]; }
        const T alpha = (T)-1 + ::log((T)1 + ::exp(wu)) - wu;
        const T npow = paper ? n2 : ::sqrt(n2);
        T ws = 0;
        for (int d = 0; d < D; ++d) {
            const T sv = U[k * D + d] + alpha * W[k * D + d] / npow;
            lc[k * D + d] = W[k * D + d];
            lc[K * D + k * D + d] = sv;
            ws += W[k * D + d] * sv;
        }
        lc[2 * K * D + k] = B[k];
        lc[2 * K * D + K + k] = paper ? ws : wu;
    }
}

// device loop workspace: [lc (2KD+2K) | res (3+2KD+2K) | count and status of the sharded exchange (3) | pad to 256 B | rows]
static inline size_t script_head_bytes(int K, int D, size_t elem) {
    const size_t nlc = (size_t)2 * K * D + 2 * K;
    return ((nlc + (3 + nlc) + 3) * elem + 255) & ~(size_t)255;
}

// the ranks of a sample-sharded device loop (nf_script_train_sharded_f32)
struct ShardDesc { void* const* peer_buffers; int world, rank; uint64_t seq0; int64_t n_global; float* status; };

// optimiser state at the start of a run: zeros, except rmsprop's running average (ones, autograd 1.3)
template <typename T>
__global__ void script_state_init_kernel(T* __restrict__ state, int P, int optimizer) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < 2 * P; i += gridDim.x * blockDim.x)
        state[i] = (optimizer == OPT_RMSPROP && i < P) ? (T)1 : (T)0;
}

// `first_iter` is the global index of the first iteration of this call (Adam's bias correction uses
// first_iter + local index + 1); `resume` != 0 keeps `state` as the previous call left it, so that a run cut
// into several calls follows the trajectory of one call; `eps_stride` is the distance in elements between
// the sample slices of consecutive iterations (0: N*D).
template <typename T>
static int script_train(const T* eps_all, int64_t n_iters, int64_t N, int D, int K, T* params, int target_id,
                        const T* tparams, int ntp, T epsv, int optimizer, T step, T h1, T h2, T h3, int64_t first_iter,
                        int resume, int64_t eps_stride, T* state, T* trace_F, void* workspace, size_t workspace_bytes,
                        void* stream_, const ShardDesc* sh = nullptr) {
    cudaStream_t st = (cudaStream_t)stream_;
    if (n_iters < 0 || N < 1 || !params || !state || !workspace || optimizer < 0 || optimizer > 2 || first_iter < 0) return NF_ERR_INVALID;
    if (eps_stride == 0) eps_stride = N * D;
    if (eps_stride < N * D) return NF_ERR_INVALID;
    // every iteration's slice is read with 16-byte vector loads (load_samples)
    if (n_iters > 1 && ((size_t)eps_stride * sizeof(T)) % 16 != 0) return NF_ERR_INVALID;
    const size_t P = (size_t)2 * K * D + K, nlc = (size_t)2 * K * D + 2 * K, nres = 3 + nlc;
    const size_t head = script_head_bytes(K, D, sizeof(T));
    const size_t need = head + nf_planar_elbo_workspace_bytes(K, D, (int)sizeof(T));
    if (workspace_bytes < need) return NF_ERR_WORKSPACE;
    T* lc = reinterpret_cast<T*>(workspace);
    T* res = lc + nlc;
    T* rows = reinterpret_cast<T*>(reinterpret_cast<unsigned char*>(workspace) + head);
    const T *w = lc, *s = lc + K * D, *b = lc + 2 * K * D, *c = lc + 2 * K * D + K;
    int rc = elbo_validate<T>(eps_all, N, D, K, w, s, b, c, target_id, tparams, ntp);
    if (rc != NF_OK) return rc;
    // wide D, float32, and a workspace sized by nf_script_train_workspace_bytes_n: matrix-product path per iteration
    bool use_gram = false;
    if constexpr (sizeof(T) == 4)
        use_gram = gram_applicable(N, D, K, target_id, ntp) && workspace_bytes >= head + gram_workspace_bytes(N, D, K);
    if (!resume)
        NF_CUDA_OK(launch_pdl(script_state_init_kernel<T>, dim3(8), dim3(256), 0, st, state, (int)P, optimizer));
    if (sh) {
        if (sizeof(T) != 4 || use_gram || (int)nres + 2 > pr::kArStride || !sh->status || sh->n_global < N) return NF_ERR_UNSUPPORTED;
        NF_CUDA_OK(cudaMemsetAsync(sh->status, 0, sizeof(float), st));
    }
    const long long n_norm = sh ? (long long)sh->n_global : (long long)N;      // the mean of get_data.py:113 is over ALL ranks' samples
    for (int64_t it = 0; it <= n_iters; ++it) {
        // apply update of iteration it-1 (if any), then prepare constants for iteration it
        NF_CUDA_OK(launch_pdl(script_step_kernel<T>, dim3(1), dim3(256), 0, st, params, lc, (const T*)res, state, K, D,
                              n_norm, it > 0 ? 1 : 0, (long long)(first_iter + it), (long long)(it - 1), optimizer, step,
                              h1, h2, h3, trace_F, 0));
        if (it == n_iters) break;
        const T* eps_it = eps_all + (size_t)it * (size_t)eps_stride;
        if (use_gram) {
            if constexpr (sizeof(T) == 4)
                rc = gram_elbo_f32(true, eps_it, N, D, K, w, s, b, c, target_id, tparams, ntp, epsv, res, res + 3,
                                   res + 3 + K * D, res + 3 + 2 * K * D, res + 3 + 2 * K * D + K, rows, st);
            if (rc != NF_OK) return rc;
            continue;
        }
        if constexpr (sizeof(T) == 4) {
            if (sh) {
                // this rank's sums are exchanged with the peers before the update kernel reads them (every rank then applies
                // the same update to its copy of the parameters); a missing peer raises the sticky status word
                pr::ArDesc ar;
                rc = make_ar_desc(&ar, sh->peer_buffers, sh->world, sh->rank, sh->seq0 + (uint64_t)it, N, res + nres);
                if (rc != NF_OK) return rc;
                ar.sticky_status = sh->status;
                bool exchanged = false;
                rc = elbo_fused<T>(true, eps_it, N, D, K, w, s, b, c, target_id, tparams, ntp, epsv, res, res + 3, res + 3 + K * D,
                                   res + 3 + 2 * K * D, res + 3 + 2 * K * D + K, rows, st, &ar, &exchanged, nullptr, workspace_bytes - head);
                if (rc == NF_OK && !exchanged) rc = launch_allreduce_small(ar, res, res, (int)nres, nullptr, st);
                if (rc != NF_OK) return rc;
                continue;
            }
        }
        rc = elbo_fused<T>(true, eps_it, N, D, K, w, s, b, c, target_id, tparams, ntp, epsv, res, res + 3, res + 3 + K * D,
                           res + 3 + 2 * K * D, res + 3 + 2 * K * D + K, rows, st, nullptr, nullptr, nullptr, workspace_bytes - head);
        if (rc != NF_OK) return rc;
    }
    return NF_OK;
}

}  // namespace nf

