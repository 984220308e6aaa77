// Packed-pair variant of the shared-parameter free-energy kernel (float32, D <= 2, K <= 64).
//
// Same arithmetic as elbo_shared_kernel (nf_elbo_impl.cuh; get_data.py:97-115), laid out for the
// instruction-issue limit that kernel runs into: every thread carries its samples as PAIRS in 64-bit
// registers and all per-layer multiply-adds are fma/mul/add.f32x2 (SASS FFMA2/FMUL2/FADD2: two results
// per issued instruction, one instruction per two clocks per scheduler when at most two operands are
// sample data -- tools/microbench.cu).  Only the transcendentals (MUFU ex2/rcp/lg2) and the two sign
// manipulations stay scalar.  Signs are arranged so that no packed negation is needed:
//     nhp = t*t - 1 (= -h'),  x = nhp*(-c) + 1,  ngx = copysign(1/(eps+|x|), x) (= -d(-log y)/dx),
//     g_c += ngx*nhp,  ngt = ngx*(-2c)*t + gz.(-s) (= -g_t),  g_a = ngt*nhp.
// Structure:
//   * one persistent block per SM; warp w of the grid owns the w-th slice of every round of samples.
//     The samples are dealt in rounds of SMAX per thread plus one shorter round, so that every warp of
//     the grid does the same amount of work for any N (no tail wave);
//   * tanh history: the first ST samples of a thread in tensor memory (tcgen05.st/ld, fetched one layer
//     ahead), the next ones in shared memory as 64-bit pairs, and -- when the caller's workspace has room for
//     it (nf_planar_elbo_workspace_bytes_n) -- the last SG in an L2-resident slice of global memory, fetched
//     one layer ahead as well: 8 samples per thread at K = 32, so 2^20 samples take two rounds instead of three;
//   * the 2D+2 gradient partials of a layer go to a per-warp staging area in shared memory (one store per value,
//     nothing waits on it); after every fourth layer the lanes of the warp each add up one staged row (one value of one
//     layer over the 32 lanes: eight 128-bit loads, rows padded so that the loads do not collide) into the warp's
//     table of sums.  No chain of dependent shuffles sits in the layer loop; the block writes one row of sums;
//   * the last block to finish (ticket in the workspace header) adds the rows in a fixed order and writes
//     the caller's outputs: one launch per call, bitwise reproducible for a fixed grid.
#pragma once
#include "nf_common.cuh"
#include "nf_launch.cuh"
#include "nf_targets.cuh"
#include "nf_philox.cuh"

namespace nf {
namespace pr {

typedef unsigned long long u64;

__device__ __forceinline__ u64 pk(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk(u64 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
// (|r| with the sign of x) in one LOP3: (r & 0x7fffffff) | (x & 0x80000000)
__device__ __forceinline__ float copysign_bits(float r, float x) {
    unsigned o;
    asm("lop3.b32 %0, %1, 0x80000000, %2, 0xb8;" : "=r"(o) : "r"(__float_as_uint(r)), "r"(__float_as_uint(x)));
    return __uint_as_float(o);
}

template <int N> struct Tm;
template <> struct Tm<2> {
    static __device__ __forceinline__ void st(uint32_t a, const float* r) {
        asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1, %2};" ::"r"(a), "r"(__float_as_uint(r[0])), "r"(__float_as_uint(r[1])) : "memory");
    }
    static __device__ __forceinline__ void ld(uint32_t a, float* r) {
        uint32_t v0, v1;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x2.b32 {%0, %1}, [%2];" : "=r"(v0), "=r"(v1) : "r"(a) : "memory");
        r[0] = __uint_as_float(v0); r[1] = __uint_as_float(v1);
    }
};
template <> struct Tm<4> {
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
template <> struct Tm<8> {
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
__device__ __forceinline__ void tm_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tm_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__host__ __device__ constexpr int tm_columns(int n) { int c = 32; while (c < n) c <<= 1; return c; }

// per-layer constant records in shared memory
//   forward : w[D]*kappa | b*kappa | s[D]            (2D+1 floats, padded to a multiple of four)
//   backward: -s[D] | -c | -2c | w[D]                (2D+2 floats, padded to a multiple of four)
template <int D> __host__ __device__ constexpr int recf_pairs() { return 2 * D + 1; }
template <int D> __host__ __device__ constexpr int recb_pairs() { return 2 * D + 2; }
// the records hold each constant once (padded to a multiple of four floats); the packed instructions take them as
// scalar operands broadcast to both halves
template <int D> __host__ __device__ constexpr int recf_floats() { return (2 * D + 1 + 3) & ~3; }
template <int D> __host__ __device__ constexpr int recb_floats() { return (2 * D + 2 + 3) & ~3; }

constexpr int kStageLayers = 4;            // layers whose partials are staged before the lanes add them up
constexpr int kStagePitch = 36;            // floats per staged row: 32 lanes + 4 (128-bit reads of 8 rows hit 8 distinct bank groups)

struct Smem {
    int lcf, lcb, tp, acc, stg, hist, total;      // offsets in floats
    __host__ __device__ Smem(int K, int D, int ssp /*pairs per thread kept in shared memory*/, int ntp_s, int nthreads) {
        const int V = 2 * D + 2;
        const int rf = (2 * D + 1 + 3) & ~3, rb = (2 * D + 2 + 3) & ~3;
        lcf = 0;
        lcb = lcf + K * rf;
        tp = lcb + K * rb;
        acc = (tp + ntp_s + 3) & ~3;
        stg = (acc + (nthreads / 32) * K * V + 3) & ~3;
        hist = stg + (nthreads / 32) * kStageLayers * V * kStagePitch;
        total = hist + 2 * K * ssp * nthreads;
    }
};

// One-shot exchange over NVLink peer memory, done by the last block after the row sum (sample-sharded runs, one
// process per GPU).  Every rank PUSHES its result vector [F parts | g_w | g_s | g_b | g_c | n_hi | n_lo] into its slot
// of every peer's exchange buffer with plain stores through the NVLink aperture, each float travelling in one 8-byte
// word together with the step's sequence number.  An 8-byte store lands whole, so a receiver that sees the sequence
// number in a word has the value too: no fence and no separate flag round trip, the latency is one NVLink crossing.
// Receivers poll the words of their own buffer and add the ranks' values in rank order -- the same order on every
// rank, so all ranks hold bitwise identical sums.  Two slot sets alternate by sequence number: a rank can be at most
// one step ahead of the slowest peer, so a set is never overwritten while a peer may still read it.  The wait is
// bounded (20 s).
// Buffer layout per rank (8-byte words): set s in {0,1}: [world][kArStride] words = {float value, uint32 sequence}.
constexpr int kArMaxWorld = 8;
constexpr int kArStride = 4096;            // words per slot
struct ArDesc {
    float* buf[kArMaxWorld];               // exchange buffer of every rank, mapped in this process
    int world, rank;                       // world <= 1: no exchange
    unsigned long long seq;                // 1, 2, 3, ...: the same on every rank for the same step
    float n_hi, n_lo;                      // this rank's sample count, split so that sums over 8 ranks stay exact
    float* xout;                           // [3]: sum of n_hi, sum of n_lo, status (0 ok, 1 a peer did not arrive)
    float* sticky_status;                  // optional: raised to 1 when a peer did not arrive, never cleared (device loops)
};
__host__ __device__ inline size_t ar_buffer_bytes(int world) { return (size_t)2 * world * kArStride * sizeof(unsigned long long); }
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }

// vec[0..nx) (shared memory) := sums over ranks, by all NT threads of the block; `s_flag` is a shared int.
// Returns false (vec unspecified) when a peer did not arrive in time.
__device__ __forceinline__ bool ar_exchange(const ArDesc& ar, float* vec, int nx, int tid, int NT, int* s_flag) {
    const int set = (int)(ar.seq & 1);
    const unsigned seq32 = (unsigned)ar.seq;
    {   // push: NT / world threads per peer
        const int per_peer = NT / ar.world, p = tid / per_peer, t = tid - p * per_peer;
        if (p < ar.world) {
            unsigned long long* slot = reinterpret_cast<unsigned long long*>(ar.buf[p]) + ((size_t)set * ar.world + ar.rank) * kArStride;
            for (int i = t; i < nx; i += per_peer) {
                const unsigned long long word = ((unsigned long long)seq32 << 32) | (unsigned long long)__float_as_uint(vec[i]);
                asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(slot + i), "l"(word) : "memory");
            }
        }
    }
    if (tid == 0) *s_flag = 0;
    __syncthreads();                      // every thread has read its part of vec before anyone overwrites it
    const unsigned long long* mine = reinterpret_cast<const unsigned long long*>(ar.buf[ar.rank]) + (size_t)set * ar.world * kArStride;
    const unsigned long long t0 = gtime();
    bool late = false;
    for (int i = tid; i < nx; i += NT) {
        float a = 0.f;
        for (int r = 0; r < ar.world; ++r) {
            unsigned long long word;
            for (;;) {
                asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(word) : "l"(mine + (size_t)r * kArStride + i) : "memory");
                if ((unsigned)(word >> 32) == seq32) break;
                if (gtime() - t0 > 20000000000ull) { late = true; break; }
            }
            a += __uint_as_float((unsigned)word);
        }
        vec[i] = a;
    }
    if (late) *s_flag = 1;
    __syncthreads();
    return *s_flag == 0;
}

// workspace header of the single-launch mode (16 bytes): a cookie that tells an initialised ticket counter from
// whatever the caller's memory held, and the ticket counter itself (zero between launches)
struct SyncHeader { u64 cookie; unsigned counter; unsigned pad; };
constexpr u64 kCookieBusy = 0x6e66636f72655f42ull;       // "nfcore_B": a block is clearing the counter
__host__ __device__ inline u64 cookie_for(const void* p) { return 0x6e66636f72653032ull ^ (u64)(uintptr_t)p; }

// tuning aid (flags bit 1): per-block phase timestamps (globaltimer, ns): entry, set-up done, samples done, row
// written and ticket taken, finalize done (last block only; 0 elsewhere), and the SM the block ran on
// One round of one warp: S samples per thread (S/2 pairs) through all K layers and back.
template <int D, int S, int SMAX, int ST, int SG, int NT, int TGT, bool MASKED>
__device__ __forceinline__ void pair_tile(const float* __restrict__ eps, long long n0, long long N, int K,
                                          const float* __restrict__ lcf, const float* __restrict__ lcb,
                                          const float* __restrict__ tp, int ntp, float epsv,
                                          u64* __restrict__ hist_t, uint32_t taddr, float* __restrict__ acc_w, int lane,
                                          float* __restrict__ stg_w /* staging area of this warp */, float& sum_q, float& sum_p,
                                          const RngDesc& rng /* eps == NULL: the samples are drawn here (nf_philox.cuh) */,
                                          u64* __restrict__ gst_t /* this thread's column of the global-memory history [K][SG/2][NT] */,
                                          unsigned long long* ph = nullptr /* phase stamps (tuning aid), else NULL */) {
    using M = Mth<float>;
    constexpr int V = 2 * D + 2;
    constexpr int P = S / 2;                        // pairs this round
    constexpr int STP = ST / 2;                     // pairs per thread that tensor memory can hold
    constexpr int PT = P < STP ? P : STP;           // pairs of this round whose history goes to tensor memory
    constexpr int SSP = (SMAX - ST - SG) / 2;       // shared-memory pair slots per thread and layer
    constexpr int SGP = SG / 2;                     // global-memory (L2-resident) pair slots per thread and layer
    constexpr int PS = (P - PT) < SSP ? (P - PT) : SSP;      // pairs of this round in shared memory
    constexpr int PG = P - PT - PS;                          // ... and in global memory
    constexpr int RF = recf_floats<D>(), RB = recb_floats<D>();
    static_assert(S % 2 == 0 && S <= SMAX && PG <= SGP, "round size");
    const float LOG2PI = 1.8378770664093453f;
    const u64 one2 = pk(1.0f, 1.0f), mone2 = pk(-1.0f, -1.0f), mtwo2 = pk(-2.0f, -2.0f);
    const u64 nln2 = pk(-0.6931471805599453f, -0.6931471805599453f);

    u64 z[P][D], gz[P][D], lq[P], prod[P], ep[MASKED ? P : 1];
    bool ok[S];
    {
        float zs[S][D];
        if (eps == nullptr) {
            // n0 is even: pair p of this thread is Philox call number n0/2 + p
#pragma unroll
            for (int p = 0; p < P; ++p) rng_pair<D>(rng, (unsigned long long)(n0 >> 1) + p, zs[2 * p], zs[2 * p + 1]);
#pragma unroll
            for (int j = 0; j < S; ++j) ok[j] = !MASKED || n0 + j < N;
        } else if (!MASKED) {
            constexpr int E = S * D;
            float buf[E];
            const float* src = eps + n0 * D;
            if ((E * 4) % 16 == 0) {
                const int4* p = reinterpret_cast<const int4*>(src);
                int4* q = reinterpret_cast<int4*>(buf);
#pragma unroll
                for (int i = 0; i < E / 4; ++i) q[i] = __ldg(p + i);
            } else {
                const int2* p = reinterpret_cast<const int2*>(src);
                int2* q = reinterpret_cast<int2*>(buf);
#pragma unroll
                for (int i = 0; i < E / 2; ++i) q[i] = __ldg(p + i);
            }
#pragma unroll
            for (int j = 0; j < S; ++j) {
                ok[j] = true;
#pragma unroll
                for (int d = 0; d < D; ++d) zs[j][d] = buf[j * D + d];
            }
        } else {
#pragma unroll
            for (int j = 0; j < S; ++j) {
                ok[j] = n0 + j < N;
#pragma unroll
                for (int d = 0; d < D; ++d) zs[j][d] = ok[j] ? eps[(n0 + j) * D + d] : 0.5f;
            }
        }
#pragma unroll
        for (int p = 0; p < P; ++p) {
            float q0 = 0.f, q1 = 0.f;
#pragma unroll
            for (int d = 0; d < D; ++d) {
                q0 = fmaf(zs[2 * p][d], zs[2 * p][d], q0);
                q1 = fmaf(zs[2 * p + 1][d], zs[2 * p + 1][d], q1);
                z[p][d] = pk(zs[2 * p][d], zs[2 * p + 1][d]);
            }
            lq[p] = pk(-0.5f * (D * LOG2PI + q0), -0.5f * (D * LOG2PI + q1));      // get_data.py:111, first term
            prod[p] = one2;
            // a sample past N gets eps = +inf: y = inf, 1/y = 0, so every gradient contribution of the layer chain
            // vanishes without a per-layer select (its log q and log p are dropped by selects at the end)
            if (MASKED) ep[p] = pk(ok[2 * p] ? epsv : __int_as_float(0x7f800000), ok[2 * p + 1] ? epsv : __int_as_float(0x7f800000));
        }
        if (!MASKED) ep[0] = pk(epsv, epsv);
    }
    if (ph) ph[0] = gtime();
    // ---- forward over the K layers (get_data.py:104-109); the log-det is left to the reverse sweep ----
    for (int k = 0; k < K; ++k) {
        u64 rf[recf_pairs<D>()];
        {
            float f[RF];
            const float4* src = reinterpret_cast<const float4*>(lcf + k * RF);
#pragma unroll
            for (int i = 0; i < RF / 4; ++i) { const float4 v = src[i]; f[4 * i] = v.x; f[4 * i + 1] = v.y; f[4 * i + 2] = v.z; f[4 * i + 3] = v.w; }
#pragma unroll
            for (int i = 0; i < recf_pairs<D>(); ++i) rf[i] = pk(f[i], f[i]);
        }
        float tt[2 * (PT > 0 ? PT : 1)];
#pragma unroll
        for (int p = 0; p < P; ++p) {
            u64 a = rf[D];
#pragma unroll
            for (int d = 0; d < D; ++d) a = fma2(rf[d], z[p][d], a);
            float a0, a1;
            upk(a, a0, a1);
            const u64 den = add2(pk(M::ex2(a0), M::ex2(a1)), one2);
            float d0, d1;
            upk(den, d0, d1);
            const u64 t = fma2(pk(M::rcp(d0), M::rcp(d1)), mtwo2, one2);         // tanh = 1 - 2/(1+e^{2a})
#pragma unroll
            for (int d = 0; d < D; ++d) z[p][d] = fma2(rf[D + 1 + d], t, z[p][d]);
            if (p < PT) upk(t, tt[2 * p], tt[2 * p + 1]);
            else if (p < PT + PS) hist_t[(k * SSP + (p - PT)) * NT] = t;
            else gst_t[(size_t)(k * SGP + (p - PT - PS)) * NT] = t;
        }
        if (PT > 0) Tm<2 * (PT > 0 ? PT : 1)>::st(taddr + k * ST, tt);
    }
    if (PT > 0) tm_wait_st();
    if (ph) ph[1] = gtime();
    // ---- target: log(eps + target(z_K)) and the reverse seed gz = -d log p / dz ----
#pragma unroll
    for (int p = 0; p < P; ++p) {
        float za[D], zb[D], ga[D], gb[D], lpa, lpb;
#pragma unroll
        for (int d = 0; d < D; ++d) upk(z[p][d], za[d], zb[d]);
        Target<float, D, TGT, true>::eval(za, tp, ntp, epsv, lpa, ga);
        Target<float, D, TGT, true>::eval(zb, tp, ntp, epsv, lpb, gb);
        if (MASKED) {
            sum_p += (ok[2 * p] ? lpa : 0.f) + (ok[2 * p + 1] ? lpb : 0.f);
#pragma unroll
            for (int d = 0; d < D; ++d) gz[p][d] = pk(ok[2 * p] ? ga[d] : 0.f, ok[2 * p + 1] ? gb[d] : 0.f);
        } else {
            sum_p += lpa + lpb;
#pragma unroll
            for (int d = 0; d < D; ++d) gz[p][d] = pk(ga[d], gb[d]);
        }
    }
    if (ph) ph[2] = gtime();
    // ---- reverse sweep: restores z, accumulates log|det| and the gradient partials ----
    // Layer k+1's partials are merged (pair halves added) at the end of its iteration and reduced across the warp
    // inside iteration k, next to arithmetic they do not depend on.  History held in tensor memory is fetched one
    // layer ahead; the loop is unrolled by two so that the prefetch buffers alternate instead of being copied.
    constexpr int NTV = 2 * (PT > 0 ? PT : 1), NGV = PG > 0 ? PG : 1;
    float ta[NTV], tb[NTV];
    u64 ga_[NGV], gb_[NGV];                                       // global-memory history, fetched one layer ahead like ta / tb
    float* const stg_wr = stg_w + lane;                           // lane's column of the staged rows
    const float* const stg_rd = stg_w + lane * kStagePitch;       // the row this lane adds up (lane < kStageLayers * V)
    if (PT > 0) Tm<NTV>::ld(taddr + (K - 1) * ST, ta);
#pragma unroll
    for (int i = 0; i < PG; ++i) ga_[i] = __ldcg(gst_t + (size_t)((K - 1) * SGP + i) * NT);
    auto layer = [&](int k, float (&tc)[NTV], float (&tn)[NTV], u64 (&gc)[NGV], u64 (&gn)[NGV]) {
        if (PT > 0) {
            tm_wait_ld();
            if (k > 0) Tm<NTV>::ld(taddr + (k - 1) * ST, tn);
        }
        if (PG > 0 && k > 0) {
#pragma unroll
            for (int i = 0; i < PG; ++i) gn[i] = __ldcg(gst_t + (size_t)((k - 1) * SGP + i) * NT);
        }
        u64 rb[recb_pairs<D>()];
        {
            float f[RB];
            const float4* src = reinterpret_cast<const float4*>(lcb + k * RB);
#pragma unroll
            for (int i = 0; i < RB / 4; ++i) { const float4 v = src[i]; f[4 * i] = v.x; f[4 * i + 1] = v.y; f[4 * i + 2] = v.z; f[4 * i + 3] = v.w; }
#pragma unroll
            for (int i = 0; i < recb_pairs<D>(); ++i) rb[i] = pk(f[i], f[i]);
        }
        const u64 nc = rb[D], c2 = rb[D + 1];
        u64 v[V];                                   // the first pair starts the per-layer sums (no zero fill)
#pragma unroll
        for (int p = 0; p < P; ++p) {
            const u64 t = (p < PT) ? pk(tc[p < PT ? 2 * p : 0], tc[p < PT ? 2 * p + 1 : 0])
                          : ((p < PT + PS) ? hist_t[(k * SSP + (p - PT)) * NT] : gc[(p >= PT + PS) ? p - PT - PS : 0]);
#pragma unroll
            for (int d = 0; d < D; ++d) z[p][d] = fma2(rb[d], t, z[p][d]);        // z_{k-1} = z_k - s_k t_k
            const u64 nhp = fma2(t, t, mone2);                                     // -(1 - t^2)
            const u64 x = fma2(nhp, nc, one2);                                     // 1 + h' c
            float x0, x1;
            upk(x, x0, x1);
            const u64 y = add2(pk(fabsf(x0), fabsf(x1)), ep[MASKED ? p : 0]);
            prod[p] = mul2(prod[p], y);
            float y0, y1;
            upk(y, y0, y1);
            const u64 ngx = pk(copysign_bits(M::rcp(y0), x0), copysign_bits(M::rcp(y1), x1));     // sign(x)/(eps+|x|)
            v[2 * D + 1] = (p == 0) ? mul2(ngx, nhp) : fma2(ngx, nhp, v[2 * D + 1]);       // g_c
            u64 ngt = mul2(mul2(ngx, c2), t);                                      // -(d/dt of the log-det term)
#pragma unroll
            for (int d = 0; d < D; ++d) {
                v[D + d] = (p == 0) ? mul2(gz[p][d], t) : fma2(gz[p][d], t, v[D + d]);     // g_s
                ngt = fma2(gz[p][d], rb[d], ngt);
            }
            const u64 ga = mul2(ngt, nhp);
            v[2 * D] = (p == 0) ? ga : add2(v[2 * D], ga);                         // g_b
#pragma unroll
            for (int d = 0; d < D; ++d) {
                v[d] = (p == 0) ? mul2(ga, z[p][d]) : fma2(ga, z[p][d], v[d]);     // g_w (direct)
                gz[p][d] = fma2(ga, rb[D + 2 + d], gz[p][d]);
            }
        }
        if ((k % kLogEvery) == 0) {
#pragma unroll
            for (int p = 0; p < P; ++p) {
                float p0, p1;
                upk(prod[p], p0, p1);
                lq[p] = fma2(pk(M::lg2(p0), M::lg2(p1)), nln2, lq[p]);
                prod[p] = one2;
            }
        }
        // stage this layer's partials (pair halves added); row = (layer within the group of four, value)
        {
            float* dst = stg_wr + (k & (kStageLayers - 1)) * (V * kStagePitch);
#pragma unroll
            for (int i = 0; i < V; ++i) { float a, b; upk(v[i], a, b); dst[i * kStagePitch] = a + b; }
        }
        if ((k & (kStageLayers - 1)) == 0) {
            // layers k .. k+3 are staged: lane r adds up row r = 32 lane values of (layer k + r / V, value r % V).  Rows of
            // layers >= K (first, partial group) hold stale data and are skipped.
            __syncwarp();
            if (lane < kStageLayers * V && k * V + lane < K * V) {
                float4 q[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) q[j] = *reinterpret_cast<const float4*>(stg_rd + 4 * j);
                float a0 = q[0].x, a1 = q[0].y, a2 = q[0].z, a3 = q[0].w;
#pragma unroll
                for (int j = 1; j < 8; ++j) { a0 += q[j].x; a1 += q[j].y; a2 += q[j].z; a3 += q[j].w; }
                acc_w[k * V + lane] += (a0 + a1) + (a2 + a3);
            }
            __syncwarp();
        }
    };
    int k = K - 1;
    for (; k >= 1; k -= 2) {
        layer(k, ta, tb, ga_, gb_);
        layer(k - 1, tb, ta, gb_, ga_);
    }
    if (k == 0) layer(0, ta, tb, ga_, gb_);
    if (ph) ph[3] = gtime();
#pragma unroll
    for (int p = 0; p < P; ++p) {
        float a, b;
        upk(lq[p], a, b);
        sum_q += ((MASKED && !ok[2 * p]) ? 0.f : a) + ((MASKED && !ok[2 * p + 1]) ? 0.f : b);
    }
}

template <int D, int NW, int SMAX, int ST, int SG, int TGT>
__global__ void __launch_bounds__(NW * 32, 1)
elbo_pair_kernel(const float* __restrict__ eps, long long N, int K, const float* __restrict__ w,
                 const float* __restrict__ s, const float* __restrict__ b, const float* __restrict__ c,
                 const float* __restrict__ tparams, int ntp, float epsv, float* __restrict__ partials /*[gridDim.x][K*V+2]*/,
                 SyncHeader* __restrict__ sync /* NULL: rows only, the caller sums them */, float* __restrict__ out_F,
                 float* __restrict__ g_w, float* __restrict__ g_s, float* __restrict__ g_b, float* __restrict__ g_c,
                 int flags /* tuning: bit 0 = no rotation of the rounds */, unsigned long long* __restrict__ stamps /* [256][8] or NULL */,
                 const ArDesc ar, const RngDesc rng, u64* __restrict__ gstash /* [gridDim.x][K][SG/2][NT] pairs, or NULL when SG == 0 */) {
    using M = Mth<float>;
    constexpr int NT = NW * 32;
    constexpr int V = 2 * D + 2;
    constexpr int RF = recf_floats<D>(), RB = recb_floats<D>();
    constexpr int SSP = (SMAX - ST - SG) / 2;
    static_assert(SMAX % 2 == 0 && ST % 2 == 0 && SG % 2 == 0 && ST + SG <= SMAX && (ST == 0 || ST == 2 || ST == 4 || ST == 8), "configuration");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* sm = reinterpret_cast<float*>(smem_raw);
    const int ntp_s = target_smem_count(TGT, D, ntp);
    const Smem L(K, D, SSP, ntp_s, NT);
    float* lcf = sm + L.lcf;
    float* lcb = sm + L.lcb;
    float* tp = sm + L.tp;
    float* acc = sm + L.acc;                                      // [NW][K][V]
    u64* hist = reinterpret_cast<u64*>(sm + L.hist);              // [K][SSP][NT] pairs
    __shared__ uint32_t tmem_base_smem;
    __shared__ float red[2 * NW];
    __shared__ int s_last;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool stamp = stamps != nullptr && tid == 0 && blockIdx.x < 256;
    if (stamp) { stamps[blockIdx.x * 8 + 0] = gtime(); stamps[blockIdx.x * 8 + 4] = 0; }
    const int tmem_cols = tm_columns(((NW + 3) / 4) * ST * K);
    if (ST > 0 && warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                         (uint32_t)__cvta_generic_to_shared(&tmem_base_smem)), "r"(tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // programmatic dependent launch: nothing written by earlier work on the stream is touched before this returns
    asm volatile("griddepcontrol.wait;" ::: "memory");

    if (sync && tid == NT - 1) {
        // first use of this workspace (or its header was overwritten): exactly one block clears the ticket counter
        const u64 good = cookie_for(sync);
        volatile u64* ck = &sync->cookie;
        u64 seen = *ck;
        if (seen != good) {
            if (seen != kCookieBusy && atomicCAS(&sync->cookie, seen, kCookieBusy) == seen) {
                sync->counter = 0;
                __threadfence();
                atomicExch(&sync->cookie, good);
            }
            while (*ck != good) { }
            __threadfence();
        }
    }

    for (int k = tid; k < K; k += NT) {
        const float kap = M::kArgScale;
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const float wv = w[k * D + d], sv = s[k * D + d];
            lcf[k * RF + d] = wv * kap;
            lcf[k * RF + D + 1 + d] = sv;
            lcb[k * RB + d] = -sv;
            lcb[k * RB + D + 2 + d] = wv;
        }
        lcf[k * RF + D] = b[k] * kap;
        lcb[k * RB + D] = -c[k];
        lcb[k * RB + D + 1] = -2.f * c[k];
    }
    target_prepare<float>(TGT, D, tparams, ntp, tp);
    for (int i = tid; i < NW * K * V; i += NT) acc[i] = 0.f;
    if (ST > 0) asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    uint32_t taddr = 0;
    if (ST > 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        taddr = tmem_base_smem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)((warp >> 2) * ST * K);
    }
    float* acc_w = acc + warp * K * V;
    float* stg_w = sm + L.stg + warp * (kStageLayers * V * kStagePitch);
    if (stamp) stamps[blockIdx.x * 8 + 1] = gtime();
    u64* hist_t = hist + tid;
    u64* gst_t = SG > 0 ? gstash + (size_t)blockIdx.x * K * (SG / 2) * NT + tid : nullptr;

    float sum_q = 0.f, sum_p = 0.f;
    // rounds: every thread of the grid takes SMAX samples per full round and `srem` in the last one
    const long long T = (long long)gridDim.x * NT;
    const long long per = ((N + T - 1) / T + 1) & ~1LL;                 // samples per thread, rounded up to a pair
    const long long nfull = per / SMAX;
    const int srem = (int)(per - nfull * SMAX);
    const long long gw = (long long)blockIdx.x * NW + warp;
    // The warps that share a scheduler (warp, warp + 4, ...) visit the rounds in rotated order: their forward sweeps
    // (transcendental-bound) and reverse sweeps (multiply-add-bound) then overlap instead of all warps of the SM
    // competing for the same pipe at the same time.  Which warp owns which samples does not change.
    const int nrounds = (int)nfull + (srem > 0 ? 1 : 0);
    const int rot = (nrounds > 1 && !(flags & 1)) ? (warp >> 2) % nrounds : 0;
    // A warp's first round may be cut in two sub-rounds (`split` samples of every lane first, the rest after): the warps
    // of a scheduler then fall out of step, so that forward sweeps (transcendental-bound) of some run beside reverse
    // sweeps (issue-bound) of the others.  flags bits 8.. : split of warp group (warp >> 2) & 3 = 1, 2, 3 (a nibble each).
    const int wgrp = (warp >> 2) & 3;
    const int split = wgrp ? (flags >> (4 + 4 * wgrp)) & 15 : 0;
    for (int rr = 0; rr < nrounds; ++rr) {
        const long long r = (rr + rot) % nrounds;
        if (rr + 1 < nrounds) {
            // pull the next round's samples of this thread into L2 while this round computes
            const long long rn = (rr + 1 + rot) % nrounds;
            const long long sn = rn < nfull ? rn * T * SMAX + gw * 32 * SMAX + (long long)lane * SMAX
                                            : nfull * T * SMAX + gw * 32 * srem + (long long)lane * srem;
            if (eps && sn < N) asm volatile("prefetch.global.L2 [%0];" ::"l"(eps + sn * D));
        }
        const int sr = r < nfull ? SMAX : srem;                   // samples per lane in this round
        const long long start = (r < nfull ? r : nfull) * T * SMAX + gw * 32 * sr;
        if (start >= N) continue;
        const bool full = start + 32LL * sr <= N;
        const long long nl = start + (long long)lane * sr;
        const int first = (rr == 0 && split > 0 && split < sr) ? split : sr;
        // tuning: phase stamps of the warps of block 0 (rows 216.. of the stamp buffer: [warp][round < 2][start, forward, target, reverse])
        unsigned long long* ph = (stamps && blockIdx.x == 0 && lane == 0 && rr < 2) ? stamps + (216 + warp) * 8 + 4 * rr : nullptr;
        for (int off = 0; off < sr;) {
            const int sc = off == 0 ? first : sr - off;
            const long long n0 = nl + off;
#define NF_PR_RUN(S_)                                                                                                        \
    if (SMAX >= S_ && sc == S_) {                                                                                            \
        if (full) pair_tile<D, (S_ <= SMAX ? S_ : 2), SMAX, ST, SG, NT, TGT, false>(eps, n0, N, K, lcf, lcb, tp, ntp, epsv, hist_t, taddr, acc_w, lane, stg_w, sum_q, sum_p, rng, gst_t, ph); \
        else pair_tile<D, (S_ <= SMAX ? S_ : 2), SMAX, ST, SG, NT, TGT, true>(eps, n0, N, K, lcf, lcb, tp, ntp, epsv, hist_t, taddr, acc_w, lane, stg_w, sum_q, sum_p, rng, gst_t, ph);      \
    }
            NF_PR_RUN(2)
            NF_PR_RUN(4)
            NF_PR_RUN(6)
            NF_PR_RUN(8)
#undef NF_PR_RUN
            off += sc;
        }
    }

    // ---- block row ----
    if (stamps && lane == 0 && blockIdx.x < 4) stamps[(200 + 4 * blockIdx.x) * 8 + warp] = gtime();      // per-warp finish times (rows 200..215)
    if (stamps) __syncthreads();
    if (stamp) stamps[blockIdx.x * 8 + 2] = gtime();
    sum_q = warp_sum(sum_q);
    sum_p = warp_sum(sum_p);
    if (lane == 0) { red[warp] = sum_q; red[NW + warp] = sum_p; }
    __syncthreads();
    const int row = K * V + 2;
    float* out = partials + (size_t)blockIdx.x * row;
    for (int i = tid; i < K * V; i += NT) {
        float a = 0.f;
#pragma unroll
        for (int wi = 0; wi < NW; ++wi) a += acc[wi * K * V + i];
        out[i] = a;
    }
    if (tid == 0) {
        float q = 0.f, p = 0.f;
#pragma unroll
        for (int wi = 0; wi < NW; ++wi) { q += red[wi]; p += red[NW + wi]; }
        out[K * V] = q;
        out[K * V + 1] = p;
    }
    if (ST > 0) asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    if (sync) {
        // the last block to arrive adds all rows in row order and scatters them into the caller's arrays
        __syncthreads();
        if (tid == 0) {
            // release: the row this block wrote (ordered before the barrier) is visible to whoever sees the ticket;
            // acquire: the last block sees every other block's row
            unsigned ticket;
            asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], 1;" : "=r"(ticket) : "l"(&sync->counter) : "memory");
            s_last = (ticket == gridDim.x - 1) ? 1 : 0;
        }
        __syncthreads();
        if (stamp) { stamps[blockIdx.x * 8 + 3] = gtime(); unsigned smid; asm("mov.u32 %0, %smid;" : "=r"(smid)); stamps[blockIdx.x * 8 + 5] = smid; }
        if (s_last) {
            const int nrows = gridDim.x;
            float* stage = acc;                                   // reuse (the staging areas follow): [groups][row]
            // A warp takes a group of 8 consecutive rows at a time, its lanes stride over the row's values two at a time
            // (a row is K*V+2 floats: an even count, 8-byte aligned).  A lane issues the loads of a whole row group before it
            // adds anything, so the pass costs a few L2 round trips; no integer division on the way.
            constexpr int GR = NW >= 16 ? 10 : 8, U = NW >= 16 ? 4 : 2;      // 16 warps x 10 rows cover 148 SMs in one batch per warp
            const int groups = (nrows + GR - 1) / GR;
            const int row2 = row >> 1;
            for (int g = warp; g < groups; g += NW) {
                const float2* base = reinterpret_cast<const float2*>(partials + (size_t)g * GR * row);
                const int nr = (nrows - g * GR) < GR ? (nrows - g * GR) : GR;
                for (int i0 = 0; i0 < row2; i0 += 32 * U) {
                    float2 t[U][GR];
#pragma unroll
                    for (int u = 0; u < U; ++u) {
                        const int i = i0 + u * 32 + lane;
#pragma unroll
                        for (int j = 0; j < GR; ++j)
                            t[u][j] = (i < row2 && j < nr) ? __ldcg(base + (size_t)j * row2 + i) : make_float2(0.f, 0.f);
                    }
#pragma unroll
                    for (int u = 0; u < U; ++u) {
                        const int i = i0 + u * 32 + lane;
                        if (i < row2) {
                            float2 r = t[u][0];
#pragma unroll
                            for (int j = 1; j < GR; ++j) { r.x += t[u][j].x; r.y += t[u][j].y; }
                            reinterpret_cast<float2*>(stage + g * row)[i] = r;
                        }
                    }
                }
            }
            __syncthreads();
            if (stamps && tid == 0) stamps[(blockIdx.x < 256 ? blockIdx.x : 0) * 8 + 6] = gtime();
            // result vector in shared memory, laid out as the caller's arrays follow one another:
            // [F, sum log q, sum log p | g_w (K*D) | g_s (K*D) | g_b (K) | g_c (K) | n_hi, n_lo]
            float* vec = stage + groups * row;
            const int KD = K * D, nvec = 3 + 2 * KD + 2 * K;
            for (int i = tid; i < row; i += NT) {
                float a = stage[i];
                for (int g = 1; g < groups; ++g) a += stage[g * row + i];
                if (i >= K * V) {
                    vec[1 + (i - K * V)] = a;
                } else {
                    const int k = i / V, j = i - k * V;
                    const int pos = j < D ? 3 + k * D + j : (j < 2 * D ? 3 + KD + k * D + (j - D) : (j == 2 * D ? 3 + 2 * KD + k : 3 + 2 * KD + K + k));
                    vec[pos] = a;
                }
            }
            __syncthreads();
            if (tid == 0) { vec[0] = vec[1] - vec[2]; vec[nvec] = ar.n_hi; vec[nvec + 1] = ar.n_lo; }
            __syncthreads();
            if (stamps && tid == 0) stamps[(blockIdx.x < 256 ? blockIdx.x : 0) * 8 + 7] = gtime();
            bool failed = false;
            if (ar.world > 1) failed = !ar_exchange(ar, vec, nvec + 2, tid, NT, &s_last);
            for (int i = tid; i < nvec; i += NT) {
                const float a = vec[i];
                if (i < 3) out_F[i] = a;
                else if (g_w) {
                    if (i < 3 + KD) g_w[i - 3] = a;
                    else if (i < 3 + 2 * KD) g_s[i - 3 - KD] = a;
                    else if (i < 3 + 2 * KD + K) g_b[i - 3 - 2 * KD] = a;
                    else g_c[i - 3 - 2 * KD - K] = a;
                }
            }
            if (tid == 0) {
                if (ar.xout) { ar.xout[0] = vec[nvec]; ar.xout[1] = vec[nvec + 1]; ar.xout[2] = failed ? 1.f : 0.f; }
                if (failed && ar.sticky_status) ar.sticky_status[0] = 1.f;
                sync->counter = 0;
                if (stamps) stamps[(blockIdx.x < 256 ? blockIdx.x : 0) * 8 + 4] = gtime();
            }
        }
    }
    if (ST > 0) {
        __syncthreads();
        if (warp == 0)
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base_smem), "r"(tmem_cols) : "memory");
    }
}

}  // namespace pr
}  // namespace nf
