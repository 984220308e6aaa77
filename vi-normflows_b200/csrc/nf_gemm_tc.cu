// Tensor-core GEMM for the flow-VAE MLP (nn_models.py:59-84 and its reverse pass), sm_100a:
// TMA tile loads -> shared memory -> tcgen05.mma (kind::tf32) -> TMEM accumulator -> fused epilogue.
//
//   C[m][n] = epi( sum_k A[m][k] * B[n][k] )       A: [M][K], B: [N][K], both row-major (K contiguous)
//
// FP32 parity on TF32 tensor cores: the reference computes in float64 and the float32 mode must
// stay within 1e-4, which one TF32 pass (10-bit mantissa) does not.  Every operand tile is therefore
// split in shared memory into hi = x with the low 13 mantissa bits cleared (exact in TF32) and
// lo = x - hi, and each k-step issues three MMAs into the same accumulator:
//   hi_a*hi_b + lo_a*hi_b + hi_a*lo_b     (the dropped lo*lo term is ~2^-22 relative).
// The split is elementwise, so it is indifferent to the 128-byte swizzle TMA applied to the tile.
//
// Warp roles (256 threads, one 128 x 64 output tile per CTA):
//   warp 0    TMA producer: per stage one A box (128 rows x 32 floats) and one B box (64 x 32)
//   warp 1    TMEM allocation + MMA issue (one elected lane), tcgen05.commit to free stages
//   warps 4-7 split the landed tiles (generic proxy -> fence.proxy.async), then the epilogue:
//             tcgen05.ld of the accumulator, bias / activation / activation-derivative mask, stores of
//             C and (optionally) of its transpose for the weight-gradient GEMMs.
#include <cstdio>
#include <cstdlib>
#include <cuda.h>
#include <cudaTypedefs.h>
#include "nf_common.cuh"
#include "nf_launch.cuh"

namespace nf {
namespace tc {

constexpr int BM = 128, BN = 64, BK = 32;          // tile (BK floats = 128 bytes = one swizzle row)
constexpr int MAX_STAGES = 3;                       // ring depth is a launch parameter (1..3): shallow rings let 2-4 CTAs share an SM
constexpr int A_BYTES = BM * BK * 4, B_BYTES = BN * BK * 4;
constexpr int STAGE_BYTES = 2 * A_BYTES + 2 * B_BYTES;      // hi + lo of both operands
constexpr int THREADS = 256;
constexpr int SPLIT_THREADS = 192;                  // warps 2-7 split; warps 4-7 also run the epilogue
constexpr int ACC_COLS = 64;                        // accumulator; with the A operand in tensor memory each stage adds 64 (hi | lo)
constexpr int STAGE_BYTES_TS = A_BYTES + 2 * B_BYTES;   // A raw | B hi | B lo

enum { EPI_ACT_NONE = 0, EPI_ACT_RELU = 1, EPI_ACT_SIGMOID = 2 };
enum { EPI_MASK_NONE = 0, EPI_MASK_RELU = 1, EPI_MASK_SIGMOID = 2 };   // multiply by act'(saved output)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {}
}
// the same with a pause between probes: for roles that share an SM with working warps for a long time (a spinning warp
// takes issue slots and shared-memory pipe slots from them)
__device__ __forceinline__ void mbar_wait_relaxed(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) __nanosleep(40);
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                     smem_u32(dst)), "l"(map), "r"(c0), "r"(c1), "r"(smem_u32(bar)) : "memory");
}
// K-major, 128-byte swizzle: rows are 128 B apart inside an 8-row atom, atoms 1024 B apart
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);         // start address
    d |= (uint64_t)1 << 16;                         // leading byte offset (unused for swizzled K-major)
    d |= (uint64_t)(1024 >> 4) << 32;               // stride byte offset: next 8-row atom
    d |= (uint64_t)1 << 46;                         // descriptor version (Blackwell)
    d |= (uint64_t)2 << 61;                         // SWIZZLE_128B
    return d;
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, bool accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc),
        "r"((uint32_t)accumulate) : "memory");
}
// A operand from tensor memory (128 lanes = rows, 8 columns = the k-slice), B from shared memory
__device__ __forceinline__ void umma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, bool accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc),
        "r"((uint32_t)accumulate) : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
        "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

struct Params {
    int M, N, K;
    float* C; long long ldc;             // [M][N] or null
    float* CT; long long ldct;           // [N][M] or null
    const float* bias;                   // [N] or null
    int act;                             // EPI_ACT_*
    const float* mask_src; long long ldm; int mask;     // multiply by act'(mask_src[m][n])
    long long* dbg;                      // optional phase timestamps of CTA (0,0) (bring-up aid), else null
    int stages;                          // depth of the shared-memory ring
    int m_split;                         // output rows >= m_split (a multiple of 64) take A from the second, sample-major operand
    int kb_per_split;                    // split-K: blockIdx.z owns k-blocks [z*kb_per_split, (z+1)*kb_per_split)
    long long split_stride;              // and writes its partial tile to C + z*split_stride
};

template <bool ATMEM>
__global__ void __launch_bounds__(THREADS)
gemm_tf32x3_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                   const __grid_constant__ CUtensorMap map_a2, const __grid_constant__ CUtensorMap map_a_half, Params p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t full_bar[MAX_STAGES], split_bar[MAX_STAGES], empty_bar[MAX_STAGES], acc_bar;
    const int STAGES = p.stages;
    // ATMEM: the hi/lo halves of A are written to tensor memory by the four warps that own its lanes, and the
    // MMAs read A from there -- A then crosses shared memory twice (TMA in, split read) instead of six times.
    constexpr int STAGE_BYTES_ = ATMEM ? STAGE_BYTES_TS : STAGE_BYTES;
    constexpr int B_OFF = ATMEM ? A_BYTES : 2 * A_BYTES;
    const uint32_t tmem_cols = ATMEM ? (STAGES == 1 ? 128u : 256u) : 64u;
    __shared__ uint32_t tmem_base;
    __shared__ float sbias[BN];

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const bool dbg = p.dbg && blockIdx.x == 0 && blockIdx.y == 0;
    if (dbg && threadIdx.x == 0) p.dbg[0] = clock64();
    const int nkb_all = (p.K + BK - 1) / BK;
    const int kb0 = blockIdx.z * p.kb_per_split;
    const int nkb = max(0, min(p.kb_per_split, nkb_all - kb0));
    if (p.C) p.C += (long long)blockIdx.z * p.split_stride;

    if (threadIdx.x == 32) {       // descriptor fetch off the critical path of the first loads
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a2) : "memory");
    }
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&split_bar[s], SPLIT_THREADS);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&acc_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2 || warp == 3) {      // bias slice of this tile (zero when absent / out of range)
        const int i = threadIdx.x - 64;
        sbias[i] = (p.bias && n0 + i < p.N) ? p.bias[n0 + i] : 0.f;
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base)), "r"(tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = tmem_base;
    if (dbg && threadIdx.x == 0) p.dbg[1] = clock64();

    if (warp == 0) {
        // ---------------- TMA producer ----------------
        if (lane == 0) {
            for (int kb = 0; kb < nkb; ++kb) {
                const int s = kb % STAGES, it = kb / STAGES;
                if (it > 0) mbar_wait(&empty_bar[s], (it - 1) & 1);
                uint8_t* st = smem + (size_t)s * STAGE_BYTES_;
                mbar_expect_tx(&full_bar[s], A_BYTES + B_BYTES);
                const int a1_rows = p.m_split - m0;       // rows of this tile that come from A1: >= 128 all, 64 half, <= 0 none
                if (ATMEM && a1_rows < BM) {
                    // A2 is stored [k][m] (contraction index outermost): 32 x 32 boxes, one per 32 output rows; a boundary
                    // tile takes its first 64 rows from A1 through a 64-row box
                    if (a1_rows > 0) tma_load_2d(st, &map_a_half, (kb0 + kb) * BK, m0, &full_bar[s]);
                    for (int q = a1_rows > 0 ? 2 : 0; q < 4; ++q)
                        tma_load_2d(st + q * 4096, &map_a2, m0 + 32 * q - p.m_split, (kb0 + kb) * BK, &full_bar[s]);
                } else
                tma_load_2d(st, &map_a, (kb0 + kb) * BK, m0, &full_bar[s]);                  // A hi slot (raw data)
                tma_load_2d(st + B_OFF, &map_b, (kb0 + kb) * BK, n0, &full_bar[s]);          // B hi slot (raw data)
            }
        }
    } else if (warp == 1) {
        // ---------------- MMA issuer ----------------
        // instruction descriptor: D=f32, A=B=tf32, both K-major, N=64, M=128
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
        if (nkb == 0 && lane == 0) umma_commit(&acc_bar);
        for (int kb = 0; kb < nkb; ++kb) {
            const int s = kb % STAGES, it = kb / STAGES;
            mbar_wait(&split_bar[s], it & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (lane == 0) {
                const uint32_t a_hi = smem_u32(smem + (size_t)s * STAGE_BYTES_);
                const uint32_t a_lo = a_hi + A_BYTES;
                const uint32_t b_hi = a_hi + B_OFF;
                const uint32_t b_lo = b_hi + B_BYTES;
                const uint32_t ta_hi = tmem_d + ACC_COLS + (uint32_t)s * 64u, ta_lo = ta_hi + 32u;
#pragma unroll
                for (int k = 0; k < BK / 8; ++k) {
                    const uint32_t off = k * 32;       // 8 tf32 = 32 bytes along K inside the swizzle row
                    if (ATMEM) {
                        umma_tf32_ts(tmem_d, ta_hi + 8u * k, umma_desc(b_hi + off), idesc, (kb | k) != 0);
                        umma_tf32_ts(tmem_d, ta_lo + 8u * k, umma_desc(b_hi + off), idesc, true);
                        umma_tf32_ts(tmem_d, ta_hi + 8u * k, umma_desc(b_lo + off), idesc, true);
                    } else {
                        umma_tf32(tmem_d, umma_desc(a_hi + off), umma_desc(b_hi + off), idesc, (kb | k) != 0);
                        umma_tf32(tmem_d, umma_desc(a_lo + off), umma_desc(b_hi + off), idesc, true);
                        umma_tf32(tmem_d, umma_desc(a_hi + off), umma_desc(b_lo + off), idesc, true);
                    }
                }
                if (dbg && kb == 0) p.dbg[4] = clock64();
                umma_commit(&empty_bar[s]);                 // stage reusable once these MMAs have read it
                if (kb == nkb - 1) umma_commit(&acc_bar);   // accumulator complete
            }
            __syncwarp();
        }
    } else {
        // ---------------- split hi/lo (warps 2-7), then epilogue (warps 4-7) ----------------
        const int t = threadIdx.x - 64;
        for (int kb = 0; kb < nkb; ++kb) {
            const int s = kb % STAGES, it = kb / STAGES;
            mbar_wait(&full_bar[s], it & 1);
            if (dbg && kb == 0 && t == 64) p.dbg[8] = clock64();
            uint8_t* st = smem + (size_t)s * STAGE_BYTES_;
            if (ATMEM) {
                if (warp < 4) {
                    // warps 2, 3: B in shared memory, hi in place, lo behind it
#pragma unroll 4
                    for (int i = t; i < B_BYTES / 16; i += 64) {
                        uint8_t* hi = st + B_OFF + (size_t)i * 16;
                        float4 v = *reinterpret_cast<float4*>(hi);
                        float4 h, l;
                        h.x = __uint_as_float(__float_as_uint(v.x) & 0xFFFFE000u); l.x = v.x - h.x;
                        h.y = __uint_as_float(__float_as_uint(v.y) & 0xFFFFE000u); l.y = v.y - h.y;
                        h.z = __uint_as_float(__float_as_uint(v.z) & 0xFFFFE000u); l.z = v.z - h.z;
                        h.w = __uint_as_float(__float_as_uint(v.w) & 0xFFFFE000u); l.w = v.w - h.w;
                        *reinterpret_cast<float4*>(hi) = h;
                        *reinterpret_cast<float4*>(hi + B_BYTES) = l;
                    }
                } else {
                    // warps 4-7: row (32 q + lane) of A, 32 k-values, from the swizzled tile into tensor memory:
                    // columns [0, 32) of the stage hold hi, [32, 64) lo.  16-byte chunk j of row r sits at chunk j ^ (r % 8).
                    const int q = warp - 4, row = 32 * q + lane;
                    const uint8_t* rp = st + (size_t)row * 128;
                    const float* cp = reinterpret_cast<const float*>(st + q * 4096) + lane;     // sample-major box of this warp: [k][32 rows]
                    const bool km = m0 + 32 * q >= p.m_split;
                    const uint32_t ta = tmem_base + ((uint32_t)(32 * q) << 16) + ACC_COLS + (uint32_t)s * 64u;
#pragma unroll
                    for (int h2 = 0; h2 < 2; ++h2) {
                        uint32_t hi[16], lo[16];
                        float x[16];
                        if (km) {
#pragma unroll
                            for (int e = 0; e < 16; ++e) x[e] = cp[(16 * h2 + e) * 32];
                        } else {
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                const float4 v = *reinterpret_cast<const float4*>(rp + 16 * ((4 * h2 + j) ^ (row & 7)));
                                x[4 * j] = v.x; x[4 * j + 1] = v.y; x[4 * j + 2] = v.z; x[4 * j + 3] = v.w;
                            }
                        }
#pragma unroll
                        for (int e = 0; e < 16; ++e) {
                            const uint32_t hb = __float_as_uint(x[e]) & 0xFFFFE000u;
                            hi[e] = hb;
                            lo[e] = __float_as_uint(x[e] - __uint_as_float(hb));
                        }
                        tmem_st16(ta + 16u * h2, hi);
                        tmem_st16(ta + 32u + 16u * h2, lo);
                    }
                    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                }
            } else {
            // A: hi in place at [0, A_BYTES), lo at [A_BYTES, 2 A_BYTES); B likewise after 2 A_BYTES
#pragma unroll 4
            for (int i = t; i < (A_BYTES + B_BYTES) / 16; i += SPLIT_THREADS) {
                const bool isA = i < A_BYTES / 16;
                uint8_t* hi = isA ? st + (size_t)i * 16 : st + 2 * A_BYTES + (size_t)(i - A_BYTES / 16) * 16;
                uint8_t* lo = hi + (isA ? A_BYTES : B_BYTES);
                float4 v = *reinterpret_cast<float4*>(hi);
                float4 h, l;
                h.x = __uint_as_float(__float_as_uint(v.x) & 0xFFFFE000u); l.x = v.x - h.x;
                h.y = __uint_as_float(__float_as_uint(v.y) & 0xFFFFE000u); l.y = v.y - h.y;
                h.z = __uint_as_float(__float_as_uint(v.z) & 0xFFFFE000u); l.z = v.z - h.z;
                h.w = __uint_as_float(__float_as_uint(v.w) & 0xFFFFE000u); l.w = v.w - h.w;
                *reinterpret_cast<float4*>(hi) = h;
                *reinterpret_cast<float4*>(lo) = l;
            }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic-proxy writes -> visible to UMMA
            mbar_arrive(&split_bar[s]);
            if (dbg && kb == 0 && t == 64) p.dbg[9] = clock64();
        }
        // epilogue: warp q = warp - 4 owns TMEM lanes 32q .. 32q+31 = rows m0 + 32q + lane
        if (warp >= 4) {
        mbar_wait(&acc_bar, 0);
        if (dbg && t == 64) p.dbg[5] = clock64();
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int q = warp - 4;
        const int m = m0 + 32 * q + lane;
        // The accumulator arrives one ROW per thread; storing it like that would touch 32 different
        // cache lines per instruction.  Each warp therefore parks its 32 x 64 slab in shared memory
        // (the pipeline stages are free by now; pitch 65 keeps both access patterns conflict-free)
        // and writes C row by row, 128 contiguous bytes per instruction.  C^T is written straight
        // from registers: consecutive lanes hold consecutive m, which is the contiguous index there.
        float* slab = reinterpret_cast<float*>(smem) + (size_t)q * 32 * 65;
#pragma unroll
        for (int c0 = 0; c0 < BN; c0 += 16) {
            uint32_t r[16];
            const uint32_t taddr = tmem_d + ((uint32_t)(32 * q) << 16) + (uint32_t)c0;
            asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                  "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                : "r"(taddr) : "memory");
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            // bias and activation with the decisions outside the element loops.  Columns past N: bias slots are zero and the
            // values are never stored; relu(0) = 0 as before, sigmoid is applied only below N
            float v[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = (nkb > 0 ? __uint_as_float(r[i]) : 0.f) + sbias[c0 + i];
            if (p.act == EPI_ACT_RELU) {
#pragma unroll
                for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], 0.f);
            } else if (p.act == EPI_ACT_SIGMOID) {
#pragma unroll
                for (int i = 0; i < 16; ++i) v[i] = (n0 + c0 + i < p.N) ? 1.f / (1.f + __expf(-v[i])) : v[i];
            }
#pragma unroll
            for (int i = 0; i < 16; ++i) slab[lane * 65 + c0 + i] = v[i];
        }
        __syncwarp();
        if (dbg && t == 64) p.dbg[2] = clock64();
        // batches of 8 rows: all shared-memory reads of a batch are issued before its stores, so the lone
        // epilogue warp of each scheduler keeps several memory operations in flight
        const int rows_here = min(32, p.M - (m0 + 32 * q));
#pragma unroll 1
        for (int r0 = 0; r0 < 32; r0 += 8) {
            float v[8][BN / 32];
#pragma unroll
            for (int j = 0; j < 8; ++j)
#pragma unroll
                for (int h = 0; h < BN / 32; ++h) v[j][h] = slab[(r0 + j) * 65 + lane + 32 * h];
            if (p.mask != EPI_MASK_NONE) {
                float y[8][BN / 32];
#pragma unroll
                for (int j = 0; j < 8; ++j)
#pragma unroll
                    for (int h = 0; h < BN / 32; ++h) {
                        const int n = n0 + lane + 32 * h;
                        y[j][h] = (r0 + j < rows_here && n < p.N) ? __ldg(p.mask_src + (long long)(m0 + 32 * q + r0 + j) * p.ldm + n) : 0.f;
                    }
#pragma unroll
                for (int j = 0; j < 8; ++j)
#pragma unroll
                    for (int h = 0; h < BN / 32; ++h) {
                        v[j][h] *= (p.mask == EPI_MASK_RELU) ? (y[j][h] > 0.f ? 1.f : 0.f) : y[j][h] * (1.f - y[j][h]);
                        slab[(r0 + j) * 65 + lane + 32 * h] = v[j][h];
                    }
            }
            if (p.C) {
#pragma unroll
                for (int j = 0; j < 8; ++j)
#pragma unroll
                    for (int h = 0; h < BN / 32; ++h) {
                        const int n = n0 + lane + 32 * h;
                        if (r0 + j < rows_here && n < p.N) p.C[(long long)(m0 + 32 * q + r0 + j) * p.ldc + n] = v[j][h];
                    }
            }
        }
        if (dbg && t == 64) p.dbg[3] = clock64();
        if (p.CT) {
            __syncwarp();
            if (m < p.M) {
#pragma unroll 1
                for (int i0 = 0; i0 < BN; i0 += 16) {
                    float v[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] = slab[lane * 65 + i0 + i];
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                        if (n0 + i0 + i < p.N) p.CT[(long long)(n0 + i0 + i) * p.ldct + m] = v[i];
                }
            }
        }
        }
    }
    if (dbg && threadIdx.x == 128) p.dbg[6] = clock64();
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (dbg && threadIdx.x == 0) p.dbg[7] = clock64();
    if (warp == 1)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(tmem_cols) : "memory");
}

// ------------------------------------------------------------------------------------------------
// Persistent variant for products with a short contraction and many output tiles (the [N][D] x [D][K] products of the
// matrix-product free-energy chain: two to eight k-blocks per tile, thousands of tiles).  In the kernel above a CTA lives
// for one tile, so its prologue (barriers, TMEM allocation), the latency of its first loads and its epilogue are paid per
// tile and covered only by the other CTAs of the SM; such a product runs at half of HBM speed.  Here a CTA walks over
// tiles blockIdx.x, blockIdx.x + gridDim.x, ...: the shared-memory ring keeps running across tile boundaries, the
// accumulator is double-buffered in tensor memory, and four more warps run the epilogue of tile j while the loads,
// splits and MMAs of tile j+1 proceed.  A operand through tensor memory only; no split-K.
//   warp 0 TMA producer | warp 1 MMA issue | warps 2-3 split B | warps 4-7 split A into TMEM | warps 8-11 epilogue
// TMEM columns: [0,64) accumulator 0 | [64,128) accumulator 1 | [128 + 64 s, +64) A of stage s (hi | lo)
constexpr int P_THREADS = 384;
constexpr int P_STAGES = 2;
constexpr int P_ACC_COLS = 128;
constexpr int P_SLAB_BYTES = BM * BN * 4;            // staging area of the epilogue's bulk stores
constexpr int P_SMEM_BYTES = P_STAGES * STAGE_BYTES_TS + P_SLAB_BYTES + 1024;

__global__ void __launch_bounds__(P_THREADS, 2)
gemm_tf32x3_persist_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                           const __grid_constant__ CUtensorMap map_a2, const __grid_constant__ CUtensorMap map_a_half,
                           const __grid_constant__ CUtensorMap map_out /* C: {32 x 128} boxes, 128-byte swizzle; else C^T: {128 x 64} boxes */, Params p) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t full_bar[P_STAGES], split_bar[P_STAGES], empty_bar[P_STAGES], acc_full[2], acc_empty[2];
    __shared__ uint32_t tmem_base;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tiles_n = (p.N + BN - 1) / BN, tiles_m = (p.M + BM - 1) / BM, tiles = tiles_m * tiles_n;
    const int nkb = (p.K + BK - 1) / BK;
    constexpr uint32_t tmem_cols = 256;

    if (threadIdx.x == 32) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a2) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_out) : "memory");
    }
    if (threadIdx.x == 0) {
        for (int s = 0; s < P_STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&split_bar[s], SPLIT_THREADS);
            mbar_init(&empty_bar[s], 1);
        }
        for (int b = 0; b < 2; ++b) { mbar_init(&acc_full[b], 1); mbar_init(&acc_empty[b], 128); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base)), "r"(tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = tmem_base;
    // bring-up aid (NF_GEMM_DBG): cycles CTA 0's role leaders spend waiting on barriers, and their total
    const bool dbg = p.dbg && blockIdx.x == 0 && lane == 0;
    long long t_wait = 0;
    const long long t_begin = dbg ? clock64() : 0;
#define NF_PWAIT(bar, par) do { if (dbg) { const long long c0_ = clock64(); mbar_wait_relaxed(bar, par); t_wait += clock64() - c0_; } else mbar_wait_relaxed(bar, par); } while (0)

    if (warp == 0) {
        // ---------------- TMA producer ----------------
        if (lane == 0) {
            int g = 0;
            for (int t = blockIdx.x; t < tiles; t += gridDim.x) {
                const int m0 = (t / tiles_n) * BM, n0 = (t % tiles_n) * BN;
                for (int kb = 0; kb < nkb; ++kb, ++g) {
                    const int s = g % P_STAGES, it = g / P_STAGES;
                    if (it > 0) NF_PWAIT(&empty_bar[s], (it - 1) & 1);
                    uint8_t* st = smem + (size_t)s * STAGE_BYTES_TS;
                    mbar_expect_tx(&full_bar[s], A_BYTES + B_BYTES);
                    const int a1_rows = p.m_split - m0;
                    if (a1_rows < BM) {
                        if (a1_rows > 0) tma_load_2d(st, &map_a_half, kb * BK, m0, &full_bar[s]);
                        for (int q = a1_rows > 0 ? 2 : 0; q < 4; ++q)
                            tma_load_2d(st + q * 4096, &map_a2, m0 + 32 * q - p.m_split, kb * BK, &full_bar[s]);
                    } else {
                        tma_load_2d(st, &map_a, kb * BK, m0, &full_bar[s]);
                    }
                    tma_load_2d(st + A_BYTES, &map_b, kb * BK, n0, &full_bar[s]);
                }
            }
        }
    } else if (warp == 1) {
        // ---------------- MMA issuer ----------------
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
        int g = 0, j = 0;
        for (int t = blockIdx.x; t < tiles; t += gridDim.x, ++j) {
            const int b = j & 1, u = j >> 1;
            if (u > 0) NF_PWAIT(&acc_empty[b], (u - 1) & 1);        // the epilogue has drained this accumulator
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t acc = tmem_d + 64u * (uint32_t)b;
            for (int kb = 0; kb < nkb; ++kb, ++g) {
                const int s = g % P_STAGES, it = g / P_STAGES;
                NF_PWAIT(&split_bar[s], it & 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                if (lane == 0) {
                    const uint32_t b_hi = smem_u32(smem + (size_t)s * STAGE_BYTES_TS) + A_BYTES;
                    const uint32_t b_lo = b_hi + B_BYTES;
                    const uint32_t ta_hi = tmem_d + P_ACC_COLS + (uint32_t)s * 64u, ta_lo = ta_hi + 32u;
#pragma unroll
                    for (int k = 0; k < BK / 8; ++k) {
                        const uint32_t off = k * 32;
                        umma_tf32_ts(acc, ta_hi + 8u * k, umma_desc(b_hi + off), idesc, (kb | k) != 0);
                        umma_tf32_ts(acc, ta_lo + 8u * k, umma_desc(b_hi + off), idesc, true);
                        umma_tf32_ts(acc, ta_hi + 8u * k, umma_desc(b_lo + off), idesc, true);
                    }
                    umma_commit(&empty_bar[s]);
                    if (kb == nkb - 1) umma_commit(&acc_full[b]);
                }
                __syncwarp();
            }
        }
    } else if (warp < 8) {
        // ---------------- split hi/lo: warps 2-3 B in shared memory, warps 4-7 A into tensor memory ----------------
        const int t64 = threadIdx.x - 64;
        int g = 0;
        for (int t = blockIdx.x; t < tiles; t += gridDim.x) {
            const int m0 = (t / tiles_n) * BM;
            for (int kb = 0; kb < nkb; ++kb, ++g) {
                const int s = g % P_STAGES, it = g / P_STAGES;
                NF_PWAIT(&full_bar[s], it & 1);
                uint8_t* st = smem + (size_t)s * STAGE_BYTES_TS;
                if (warp < 4) {
#pragma unroll 4
                    for (int i = t64; i < B_BYTES / 16; i += 64) {
                        uint8_t* hi = st + A_BYTES + (size_t)i * 16;
                        float4 v = *reinterpret_cast<float4*>(hi);
                        float4 h, l;
                        h.x = __uint_as_float(__float_as_uint(v.x) & 0xFFFFE000u); l.x = v.x - h.x;
                        h.y = __uint_as_float(__float_as_uint(v.y) & 0xFFFFE000u); l.y = v.y - h.y;
                        h.z = __uint_as_float(__float_as_uint(v.z) & 0xFFFFE000u); l.z = v.z - h.z;
                        h.w = __uint_as_float(__float_as_uint(v.w) & 0xFFFFE000u); l.w = v.w - h.w;
                        *reinterpret_cast<float4*>(hi) = h;
                        *reinterpret_cast<float4*>(hi + B_BYTES) = l;
                    }
                } else {
                    const int q = warp - 4, row = 32 * q + lane;
                    const uint8_t* rp = st + (size_t)row * 128;
                    const float* cp = reinterpret_cast<const float*>(st + q * 4096) + lane;
                    const bool km = m0 + 32 * q >= p.m_split;
                    const uint32_t ta = tmem_base + ((uint32_t)(32 * q) << 16) + P_ACC_COLS + (uint32_t)s * 64u;
#pragma unroll
                    for (int h2 = 0; h2 < 2; ++h2) {
                        uint32_t hi[16], lo[16];
                        float x[16];
                        if (km) {
#pragma unroll
                            for (int e = 0; e < 16; ++e) x[e] = cp[(16 * h2 + e) * 32];
                        } else {
#pragma unroll
                            for (int jj = 0; jj < 4; ++jj) {
                                const float4 v = *reinterpret_cast<const float4*>(rp + 16 * ((4 * h2 + jj) ^ (row & 7)));
                                x[4 * jj] = v.x; x[4 * jj + 1] = v.y; x[4 * jj + 2] = v.z; x[4 * jj + 3] = v.w;
                            }
                        }
#pragma unroll
                        for (int e = 0; e < 16; ++e) {
                            const uint32_t hb = __float_as_uint(x[e]) & 0xFFFFE000u;
                            hi[e] = hb;
                            lo[e] = __float_as_uint(x[e] - __uint_as_float(hb));
                        }
                        tmem_st16(ta + 16u * h2, hi);
                        tmem_st16(ta + 32u + 16u * h2, lo);
                    }
                    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_arrive(&split_bar[s]);
            }
        }
    } else {
        // ---------------- epilogue: warp q = warp - 8 owns TMEM lanes 32q .. 32q+31 ----------------
        // The tile goes to global memory as bulk tensor stores: the four warps lay it out in shared memory (C: two
        // 128-byte-swizzled boxes of 32 columns; C^T: [64 n][128 m], consecutive lanes = consecutive m), one thread
        // issues the store(s).  Rows and columns past M / N are clipped by the tensor map.
        const int q = warp - 8;
        uint8_t* stg = smem + (size_t)P_STAGES * STAGE_BYTES_TS;
        const bool to_ct = p.CT != nullptr;
        const int row = 32 * q + lane;                    // row of the tile this thread holds
        int j = 0;
        for (int t = blockIdx.x; t < tiles; t += gridDim.x, ++j) {
            const int m0 = (t / tiles_n) * BM, n0 = (t % tiles_n) * BN;
            const int b = j & 1, u = j >> 1;
            NF_PWAIT(&acc_full[b], u & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const long long e0_ = dbg ? clock64() : 0;
            if (j > 0) {
                // the previous tile's store must have read the staging area before it is overwritten
                if (threadIdx.x == 8 * 32) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                asm volatile("bar.sync 1, 128;" ::: "memory");
            }
            const long long e1_ = dbg ? clock64() : 0;
#pragma unroll
            for (int c0 = 0; c0 < BN; c0 += 16) {
                uint32_t r[16];
                const uint32_t taddr = tmem_d + ((uint32_t)(32 * q) << 16) + 64u * (uint32_t)b + (uint32_t)c0;
                asm volatile(
                    "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                    : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                      "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                    : "r"(taddr) : "memory");
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                // bias and activation with the decisions outside the element loops (columns past N are clipped by the store)
                float v[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
                if (p.bias) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] += (n0 + c0 + i < p.N) ? __ldg(p.bias + n0 + c0 + i) : 0.f;
                }
                if (p.act == EPI_ACT_RELU) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], 0.f);
                } else if (p.act == EPI_ACT_SIGMOID) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] = 1.f / (1.f + __expf(-v[i]));
                }
                if (to_ct) {
                    float* dst = reinterpret_cast<float*>(stg) + (size_t)c0 * BM + row;
#pragma unroll
                    for (int i = 0; i < 16; ++i) dst[(size_t)i * BM] = v[i];
                } else {
                    // box = columns [32 h, 32 h + 32) of the tile: 128 rows of 128 bytes, 16-byte chunk c of row r at c ^ (r % 8)
                    uint8_t* box = stg + (size_t)(c0 / 32) * (BM * 128) + (size_t)row * 128;
#pragma unroll
                    for (int c4 = 0; c4 < 4; ++c4) {
                        const int chunk = (c0 % 32) / 4 + c4;
                        *reinterpret_cast<float4*>(box + 16 * (chunk ^ (row & 7))) = make_float4(v[4 * c4], v[4 * c4 + 1], v[4 * c4 + 2], v[4 * c4 + 3]);
                    }
                }
            }
            const long long e2_ = dbg ? clock64() : 0;
            // the accumulator has been read: hand the buffer back, publish the staged tile to the async proxy
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            mbar_arrive(&acc_empty[b]);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (threadIdx.x == 8 * 32) {
                if (to_ct) {
                    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(&map_out), "r"(m0), "r"(n0),
                                 "r"(smem_u32(stg)) : "memory");
                } else {
                    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(&map_out), "r"(n0), "r"(m0),
                                 "r"(smem_u32(stg)) : "memory");
                    if (n0 + 32 < p.N)
                        asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(&map_out), "r"(n0 + 32),
                                     "r"(m0), "r"(smem_u32(stg + BM * 128)) : "memory");
                }
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
            if (dbg && warp == 8) { const long long e3_ = clock64(); p.dbg[10] += e1_ - e0_; p.dbg[11] += e2_ - e1_; p.dbg[12] += e3_ - e2_; }
        }
        if (threadIdx.x == 8 * 32) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
#undef NF_PWAIT
    if (dbg && (warp == 0 || warp == 1 || warp == 2 || warp == 4 || warp == 8)) {
        const int role = warp == 0 ? 0 : (warp == 1 ? 1 : (warp == 2 ? 2 : (warp == 4 ? 3 : 4)));
        p.dbg[2 * role] = t_wait;
        p.dbg[2 * role + 1] = clock64() - t_begin;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(tmem_cols) : "memory");
}

static PFN_cuTensorMapEncodeTiled get_encoder() {
    static PFN_cuTensorMapEncodeTiled fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) == cudaSuccess && qr == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled>(p);
    }
    return fn;
}

// row-major [rows][cols] float matrix with leading dimension ld (elements); box = box_rows x 32 columns.
// Encoding a map costs several microseconds on the host; the MLP re-uses the same buffers step after
// step (the framework's caching allocator hands the same addresses back), so maps are memoised.
struct MapKey { const float* base; long long rows, cols, ld; int box_rows; bool swz; int box_cols; };
static bool make_map(CUtensorMap* map, const float* base, long long rows, long long cols, long long ld, int box_rows, bool swz = true,
                     int box_cols = BK) {
    constexpr int kCache = 128;
    static thread_local MapKey keys[kCache];
    static thread_local CUtensorMap vals[kCache];
    static thread_local int used = 0, next = 0;
    for (int i = 0; i < used; ++i) {
        const MapKey& k = keys[i];
        if (k.base == base && k.rows == rows && k.cols == cols && k.ld == ld && k.box_rows == box_rows && k.swz == swz && k.box_cols == box_cols) { *map = vals[i]; return true; }
    }
    PFN_cuTensorMapEncodeTiled enc = get_encoder();
    if (!enc) return false;
    cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t gstride[1] = {(cuuint64_t)ld * 4};
    cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    if (enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), gdim, gstride, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, swz ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return false;
    const int slot = used < kCache ? used++ : (next++ % kCache);
    keys[slot] = MapKey{base, rows, cols, ld, box_rows, swz, box_cols};
    vals[slot] = *map;
    return true;
}

}  // namespace tc

// C = epi(A B^T); returns NF_ERR_UNSUPPORTED when TMA cannot describe the operands (the caller then
// uses the CUDA-core GEMM).  A: [M][K] ld lda, B: [N][K] ld ldb.
// Block (32, 16): thread (x, y) adds slices y, y+16, ... of output element x in ascending order, then the
// sixteen group sums are added in ascending order -- a fixed tree, with sixteen loads in flight per element.
__global__ void tc_splitk_reduce_kernel(const float* __restrict__ part, int nz, long long count, float* __restrict__ out, int accumulate) {
    __shared__ float sm[16][33];
    const long long i = (long long)blockIdx.x * 32 + threadIdx.x;
    float s = 0.f;
    if (i < count)
        for (int z = threadIdx.y; z < nz; z += 16) s += part[(long long)z * count + i];
    sm[threadIdx.y][threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.y == 0 && i < count) {
        float t = 0.f;
#pragma unroll
        for (int y = 0; y < 16; ++y) t += sm[y][threadIdx.x];
        out[i] = accumulate ? out[i] + t : t;
    }
}

// splits > 1: the contraction is cut into `splits` ranges, each CTA layer writes a partial [M][N] tile set
// into split_ws (splits*M*N floats) and a second kernel adds them in a fixed order (no bias/activation/C^T);
// with `accumulate` that sum is added to C instead of replacing it (split mode only).
// General form: output rows [0, M1) take A1 ([M1][K], K contiguous); when A2 != nullptr, rows [m_split, m_split + M2) with
// m_split = M1 rounded up to 64 take A2, stored sample-major: A2[k][m] ([K][M2], ld lda2) -- the layout in which the
// free-energy chain holds z_0 and gz_K -- and rows [M1, m_split) of C are zero padding.  C has m_split + M2 rows then.
int tc_gemm_stacked_f32(const float* A, long long lda, int M1, const float* A2, long long lda2, int M2, const float* B, long long ldb,
                        int N, int K, float* C, long long ldc, float* CT, long long ldct, const float* bias, int act,
                        const float* mask_src, long long ldm, int mask, cudaStream_t st, int splits, float* split_ws, bool accumulate) {
    if (M1 < 0 || M2 < 0 || (M1 < 1 && M2 < 1) || N < 1 || K < 1) return NF_ERR_INVALID;
    if ((ldb % 4) || ((uintptr_t)B & 15)) return NF_ERR_UNSUPPORTED;                                   // 16-byte strides
    if (M1 > 0 && ((lda % 4) || ((uintptr_t)A & 15))) return NF_ERR_UNSUPPORTED;
    if (M2 > 0 && (!A2 || (lda2 % 4) || ((uintptr_t)A2 & 15))) return NF_ERR_UNSUPPORTED;
    static const bool atmem = []() { const char* e = getenv("NF_GEMM_ATMEM"); return !(e && atoi(e) == 0); }();
    if (M2 > 0 && !atmem) return NF_ERR_UNSUPPORTED;          // the sample-major operand needs the tensor-memory path
    const int m_split = M2 > 0 ? (M1 + 63) / 64 * 64 : (1 << 30);
    const int M = M2 > 0 ? m_split + M2 : M1;
    CUtensorMap ma, mb, ma2, mah;
    if (M1 > 0) { if (!tc::make_map(&ma, A, M1, K, lda, tc::BM)) return NF_ERR_UNSUPPORTED; }
    if (!tc::make_map(&mb, B, N, K, ldb, tc::BN)) return NF_ERR_UNSUPPORTED;
    if (M2 > 0) { if (!tc::make_map(&ma2, A2, K, M2, lda2, 32, false)) return NF_ERR_UNSUPPORTED; }
    else ma2 = mb;
    if (M1 < 1) ma = mb;
    if (M2 > 0 && (m_split % tc::BM) != 0) { if (!tc::make_map(&mah, A, M1, K, lda, 64)) return NF_ERR_UNSUPPORTED; }
    else mah = mb;
    static long long* dbg_buf = nullptr;
    if (getenv("NF_GEMM_DBG") && !dbg_buf) cudaMallocManaged(&dbg_buf, 128);
    const int nkb_all = (K + tc::BK - 1) / tc::BK;
    if (splits > nkb_all) splits = nkb_all;
    if (splits < 1 || !split_ws || bias || act || CT || mask || ldc != N) {
        if (accumulate) return NF_ERR_UNSUPPORTED;
        splits = 1;
    }
    const bool via_ws = splits > 1 || accumulate;
    const int kbps = (nkb_all + splits - 1) / splits;
    // A through tensor memory (default) or split in shared memory (NF_GEMM_ATMEM=0, the first version of this kernel).
    // Ring depth: short contractions are dominated by the per-tile prologue and epilogue, so they get a one-stage
    // ring and four CTAs per SM that cover for one another; longer ones a deeper ring and two CTAs per SM
    // (tensor-memory path: 3 stages, 256 columns per CTA; shared-memory path: 2 stages).  NF_GEMM_STAGES=1..3 overrides.
    int stages = kbps <= 2 ? 1 : (atmem ? 3 : 2);
    if (const char* e = getenv("NF_GEMM_STAGES")) { const int v = atoi(e); if (v >= 1 && v <= tc::MAX_STAGES) stages = v; }
    tc::Params p{M, N, K, via_ws ? split_ws : C, ldc, CT, ldct, bias, act, mask_src, ldm, mask, dbg_buf, stages, m_split, kbps, (long long)M * N};
    constexpr size_t kSlab = 4 * 32 * 65 * sizeof(float);       // the epilogue parks the accumulator in the ring's memory
    size_t ring = (size_t)stages * (atmem ? tc::STAGE_BYTES_TS : tc::STAGE_BYTES);
    if (ring < kSlab) ring = kSlab;
    const size_t smem = ring + 1024;
    static bool attr_done_dev[64] = {};           // the attribute is per device
    int dev_id = 0;
    NF_CUDA_OK(cudaGetDevice(&dev_id));
    bool& attr_done = attr_done_dev[dev_id & 63];
    if (!attr_done) {
        NF_CUDA_OK(cudaFuncSetAttribute(tc::gemm_tf32x3_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)((size_t)tc::MAX_STAGES * tc::STAGE_BYTES + 1024)));
        NF_CUDA_OK(cudaFuncSetAttribute(tc::gemm_tf32x3_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)((size_t)tc::MAX_STAGES * tc::STAGE_BYTES_TS + 1024)));
        attr_done = true;
    }
    dim3 grid((N + tc::BN - 1) / tc::BN, (M + tc::BM - 1) / tc::BM, splits);
    // short contraction, many tiles, no split: the persistent kernel (NF_GEMM_PERSIST=0 keeps the per-tile one)
    const char* pe = getenv("NF_GEMM_PERSIST");                  // read per call: the tests compare the two kernels in one process
    const bool persist_on = !(pe && atoi(pe) == 0);
    const long long n_tiles = (long long)grid.x * grid.y;
    const DevInfo di = dev_info();
    const bool one_out = (C != nullptr) != (CT != nullptr);      // the persistent kernel stores one of C / C^T, by bulk tensor stores
    const float* outp = C ? C : CT;
    const long long ldo = C ? ldc : ldct;
    CUtensorMap mo;
    bool persist = persist_on && atmem && !via_ws && splits == 1 && di.ok && nkb_all <= 16 && n_tiles >= 4LL * di.sms && n_tiles < (1LL << 30) &&
                   one_out && mask == 0 && (ldo % 4) == 0 && ((uintptr_t)outp & 15) == 0 &&
                   ((C ? N : M) % 4) == 0;       // bulk tensor stores write whole 16-byte units of a row
    if (persist) persist = C ? tc::make_map(&mo, C, M, N, ldc, tc::BM, true, 32) : tc::make_map(&mo, CT, N, M, ldct, tc::BN, false, tc::BM);
    if (persist) {
        static bool pattr_done_dev[64] = {};
        bool& pattr_done = pattr_done_dev[dev_id & 63];
        if (!pattr_done) {
            NF_CUDA_OK(cudaFuncSetAttribute(tc::gemm_tf32x3_persist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, tc::P_SMEM_BYTES));
            pattr_done = true;
        }
        tc::gemm_tf32x3_persist_kernel<<<2 * di.sms, tc::P_THREADS, tc::P_SMEM_BYTES, st>>>(ma, mb, ma2, mah, mo, p);
        NF_CUDA_OK(cudaGetLastError());
        if (dbg_buf) {
            cudaStreamSynchronize(st);
            fprintf(stderr, "[gemm persist dbg] M=%d N=%d K=%d tiles=%lld (per CTA %.1f) | waiting/total clk of CTA 0: producer %lld/%lld, mma %lld/%lld, "
                    "split-B %lld/%lld, split-A %lld/%lld, epilogue %lld/%lld\n", M, N, K, n_tiles, (double)n_tiles / (2 * di.sms), dbg_buf[0], dbg_buf[1],
                    dbg_buf[2], dbg_buf[3], dbg_buf[4], dbg_buf[5], dbg_buf[6], dbg_buf[7], dbg_buf[8], dbg_buf[9]);
            fprintf(stderr, "[gemm persist dbg]   epilogue of CTA 0: staging free %lld clk, TMEM -> staging %lld, hand-over + store issue %lld\n", dbg_buf[10],
                    dbg_buf[11], dbg_buf[12]);
            dbg_buf[10] = dbg_buf[11] = dbg_buf[12] = 0;
        }
        return NF_OK;
    }
    if (atmem) tc::gemm_tf32x3_kernel<true><<<grid, tc::THREADS, smem, st>>>(ma, mb, ma2, mah, p);
    else tc::gemm_tf32x3_kernel<false><<<grid, tc::THREADS, smem, st>>>(ma, mb, ma2, mah, p);
    NF_CUDA_OK(cudaGetLastError());
    if (via_ws) {
        const long long count = (long long)M * N;
        tc_splitk_reduce_kernel<<<(unsigned)((count + 31) / 32), dim3(32, 16), 0, st>>>(split_ws, splits, count, C, accumulate ? 1 : 0);
        NF_CUDA_OK(cudaGetLastError());
    }
    if (dbg_buf) {
        cudaStreamSynchronize(st);
        fprintf(stderr, "[gemm dbg] M=%d N=%d K=%d stages=%d kb/cta=%d setup=%lld tile-landed=%lld split=%lld first-mma=%lld rest-of-k=%lld | epilogue: tmem->smem %lld, C rows %lld, CT %lld | total %lld clk\n",
                M, N, K, stages, kbps, dbg_buf[1] - dbg_buf[0], dbg_buf[8] - dbg_buf[1], dbg_buf[9] - dbg_buf[8], dbg_buf[4] - dbg_buf[1], dbg_buf[5] - dbg_buf[4], dbg_buf[2] - dbg_buf[5],
                dbg_buf[3] - dbg_buf[2], dbg_buf[6] - dbg_buf[3], dbg_buf[7] - dbg_buf[0]);
    }
    return NF_OK;
}

int tc_gemm_f32(const float* A, long long lda, const float* B, long long ldb, int M, int N, int K, float* C, long long ldc,
                float* CT, long long ldct, const float* bias, int act, const float* mask_src, long long ldm, int mask,
                cudaStream_t st, int splits = 1, float* split_ws = nullptr, bool accumulate = false) {
    return tc_gemm_stacked_f32(A, lda, M, nullptr, 0, 0, B, ldb, N, K, C, ldc, CT, ldct, bias, act, mask_src, ldm, mask, st, splits,
                               split_ws, accumulate);
}

}  // namespace nf

extern "C" int nf_gemm_tn_tf32x3(const float* A, int64_t lda, const float* B, int64_t ldb, int M, int N, int K, float* C,
                                 int64_t ldc, float* CT, int64_t ldct, const float* bias, int act, void* stream) {
    return nf::tc_gemm_f32(A, lda, B, ldb, M, N, K, C, ldc, CT, ldct, bias, act, nullptr, 0, 0, (cudaStream_t)stream);
}
