// Common device helpers for libnfcore (sm_100a).
// Math traits: float uses MUFU-backed approximations chosen so that a K<=64 layer
// chain stays inside the 1e-4 relative budget (SURVEY.md §7 "hard parts");
// double uses the CUDA libm double routines (correctness mode, 1e-10).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/nfcore.h"

#define NF_CUDA_OK(expr)                                  \
    do {                                                  \
        cudaError_t _e = (expr);                          \
        if (_e != cudaSuccess) { nf_set_last_cuda_error(_e); return NF_ERR_CUDA; } \
    } while (0)

void nf_set_last_cuda_error(cudaError_t e);

namespace nf {

constexpr int kWarp = 32;

template <typename T> struct Mth;

template <> struct Mth<float> {
    static __device__ __forceinline__ float ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
    static __device__ __forceinline__ float lg2(float x) { float y; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
    static __device__ __forceinline__ float rcp(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
    static __device__ __forceinline__ float rsqrt(float x) { float y; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
    static __device__ __forceinline__ float sqrt(float x) { return sqrtf(x); }
    // tanh(a) = 1 - 2/(1+e^{2a}); saturates correctly for e -> 0 / inf.  2 MUFU.
    static __device__ __forceinline__ float tanh(float a) {
        float e = ex2(a * 2.885390081777927f);          // 2*log2(e)
        return fmaf(-2.0f, rcp(1.0f + e), 1.0f);
    }
    static __device__ __forceinline__ float log(float x) { return lg2(x) * 0.6931471805599453f; }
    // Pre-scaled form used by the layer loop: the caller folds 2*log2(e) into w and b once
    // per layer, so tanh costs FADD + FFMA + 2 MUFU per sample.
    static constexpr float kArgScale = 2.885390081777927f;
    static __device__ __forceinline__ float tanh_pre(float a2) { return fmaf(-2.0f, rcp(1.0f + ex2(a2)), 1.0f); }
    // -sign(x) * |r|  in one LOP3: (r & 0x7fffffff) | (~x & 0x80000000)
    static __device__ __forceinline__ float neg_copysign(float r, float x) {
        unsigned o;
        asm("lop3.b32 %0, %1, 0x80000000, %2, 0x74;" : "=r"(o) : "r"(__float_as_uint(r)), "r"(__float_as_uint(x)));
        return __uint_as_float(o);
    }
    // MUFU-based variants for the once-per-sample target code of the hot kernels (rel. error ~1e-6)
    static __device__ __forceinline__ float fexp(float x) { return ex2(x * 1.4426950408889634f); }
    static __device__ __forceinline__ float flog(float x) { return lg2(x) * 0.6931471805599453f; }
    static __device__ __forceinline__ float flog1p(float x) { return lg2(1.0f + x) * 0.6931471805599453f; }   // x in [0,1]
    static __device__ __forceinline__ float fsqrt(float x) { float y; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
    // accurate variants
    static __device__ __forceinline__ float exp(float x) { return expf(x); }
    static __device__ __forceinline__ float logp(float x) { return logf(x); }
    static __device__ __forceinline__ float log1p(float x) { return log1pf(x); }
    static __device__ __forceinline__ float sin(float x) { return sinf(x); }
    static __device__ __forceinline__ float cos(float x) { return cosf(x); }
    static __device__ __forceinline__ float div(float a, float b) { return __fdividef(a, b); }
    static __device__ __forceinline__ float fabs(float x) { return fabsf(x); }
    static __device__ __forceinline__ float copysign(float m, float s) { return copysignf(m, s); }
    static __device__ __forceinline__ float fma(float a, float b, float c) { return fmaf(a, b, c); }
    static __device__ __forceinline__ float fmax(float a, float b) { return fmaxf(a, b); }
};

template <> struct Mth<double> {
    static __device__ __forceinline__ double rcp(double x) { return 1.0 / x; }
    static __device__ __forceinline__ double sqrt(double x) { return ::sqrt(x); }
    static __device__ __forceinline__ double tanh(double a) { return ::tanh(a); }
    static __device__ __forceinline__ double log(double x) { return ::log(x); }
    static constexpr double kArgScale = 1.0;
    static __device__ __forceinline__ double tanh_pre(double a) { return ::tanh(a); }
    static __device__ __forceinline__ double neg_copysign(double r, double x) { return ::copysign(r, -x); }
    static __device__ __forceinline__ double fexp(double x) { return ::exp(x); }
    static __device__ __forceinline__ double flog(double x) { return ::log(x); }
    static __device__ __forceinline__ double flog1p(double x) { return ::log1p(x); }
    static __device__ __forceinline__ double fsqrt(double x) { return ::sqrt(x); }
    static __device__ __forceinline__ double exp(double x) { return ::exp(x); }
    static __device__ __forceinline__ double logp(double x) { return ::log(x); }
    static __device__ __forceinline__ double log1p(double x) { return ::log1p(x); }
    static __device__ __forceinline__ double sin(double x) { return ::sin(x); }
    static __device__ __forceinline__ double cos(double x) { return ::cos(x); }
    static __device__ __forceinline__ double div(double a, double b) { return a / b; }
    static __device__ __forceinline__ double fabs(double x) { return ::fabs(x); }
    static __device__ __forceinline__ double copysign(double m, double s) { return ::copysign(m, s); }
    static __device__ __forceinline__ double fma(double a, double b, double c) { return ::fma(a, b, c); }
    static __device__ __forceinline__ double fmax(double a, double b) { return ::fmax(a, b); }
};

template <typename T>
__device__ __forceinline__ T shfl_xor(T v, int mask) { return __shfl_xor_sync(0xffffffffu, v, mask); }

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) v += shfl_xor(v, m);
    return v;
}

// Multi-value warp reduction ("halving butterfly").  On entry every lane holds V
// partial sums v[0..V).  On exit v[j], j < ceil(V/32), of lane L holds the warp-wide total of
// original index warp_slot<V>(L, j); other v[] entries are garbage.  At each of the
// five xor stages the surviving values are paired: the lane whose stage bit is
// set keeps the upper member of the pair and ships the lower one, so the count
// of live values halves while the number of lanes covered doubles.  Cost
// ~4 instructions per value instead of 10 for V independent butterflies, and the
// summation tree is fixed (deterministic).
template <typename T, int V>
__device__ __forceinline__ void warp_reduce_multi(T (&v)[V], int lane) {
    int n = V;
#pragma unroll
    for (int m = 16; m > 0; m >>= 1) {
        const bool up = (lane & m) != 0;
        const int h = (n + 1) / 2;
#pragma unroll
        for (int i = 0; i < V; ++i) {
            if (i < n - h) {                 // paired (i, i+h)
                T keep = up ? v[i + h] : v[i];
                T send = up ? v[i] : v[i + h];
                v[i] = keep + shfl_xor(send, m);
            } else if (i < h) {              // unpaired middle value: plain butterfly
                v[i] += shfl_xor(v[i], m);
            }
        }
        n = h;
    }
}

// Which original index ends up in v[0] of `lane` after warp_reduce_multi<T,V>, and whether `lane` is the
// lowest lane holding that index (the one that publishes it).  Evaluated at compile time into a 32-entry
// table: a kernel prologue only does one table lookup per thread.
template <int V>
struct WarpSlotTable {
    static constexpr int KEEP = (V + 31) / 32;      // values left per lane after the five stages (1 for V <= 32)
    int slot[32][KEEP];
    bool own[32][KEEP];
    constexpr WarpSlotTable() : slot{}, own{} {
        for (int lane = 0; lane < 32; ++lane) {
            int idx[V] = {};
            for (int i = 0; i < V; ++i) idx[i] = i;
            int n = V;
            for (int m = 16; m > 0; m >>= 1) {
                const bool up = (lane & m) != 0;
                const int h = (n + 1) / 2;
                for (int i = 0; i < n - h; ++i) idx[i] = up ? idx[i + h] : idx[i];
                n = h;
            }
            for (int j = 0; j < KEEP; ++j) slot[lane][j] = j < n ? idx[j] : -1;
        }
        for (int lane = 0; lane < 32; ++lane)
            for (int j = 0; j < KEEP; ++j) {
                bool first = slot[lane][j] >= 0;
                for (int l = 0; l < lane; ++l)
                    for (int jj = 0; jj < KEEP; ++jj)
                        if (slot[l][jj] == slot[lane][j]) first = false;
                own[lane][j] = first;
            }
    }
};
template <int V> __device__ constexpr WarpSlotTable<V> kWarpSlotTable{};

template <int V>
__device__ __forceinline__ int warp_slot(int lane, int j = 0) { return kWarpSlotTable<V>.slot[lane][j]; }
template <int V>
__device__ __forceinline__ bool warp_slot_owner(int lane, int j = 0) { return kWarpSlotTable<V>.own[lane][j]; }

}  // namespace nf
