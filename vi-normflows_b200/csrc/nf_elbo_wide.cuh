// Shared-parameter free energy for larger latent dimension (5 <= D <= 512), sm_100a.
//
// Same math as nf_elbo_impl.cuh (get_data.py:97-115) but a sample no longer fits one
// thread: L lanes (2..32) share a sample, each holding C=8 (or 16) coordinates of z and
// of the running gradient in registers.  The two dot products per layer (w.z forward,
// gz.s backward) are finished with log2(L) xor shuffles; everything that is a scalar per
// sample (tanh, h', the log-det factor, its reciprocal) is computed redundantly on the L
// lanes.  Gradient partials for w_k and s_k are lane-local (a lane owns its coordinates);
// after the S interleaved sample groups of a warp have passed layer k they are
// reduce-scattered across the 32/L sample groups of the warp and added to the warp's
// private row of the partial-sum workspace in global memory (L1/L2 resident, touched by
// one warp only => deterministic).  The row kernel in nf_elbo_impl.cuh then adds rows.
// Layer constants are read straight from global memory through L1 (every warp of the SM
// reads the same 2D+2 numbers per layer), activations t_k go to shared memory.
#pragma once
#include "nf_common.cuh"

namespace nf {

constexpr int kWideThreads = 128;
constexpr int kWideWarps = kWideThreads / kWarp;

__host__ __device__ constexpr int wide_vk(int D) { return 2 * ((D + 3) & ~3) + 4; }   // per-layer row stride

// sum over the L lanes that share a sample (xor masks 1 .. L/2); result on all of them
template <typename T, int L>
__device__ __forceinline__ T group_sum(T v) {
#pragma unroll
    for (int m = 1; m < L; m <<= 1) v += shfl_xor(v, m);
    return v;
}
// sum over the 32/L sample groups of the warp (xor masks L .. 16); result on all lanes
template <typename T, int L>
__device__ __forceinline__ T cross_group_sum(T v) {
#pragma unroll
    for (int m = L; m < 32; m <<= 1) v += shfl_xor(v, m);
    return v;
}

// Reduce-scatter of N values per lane across the sample groups: at xor mask M the lane whose
// bit is set keeps the upper half.  On exit v[0 .. N*L/32) holds slice number (lane / L) --
// read with bit 16 as the most significant bit -- of the group-summed array.
template <typename T, int N, int M, int L>
struct ReduceScatter {
    static __device__ __forceinline__ void run(T* v, int lane) {
        if constexpr (M >= L && N >= 2) {
            constexpr int H = N / 2;
            const bool up = (lane & M) != 0;
#pragma unroll
            for (int i = 0; i < H; ++i) {
                const T keep = up ? v[H + i] : v[i];
                const T send = up ? v[i] : v[H + i];
                v[i] = keep + shfl_xor(send, M);
            }
            ReduceScatter<T, H, M / 2, L>::run(v, lane);
        }
    }
};

template <typename T, int C, int L, int S, int TGT, bool BWD>
__global__ void __launch_bounds__(kWideThreads)
elbo_wide_kernel(const T* __restrict__ eps, long long N, int D, int K,
                 const T* __restrict__ w, const T* __restrict__ s,
                 const T* __restrict__ b, const T* __restrict__ c,
                 const T* __restrict__ tparams, int ntp, T epsv,
                 T* __restrict__ partials /* [gridDim.x * 4 warps][K*Vk + 2] */) {
    using M = Mth<T>;
    constexpr int G = 32 / L;                 // sample groups per warp
    constexpr int NV = 2 * C;                 // lane-local gradient values per layer (g_w | g_s)
    constexpr int KEEP = NV / G;              // values a lane owns after the reduce-scatter
    static_assert(NV >= G, "need at least one value per lane after the reduce-scatter");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* hist = reinterpret_cast<T*>(smem_raw);              // [K][S][warps][G]

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int l = lane % L, g = lane / L;
    const int d0 = l * C;                                  // first coordinate of this lane
    const int Dp = (D + 3) & ~3;
    const int Vk = 2 * Dp + 4;
    const T LOG2PI = (T)1.8378770664093453;
    T* row = partials + ((size_t)blockIdx.x * kWideWarps + warp) * ((size_t)K * Vk + 2);

    // target parameters for this lane's coordinates (GAUSSDIAG: mu, 1/var, const; MIXDIAG: per component)
    T g_mu[C], g_iv[C];
    T g_const = (T)0;
    if (TGT == NF_TARGET_GAUSSDIAG) {
        T ls_sum = (T)0;
#pragma unroll
        for (int i = 0; i < C; ++i) {
            const int d = d0 + i;
            const bool in = d < D;
            g_mu[i] = in ? tparams[d] : (T)0;
            const T ls = in ? tparams[D + d] : (T)0;
            g_iv[i] = in ? M::exp(-ls) : (T)0;                 // distributions.py:49 (D > 1 branch)
            ls_sum += ls;
        }
        ls_sum = group_sum<T, L>(ls_sum);
        g_const = (T)-0.5 * (D * LOG2PI + ls_sum);             // distributions.py:50
    }

    if (BWD) {   // the row is accumulated into: clear it first (workspace contents on entry are ignored)
        for (size_t i = lane; i < (size_t)K * Vk; i += 32) row[i] = (T)0;
        __syncwarp();
    }
    T sum_q = (T)0, sum_p = (T)0;
    const long long tile_samples = (long long)kWideWarps * S * G;
    const long long ntiles = (N + tile_samples - 1) / tile_samples;

    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        T z[S][C], lq[S], prod[S];
        bool ok[S];
#pragma unroll
        for (int j = 0; j < S; ++j) {
            const long long n = tile * tile_samples + ((long long)warp * S + j) * G + g;
            ok[j] = n < N;
            T zz = (T)0;
#pragma unroll
            for (int i = 0; i < C; ++i) {
                const int d = d0 + i;
                z[j][i] = (ok[j] && d < D) ? eps[n * D + d] : (T)0;
                zz = M::fma(z[j][i], z[j][i], zz);
            }
            zz = group_sum<T, L>(zz);
            lq[j] = (T)-0.5 * (D * LOG2PI + zz);
            prod[j] = (T)1;
        }
        // ---- forward ----  (layer constants are fetched one layer ahead: they come from L1/L2, and a
        // warp that owns only S sample groups has nothing else to hide that latency behind)
        T wn[C], sn[C];
#pragma unroll
        for (int i = 0; i < C; ++i) {
            const int d = d0 + i;
            wn[i] = d < D ? __ldg(w + d) : (T)0;
            sn[i] = d < D ? __ldg(s + d) : (T)0;
        }
        T bn = __ldg(b), cn = __ldg(c);
        for (int k = 0; k < K; ++k) {
            T wk[C], sk[C];
#pragma unroll
            for (int i = 0; i < C; ++i) { wk[i] = wn[i]; sk[i] = sn[i]; }
            const T bk = bn, ck = cn;
            if (k + 1 < K) {
#pragma unroll
                for (int i = 0; i < C; ++i) {
                    const int d = d0 + i;
                    wn[i] = d < D ? __ldg(w + (size_t)(k + 1) * D + d) : (T)0;
                    sn[i] = d < D ? __ldg(s + (size_t)(k + 1) * D + d) : (T)0;
                }
                bn = __ldg(b + k + 1); cn = __ldg(c + k + 1);
            }
#pragma unroll
            for (int j = 0; j < S; ++j) {
                T a = (T)0;
#pragma unroll
                for (int i = 0; i < C; ++i) a = M::fma(wk[i], z[j][i], a);
                a = group_sum<T, L>(a) + bk;
                const T t = M::tanh(a);
#pragma unroll
                for (int i = 0; i < C; ++i) z[j][i] = M::fma(sk[i], t, z[j][i]);
                if (BWD) {
                    if (l == 0) hist[((k * S + j) * kWideWarps + warp) * G + g] = t;
                } else {
                    const T hp = M::fma(-t, t, (T)1);
                    const T x = M::fma(hp, ck, (T)1);
                    prod[j] *= epsv + M::fabs(x);
                }
            }
            if (!BWD && ((k & 3) == 3 || k == K - 1)) {
#pragma unroll
                for (int j = 0; j < S; ++j) { lq[j] -= M::logp(prod[j]); prod[j] = (T)1; }
            }
        }
        if (BWD) __syncwarp();
        // ---- target ----
        T gz[S][C];
#pragma unroll
        for (int j = 0; j < S; ++j) {
            T lp;
            if (TGT == NF_TARGET_GAUSSDIAG) {
                T q = (T)0;
                T sc[C];
#pragma unroll
                for (int i = 0; i < C; ++i) {
                    const T df = z[j][i] - g_mu[i];
                    sc[i] = df * g_iv[i];
                    q = M::fma(df, sc[i], q);
                }
                q = group_sum<T, L>(q);
                const T dens = M::exp(g_const - (T)0.5 * q);
                const T den = epsv + dens;
                lp = M::logp(den);
                const T rho = M::div(dens, den);
#pragma unroll
                for (int i = 0; i < C; ++i) gz[j][i] = ok[j] ? rho * sc[i] : (T)0;
            } else {   // NF_TARGET_MIXDIAG: raw params [pi (Gc) | mu (Gc*D) | var (Gc*D)]
                const int Gc = ntp / (1 + 2 * D);
                T dens = (T)0;
                T dd[C];
#pragma unroll
                for (int i = 0; i < C; ++i) dd[i] = (T)0;
                for (int gc = 0; gc < Gc; ++gc) {
                    T q = (T)0, lsum = (T)0;
                    T sc[C];
#pragma unroll
                    for (int i = 0; i < C; ++i) {
                        const int d = d0 + i;
                        const bool in = d < D;
                        const T mu = in ? tparams[Gc + gc * D + d] : (T)0;
                        const T var = in ? tparams[Gc + Gc * D + gc * D + d] : (T)1;
                        const T df = z[j][i] - mu;
                        sc[i] = in ? M::div(df, var) : (T)0;
                        q = M::fma(df, sc[i], q);
                        lsum += M::logp(var);
                    }
                    q = group_sum<T, L>(q);
                    lsum = group_sum<T, L>(lsum);
                    const T e = tparams[gc] * M::exp((T)-0.5 * (D * LOG2PI + lsum + q));    // distributions.py:29-31
                    dens += e;
#pragma unroll
                    for (int i = 0; i < C; ++i) dd[i] = M::fma(e, sc[i], dd[i]);
                }
                const T den = epsv + dens;
                lp = M::logp(den);
                const T inv = M::div((T)1, den);
#pragma unroll
                for (int i = 0; i < C; ++i) gz[j][i] = ok[j] ? dd[i] * inv : (T)0;
            }
            if (l == 0 && ok[j]) sum_p += lp;
        }
        if (BWD) {
            // where this lane's KEEP reduced values live inside a layer record, and whether they are real coordinates
            int roff[KEEP];
            bool rok[KEEP];
#pragma unroll
            for (int i = 0; i < KEEP; ++i) {
                const int e = g * KEEP + i;                 // element of [g_w(C) | g_s(C)]; slice number == g
                const int d = d0 + (e < C ? e : e - C);
                rok[i] = d < D;
                roff[i] = e < C ? d : Dp + d;
            }
#pragma unroll
            for (int i = 0; i < C; ++i) {
                const int d = d0 + i;
                wn[i] = d < D ? __ldg(w + (size_t)(K - 1) * D + d) : (T)0;
                sn[i] = d < D ? __ldg(s + (size_t)(K - 1) * D + d) : (T)0;
            }
            cn = __ldg(c + K - 1);
            for (int k = K - 1; k >= 0; --k) {
                T wk[C], sk[C];
#pragma unroll
                for (int i = 0; i < C; ++i) { wk[i] = wn[i]; sk[i] = sn[i]; }
                const T ck = cn;
                if (k > 0) {
#pragma unroll
                    for (int i = 0; i < C; ++i) {
                        const int d = d0 + i;
                        wn[i] = d < D ? __ldg(w + (size_t)(k - 1) * D + d) : (T)0;
                        sn[i] = d < D ? __ldg(s + (size_t)(k - 1) * D + d) : (T)0;
                    }
                    cn = __ldg(c + k - 1);
                }
                // the running totals of this layer are read now and written back after the layer math
                T* rk = row + (size_t)k * Vk;
                T old[KEEP];
#pragma unroll
                for (int i = 0; i < KEEP; ++i) old[i] = rok[i] ? rk[roff[i]] : (T)0;
                T old_b = (T)0, old_c = (T)0;
                if (lane == 0) { old_b = rk[2 * Dp]; old_c = rk[2 * Dp + 1]; }
                T v[NV];
#pragma unroll
                for (int i = 0; i < NV; ++i) v[i] = (T)0;
                T gb = (T)0, gcc = (T)0;
#pragma unroll
                for (int j = 0; j < S; ++j) {
                    const T t = hist[((k * S + j) * kWideWarps + warp) * G + g];
                    T gt = (T)0;
#pragma unroll
                    for (int i = 0; i < C; ++i) {
                        z[j][i] = M::fma(-sk[i], t, z[j][i]);             // z_{k-1}
                        v[C + i] = M::fma(gz[j][i], t, v[C + i]);         // g_s
                        gt = M::fma(gz[j][i], sk[i], gt);
                    }
                    gt = group_sum<T, L>(gt);
                    const T hp = M::fma(-t, t, (T)1);
                    const T x = M::fma(hp, ck, (T)1);
                    const T y = epsv + M::fabs(x);
                    prod[j] *= y;
                    T gx = M::neg_copysign(M::rcp(y), x);
                    gx = ok[j] ? gx : (T)0;
                    gcc = M::fma(gx, hp, gcc);
                    gt = M::fma((T)-2 * ck * gx, t, gt);
                    const T ga = gt * hp;
                    gb += ga;
#pragma unroll
                    for (int i = 0; i < C; ++i) {
                        v[i] = M::fma(ga, z[j][i], v[i]);                 // g_w (direct)
                        gz[j][i] = M::fma(ga, wk[i], gz[j][i]);
                    }
                }
                if ((k & 3) == 0) {
#pragma unroll
                    for (int j = 0; j < S; ++j) { lq[j] -= M::logp(prod[j]); prod[j] = (T)1; }
                }
                // reduce over the sample groups of the warp and add to the warp's row
                ReduceScatter<T, NV, 16, L>::run(v, lane);
#pragma unroll
                for (int i = 0; i < KEEP; ++i)
                    if (rok[i]) rk[roff[i]] = old[i] + v[i];
                gb = cross_group_sum<T, L>(gb);
                gcc = cross_group_sum<T, L>(gcc);
                if (lane == 0) { rk[2 * Dp] = old_b + gb; rk[2 * Dp + 1] = old_c + gcc; }
            }
        }
#pragma unroll
        for (int j = 0; j < S; ++j)
            if (l == 0 && ok[j]) sum_q += lq[j];
        if (BWD) __syncwarp();
    }
    sum_q = warp_sum(sum_q);
    sum_p = warp_sum(sum_p);
    if (lane == 0) {
        row[(size_t)K * Vk] = sum_q;
        row[(size_t)K * Vk + 1] = sum_p;
    }
}

}  // namespace nf
