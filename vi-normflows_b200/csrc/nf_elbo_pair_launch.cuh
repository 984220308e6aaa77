// Launchers of the packed-pair free-energy kernel (nf_elbo_pair.cuh).  The instantiations are spread over several
// translation units (nf_elbo_pair_*.cu) so that they compile in parallel.
#pragma once
#include <cstdio>
#include <cstdlib>
#include "nf_launch.cuh"
#include "nf_elbo_pair.cuh"

namespace nf {

struct PairCfg { int nw, smax, st, sg, flags; unsigned long long* stamps; };
const PairCfg& pair_override();

#define NF_PAIR_ARGS const float* eps, long long N, int K, const float* w, const float* s, const float* b, const float* c, \
                     const float* tp, int ntp, float epsv, float* partials, pr::SyncHeader* sync, float* out_F, float* g_w, \
                     float* g_s, float* g_b, float* g_c, int* nb, cudaStream_t st, const pr::ArDesc& ar, const RngDesc& rng, void* stash, size_t stash_bytes
#define NF_PAIR_PASS eps, N, K, w, s, b, c, tp, ntp, epsv, partials, sync, out_F, g_w, g_s, g_b, g_c, nb, st, ar, rng, stash, stash_bytes

// bytes of L2-resident history a launch with SG samples per thread in global memory needs (grid <= `sms` blocks)
static inline size_t pair_stash_bytes(int sms, int nt, int K, int sg) { return (size_t)sms * nt * K * (sg / 2) * sizeof(unsigned long long); }

template <int D, int NW, int SMAX, int ST, int SG, int TGT>
static int launch_pair(const float* eps, long long N, int K, const float* w, const float* s, const float* b, const float* c,
                       const float* tparams, int ntp, float epsv, float* partials, pr::SyncHeader* sync, float* out_F,
                       float* g_w, float* g_s, float* g_b, float* g_c, int* nblocks_out, cudaStream_t stream, const pr::ArDesc& ar, const RngDesc& rng,
                       void* stash, size_t stash_bytes) {
    auto kern = pr::elbo_pair_kernel<D, NW, SMAX, ST, SG, TGT>;
    const int ntp_s = target_smem_count(TGT, D, ntp);
    const pr::Smem L(K, D, (SMAX - ST - SG) / 2, ntp_s, NW * 32);
    const size_t smem = (size_t)L.total * sizeof(float);
    if (smem > 227 * 1024 - 256 || ((NW + 3) / 4) * ST * K > 512) return NF_ERR_UNSUPPORTED;
    int dev = 0;
    NF_CUDA_OK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return NF_ERR_UNSUPPORTED;
    static size_t smem_set[kMaxDevices] = {};          // per instantiation and device: the attribute is raised once
    static std::mutex mu;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (smem_set[dev] < smem) {
            NF_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            smem_set[dev] = smem;
        }
    }
    const DevInfo di = dev_info();
    if (SG > 0 && (!stash || stash_bytes < pair_stash_bytes(di.sms, NW * 32, K, SG))) return NF_ERR_WORKSPACE;
    long long grid = di.sms;
    const long long tiles = (N + (long long)NW * 32 * 2 - 1) / ((long long)NW * 32 * 2);
    if (grid > tiles) grid = tiles;
    if (grid < 1) grid = 1;
    NF_CUDA_OK(launch_pdl(kern, dim3((unsigned)grid), dim3(NW * 32), smem, stream, eps, N, K, w, s, b, c, tparams, ntp, epsv, partials,
                          sync, out_F, g_w, g_s, g_b, g_c, pair_override().flags, pair_override().stamps, ar, rng, reinterpret_cast<unsigned long long*>(stash)));
    *nblocks_out = (int)grid;
    return NF_OK;
}


// production configurations by chain depth.  With room for an L2-resident part of the history in the caller's workspace
// (nf_planar_elbo_workspace_bytes_n) K <= 32 runs 8 samples per thread (4 in tensor memory, 2 in shared memory, 2 in L2):
// 2^20 samples are two rounds instead of three (measured 61.8 vs 62.8 us per step; 14 warps x 8 samples: 72.7).
template <int D, int TGT>
static int dispatch_pair(NF_PAIR_ARGS) {
    if (K <= 16) return launch_pair<D, 16, 8, 4, 0, TGT>(NF_PAIR_PASS);
    if (K <= 32) {
        const int rc = launch_pair<D, 16, 8, 4, 2, TGT>(NF_PAIR_PASS);
        if (rc != NF_ERR_WORKSPACE) return rc;
        return launch_pair<D, 16, 6, 4, 0, TGT>(NF_PAIR_PASS);
    }
    return launch_pair<D, 8, 6, 4, 0, TGT>(NF_PAIR_PASS);
}

// one function per translation unit; NF_ERR_UNSUPPORTED when the target is not one of the unit's
int pair_group_ring(int tgt, int D, NF_PAIR_ARGS);       // U1, TRIAL1
int pair_group_u2(int tgt, int D, NF_PAIR_ARGS);         // U2
int pair_group_ridges(int tgt, int D, NF_PAIR_ARGS);     // U3, U4
int pair_group_gauss(int tgt, int D, NF_PAIR_ARGS);      // MIXDIAG, GAUSSDIAG (D = 1, 2)
int pair_group_sweep(NF_PAIR_ARGS);                      // U1 with the NF_PAIR_CFG configuration

}  // namespace nf
