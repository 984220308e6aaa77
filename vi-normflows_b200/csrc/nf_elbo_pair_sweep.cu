// Instantiations of the packed-pair free-energy kernel (see nf_elbo_pair_launch.cuh).
#include "nf_elbo_pair_launch.cuh"

namespace nf {

// tuning knob NF_PAIR_CFG="warps,samples per thread,of which in tensor memory,of which in global memory" (read once); NF_PAIR=0 switches the
// packed-pair kernel off (the launch then takes the scalar kernel of nf_elbo_impl.cuh)
const PairCfg& pair_override() {
    static const PairCfg cfg = []() {
        PairCfg c = {0, 0, 0, 0, 0, nullptr};
        if (const char* e = getenv("NF_PAIR_CFG")) sscanf(e, "%d,%d,%d,%d", &c.nw, &c.smax, &c.st, &c.sg);
        if (const char* e = getenv("NF_PAIR_FLAGS")) c.flags = atoi(e);
        if (c.flags & 2) {      // tuning aid: phase timestamps of every launch (of the device current at the first call)
            if (cudaMalloc(&c.stamps, sizeof(unsigned long long) * 256 * 8) != cudaSuccess) c.stamps = nullptr;
            else cudaMemset(c.stamps, 0, sizeof(unsigned long long) * 256 * 8);
        }
        if (const char* e = getenv("NF_PAIR")) if (e[0] == '0') c.nw = -1;
        return c;
    }();
    return cfg;
}

size_t pair_stash_bytes_max(int K) {
    const DevInfo di = dev_info();
    return pair_stash_bytes(di.sms > 0 ? di.sms : 148, 512, K, 2);
}

bool pair_applicable(size_t elem, bool bwd, long long N, int D, int K) {
    return elem == 4 && bwd && D >= 1 && D <= 2 && K <= 64 && N >= (long long)148 * 128 * 8 && pair_override().nw >= 0;
}

int pair_group_sweep(NF_PAIR_ARGS) {
    const PairCfg& o = pair_override();
#define NF_SW(NW_, SM_, ST_, SG_) if (o.nw == NW_ && o.smax == SM_ && o.st == ST_ && o.sg == SG_) return launch_pair<2, NW_, SM_, ST_, SG_, NF_TARGET_U1>(NF_PAIR_PASS)
    NF_SW(16, 6, 4, 0);
    NF_SW(14, 8, 4, 2);
    NF_SW(16, 8, 4, 2);
    NF_SW(12, 8, 4, 2);
    NF_SW(12, 8, 4, 0);
    NF_SW(16, 4, 4, 0);
    NF_SW(8, 6, 4, 0);
    NF_SW(8, 8, 4, 2);
    NF_SW(12, 6, 4, 0);
#undef NF_SW
    return NF_ERR_UNSUPPORTED;
}

int pair_elbo(int tgt, const float* eps, long long N, int D, int K, const float* w, const float* s, const float* b,
              const float* c, const float* tp, int ntp, float epsv, float* partials, pr::SyncHeader* sync,
              float* out_F, float* g_w, float* g_s, float* g_b, float* g_c, int* nb, cudaStream_t st, const pr::ArDesc& ar, const RngDesc& rng, void* stash, size_t stash_bytes) {
    if (tgt == NF_TARGET_U1 && D == 2 && pair_override().nw > 0) {
        const int rc = pair_group_sweep(NF_PAIR_PASS);
        if (rc != NF_ERR_UNSUPPORTED) return rc;
    }
    int rc = pair_group_ring(tgt, D, NF_PAIR_PASS);
    if (rc == NF_ERR_UNSUPPORTED) rc = pair_group_u2(tgt, D, NF_PAIR_PASS);
    if (rc == NF_ERR_UNSUPPORTED) rc = pair_group_ridges(tgt, D, NF_PAIR_PASS);
    if (rc == NF_ERR_UNSUPPORTED) rc = pair_group_gauss(tgt, D, NF_PAIR_PASS);
    return rc;
}

}  // namespace nf

// tuning aid: phase timestamps of the last packed-pair launch made with NF_PAIR_FLAGS bit 1 (see nf_elbo_pair.cuh)
extern "C" int nf_debug_pair_stamps(unsigned long long* out /*[256][8]*/) {
    if (!nf::pair_override().stamps) return NF_ERR_UNSUPPORTED;
    NF_CUDA_OK(cudaMemcpy(out, nf::pair_override().stamps, sizeof(unsigned long long) * 256 * 8, cudaMemcpyDeviceToHost));
    return NF_OK;
}
