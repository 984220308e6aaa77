// Host-side launch helpers shared by the translation units of libnfcore.
#pragma once
#include <cstdlib>
#include <mutex>
#include "nf_common.cuh"

namespace nf {

constexpr int kLogEvery = 4;               // layers per MUFU.LG2 (product of 4 factors >= eps^4 = 1e-28: no underflow)
constexpr size_t kSyncHeaderBytes = 256;   // tail of the workspace: ticket of the single-launch kernel (nf_elbo_pair.cuh)
constexpr int kMaxDevices = 64;

// Launch with the programmatic-stream-serialization attribute: the kernel's launch latency and prologue
// overlap the tail of its predecessor; the kernel itself waits (griddepcontrol.wait) before touching memory.
template <typename... KArgs, typename... Args>
static inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    static const bool plain = getenv("NF_NO_PDL") != nullptr;      // tuning knob: ordinary stream-ordered launches
    cfg.attrs = attr; cfg.numAttrs = plain ? 0 : 1;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

struct DevInfo { int sms; int ok; };
static inline DevInfo dev_info() {
    static thread_local int cached_dev = -1;
    static thread_local DevInfo info = {0, 0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return {0, 0};
    if (dev != cached_dev) {
        int sms = 0, major = 0;
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return {0, 0};
        if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return {0, 0};
        info.sms = sms;
        info.ok = (major == 10);
        cached_dev = dev;
    }
    return info;
}


}  // namespace nf
