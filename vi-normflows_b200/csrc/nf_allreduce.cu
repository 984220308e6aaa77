// One-shot all-reduce of a small float vector over NVLink peer memory (sm_100a, <= 8 GPUs of one box).
//
// The free-energy step ends in a 3+2KD+2K-float result vector per GPU (195 floats at the headline
// config).  A ring/tree collective is pure latency at that size; here every rank PUSHES its vector into a
// slot of every peer's exchange buffer (plain stores through the NVLink aperture), publishes a sequence
// flag with system scope, waits for the peers' flags and adds the world slots in rank order -- the same
// order on every rank, so all ranks hold bitwise identical sums.  Two slot sets alternate by sequence
// number: a rank can run at most one step ahead of the slowest peer, so a set is never overwritten while a
// peer may still read it.  The spin is bounded; on expiry the kernel raises `status[0]` instead of hanging.
//
// Buffer layout per rank (floats): set s in {0,1}: [world][stride] data, then flags as uint64 [2][world].
#include "nf_common.cuh"

namespace nf {

constexpr int kArMaxWorld = 8;
constexpr int kArStride = 1024;          // floats per slot
struct ArPeers { float* buf[kArMaxWorld]; };

__host__ __device__ inline size_t ar_flag_offset_floats(int world) { return (size_t)2 * world * kArStride; }

__global__ void __launch_bounds__(256)
allreduce_small_kernel(ArPeers peers, int world, int rank, const float* __restrict__ src, float* __restrict__ dst, int n,
                       unsigned long long seq, int* __restrict__ status) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int set = (int)(seq & 1);
    const int tid = threadIdx.x;
    // 1. push my vector into slot `rank` of every peer (including myself)
    for (int p = 0; p < world; ++p) {
        float* slot = peers.buf[p] + ((size_t)set * world + rank) * kArStride;
        for (int i = tid; i < n; i += blockDim.x) slot[i] = src[i];
    }
    __threadfence_system();
    __syncthreads();
    // 2. publish: flag[set][rank] on every peer := seq
    if (tid < world) {
        unsigned long long* flags = reinterpret_cast<unsigned long long*>(peers.buf[tid] + ar_flag_offset_floats(world));
        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(flags + (size_t)set * world + rank), "l"(seq) : "memory");
    }
    // 3. wait for every peer's flag in MY buffer
    __shared__ int failed;
    if (tid == 0) failed = 0;
    __syncthreads();
    if (tid < world) {
        const unsigned long long* f = reinterpret_cast<const unsigned long long*>(peers.buf[rank] + ar_flag_offset_floats(world)) +
                                      (size_t)set * world + tid;
        unsigned long long v = 0;
        long long spins = 0;
        do {
            asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(f) : "memory");
        } while (v < seq && ++spins < 400000000LL);
        if (v < seq) failed = 1;
    }
    __syncthreads();
    if (failed) {
        if (tid == 0) status[0] = 1;
        return;
    }
    // 4. add the slots in rank order
    const float* mine = peers.buf[rank] + (size_t)set * world * kArStride;
    for (int i = tid; i < n; i += blockDim.x) {
        float s = 0.f;
        for (int r = 0; r < world; ++r) s += __ldcv(mine + (size_t)r * kArStride + i);
        dst[i] = s;
    }
}

}  // namespace nf

extern "C" {

size_t nf_allreduce_small_buffer_bytes(int world) {
    if (world < 1 || world > nf::kArMaxWorld) return 0;
    return nf::ar_flag_offset_floats(world) * sizeof(float) + (size_t)2 * world * sizeof(unsigned long long);
}
int nf_allreduce_small_max_count(void) { return nf::kArStride; }

int nf_allreduce_small_f32(void* const* peer_buffers, int world, int rank, const float* src, float* dst, int n,
                           uint64_t seq, int* status, void* stream) {
    if (!peer_buffers || world < 1 || world > nf::kArMaxWorld || rank < 0 || rank >= world || !src || !dst || !status) return NF_ERR_INVALID;
    if (n < 0 || n > nf::kArStride || seq == 0) return NF_ERR_INVALID;
    nf::ArPeers peers;
    for (int i = 0; i < nf::kArMaxWorld; ++i) peers.buf[i] = i < world ? reinterpret_cast<float*>(peer_buffers[i]) : nullptr;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(1); cfg.blockDim = dim3(256); cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    NF_CUDA_OK(cudaLaunchKernelEx(&cfg, nf::allreduce_small_kernel, peers, world, rank, src, dst, n, (unsigned long long)seq, status));
    return NF_OK;
}

}  // extern "C"
