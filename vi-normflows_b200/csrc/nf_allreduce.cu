// One-shot all-reduce of a small float vector over NVLink peer memory (sm_100a, <= 8 GPUs of one box).
//
// The free-energy step ends in a 3+2KD+2K-float result vector per GPU (195 floats at the headline
// config).  A ring/tree collective is pure latency at that size; here every rank PUSHES its vector into a
// slot of every peer's exchange buffer (plain stores through the NVLink aperture), publishes a sequence
// flag with system scope, waits for the peers' flags and adds the world slots in rank order -- the same
// order on every rank, so all ranks hold bitwise identical sums.  Two slot sets alternate by sequence
// number: a rank can run at most one step ahead of the slowest peer, so a set is never overwritten while a
// peer may still read it.  The wait is bounded (20 s); on expiry the kernel reports it instead of hanging.
//
// The packed-pair free-energy kernel does this exchange in its last block (nf_elbo_pair.cuh, ArDesc); the
// kernel here is the same exchange as a launch of its own, for the shapes that kernel does not cover.
//
// Buffer layout per rank (floats): set s in {0,1}: [world][kArStride] data, then flags as uint64 [2][world].
#include "nf_elbo_pair.cuh"

namespace nf {

using pr::ArDesc;
using pr::kArMaxWorld;
using pr::kArStride;
using pr::ar_flag_offset_floats;

// src[0..n) (+ n_hi, n_lo when ar.xout != NULL) -> dst[0..n) = sums over ranks, ar.xout = {sum n_hi, sum n_lo, status};
// status_i (legacy interface) is raised to 1 when a peer did not arrive.
__global__ void __launch_bounds__(512)
allreduce_small_kernel(const ArDesc ar, const float* src, float* dst /* may be src */, int n, int* __restrict__ status_i) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int set = (int)(ar.seq & 1);
    const int tid = threadIdx.x, NT = blockDim.x;
    const int nx = n + (ar.xout ? 2 : 0);
    // 1. push my vector into slot `rank` of every peer (including myself): NT / world threads per peer
    {
        const int per_peer = NT / ar.world, p = tid / per_peer, t = tid - p * per_peer;
        if (p < ar.world) {
            float* slot = ar.buf[p] + ((size_t)set * ar.world + ar.rank) * kArStride;
            for (int i = t; i < nx; i += per_peer) slot[i] = i < n ? src[i] : (i == n ? ar.n_hi : ar.n_lo);
        }
    }
    __threadfence_system();
    __syncthreads();
    // 2. publish: flag[set][rank] on every peer := seq
    if (tid < ar.world) {
        unsigned long long* flags = reinterpret_cast<unsigned long long*>(ar.buf[tid] + ar_flag_offset_floats(ar.world));
        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(flags + (size_t)set * ar.world + ar.rank), "l"(ar.seq) : "memory");
    }
    // 3. wait for every peer's flag in MY buffer
    __shared__ int failed;
    if (tid == 0) failed = 0;
    __syncthreads();
    if (tid < ar.world) {
        const unsigned long long* f = reinterpret_cast<const unsigned long long*>(ar.buf[ar.rank] + ar_flag_offset_floats(ar.world)) +
                                      (size_t)set * ar.world + tid;
        unsigned long long v = 0;
        const unsigned long long t0 = pr::gtime();
        do {
            asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(f) : "memory");
        } while (v < ar.seq && pr::gtime() - t0 < 20000000000ull);
        if (v < ar.seq) failed = 1;
    }
    __syncthreads();
    if (failed) {
        if (tid == 0) {
            if (status_i) status_i[0] = 1;
            if (ar.xout) ar.xout[2] = 1.f;
        }
        return;
    }
    // 4. add the slots in rank order
    const float* mine = ar.buf[ar.rank] + (size_t)set * ar.world * kArStride;
    for (int i = tid; i < nx; i += NT) {
        float a = 0.f;
        for (int r = 0; r < ar.world; ++r) a += __ldcv(mine + (size_t)r * kArStride + i);
        if (i < n) dst[i] = a;
        else ar.xout[i - n] = a;
    }
    if (tid == 0 && ar.xout) ar.xout[2] = 0.f;
}

int launch_allreduce_small(const ArDesc& ar, const float* src, float* dst, int n, int* status_i, cudaStream_t stream) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(1); cfg.blockDim = dim3(512); cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    NF_CUDA_OK(cudaLaunchKernelEx(&cfg, allreduce_small_kernel, ar, src, dst, n, status_i));
    return NF_OK;
}

int make_ar_desc(ArDesc* ar, void* const* peer_buffers, int world, int rank, uint64_t seq, int64_t n_local, float* xout) {
    if (!peer_buffers || world < 1 || world > kArMaxWorld || rank < 0 || rank >= world || seq == 0 || n_local < 0) return NF_ERR_INVALID;
    for (int i = 0; i < kArMaxWorld; ++i) ar->buf[i] = i < world ? reinterpret_cast<float*>(peer_buffers[i]) : nullptr;
    for (int i = 0; i < world; ++i) if (!ar->buf[i]) return NF_ERR_INVALID;
    ar->world = world; ar->rank = rank; ar->seq = (unsigned long long)seq;
    ar->n_hi = (float)(n_local >> 20); ar->n_lo = (float)(n_local & ((1 << 20) - 1));      // both exact; so are their sums over 8 ranks
    ar->xout = xout;
    return NF_OK;
}

}  // namespace nf

extern "C" {

size_t nf_allreduce_small_buffer_bytes(int world) {
    if (world < 1 || world > nf::kArMaxWorld) return 0;
    return nf::ar_flag_offset_floats(world) * sizeof(float) + (size_t)2 * world * sizeof(unsigned long long);
}
int nf_allreduce_small_max_count(void) { return nf::kArStride - 2; }

int nf_allreduce_small_f32(void* const* peer_buffers, int world, int rank, const float* src, float* dst, int n,
                           uint64_t seq, int* status, void* stream) {
    if (!src || !dst || !status || n < 0 || n > nf::kArStride - 2) return NF_ERR_INVALID;
    nf::ArDesc ar;
    const int rc = nf::make_ar_desc(&ar, peer_buffers, world, rank, seq, 0, nullptr);
    if (rc != NF_OK) return rc;
    return nf::launch_allreduce_small(ar, src, dst, n, status, (cudaStream_t)stream);
}

/* In-place form that also carries the caller's sample count: vec[0..n) := sums over ranks, vec[n..n+3) := n_hi, n_lo
 * (global count = n_hi * 2^20 + n_lo) and status (0, or 1 when a peer did not arrive). */
int nf_allreduce_small_counted_f32(void* const* peer_buffers, int world, int rank, float* vec, int n, int64_t n_local,
                                   uint64_t seq, void* stream) {
    if (!vec || n < 0 || n > nf::kArStride - 2) return NF_ERR_INVALID;
    nf::ArDesc ar;
    const int rc = nf::make_ar_desc(&ar, peer_buffers, world, rank, seq, n_local, vec + n);
    if (rc != NF_OK) return rc;
    return nf::launch_allreduce_small(ar, vec, vec, n, nullptr, (cudaStream_t)stream);
}

}  // extern "C"
