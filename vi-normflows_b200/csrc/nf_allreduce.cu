// One-shot all-reduce of a small float vector over NVLink peer memory (sm_100a, <= 8 GPUs of one box).
//
// The free-energy step ends in a 3+2KD+2K-float result vector per GPU (195 floats at the headline
// config).  A ring/tree collective is pure latency at that size; the exchange here is one NVLink crossing:
// see ArDesc / ar_exchange in nf_elbo_pair.cuh (values pushed into every peer's buffer in 8-byte words that
// carry the step's sequence number; receivers poll their own buffer and add in rank order, so all ranks hold
// bitwise identical sums; two alternating slot sets; bounded wait).
//
// The packed-pair free-energy kernel does this exchange in its last block; the kernel here is the same
// exchange as a launch of its own, for the shapes that kernel does not cover.
#include "nf_elbo_pair.cuh"

namespace nf {

using pr::ArDesc;
using pr::kArMaxWorld;
using pr::kArStride;

// src[0..n) (+ n_hi, n_lo when ar.xout != NULL) -> dst[0..n) = sums over ranks, ar.xout = {sum n_hi, sum n_lo, status};
// status_i (legacy interface) is raised to 1 when a peer did not arrive.
__global__ void __launch_bounds__(512)
allreduce_small_kernel(const ArDesc ar, const float* src, float* dst /* may be src */, int n, int* __restrict__ status_i) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    __shared__ float vec[kArStride];
    __shared__ int s_flag;
    const int tid = threadIdx.x, NT = blockDim.x;
    const int nx = n + (ar.xout ? 2 : 0);
    for (int i = tid; i < nx; i += NT) vec[i] = i < n ? src[i] : (i == n ? ar.n_hi : ar.n_lo);
    __syncthreads();
    const bool ok = pr::ar_exchange(ar, vec, nx, tid, NT, &s_flag);
    if (!ok) {
        if (tid == 0) {
            if (status_i) status_i[0] = 1;
            if (ar.xout) ar.xout[2] = 1.f;
            if (ar.sticky_status) ar.sticky_status[0] = 1.f;
        }
        return;
    }
    for (int i = tid; i < nx; i += NT) {
        if (i < n) dst[i] = vec[i];
        else ar.xout[i - n] = vec[i];
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
    ar->sticky_status = nullptr;
    return NF_OK;
}

}  // namespace nf

extern "C" {

size_t nf_allreduce_small_buffer_bytes(int world) {
    if (world < 1 || world > nf::kArMaxWorld) return 0;
    return nf::pr::ar_buffer_bytes(world);
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
