// float32 instantiation of the shared-parameter free-energy kernels + the C entry points
// that do not depend on the element type.
#include "nf_elbo_impl.cuh"

extern "C" {

size_t nf_planar_elbo_workspace_bytes(int K, int D, int elem_size) {
    if (K < 1 || D < 1 || elem_size < 1) return 0;
    if (nf::elbo_thread_per_sample(D)) {
        const size_t row = (size_t)K * (2 * D + 2) + 2;
        return row * (size_t)nf::max_grid_blocks() * (size_t)elem_size + nf::kSyncHeaderBytes;
    }
    const size_t row = (size_t)K * nf::wide_vk(D) + 2;      // one row per warp of the wide kernel
    const nf::DevInfo di = nf::dev_info();
    return row * (size_t)((di.sms > 0 ? di.sms : 148) * nf::wide_blocks_per_sm(K, D, (size_t)elem_size) * nf::kWideWarps) * (size_t)elem_size +
           nf::kSyncHeaderBytes;
}

size_t nf_planar_elbo_workspace_bytes_n(int64_t N, int K, int D, int elem_size) {
    size_t need = nf_planar_elbo_workspace_bytes(K, D, elem_size);
    if (need && elem_size == 4 && nf::gram_applicable(N, D, K, -1, 0)) {
        const size_t g = nf::gram_workspace_bytes(N, D, K);
        if (g > need) need = g;
    }
    // packed-pair kernel: room behind the rows for the L2-resident part of the tanh history (8 samples per thread)
    if (need && nf::pair_applicable((size_t)elem_size, true, N, D, K)) need = ((need + 255) & ~(size_t)255) + nf::pair_stash_bytes_max(K);
    return need;
}

int nf_planar_elbo_fwd_bwd_f32(const float* eps, int64_t N, int D, int K, const float* w, const float* s,
                               const float* b, const float* c, int target_id, const float* target_params,
                               int n_target_params, float log_eps, float* out_F, float* g_w, float* g_s,
                               float* g_b, float* g_c, void* workspace, size_t workspace_bytes, void* stream) {
    // wide D with a workspace large enough for the intermediates of a chunk: matrix-product path
    if (out_F && workspace && nf::gram_applicable(N, D, K, target_id, n_target_params) &&
        workspace_bytes >= nf::gram_workspace_bytes(N, D, K) &&
        ((reinterpret_cast<uintptr_t>(workspace) | reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(s)) & 15) == 0) {
        const bool any_g = g_w || g_s || g_b || g_c, all_g = g_w && g_s && g_b && g_c;
        if (any_g && !all_g) return NF_ERR_INVALID;
        const int rc = nf::elbo_validate<float>(eps, N, D, K, w, s, b, c, target_id, target_params, n_target_params);
        if (rc != NF_OK) return rc;
        return nf::gram_elbo_f32(all_g, eps, N, D, K, w, s, b, c, target_id, target_params, n_target_params, log_eps, out_F,
                                 g_w, g_s, g_b, g_c, workspace, (cudaStream_t)stream);
    }
    return nf::elbo_entry<float>(eps, N, D, K, w, s, b, c, target_id, target_params, n_target_params, log_eps,
                                 out_F, g_w, g_s, g_b, g_c, workspace, workspace_bytes, stream);
}

// sample-sharded form: the local sums are exchanged over NVLink peer memory before they reach `out`
int nf_planar_elbo_fwd_bwd_exchange_f32(const float* eps, int64_t N, int D, int K, const float* w, const float* s,
                                        const float* b, const float* c, int target_id, const float* target_params,
                                        int n_target_params, float log_eps, float* out, void* workspace, size_t workspace_bytes,
                                        void* const* peer_buffers, int world, int rank, uint64_t seq, void* stream) {
    if (!out || !workspace) return NF_ERR_INVALID;
    const int nvec = 3 + 2 * K * D + 2 * K;
    if (nvec + 2 > nf::pr::kArStride) return NF_ERR_UNSUPPORTED;
    int rc = nf::elbo_validate<float>(eps, N, D, K, w, s, b, c, target_id, target_params, n_target_params);
    if (rc != NF_OK) return rc;
    if (workspace_bytes < nf_planar_elbo_workspace_bytes(K, D, 4)) return NF_ERR_WORKSPACE;
    nf::pr::ArDesc ar;
    rc = nf::make_ar_desc(&ar, peer_buffers, world, rank, seq, N, out + nvec);
    if (rc != NF_OK) return rc;
    float *g_w = out + 3, *g_s = g_w + K * D, *g_b = g_s + K * D, *g_c = g_b + K;
    bool exchanged = false;
    rc = nf::elbo_fused<float>(true, eps, N, D, K, w, s, b, c, target_id, target_params, n_target_params, log_eps, out, g_w, g_s, g_b,
                               g_c, reinterpret_cast<float*>(workspace), (cudaStream_t)stream, &ar, &exchanged, nullptr, workspace_bytes);
    if (rc != NF_OK || exchanged) return rc;
    return nf::launch_allreduce_small(ar, out, out, nvec, nullptr, (cudaStream_t)stream);
}

// the samples are drawn inside the kernel (nf_philox.cuh): nothing is read from memory but the layer constants
int nf_planar_elbo_fwd_bwd_rng_f32(uint64_t seed, uint64_t step, int64_t N, int D, int K, const float* w, const float* s,
                                   const float* b, const float* c, int target_id, const float* target_params, int n_target_params,
                                   float log_eps, float* out_F, float* g_w, float* g_s, float* g_b, float* g_c, void* workspace,
                                   size_t workspace_bytes, void* stream) {
    if (!out_F || !workspace || !g_w || !g_s || !g_b || !g_c) return NF_ERR_INVALID;
    if (D != 1 && D != 2) return NF_ERR_UNSUPPORTED;
    // validation of everything but the (absent) sample pointer
    const int rc = nf::elbo_validate<float>(reinterpret_cast<const float*>(workspace), N, D, K, w, s, b, c, target_id, target_params,
                                            n_target_params);
    if (rc != NF_OK) return rc;
    if (workspace_bytes < nf_planar_elbo_workspace_bytes(K, D, 4)) return NF_ERR_WORKSPACE;
    const nf::RngDesc rng = {seed, step};
    return nf::elbo_fused<float>(true, nullptr, N, D, K, w, s, b, c, target_id, target_params, n_target_params, log_eps, out_F, g_w,
                                 g_s, g_b, g_c, reinterpret_cast<float*>(workspace), (cudaStream_t)stream, nullptr, nullptr, &rng, workspace_bytes);
}

// ---- host-buffer sessions (see nfcore.h) ----
int nf_session_create(nf_session** out, int64_t max_samples, int D, int K, int elem_size) {
    return nf::session_create(reinterpret_cast<nf::Session**>(out), max_samples, D, K, elem_size);
}
void nf_session_destroy(nf_session* s) { nf::session_free(reinterpret_cast<nf::Session*>(s)); }
int nf_session_elbo_host_f32(nf_session* s, const float* eps_host, int64_t N, const float* layer_consts_host, int target_id,
                             const float* target_params_host, int n_target_params, float log_eps, int backward,
                             float* out_host) {
    return nf::session_elbo_host<float>(reinterpret_cast<nf::Session*>(s), eps_host, N, layer_consts_host, target_id,
                                        target_params_host, n_target_params, log_eps, backward, out_host);
}

int nf_session_objective_host_f32(nf_session* s, const float* eps_host, int64_t N, const double* params, int semantics, int target_id,
                                  const float* target_params_host, int n_target_params, float log_eps, int backward, double* F_parts,
                                  double* grad) {
    return nf::session_objective_host<float>(reinterpret_cast<nf::Session*>(s), eps_host, N, params, semantics, target_id,
                                              target_params_host, n_target_params, log_eps, backward, F_parts, grad);
}

int nf_session_objective_host_exchange_f32(nf_session* s, const float* eps_host, int64_t N_local, const double* params, int semantics,
                                           int target_id, const float* target_params_host, int n_target_params, float log_eps,
                                           void* const* peer_buffers, int world, int rank, uint64_t seq, double* F_parts,
                                           double* grad, int64_t* n_global) {
    nf::pr::ArDesc ar;
    const int rc = nf::make_ar_desc(&ar, peer_buffers, world, rank, seq, N_local, nullptr);
    if (rc != NF_OK) return rc;
    return nf::session_objective_host<float>(reinterpret_cast<nf::Session*>(s), eps_host, N_local, params, semantics, target_id,
                                             target_params_host, n_target_params, log_eps, 1, F_parts, grad, &ar, n_global);
}

int nf_session_objective_rng_f32(nf_session* s, uint64_t seed, uint64_t step, int64_t N, const double* params, int semantics,
                                 int target_id, const float* target_params_host, int n_target_params, float log_eps, double* F_parts,
                                 double* grad) {
    return nf::session_objective_rng(reinterpret_cast<nf::Session*>(s), seed, step, N, params, semantics, target_id, target_params_host,
                                     n_target_params, log_eps, F_parts, grad);
}

int nf_script_train_f32(const float* eps_all, int64_t n_iters, int64_t N, int D, int K, float* params, int target_id,
                        const float* target_params, int n_target_params, float log_eps, int optimizer, float step_size, float h1,
                        float h2, float h3, int64_t first_iter, int resume, int64_t eps_stride, float* opt_state, float* trace_F,
                        void* workspace, size_t workspace_bytes, void* stream) {
    return nf::script_train<float>(eps_all, n_iters, N, D, K, params, target_id, target_params, n_target_params, log_eps, optimizer,
                                step_size, h1, h2, h3, first_iter, resume, eps_stride, opt_state, trace_F, workspace,
                                workspace_bytes, stream);
}

size_t nf_script_train_workspace_bytes(int K, int D, int elem_size) {
    if (K < 1 || D < 1 || elem_size < 1) return 0;
    return nf::script_head_bytes(K, D, (size_t)elem_size) + nf_planar_elbo_workspace_bytes(K, D, elem_size);
}

size_t nf_script_train_workspace_bytes_n(int64_t N, int K, int D, int elem_size) {
    if (K < 1 || D < 1 || elem_size < 1) return 0;
    return nf::script_head_bytes(K, D, (size_t)elem_size) + nf_planar_elbo_workspace_bytes_n(N, K, D, elem_size);
}

int nf_script_train_sharded_f32(const float* eps_all, int64_t n_iters, int64_t N, int64_t n_global, int D, int K, float* params,
                                int target_id, const float* target_params, int n_target_params, float log_eps, int optimizer,
                                float step_size, float h1, float h2, float h3, int64_t first_iter, int resume, int64_t eps_stride,
                                float* opt_state, float* trace_F, void* workspace, size_t workspace_bytes,
                                void* const* peer_buffers, int world, int rank, uint64_t seq0, float* status, void* stream) {
    if (!peer_buffers || !status || seq0 == 0) return NF_ERR_INVALID;
    const nf::ShardDesc sh = {peer_buffers, world, rank, seq0, n_global, status};
    return nf::script_train<float>(eps_all, n_iters, N, D, K, params, target_id, target_params, n_target_params, log_eps, optimizer,
                                   step_size, h1, h2, h3, first_iter, resume, eps_stride, opt_state, trace_F, workspace,
                                   workspace_bytes, stream, &sh);
}

}  // extern "C"

// ---------------------------------------------------------------------------------------
// Allocation-free variant of nf_planar_elbo_fwd_bwd_f32.
//
// Who needs this: host programs that are not allowed to ask the CUDA driver for memory at
// the point where they evaluate the free energy -- an optimiser loop inside an application
// that handed the whole device to its own pool at start-up (a later cudaMalloc would fail
// or fight that pool), and latency-sensitive callers for whom cudaMalloc/cudaFree are off
// limits because both synchronise the device.  Such callers still own their inputs and
// outputs; the one thing they cannot provide is the partial-sum workspace, so the library
// carries one: a fixed array in device global memory that is part of the loaded module.
// Every GPU that loads libnfcore therefore holds NF_NOALLOC_SCRATCH_BYTES (4 MiB) until the
// process exits, whether or not this entry point is ever called.
//
// Rules that follow from sharing one array:
//   * problems whose workspace (nf_planar_elbo_workspace_bytes(K, D, 4)) exceeds the array
//     are refused with NF_ERR_SCRATCH_LIMIT -- nothing is launched;
//   * the array keeps whatever the previous call left in it.  Nothing in it has to start at
//     zero: every block (D <= 4) overwrites its whole row before the row kernel reads it, and
//     the wide-D kernel clears each warp's row itself before accumulating;
//   * two calls that overlap on the GPU (different streams, or different host threads on one
//     device) use the same rows and race.  Serialise them on one stream;
//   * the first call on each device looks up the array's device address (set-up, under a
//     lock).  Make that first call before any stream capture begins.
// The body below calls no cudaMalloc / cudaFree / allocating library routine.
// ---------------------------------------------------------------------------------------
namespace nf {

static __device__ __align__(256) unsigned char g_noalloc_scratch[NF_NOALLOC_SCRATCH_BYTES];

static void* g_noalloc_ptr[kMaxDevices];          // device address of the array, per device ordinal
static bool g_noalloc_ready[kMaxDevices];
static std::mutex g_noalloc_mutex;

// In host code `g_noalloc_scratch` is only a handle for the symbol, not an address a kernel
// can use; the address comes from cudaGetSymbolAddress on the device that is current.
static int noalloc_scratch(void** out) {
    int dev = 0;
    NF_CUDA_OK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return NF_ERR_UNSUPPORTED;
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ready[dev]) {
        void* p = nullptr;
        NF_CUDA_OK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        g_noalloc_ptr[dev] = p;
        g_noalloc_ready[dev] = true;
    }
    *out = g_noalloc_ptr[dev];
    return NF_OK;
}

}  // namespace nf

extern "C" int nf_planar_elbo_fwd_bwd_noalloc_f32(const float* eps, int64_t N, int D, int K, const float* w,
                                                  const float* s, const float* b, const float* c, int target_id,
                                                  const float* target_params, int n_target_params, float log_eps,
                                                  float* out_F, float* g_w, float* g_s, float* g_b, float* g_c,
                                                  void* stream) {
    if (D < 1 || K < 1) return NF_ERR_INVALID;
    const size_t need = nf_planar_elbo_workspace_bytes(K, D, (int)sizeof(float));
    if (need == 0 || need > (size_t)NF_NOALLOC_SCRATCH_BYTES) return NF_ERR_SCRATCH_LIMIT;
    void* scratch = nullptr;
    const int rc = nf::noalloc_scratch(&scratch);
    if (rc != NF_OK) return rc;
    return nf::elbo_entry<float>(eps, N, D, K, w, s, b, c, target_id, target_params, n_target_params, log_eps, out_F,
                                 g_w, g_s, g_b, g_c, scratch, (size_t)NF_NOALLOC_SCRATCH_BYTES, stream);
}
