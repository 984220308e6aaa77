/* libnfcore -- C ABI of the B200 (sm_100a) planar-flow free-energy hot path.
 *
 * The reference (benlevyx/vi-normflows) is pure Python and has no FFI of its own;
 * these entry points are what a binding for its hot path would call.  Each one
 * names the reference code it replaces (paths relative to the reference checkout).
 *
 * Conventions
 *   - plain pointers and sizes only; every `*_f32`/`*_f64` pair differs only in the
 *     element type.  f32 is the throughput mode, f64 the reference-precision mode.
 *   - unless the name says `host`, pointers are DEVICE pointers owned by the caller;
 *     the call enqueues work on `stream` (a cudaStream_t passed as void*) and returns
 *     without synchronising.
 *   - return value: NF_OK (0) or a negative nf_status; nf_status_string() explains it,
 *     nf_last_cuda_error() returns the CUDA error string behind NF_ERR_CUDA.
 *   - reductions over samples are returned as SUMS (not means) so that sample-sharded
 *     callers need one all-reduce followed by a scale by 1/N_global.
 *   - matrices are row-major: eps[N][D], w[K][D], s[K][D].
 */
#ifndef NFCORE_H_
#define NFCORE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    NF_OK = 0,
    NF_ERR_INVALID = -1,      /* bad argument (null pointer, N<0, K<1, D<1, ...) */
    NF_ERR_UNSUPPORTED = -2,  /* shape/target combination this build does not implement */
    NF_ERR_WORKSPACE = -3,    /* workspace_bytes smaller than nf_planar_elbo_workspace_bytes() */
    NF_ERR_CUDA = -4,         /* a CUDA runtime call failed; see nf_last_cuda_error() */
    NF_ERR_NO_DEVICE = -5,    /* no CUDA device / not an sm_100 device */
    NF_ERR_SCRATCH_LIMIT = -6, /* nf_planar_elbo_fwd_bwd_noalloc_f32: problem needs more than NF_NOALLOC_SCRATCH_BYTES */
    NF_ERR_PEER = -7          /* multi-GPU exchange: a rank did not arrive within the time limit; the sums are not valid */
} nf_status;

/* Target densities exp(-U(z)) the free energy is taken against.  The objective
 * always forms log(eps + target(z_K)) as get_data.py:112 does.
 *   U1..U4, TRIAL1 : get_data.py:28-59, 20-26 (p1..p4, trial1); D must be 2; no params.
 *   MIXDIAG        : distributions.py:60-69 prob_gm with diagonal covariances, and
 *                    get_data.py:61-66 / experimentation.py:18-23 gmm (D=1, var=sigma^2).
 *                    params = [pi_0..pi_{G-1} | mu (G*D) | var (G*D)], G = n/(1+2D);
 *                    pi_{G-1} is passed explicitly (= 1 - sum of the others).
 *   GAUSSDIAG      : exp(distributions.py:35-54 log_mvn); params = [mu (D) | log_sigma_diag (D)];
 *                    for D == 1 the inverse variance is exp(1 - log_sigma) as at :46-47.
 *   MIXFULL        : distributions.py:60-69 prob_gm with FULL covariance matrices, which is what mvn takes when
 *                    len(sigma_diag) > 1 (:20-25: np.linalg.inv / np.linalg.det of a D x D matrix);
 *                    params = [pi (G) | mu (G*D) | Sigma (G*D*D, row-major)], G = n/(1+D+D*D); D <= 4.
 *                    The kernels invert the matrices once per block (Gauss-Jordan in float64).
 */
typedef enum {
    NF_TARGET_U1 = 0,
    NF_TARGET_U2 = 1,
    NF_TARGET_U3 = 2,
    NF_TARGET_U4 = 3,
    NF_TARGET_TRIAL1 = 4,
    NF_TARGET_MIXDIAG = 5,
    NF_TARGET_GAUSSDIAG = 6,
    NF_TARGET_MIXFULL = 7
} nf_target;

const char* nf_status_string(int status);
const char* nf_last_cuda_error(void);
/* Library version, and the compute capability it was built for (100 = sm_100a). */
int nf_version(void);
int nf_built_for_sm(void);
/* NF_OK when the current device can run the kernels (compute capability 10.x). */
int nf_check_device(void);

/* ---------------------------------------------------------------------------
 * Shared-parameter Monte-Carlo free energy, forward + backward in one launch.
 * Replaces get_data.py:97-113 `variational_objective` (= experimentation.py:54-70)
 * and the autograd tape replay `grad(variational_objective)` at get_data.py:115.
 *
 * Per layer k the caller supplies the direction w_k, the step vector s_k, the
 * offset b_k and the log-det coefficient c_k; per sample the kernel evaluates
 *     a = w_k.z + b_k ; t = tanh(a) ; ld += log(eps + |1 + (1 - t^2) c_k|) ; z += s_k t
 * which is get_data.py:106-109 with s_k = u_hat (:106) and c_k = w_k.u_k (:107-108).
 * Outputs (all sums over the N samples given):
 * `log_eps` is the eps INSIDE the logs (1e-7 at get_data.py:121), not a logarithm.
 *     out_F[0] = sum_n (log q_K - log p)      (get_data.py:113 before the mean)
 *     out_F[1] = sum_n log q_K                (get_data.py:94, `other_metrics`)
 *     out_F[2] = sum_n log(eps + target(z_K)) (get_data.py:95)
 *     g_w[K][D], g_s[K][D], g_b[K], g_c[K] = d out_F[0] / d (w, s, b, c)
 * g_* may all be NULL: then only out_F is produced (forward only; what the
 * reference's callbacks use, get_data.py:128-138).
 * `workspace` holds per-block partial sums; its contents on entry are ignored.
 * ------------------------------------------------------------------------- */
size_t nf_planar_elbo_workspace_bytes(int K, int D, int elem_size);
/* Workspace that also lets a call over N samples take the fastest path the library has for the shape:
 * for float32, 32 <= D <= 512 (D % 4 == 0), K <= 64 and the GAUSSDIAG / MIXDIAG targets that can be the
 * matrix-product formulation (csrc/nf_elbo_gram.cu; the library picks it where its fitted cost model says it wins,
 * roughly from 2^15 samples at D = 64 down to 2^10 at D = 256), which keeps the intermediates of a chunk of
 * samples (about 3 * 64 + D floats per sample, chunks of up to 32M/D samples) in the workspace.  Never smaller than
 * nf_planar_elbo_workspace_bytes(); a call given less than this runs the direct kernel.  Same quantity
 * (get_data.py:97-115), same outputs.  No counterpart in the reference (NumPy allocates as it goes). */
size_t nf_planar_elbo_workspace_bytes_n(int64_t N, int K, int D, int elem_size);

int nf_planar_elbo_fwd_bwd_f32(const float* eps, int64_t N, int D, int K,
                               const float* w, const float* s, const float* b, const float* c,
                               int target_id, const float* target_params, int n_target_params,
                               float log_eps,
                               float* out_F, float* g_w, float* g_s, float* g_b, float* g_c,
                               void* workspace, size_t workspace_bytes, void* stream);

int nf_planar_elbo_fwd_bwd_f64(const double* eps, int64_t N, int D, int K,
                               const double* w, const double* s, const double* b, const double* c,
                               int target_id, const double* target_params, int n_target_params,
                               double log_eps,
                               double* out_F, double* g_w, double* g_s, double* g_b, double* g_c,
                               void* workspace, size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------
 * The same free energy with the samples drawn inside the kernel.  Replaces the draw the reference makes in every
 * objective call, z0 = np.random.multivariate_normal(0, I, N) (get_data.py:100-101), together with the pass over z0:
 * nothing crosses PCIe and nothing is read from HBM but the layer constants.  Opt-in: the stream is this library's own
 * (Philox-4x32-10 keyed by `seed`, counter = (sample pair, `step`), Box-Muller; csrc/nf_philox.cuh), not NumPy's, so
 * runs that must reproduce the reference's draws pass their samples to nf_planar_elbo_fwd_bwd_f32 instead.
 * Sample n of step t depends on (seed, t, n) only; nf_philox_normal_f32 writes the same stream to memory
 * ([N][D], D = 1 or 2).  float32, D <= 2, K <= 64; all five output arrays are required.
 * ------------------------------------------------------------------------- */
int nf_philox_normal_f32(uint64_t seed, uint64_t step, int64_t N, int D, float* out, void* stream);
int nf_planar_elbo_fwd_bwd_rng_f32(uint64_t seed, uint64_t step, int64_t N, int D, int K,
                                   const float* w, const float* s, const float* b, const float* c,
                                   int target_id, const float* target_params, int n_target_params, float log_eps,
                                   float* out_F, float* g_w, float* g_s, float* g_b, float* g_c,
                                   void* workspace, size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------
 * Host-buffer form of the shared-parameter free energy: the call the reference's
 * `objective(params, t)` / `gradient(params, t)` make, with everything in host memory
 * (get_data.py:97-115: NumPy parameters, freshly drawn NumPy z0).  A session owns the device
 * staging for one (max_samples, D, K, element size); nf_session_elbo_host_* sends the samples in
 * chunks on a copy stream, runs the objective kernel on each chunk as it lands, sums all partial
 * rows once, copies the result back and returns when `out_host` is valid (it synchronises).
 *   eps_host          : [N][D]; pinned memory lets the copies overlap the kernels
 *   layer_consts_host : [w (K*D) | s (K*D) | b (K) | c (K)]
 *   out_host          : [out_F (3) | g_w (K*D) | g_s (K*D) | g_b (K) | g_c (K)]  (gradients only
 *                       when backward != 0)
 * Session creation / destruction allocate and free device and pinned memory; the per-call entry
 * points do not.  A session is not thread-safe.
 * ------------------------------------------------------------------------- */
typedef struct nf_session nf_session;
int nf_session_create(nf_session** out, int64_t max_samples, int D, int K, int elem_size);
void nf_session_destroy(nf_session* s);
int nf_session_elbo_host_f32(nf_session* s, const float* eps_host, int64_t N, const float* layer_consts_host,
                             int target_id, const float* target_params_host, int n_target_params,
                             float log_eps, int backward, float* out_host);
int nf_session_elbo_host_f64(nf_session* s, const double* eps_host, int64_t N, const double* layer_consts_host,
                             int target_id, const double* target_params_host, int n_target_params,
                             double log_eps, int backward, double* out_host);
/* The literal objective(params, t) / gradient(params, t) of get_data.py:97-115 (`variational_objective` and
 * `grad(variational_objective)`) in one call: `params` is the reference's float64 parameter vector [W | U | B]
 * (get_data.py:74-78), `eps_host` the N x D draws of :100-101 in host memory.  Internally nf_script_constants_host,
 * nf_session_elbo_host_* and nf_script_chain_host.  F_parts[3] = sums over the samples of (F, log q_K, log p)
 * (divide by N for the means of :113 and :80-95); grad[2KD+K] = d(mean F)/d params (float64), untouched when backward = 0.
 * semantics: 0 = script (get_data.py:106-108), 1 = package (flows.py:37-39). */
int nf_session_objective_host_f32(nf_session* s, const float* eps_host, int64_t N, const double* params, int semantics,
                                  int target_id, const float* target_params_host, int n_target_params, float log_eps,
                                  int backward, double* F_parts, double* grad);
int nf_session_objective_host_f64(nf_session* s, const double* eps_host, int64_t N, const double* params, int semantics,
                                  int target_id, const double* target_params_host, int n_target_params, double log_eps,
                                  int backward, double* F_parts, double* grad);
/* The same objective + gradient with the draw of get_data.py:100-101 made inside the kernel (the library's counter-based
 * stream for (seed, step), see nf_planar_elbo_fwd_bwd_rng_f32): no sample buffer at all.  float32 session, D <= 2,
 * N large enough for the kernel that draws (NF_ERR_UNSUPPORTED otherwise). */
/* nf_session_objective_host_f32 over this rank's shard of the draws, summed over the `world` GPUs of one box by the
 * peer exchange of nf_allreduce_small_counted_f32 (same buffers, same seq discipline: every rank calls with the same
 * seq > the previous one).  F_parts are global sums, grad = d(mean over ALL samples of F)/d params, *n_global the
 * global sample count.  NF_ERR_PEER when a rank did not arrive. */
int nf_session_objective_host_exchange_f32(nf_session* s, const float* eps_host, int64_t N_local, const double* params,
                                           int semantics, int target_id, const float* target_params_host, int n_target_params,
                                           float log_eps, void* const* peer_buffers, int world, int rank, uint64_t seq,
                                           double* F_parts, double* grad, int64_t* n_global);
int nf_session_objective_rng_f32(nf_session* s, uint64_t seed, uint64_t step, int64_t N, const double* params, int semantics,
                                 int target_id, const float* target_params_host, int n_target_params, float log_eps,
                                 double* F_parts, double* grad);

/* ---------------------------------------------------------------------------
 * The whole optimisation loop of the script path on the device: get_data.py:119-142 `optimise`
 * (rmsprop), experimentation.py:99-109 (sgd), updated_experimentation.ipynb cell 1 (adam), with the
 * update rules of autograd 1.3's autograd.misc.optimizers.  Per iteration the library enqueues three
 * kernels -- apply the previous update + derive (s, c) from (W, U) (get_data.py:106-108), the objective
 * kernel on that iteration's samples, the row kernel -- and never copies parameters to the host.
 *   eps_all   : [n_iters][N][D] samples, iteration t uses slice t (the draw the reference makes inside
 *               `gradient(params, t)`, get_data.py:100-101)
 *   params    : [W (K*D) | U (K*D) | B (K)], updated in place (get_data.py:74-78 layout)
 *   optimizer : 0 sgd (h1 = mass), 1 rmsprop (h1 = gamma, h2 = eps), 2 adam (h1 = b1, h2 = b2, h3 = eps)
 *   first_iter: global index of the first iteration of this call (0 for a fresh run).  Adam's bias correction
 *               uses first_iter + t + 1, as autograd 1.3 does for iteration index first_iter + t.
 *   resume    : 0: the call initialises opt_state (zeros; rmsprop's running average starts at ones, autograd 1.3);
 *               != 0: opt_state is taken as a previous call left it, so that a run cut into several calls (the
 *               callback cadence of get_data.py:128-138) follows exactly the trajectory of a single call
 *   eps_stride: elements between the slices of consecutive iterations (0: N*D).  Slices are read with 16-byte
 *               vector loads: eps_stride * elem_size must be a multiple of 16 when n_iters > 1 (NF_ERR_INVALID
 *               otherwise) -- pad the slices
 *   opt_state : 2*(2KD+K) elements kept by the caller between calls
 *   trace_F   : [n_iters] free energy seen by each gradient evaluation, or NULL
 *   workspace : nf_script_train_workspace_bytes(K, D, elem_size)
 * D <= 4 or the wide-D targets, script semantics.  Asynchronous on `stream`: nothing in the call synchronises.
 * ------------------------------------------------------------------------- */
size_t nf_script_train_workspace_bytes(int K, int D, int elem_size);
/* The same with room for the matrix-product path of nf_planar_elbo_workspace_bytes_n (N samples per iteration):
 * nf_script_train_f32 given at least this much evaluates wide-D objectives through it. */
size_t nf_script_train_workspace_bytes_n(int64_t N, int K, int D, int elem_size);
int nf_script_train_f32(const float* eps_all, int64_t n_iters, int64_t N, int D, int K, float* params, int target_id,
                        const float* target_params, int n_target_params, float log_eps, int optimizer,
                        float step_size, float h1, float h2, float h3, int64_t first_iter, int resume, int64_t eps_stride,
                        float* opt_state, float* trace_F, void* workspace, size_t workspace_bytes, void* stream);
int nf_script_train_f64(const double* eps_all, int64_t n_iters, int64_t N, int D, int K, double* params, int target_id,
                        const double* target_params, int n_target_params, double log_eps, int optimizer,
                        double step_size, double h1, double h2, double h3, int64_t first_iter, int resume, int64_t eps_stride,
                        double* opt_state, double* trace_F, void* workspace, size_t workspace_bytes, void* stream);

/* nf_script_train_f32 for a sample-sharded run (one process per GPU): every iteration's sums over this rank's N samples are
 * exchanged with the peers (inside the objective kernel's last block where the shape allows it, see
 * nf_planar_elbo_fwd_bwd_exchange_f32) before the update kernel reads them, normalised by `n_global` (the samples of all
 * ranks).  Every rank applies the same update to its own copy of the parameters -- the sums are bitwise identical on all
 * ranks -- so the copies stay equal without a broadcast.  No counterpart in the reference (single process).
 *   seq0   : sequence number of the first iteration's exchange; the call uses seq0 .. seq0 + n_iters - 1
 *   status : one float in device memory; cleared at the start, raised to 1 if a peer ever failed to arrive within 20 s
 * float32, result vectors of at most nf_allreduce_small_max_count() floats, not the matrix-product path. */
int nf_script_train_sharded_f32(const float* eps_all, int64_t n_iters, int64_t N, int64_t n_global, int D, int K, float* params,
                                int target_id, const float* target_params, int n_target_params, float log_eps, int optimizer,
                                float step_size, float h1, float h2, float h3, int64_t first_iter, int resume, int64_t eps_stride,
                                float* opt_state, float* trace_F, void* workspace, size_t workspace_bytes,
                                void* const* peer_buffers, int world, int rank, uint64_t seq0, float* status, void* stream);

/* Host-side parameter algebra of the shared-parameter path (float64, K*D numbers, no GPU involved).
 * get_data.py:106-108 derives the step vector u_hat and the log-det coefficient w.u from the raw (w, u) inside
 * the objective; the entry points above take them as inputs (s, c) and return gradients with respect to them.
 *   nf_script_constants_host : params [W | U | B] (get_data.py:74-78 layout) -> lc [w | s | b | c]
 *   nf_script_chain_host     : g [g_w | g_s | g_b | g_c] -> dparams [dW | dU | dB] = scale * d/d(params)
 * semantics 0: script (get_data.py:106-108: ||w|| not squared, c = w.u); 1: paper (Rezende & Mohamed eq. 10-12). */
int nf_script_constants_host(const double* params, int K, int D, int semantics, double* lc);
int nf_script_chain_host(const double* params, const double* g, int K, int D, int semantics, double scale,
                         double* dparams);

/* nf_planar_elbo_fwd_bwd_f32 for callers that must not allocate device memory.  It has no
 * counterpart in the reference (which never touches a GPU); it exists for host programs
 * whose device memory was all handed out at start-up, or for which cudaMalloc/cudaFree --
 * both synchronise the device -- are forbidden where the objective is evaluated.
 * Same arguments and results as nf_planar_elbo_fwd_bwd_f32 without the workspace pair: the
 * partial sums go to a fixed NF_NOALLOC_SCRATCH_BYTES array that lives in the library's
 * device image (every GPU that loads the library holds it until exit).
 *   - returns NF_ERR_SCRATCH_LIMIT, launching nothing, when
 *     nf_planar_elbo_workspace_bytes(K, D, 4) > NF_NOALLOC_SCRATCH_BYTES
 *     (float32, D <= 4: K*(2D+2)+2 <= 442, e.g. K <= 73 at D = 2);
 *   - all calls share the array: calls that overlap on the GPU race -- keep them on one stream;
 *   - the first call on a device resolves the array's address; issue it before any stream
 *     capture starts.
 * The call itself performs no cudaMalloc/cudaFree and uses no allocating library routine. */
#define NF_NOALLOC_SCRATCH_BYTES (4u << 20)
int nf_planar_elbo_fwd_bwd_noalloc_f32(const float* eps, int64_t N, int D, int K,
                                       const float* w, const float* s, const float* b, const float* c,
                                       int target_id, const float* target_params, int n_target_params,
                                       float log_eps,
                                       float* out_F, float* g_w, float* g_s, float* g_b, float* g_c,
                                       void* stream);

/* ---------------------------------------------------------------------------
 * Evaluation surface of the shared-parameter path (forward only), what visualise_flow (Final (master).ipynb cell 16)
 * computes with NumPy: the target density at given points (its contour grid), and samples pushed through the
 * learned flow.
 *   nf_target_density_* : out[n] = target(z[n]), the density itself (get_data.py:20-66 p1..p4, trial1, gmm;
 *                         distributions.py:10-32 mvn, :60-69 prob_gm); target ids and parameters as above.
 *                         D <= 4 for every target; D <= 512 for MIXDIAG / GAUSSDIAG.
 *   nf_planar_chain_forward_* : z_K[n] (get_data.py:104-109 applied to the given z0[n]) and, when log_qK != NULL,
 *                         log q_K[n] = -1/2 sum_d (log 2 pi + z0_d^2) - sum_k log(eps + |1 + h'(a_k) c_k|) (:111).
 *                         Layer constants (w, s, b, c) as for nf_planar_elbo_fwd_bwd_*; D <= 512.
 * ------------------------------------------------------------------------- */
int nf_target_density_f32(const float* z, int64_t N, int D, int target_id, const float* target_params, int n_target_params,
                          float* out, void* stream);
int nf_target_density_f64(const double* z, int64_t N, int D, int target_id, const double* target_params, int n_target_params,
                          double* out, void* stream);
int nf_planar_chain_forward_f32(const float* z0, int64_t N, int D, int K, const float* w, const float* s, const float* b,
                                const float* c, float log_eps, float* zK, float* log_qK, void* stream);
int nf_planar_chain_forward_f64(const double* z0, int64_t N, int D, int K, const double* w, const double* s, const double* b,
                                const double* c, double log_eps, double* zK, double* log_qK, void* stream);

/* ---------------------------------------------------------------------------
 * flows.planar_flow (normflows/normflows/flows.py:8-34, _get_uhat :37-39, m :42-43).
 * z, out: [N][D].  per_sample != 0: w,u are [N][D] and b is [N] (the HEAD signature,
 * flows.py:26-27); per_sample == 0: w,u are [D], b is [1] (the call the stale drivers
 * make, src/learning_basic_flow.py:33,97).  u_hat uses ||w||^2 (flows.py:39).
 *   semantics 0 ("package"): out = z + (sum_d u_hat_d tanh(w.z+b)) broadcast to every
 *                            coordinate -- what flows.py:32 computes;
 *   semantics 1 ("paper")  : out = z + u_hat tanh(w.z+b) (Rezende & Mohamed eq. 10).
 * D <= 256.  Also the forward-only building block of distributions.sample_from_pz
 * (distributions.py:92-102).
 * ------------------------------------------------------------------------- */
int nf_planar_flow_f32(const float* z, const float* w, const float* u, const float* b, int64_t N, int D,
                       int per_sample, int semantics, float* out, void* stream);
int nf_planar_flow_f64(const double* z, const double* w, const double* u, const double* b, int64_t N, int D,
                       int per_sample, int semantics, double* out, void* stream);

/* ---------------------------------------------------------------------------
 * Amortised (per-sample-parameter) K-layer flow of the flow-VAE objective,
 * normflows/normflows/optimization.py:74-85 (reparameterisation :74-75, running
 * log-det :64,83 with (sum w)(sum u) and raw u, planar_flow :84).
 *   z0, mu0, log_sigma0: [N][D];  W, U: [K][N][D];  Bb: [K][N]
 * forward :  zK[N][D], logdet[N] = sum_k log(eps + |1 + (sum u_k)(sum w_k) h'(w_k.z+b_k)|),
 *            t_save[K][N] = tanh activations kept for the backward pass (may be NULL
 *            when no backward pass follows).
 * backward:  given g_zK[N][D] = dF/dzK and the scalar g_logdet = dF/dlogdet[n] (the same
 *            for every n: -1/N for optimization.py:85,92), writes dF/d(mu0, log_sigma0,
 *            W, U, Bb) with the shapes of the inputs.
 * No cross-sample reduction happens here; D <= 256.
 * ------------------------------------------------------------------------- */
int nf_planar_amortised_fwd_f32(const float* z0, const float* mu0, const float* log_sigma0, const float* W,
                                const float* U, const float* Bb, int64_t N, int D, int K, float log_eps,
                                float* zK, float* logdet, float* t_save, void* stream);
int nf_planar_amortised_fwd_f64(const double* z0, const double* mu0, const double* log_sigma0, const double* W,
                                const double* U, const double* Bb, int64_t N, int D, int K, double log_eps,
                                double* zK, double* logdet, double* t_save, void* stream);
int nf_planar_amortised_bwd_f32(const float* z0, const float* log_sigma0, const float* W, const float* U,
                                const float* zK, const float* t_save, const float* g_zK, float g_logdet,
                                int64_t N, int D, int K, float log_eps, float* g_mu0, float* g_log_sigma0,
                                float* g_W, float* g_U, float* g_Bb, void* stream);
int nf_planar_amortised_bwd_f64(const double* z0, const double* log_sigma0, const double* W, const double* U,
                                const double* zK, const double* t_save, const double* g_zK, double g_logdet,
                                int64_t N, int D, int K, double log_eps, double* g_mu0, double* g_log_sigma0,
                                double* g_W, double* g_U, double* g_Bb, void* stream);

/* The same reverse pass with dF/dlogdet given per sample in device memory (g_logdet_vec[N]): what a tape hands back
 * without a host round trip for the scalar (optimization.py:85,92 make it -1/N for every n, but it lives on the device). */
int nf_planar_amortised_bwd_v_f32(const float* z0, const float* log_sigma0, const float* W, const float* U,
                                  const float* zK, const float* t_save, const float* g_zK, const float* g_logdet_vec,
                                  int64_t N, int D, int K, float log_eps, float* g_mu0, float* g_log_sigma0,
                                  float* g_W, float* g_U, float* g_Bb, void* stream);
int nf_planar_amortised_bwd_v_f64(const double* z0, const double* log_sigma0, const double* W, const double* U,
                                  const double* zK, const double* t_save, const double* g_zK, const double* g_logdet_vec,
                                  int64_t N, int D, int K, double log_eps, double* g_mu0, double* g_log_sigma0,
                                  double* g_W, double* g_U, double* g_Bb, void* stream);

/* autograd.misc.optimizers.adam (autograd 1.3; call site optimization.py:118), one step on the device:
 *   m = (1-b1) g + b1 m;  v = (1-b2) g^2 + b2 v;  x -= step * (m / (1 - b1^(t+1))) / (sqrt(v / (1 - b2^(t+1))) + eps)
 * params, m, v: float64[n] in device memory (m, v zero before step t = 0); grad: float32 or float64 (grad_elem_size). */
int nf_adam_step(double* params, const void* grad, int grad_elem_size, double* m, double* v, int64_t n, double step_size,
                 double b1, double b2, double eps, int64_t t, void* stream);

/* ---------------------------------------------------------------------------
 * Encoder / decoder MLP: nn_models.Feedforward.forward (normflows/normflows/nn_models.py:41-84)
 * as used by src/learning_mnist.py:60-73, and its reverse pass (the autograd tape replay of
 * optimization.py:35 through those closures).
 *   weights : the reference's flat vector (nn_models.py:59-80):
 *             W1 (H x D_in, row-major) | b1 (H) | {W_l (H x H) | b_l (H)} x (L-1) | W_out (D_out x H) | b_out
 *             nf_mlp_param_count() = Feedforward.D (nn_models.py:16-19)
 *   x       : [N][D_in]  SAMPLE-major (the reference passes X.T, feature-major; same numbers)
 *   out     : [N][D_out] sample-major (the reference returns (1, D_out, N))
 *   activations: 0 none, 1 relu (transformations.py:21), 2 sigmoid (:7-8), 3 rbf exp(-x^2)
 *             (nn_models.py:175 default_architecture).  out_act 3 is not supported backward.
 *   fwd_ws  : nf_mlp_forward_workspace_bytes(); holds the hidden activations the backward pass needs
 *   backward: g_out = dF/d out -> g_weights (same layout as weights), g_x (may be NULL);
 *             bwd_ws of nf_mlp_backward_workspace_bytes().
 * ------------------------------------------------------------------------- */
int64_t nf_mlp_param_count(int D_in, int H, int L, int D_out);
size_t nf_mlp_forward_workspace_bytes(int64_t N, int D_in, int H, int L, int D_out, int hidden_act, int elem_size);
size_t nf_mlp_backward_workspace_bytes(int64_t N, int D_in, int H, int L, int D_out, int elem_size);
int nf_mlp_forward_f32(const float* weights, const float* x, int64_t N, int D_in, int H, int L, int D_out,
                       int hidden_act, int out_act, float* fwd_ws, float* out, void* stream);
int nf_mlp_forward_f64(const double* weights, const double* x, int64_t N, int D_in, int H, int L, int D_out,
                       int hidden_act, int out_act, double* fwd_ws, double* out, void* stream);
int nf_mlp_backward_f32(const float* weights, const float* x, int64_t N, int D_in, int H, int L, int D_out,
                        int hidden_act, int out_act, const float* fwd_ws, const float* out, const float* g_out,
                        float* g_weights, float* g_x, float* bwd_ws, void* stream);
int nf_mlp_backward_f64(const double* weights, const double* x, int64_t N, int D_in, int H, int L, int D_out,
                        int hidden_act, int out_act, const double* fwd_ws, const double* out, const double* g_out,
                        double* g_weights, double* g_x, double* bwd_ws, void* stream);

/* distributions.log_bern_mult (distributions.py:86-89): out_sum[0] = sum over ALL count elements of
 * x log(eps+p) + (1-x) log(eps+1-p) (one scalar for the whole batch, as the reference), and, when
 * g_p != NULL, g_p[i] = g_scale * d out_sum / d p[i].  workspace: nf_bernoulli_workspace_bytes(). */
size_t nf_bernoulli_workspace_bytes(int elem_size);
int nf_bernoulli_ll_f32(const float* x, const float* p, int64_t count, float log_eps, float g_scale,
                        float* out_sum, float* g_p, float* workspace, void* stream);
int nf_bernoulli_ll_f64(const double* x, const double* p, int64_t count, double log_eps, double g_scale,
                        double* out_sum, double* g_p, double* workspace, void* stream);

/* nf_bernoulli_ll_* with the upstream scalar dF/d out_sum read from device memory (g_scale_dev[0]): no host round trip. */
int nf_bernoulli_ll_dev_f32(const float* x, const float* p, int64_t count, float log_eps, const float* g_scale_dev,
                            float* out_sum, float* g_p, float* workspace, void* stream);
int nf_bernoulli_ll_dev_f64(const double* x, const double* p, int64_t count, double log_eps, const double* g_scale_dev,
                            double* out_sum, double* g_p, double* workspace, void* stream);

/* Tensor-core GEMM used by the MLP entry points above, exposed for testing and reuse:
 *   C[m][n] = act( sum_k A[m][k] B[n][k] + bias[n] ),  A: [M][K] (lda), B: [N][K] (ldb), float32.
 * TMA tile loads, tcgen05.mma kind::tf32; both operands are split into hi/lo halves on the fly (B in shared
 * memory, A into tensor memory by the warps that own its lanes) and every k-step issues three MMAs, so that
 * results stay at float32 accuracy; accumulator in TMEM.
 * C (ldc) and/or its transpose CT[n][m] (ldct) are written; act: 0 none, 1 relu, 2 sigmoid.
 * lda, ldb must be multiples of 4 and A, B 16-byte aligned, else NF_ERR_UNSUPPORTED.
 * (The reference's counterpart is the np.matmul chain of nn_models.py:59-84.) */
int nf_gemm_tn_tf32x3(const float* A, int64_t lda, const float* B, int64_t ldb, int M, int N, int K,
                      float* C, int64_t ldc, float* CT, int64_t ldct, const float* bias, int act, void* stream);

/* ---------------------------------------------------------------------------
 * One-shot all-reduce(sum) of a small float32 vector (n <= nf_allreduce_small_max_count()) across the
 * <= 8 GPUs of one box, over NVLink peer memory.  No counterpart in the reference (it is single-process);
 * it is the exchange step of the sample-sharded objective: one vector of
 * [F parts | gradient partials] per step.
 *   peer_buffers[r] : device pointer, valid in THIS process, to rank r's exchange buffer of
 *                     nf_allreduce_small_buffer_bytes(world) bytes (peer-mapped, zero-filled once).
 *   seq             : 1, 2, 3, ... the same on every rank for the same step.
 *   status[0]       : set to 1 by the kernel if a peer did not arrive within 20 s.
 * All ranks add the slots in rank order, so every rank ends with bitwise identical sums.
 * ------------------------------------------------------------------------- */
size_t nf_allreduce_small_buffer_bytes(int world);
int nf_allreduce_small_max_count(void);
int nf_allreduce_small_f32(void* const* peer_buffers, int world, int rank, const float* src, float* dst, int n,
                           uint64_t seq, int* status, void* stream);
/* The same exchange in place, carrying the caller's sample count along (no counterpart in the reference):
 * vec[0..n) := sums over ranks; vec[n], vec[n+1] := n_hi, n_lo with global count = n_hi * 2^20 + n_lo (both parts and
 * their sums over 8 ranks are exact in float32); vec[n+2] := 0, or 1 when a peer did not arrive within 20 s. */
int nf_allreduce_small_counted_f32(void* const* peer_buffers, int world, int rank, float* vec, int n, int64_t n_local,
                                   uint64_t seq, void* stream);

/* nf_planar_elbo_fwd_bwd_f32 for a sample-sharded run (one process per GPU): this rank's sums over its
 * N samples are exchanged with the peers through the buffers above before they are written, so `out` holds the sums
 * over ALL ranks' samples -- what get_data.py:113/115 needs before its mean.  No counterpart in the reference.
 *   out : [F parts (3) | g_w (K*D) | g_s (K*D) | g_b (K) | g_c (K) | n_hi | n_lo | status], 3+2KD+2K+3 floats:
 *         the global sample count is n_hi * 2^20 + n_lo; status is 0, or 1 when a peer did not arrive within 20 s
 *         (then the other entries are this rank's own sums and must not be used).
 * For float32, D <= 2, K <= 64 and enough samples the exchange happens inside the objective kernel (its last block
 * pushes, waits and adds: one launch per step); other shapes run the objective and nf_allreduce_small_f32's kernel
 * back to back.  3+2KD+2K+2 must not exceed nf_allreduce_small_max_count() + 2 (NF_ERR_UNSUPPORTED otherwise).
 * Every rank must make the call with the same `seq` (1, 2, 3, ...) for the same step. */
int nf_planar_elbo_fwd_bwd_exchange_f32(const float* eps, int64_t N, int D, int K,
                                        const float* w, const float* s, const float* b, const float* c,
                                        int target_id, const float* target_params, int n_target_params, float log_eps,
                                        float* out, void* workspace, size_t workspace_bytes,
                                        void* const* peer_buffers, int world, int rank, uint64_t seq, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NFCORE_H_ */
