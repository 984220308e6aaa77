// Status strings, version and device checks for libnfcore.
#include <cmath>
#include "nf_common.cuh"

static thread_local cudaError_t g_last_cuda = cudaSuccess;
void nf_set_last_cuda_error(cudaError_t e) { g_last_cuda = e; }

extern "C" {

const char* nf_status_string(int status) {
    switch (status) {
        case NF_OK: return "ok";
        case NF_ERR_INVALID: return "invalid argument";
        case NF_ERR_UNSUPPORTED: return "unsupported shape or target for this build";
        case NF_ERR_WORKSPACE: return "workspace too small (see nf_planar_elbo_workspace_bytes)";
        case NF_ERR_CUDA: return "CUDA runtime error (see nf_last_cuda_error)";
        case NF_ERR_NO_DEVICE: return "no sm_100 CUDA device is current";
        case NF_ERR_SCRATCH_LIMIT: return "problem needs more scratch than the built-in no-allocation buffer holds (NF_NOALLOC_SCRATCH_BYTES)";
        case NF_ERR_PEER: return "multi-GPU exchange: a rank did not arrive within the time limit";
        default: return "unknown status";
    }
}

const char* nf_last_cuda_error(void) { return cudaGetErrorString(g_last_cuda); }

int nf_version(void) { return 100; }   /* 0.1.0 */

int nf_built_for_sm(void) { return 100; }

int nf_check_device(void) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return NF_ERR_NO_DEVICE;
    int major = 0;     // attribute query: cheap (cudaGetDeviceProperties takes milliseconds)
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return NF_ERR_NO_DEVICE;
    return major == 10 ? NF_OK : NF_ERR_NO_DEVICE;
}

// ---- host-side parameter algebra of the shared-parameter path (K*D numbers, float64, no GPU involved) ----
// get_data.py:106-108 derives, per layer, the step vector and the log-det coefficient from the raw (w, u); the kernels take
// those as inputs and return gradients with respect to them.  These two routines are that map and its chain rule, in the
// caller's parameter layout [W | U | B] (get_data.py:74-78).  semantics 0: script (s = u + (m(w.u) - w.u) w/||w||, c = w.u),
// 1: paper (norm squared, c = w.s).  Done in NumPy this bookkeeping costs ~85 us per objective call.
static inline double nf_m(double x) { return -1.0 + std::log(1.0 + std::exp(x)); }

int nf_script_constants_host(const double* params, int K, int D, int semantics, double* lc /* [w | s | b | c] */) {
    if (!params || !lc || K < 1 || D < 1 || semantics < 0 || semantics > 1) return NF_ERR_INVALID;
    const double *W = params, *U = params + (size_t)K * D, *B = params + (size_t)2 * K * D;
    double *w = lc, *s = lc + (size_t)K * D, *b = lc + (size_t)2 * K * D, *c = b + K;
    for (int k = 0; k < K; ++k) {
        const double *wk = W + (size_t)k * D, *uk = U + (size_t)k * D;
        double wu = 0.0, n2 = 0.0;
        for (int d = 0; d < D; ++d) { wu += wk[d] * uk[d]; n2 += wk[d] * wk[d]; }
        const double alpha = nf_m(wu) - wu;
        const double npow = semantics == 0 ? std::sqrt(n2) : n2;
        double ws = 0.0;
        for (int d = 0; d < D; ++d) {
            const double sv = uk[d] + alpha * wk[d] / npow;
            w[(size_t)k * D + d] = wk[d];
            s[(size_t)k * D + d] = sv;
            ws += wk[d] * sv;
        }
        b[k] = B[k];
        c[k] = semantics == 0 ? wu : ws;
    }
    return NF_OK;
}

int nf_script_chain_host(const double* params, const double* g /* [g_w | g_s | g_b | g_c] */, int K, int D, int semantics,
                         double scale, double* dparams /* [dW | dU | dB] */) {
    if (!params || !g || !dparams || K < 1 || D < 1 || semantics < 0 || semantics > 1) return NF_ERR_INVALID;
    const double *W = params, *U = params + (size_t)K * D;
    const double *g_w = g, *g_s = g + (size_t)K * D, *g_b = g + (size_t)2 * K * D, *g_c = g_b + K;
    double *dW = dparams, *dU = dparams + (size_t)K * D, *dB = dparams + (size_t)2 * K * D;
    const double p = semantics == 0 ? 1.0 : 2.0;
    for (int k = 0; k < K; ++k) {
        const double *wk = W + (size_t)k * D, *uk = U + (size_t)k * D;
        const double *gwk = g_w + (size_t)k * D, *gsk = g_s + (size_t)k * D;
        double wu = 0.0, n2 = 0.0;
        for (int d = 0; d < D; ++d) { wu += wk[d] * uk[d]; n2 += wk[d] * wk[d]; }
        const double alpha = nf_m(wu) - wu;
        const double dalpha = 1.0 / (1.0 + std::exp(-wu)) - 1.0;
        const double n = std::sqrt(n2), npow = semantics == 0 ? n : n2;
        // paper: c = w.s, so its gradient folds into g_s and the direct g_w; script: c = w.u
        const double gc_fold = semantics == 1 ? g_c[k] : 0.0, gc_raw = semantics == 1 ? 0.0 : g_c[k];
        double gs_w = 0.0;
        for (int d = 0; d < D; ++d) gs_w += (gsk[d] + gc_fold * wk[d]) * wk[d];
        const double a1 = dalpha * gs_w / npow, a2 = alpha / npow, a3 = alpha * p * gs_w / (npow * n2);
        for (int d = 0; d < D; ++d) {
            const double sv = uk[d] + alpha * wk[d] / npow;
            const double gs = gsk[d] + gc_fold * wk[d];
            const double gw = gwk[d] + gc_fold * sv;
            dU[(size_t)k * D + d] = scale * (gs + a1 * wk[d] + gc_raw * wk[d]);
            dW[(size_t)k * D + d] = scale * (gw + a1 * uk[d] + a2 * gs - a3 * wk[d] + gc_raw * uk[d]);
        }
        dB[k] = scale * g_b[k];
    }
    return NF_OK;
}

}  // extern "C"
