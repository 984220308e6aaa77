// Status strings, version and device checks for libnfcore.
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

}  // extern "C"
