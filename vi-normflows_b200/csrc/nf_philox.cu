// The sample stream of nf_philox.cuh written to memory: what the free-energy kernel draws in place when it is
// called through nf_planar_elbo_fwd_bwd_rng_f32, for callers (and tests) that want to see the samples.
#include "nf_philox.cuh"
#include "nf_common.cuh"
#include "nf_launch.cuh"

namespace nf {

template <int D>
__global__ void __launch_bounds__(256) philox_normal_kernel(RngDesc rng, long long N, float* __restrict__ out) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const long long npairs = (N + 1) / 2;
    for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < npairs; q += (long long)gridDim.x * blockDim.x) {
        float za[D], zb[D];
        rng_pair<D>(rng, (unsigned long long)q, za, zb);
#pragma unroll
        for (int d = 0; d < D; ++d) {
            out[(2 * q) * D + d] = za[d];
            if (2 * q + 1 < N) out[(2 * q + 1) * D + d] = zb[d];
        }
    }
}

}  // namespace nf

extern "C" int nf_philox_normal_f32(uint64_t seed, uint64_t step, int64_t N, int D, float* out, void* stream) {
    if (N < 0 || (N > 0 && !out)) return NF_ERR_INVALID;
    if (D != 1 && D != 2) return NF_ERR_UNSUPPORTED;
    if (N == 0) return NF_OK;
    const nf::DevInfo di = nf::dev_info();
    if (!di.ok) return NF_ERR_NO_DEVICE;
    nf::RngDesc rng = {seed, step};
    long long blocks = ((N + 1) / 2 + 255) / 256;
    if (blocks > (long long)di.sms * 16) blocks = (long long)di.sms * 16;
    if (D == 1) NF_CUDA_OK(nf::launch_pdl(nf::philox_normal_kernel<1>, dim3((unsigned)blocks), dim3(256), 0, (cudaStream_t)stream, rng, (long long)N, out));
    else NF_CUDA_OK(nf::launch_pdl(nf::philox_normal_kernel<2>, dim3((unsigned)blocks), dim3(256), 0, (cudaStream_t)stream, rng, (long long)N, out));
    return NF_OK;
}
