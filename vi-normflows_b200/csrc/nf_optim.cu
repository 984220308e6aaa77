// Adam update of a flat parameter vector on the device: the step autograd.misc.optimizers.adam makes on the host
// (autograd 1.3; call site optimization.py:118), for the flow-VAE loop whose 210 k parameters and gradients would
// otherwise cross PCIe twice per iteration.  Master parameters and the two moment vectors are float64 (the reference's
// precision); the gradient comes in the compute type of the objective.
#include "nf_common.cuh"

namespace nf {

template <typename G>
__global__ void __launch_bounds__(256)
adam_step_kernel(double* __restrict__ x, const G* __restrict__ g, double* __restrict__ m, double* __restrict__ v, long long n,
                 double step, double b1, double b2, double eps, double c1 /* 1 - b1^(t+1) */, double c2 /* 1 - b2^(t+1) */) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const double gi = (double)g[i];
        const double mi = (1.0 - b1) * gi + b1 * m[i];
        const double vi = (1.0 - b2) * (gi * gi) + b2 * v[i];
        m[i] = mi;
        v[i] = vi;
        x[i] = x[i] - step * (mi / c1) / (::sqrt(vi / c2) + eps);
    }
}

}  // namespace nf

extern "C" int nf_adam_step(double* params, const void* grad, int grad_elem_size, double* m, double* v, int64_t n, double step_size,
                            double b1, double b2, double eps, int64_t t, void* stream) {
    if (n < 0 || t < 0 || (n > 0 && (!params || !grad || !m || !v)) || (grad_elem_size != 4 && grad_elem_size != 8)) return NF_ERR_INVALID;
    if (n == 0) return NF_OK;
    const double c1 = 1.0 - ::pow(b1, (double)(t + 1)), c2 = 1.0 - ::pow(b2, (double)(t + 1));
    long long blocks = (n + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (grad_elem_size == 4)
        nf::adam_step_kernel<float><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(params, (const float*)grad, m, v, n, step_size, b1, b2, eps, c1, c2);
    else
        nf::adam_step_kernel<double><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(params, (const double*)grad, m, v, n, step_size, b1, b2, eps, c1, c2);
    NF_CUDA_OK(cudaGetLastError());
    return NF_OK;
}
