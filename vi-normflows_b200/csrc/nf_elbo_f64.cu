// float64 (reference-precision mode) instantiation of the shared-parameter free-energy kernels.
#include "nf_elbo_impl.cuh"

extern "C" {

int nf_planar_elbo_fwd_bwd_f64(const double* eps, int64_t N, int D, int K, const double* w, const double* s,
                               const double* b, const double* c, int target_id, const double* target_params,
                               int n_target_params, double log_eps, double* out_F, double* g_w, double* g_s,
                               double* g_b, double* g_c, void* workspace, size_t workspace_bytes, void* stream) {
    return nf::elbo_entry<double>(eps, N, D, K, w, s, b, c, target_id, target_params, n_target_params, log_eps,
                                  out_F, g_w, g_s, g_b, g_c, workspace, workspace_bytes, stream);
}

int nf_session_elbo_host_f64(nf_session* s, const double* eps_host, int64_t N, const double* layer_consts_host, int target_id,
                             const double* target_params_host, int n_target_params, double log_eps, int backward,
                             double* out_host) {
    return nf::session_elbo_host<double>(reinterpret_cast<nf::Session*>(s), eps_host, N, layer_consts_host, target_id,
                                         target_params_host, n_target_params, log_eps, backward, out_host);
}

int nf_session_objective_host_f64(nf_session* s, const double* eps_host, int64_t N, const double* params, int semantics, int target_id,
                                  const double* target_params_host, int n_target_params, double log_eps, int backward, double* F_parts,
                                  double* grad) {
    return nf::session_objective_host<double>(reinterpret_cast<nf::Session*>(s), eps_host, N, params, semantics, target_id,
                                              target_params_host, n_target_params, log_eps, backward, F_parts, grad);
}

int nf_script_train_f64(const double* eps_all, int64_t n_iters, int64_t N, int D, int K, double* params, int target_id,
                        const double* target_params, int n_target_params, double log_eps, int optimizer, double step_size, double h1,
                        double h2, double h3, int64_t first_iter, int resume, int64_t eps_stride, double* opt_state, double* trace_F,
                        void* workspace, size_t workspace_bytes, void* stream) {
    return nf::script_train<double>(eps_all, n_iters, N, D, K, params, target_id, target_params, n_target_params, log_eps, optimizer,
                                step_size, h1, h2, h3, first_iter, resume, eps_stride, opt_state, trace_F, workspace,
                                workspace_bytes, stream);
}

}  // extern "C"
