"""ctypes loader for libnfcore.so -- the only way Python reaches the CUDA kernels.

There is no CPU fallback: if the library is missing, or no sm_100 device is
current when a kernel is requested, an exception is raised.
"""
import ctypes
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(HERE), 'lib', 'libnfcore.so')
HEADER = os.path.join(os.path.dirname(os.path.dirname(HERE)), 'include', 'nfcore.h')


class NFError(RuntimeError):
    pass


_lib = None

c_void_p, c_int, c_int64, c_size_t = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_size_t
c_float, c_double, c_char_p = ctypes.c_float, ctypes.c_double, ctypes.c_char_p


def _elbo_args(real):
    return [c_void_p, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int,
            real, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]


PROTOTYPES = {
    'nf_status_string': (c_char_p, [c_int]),
    'nf_last_cuda_error': (c_char_p, []),
    'nf_version': (c_int, []),
    'nf_built_for_sm': (c_int, []),
    'nf_check_device': (c_int, []),
    'nf_script_constants_host': (c_int, [c_void_p, c_int, c_int, c_int, c_void_p]),
    'nf_script_chain_host': (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, ctypes.c_double, c_void_p]),
    'nf_planar_elbo_workspace_bytes': (c_size_t, [c_int, c_int, c_int]),
    'nf_planar_elbo_workspace_bytes_n': (c_size_t, [c_int64, c_int, c_int, c_int]),
    'nf_planar_elbo_fwd_bwd_f32': (c_int, _elbo_args(c_float)),
    'nf_planar_elbo_fwd_bwd_f64': (c_int, _elbo_args(c_double)),
    'nf_planar_elbo_fwd_bwd_noalloc_f32': (c_int, _elbo_args(c_float)[:17] + [c_void_p]),
    'nf_script_train_workspace_bytes': (c_size_t, [c_int, c_int, c_int]),
    'nf_script_train_workspace_bytes_n': (c_size_t, [c_int64, c_int, c_int, c_int]),
    'nf_script_train_f32': (c_int, [c_void_p, c_int64, c_int64, c_int, c_int, c_void_p, c_int, c_void_p, c_int, c_float, c_int] + [c_float] * 4 + [c_int64, c_int, c_int64, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    'nf_script_train_f64': (c_int, [c_void_p, c_int64, c_int64, c_int, c_int, c_void_p, c_int, c_void_p, c_int, c_double, c_int] + [c_double] * 4 + [c_int64, c_int, c_int64, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    'nf_script_train_sharded_f32': (c_int, [c_void_p, c_int64, c_int64, c_int64, c_int, c_int, c_void_p, c_int, c_void_p, c_int, c_float, c_int] + [c_float] * 4
                                    + [c_int64, c_int, c_int64, c_void_p, c_void_p, c_void_p, c_size_t, ctypes.POINTER(c_void_p), c_int, c_int, ctypes.c_uint64,
                                       c_void_p, c_void_p]),
    'nf_session_create': (c_int, [ctypes.POINTER(c_void_p), c_int64, c_int, c_int, c_int]),
    'nf_session_destroy': (None, [c_void_p]),
    'nf_session_elbo_host_f32': (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int, c_void_p, c_int, c_float, c_int, c_void_p]),
    'nf_session_elbo_host_f64': (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int, c_void_p, c_int, c_double, c_int, c_void_p]),
    'nf_session_objective_host_f32': (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int, c_int, c_void_p, c_int, c_float, c_int, c_void_p, c_void_p]),
    'nf_session_objective_host_f64': (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int, c_int, c_void_p, c_int, c_double, c_int, c_void_p, c_void_p]),
    'nf_session_objective_host_exchange_f32': (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int, c_int, c_void_p, c_int, c_float,
                                               ctypes.POINTER(c_void_p), c_int, c_int, ctypes.c_uint64, c_void_p, c_void_p, c_void_p]),
    'nf_session_objective_rng_f32': (c_int, [c_void_p, ctypes.c_uint64, ctypes.c_uint64, c_int64, c_void_p, c_int, c_int, c_void_p, c_int, c_float, c_void_p, c_void_p]),
    'nf_planar_flow_f32': (c_int, [c_void_p] * 4 + [c_int64, c_int, c_int, c_int, c_void_p, c_void_p]),
    'nf_planar_flow_f64': (c_int, [c_void_p] * 4 + [c_int64, c_int, c_int, c_int, c_void_p, c_void_p]),
    'nf_planar_amortised_fwd_f32': (c_int, [c_void_p] * 6 + [c_int64, c_int, c_int, c_float] + [c_void_p] * 4),
    'nf_planar_amortised_fwd_f64': (c_int, [c_void_p] * 6 + [c_int64, c_int, c_int, c_double] + [c_void_p] * 4),
    'nf_planar_amortised_bwd_f32': (c_int, [c_void_p] * 7 + [c_float, c_int64, c_int, c_int, c_float] + [c_void_p] * 6),
    'nf_planar_amortised_bwd_f64': (c_int, [c_void_p] * 7 + [c_double, c_int64, c_int, c_int, c_double] + [c_void_p] * 6),
    'nf_planar_amortised_bwd_v_f32': (c_int, [c_void_p] * 8 + [c_int64, c_int, c_int, c_float] + [c_void_p] * 6),
    'nf_planar_amortised_bwd_v_f64': (c_int, [c_void_p] * 8 + [c_int64, c_int, c_int, c_double] + [c_void_p] * 6),
    'nf_adam_step': (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int64, c_double, c_double, c_double, c_double, c_int64, c_void_p]),
    'nf_bernoulli_ll_dev_f32': (c_int, [c_void_p, c_void_p, c_int64, c_float] + [c_void_p] * 5),
    'nf_bernoulli_ll_dev_f64': (c_int, [c_void_p, c_void_p, c_int64, c_double] + [c_void_p] * 5),
    'nf_mlp_param_count': (c_int64, [c_int] * 4),
    'nf_mlp_forward_workspace_bytes': (c_size_t, [c_int64] + [c_int] * 6),
    'nf_mlp_backward_workspace_bytes': (c_size_t, [c_int64] + [c_int] * 5),
    'nf_mlp_forward_f32': (c_int, [c_void_p, c_void_p, c_int64] + [c_int] * 6 + [c_void_p] * 3),
    'nf_mlp_forward_f64': (c_int, [c_void_p, c_void_p, c_int64] + [c_int] * 6 + [c_void_p] * 3),
    'nf_mlp_backward_f32': (c_int, [c_void_p, c_void_p, c_int64] + [c_int] * 6 + [c_void_p] * 7),
    'nf_mlp_backward_f64': (c_int, [c_void_p, c_void_p, c_int64] + [c_int] * 6 + [c_void_p] * 7),
    'nf_gemm_tn_tf32x3': (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int, c_int, c_int, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int, c_void_p]),
    'nf_allreduce_small_buffer_bytes': (c_size_t, [c_int]),
    'nf_allreduce_small_max_count': (c_int, []),
    'nf_allreduce_small_f32': (c_int, [ctypes.POINTER(c_void_p), c_int, c_int, c_void_p, c_void_p, c_int, ctypes.c_uint64, c_void_p, c_void_p]),
    'nf_target_density_f32': (c_int, [c_void_p, c_int64, c_int, c_int, c_void_p, c_int, c_void_p, c_void_p]),
    'nf_target_density_f64': (c_int, [c_void_p, c_int64, c_int, c_int, c_void_p, c_int, c_void_p, c_void_p]),
    'nf_planar_chain_forward_f32': (c_int, [c_void_p, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_void_p, c_void_p, c_void_p]),
    'nf_planar_chain_forward_f64': (c_int, [c_void_p, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_double, c_void_p, c_void_p, c_void_p]),
    'nf_philox_normal_f32': (c_int, [ctypes.c_uint64, ctypes.c_uint64, c_int64, c_int, c_void_p, c_void_p]),
    'nf_planar_elbo_fwd_bwd_rng_f32': (c_int, [ctypes.c_uint64, ctypes.c_uint64, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                                               c_void_p, c_int, c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    'nf_allreduce_small_counted_f32': (c_int, [ctypes.POINTER(c_void_p), c_int, c_int, c_void_p, c_int, c_int64, ctypes.c_uint64, c_void_p]),
    'nf_planar_elbo_fwd_bwd_exchange_f32': (c_int, [c_void_p, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int,
                                                    c_float, c_void_p, c_void_p, c_size_t, ctypes.POINTER(c_void_p), c_int, c_int, ctypes.c_uint64, c_void_p]),
    'nf_bernoulli_workspace_bytes': (c_size_t, [c_int]),
    'nf_bernoulli_ll_f32': (c_int, [c_void_p, c_void_p, c_int64, c_float, c_float] + [c_void_p] * 4),
    'nf_bernoulli_ll_f64': (c_int, [c_void_p, c_void_p, c_int64, c_double, c_double] + [c_void_p] * 4),
}


def declared_symbols():
    """Every function the public header declares (used by the CPU export test)."""
    text = open(HEADER).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(nf_[a-z0-9_]+)\s*\(', text)))


def load():
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise NFError('libnfcore.so not found at %s -- run `python __graft_entry__.py build` '
                          '(there is no CPU fallback)' % LIB_PATH)
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(status):
    if status != 0:
        lib = load()
        msg = lib.nf_status_string(status).decode()
        if status == -4:
            msg += ': ' + lib.nf_last_cuda_error().decode()
        raise NFError('libnfcore: %s (status %d)' % (msg, status))
