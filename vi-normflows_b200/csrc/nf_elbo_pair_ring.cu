// Instantiations of the packed-pair free-energy kernel (see nf_elbo_pair_launch.cuh).
#include "nf_elbo_pair_launch.cuh"

namespace nf {

int pair_group_ring(int tgt, int D, NF_PAIR_ARGS) {
    if (D != 2) return NF_ERR_UNSUPPORTED;
    if (tgt == NF_TARGET_U1) return dispatch_pair<2, NF_TARGET_U1>(NF_PAIR_PASS);
    if (tgt == NF_TARGET_TRIAL1) return dispatch_pair<2, NF_TARGET_TRIAL1>(NF_PAIR_PASS);
    return NF_ERR_UNSUPPORTED;
}

}  // namespace nf
