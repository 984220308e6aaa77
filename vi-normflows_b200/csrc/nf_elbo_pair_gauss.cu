// Instantiations of the packed-pair free-energy kernel (see nf_elbo_pair_launch.cuh).
#include "nf_elbo_pair_launch.cuh"

namespace nf {

int pair_group_gauss(int tgt, int D, NF_PAIR_ARGS) {
    if (tgt == NF_TARGET_MIXDIAG) {
        if (D == 1) return dispatch_pair<1, NF_TARGET_MIXDIAG>(NF_PAIR_PASS);
        if (D == 2) return dispatch_pair<2, NF_TARGET_MIXDIAG>(NF_PAIR_PASS);
    }
    if (tgt == NF_TARGET_GAUSSDIAG) {
        if (D == 1) return dispatch_pair<1, NF_TARGET_GAUSSDIAG>(NF_PAIR_PASS);
        if (D == 2) return dispatch_pair<2, NF_TARGET_GAUSSDIAG>(NF_PAIR_PASS);
    }
    if (tgt == NF_TARGET_MIXFULL) {
        if (D == 1) return dispatch_pair<1, NF_TARGET_MIXFULL>(NF_PAIR_PASS);
        if (D == 2) return dispatch_pair<2, NF_TARGET_MIXFULL>(NF_PAIR_PASS);
    }
    return NF_ERR_UNSUPPORTED;
}

}  // namespace nf
