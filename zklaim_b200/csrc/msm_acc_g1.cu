// msm_acc_g1.cu -- the hot kernel of the prover: G1 bucket accumulation (XYZZ mixed additions)
// with the Montgomery multiplier fully inlined.
#include "msm_acc_impl.cuh"
namespace zkb {
template int launch_accumulate<Fq>(zkb_ctx *, const Affine<Fq> *, const uint32_t *, const uint32_t *, uint32_t,
                                   uint32_t, XYZZ<Fq> *, XYZZ<Fq> *);
}
