// msm_g1.cu -- G1 instantiation of the latency-bound MSM stages (noinline multiplier).
#define ZKB_NOINLINE_MUL 1
#include "msm_finish_impl.cuh"
namespace zkb {
template int msm_finish<Fq>(zkb_ctx *, const zkb_bases *, size_t, uint32_t, uint64_t *);
template int bases_build<Fq>(zkb_ctx *, zkb_bases *, const uint64_t *);
}
