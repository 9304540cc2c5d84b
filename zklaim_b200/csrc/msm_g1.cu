// msm_g1.cu -- G1 instantiation of the latency-bound MSM stages (noinline multiplier).
#ifndef ZKB_INLINE_FINISH
#define ZKB_NOINLINE_MUL 1
#endif
#include "msm_finish_impl.cuh"
namespace zkb {
template int msm_finish<Fq>(zkb_ctx *, const MsmPlan *, const JobTab &, int, uint32_t, XYZZ<Fq> *);
template int plan_upload_job<Fq>(zkb_ctx *, MsmPlan *, int, const uint64_t *, bool);
template int results_to_host<Fq>(zkb_ctx *, const XYZZ<Fq> *, size_t, uint64_t *);
template int plan_synth_job<Fq>(zkb_ctx *, MsmPlan *, int, const Affine<Fq> &, uint64_t);
}
