// msm_g2.cu -- G2 (Fq2 coordinates) instantiation of all MSM stages (noinline multiplier: an
// XYZZ<Fq2> accumulator alone is 64 registers, the inlined code would only spill more).
#define ZKB_NOINLINE_MUL 1
#include "msm_acc_impl.cuh"
#include "msm_finish_impl.cuh"
namespace zkb {
template int launch_accumulate<Fq2>(zkb_ctx *, const Affine<Fq2> *, const uint32_t *, const uint32_t *,
                                    uint32_t, uint32_t, XYZZ<Fq2> *, XYZZ<Fq2> *);
template int msm_finish<Fq2>(zkb_ctx *, const MsmPlan *, const JobTab &, int, uint32_t, XYZZ<Fq2> *);
template int plan_upload_job<Fq2>(zkb_ctx *, MsmPlan *, int, const uint64_t *, bool);
template int results_to_host<Fq2>(zkb_ctx *, const XYZZ<Fq2> *, size_t, uint64_t *);
}
