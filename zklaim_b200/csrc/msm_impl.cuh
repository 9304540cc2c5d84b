// msm_impl.cuh -- shared declarations of the MSM translation units (see msm.cu for the design).
#pragma once
#include "common.cuh"

namespace zkb {

// One query ("job") of an MSM plan.  Trivially copyable: the whole table travels to the kernels by value.
struct MsmJob {
    uint32_t item_begin;  // first item of this job in the plan's item space
    uint32_t n;           // items of this job that take part in this run
    uint32_t base_first;  // index inside a copy of the base that belongs to item 0 (zkb_msm's `offset`)
    uint32_t n_total;     // bases per precomputed copy
    uint32_t pts_off;     // first point of copy 0 inside the plan's point array
    uint32_t scal_first;  // scalar index of item 0 when the plan has no scalar map
    uint32_t bucket_off;  // first bucket of the job inside one proof's bucket space
    uint16_t c, W, G, F;  // window bits, windows, bucket sets, precomputed copies (W <= G * F)
};

constexpr int MSM_MAX_JOBS = 4;
constexpr int MSM_MAX_C = 20; // widest window: 2^19 buckets per set

struct JobTab {
    uint32_t njobs, n_items, Bp, pad; // Bp = buckets of one proof (all jobs, all sets)
    MsmJob j[MSM_MAX_JOBS];
};

// A group's precomputed bases plus the job table: immutable after build, lives on one device.
struct MsmPlan {
    int group = 0;
    int dev = 0;
    JobTab tab = {};
    void *d_pts = nullptr;       // affine points, per job copy-major [F][n_total]
    uint32_t *d_sidx = nullptr;  // optional scalar map: item -> index into one proof's scalar vector
    size_t pts_count = 0;
    size_t bytes = 0;
};

constexpr int SCAN_T = 256, SCAN_E = 8, SCAN_BLK = SCAN_T * SCAN_E;

// window choice: minimise  n_eff * W * (mixed add) + 2^c * (full add)  field multiplications
int msm_choose_c(size_t n_eff, int group, int precompute_factor);
// Montgomery -> plain integers (fp.tcc:228-238), n elements
int msm_to_plain(zkb_ctx *ctx, const void *d_mont, uint32_t *d_plain, size_t n);
// fills job `k` of `plan` (host side only); returns 0 or ZKB_ERR_ARG
int plan_add_job(zkb_ctx *ctx, MsmPlan *plan, size_t n_total, int c, int precompute_factor);
// allocates the point array of a plan after all jobs were added
int plan_alloc(zkb_ctx *ctx, MsmPlan *plan);
void plan_free(MsmPlan *plan);

// msm.cu: digit extraction + counting sort of the (item, window) pairs of `batch` proofs
int msm_sort_pairs(zkb_ctx *ctx, const MsmPlan *plan, const JobTab &tab, const uint32_t *d_plain, size_t zstride,
                   int batch);

// msm_acc_g1.cu / msm_g2.cu: bucket accumulation over fixed-length runs of the sorted pair list
template <class F>
int launch_accumulate(zkb_ctx *ctx, const Affine<F> *pts, const uint32_t *ent, const uint32_t *start, uint32_t B,
                      uint32_t rmax, XYZZ<F> *buckets, XYZZ<F> *parts);

// msm_g1.cu / msm_g2.cu: fix-up, bucket reduction, per-job result (XYZZ, device, [batch][njobs]); base precomputation
template <class F>
int msm_finish(zkb_ctx *ctx, const MsmPlan *plan, const JobTab &tab, int batch, uint32_t rmax, XYZZ<F> *d_res);
template <class F> int plan_upload_job(zkb_ctx *ctx, MsmPlan *plan, int job, const uint64_t *affine, bool from_device = false);
template <class F> int plan_synth_job(zkb_ctx *ctx, MsmPlan *plan, int job, const Affine<F> &gen, uint64_t seed);
template <class F> int results_to_host(zkb_ctx *ctx, const XYZZ<F> *d_res, size_t count, uint64_t *out_jacobian);

// the whole pipeline: plain scalars (device, `zstride` words between proofs) -> d_res[batch][njobs]
template <class F>
int msm_run(zkb_ctx *ctx, const MsmPlan *plan, const JobTab &tab, const uint32_t *d_plain, size_t zstride, int batch,
            XYZZ<F> *d_res);

// Run length of the bucket accumulation: the sorted pair list (E pairs, known on the device only) is cut into at
// most `rmax` runs of S = 32, 64, 128, ... pairs; every kernel derives S from E the same way.  (S started at 4 in
// round 1: with few pairs most of the additions then moved into the fix-up pass, which merges run-boundary partials
// with FULL additions -- 0.48 of the 2.1 ms of a single zklaim proof.)
// G2 (Fq2 coordinates: three times the work per addition, a third of the resident threads) keeps shorter runs.
template <class F> __host__ __device__ inline uint32_t msm_run_length(uint32_t E, uint32_t rmax)
{
    uint32_t S = sizeof(F) > 32 ? 8u : 32u;
    while ((uint64_t)S * rmax < E) S <<= 1;
    return S;
}
// runs per MSM launch: four waves of accumulation threads
inline uint32_t msm_max_runs(const zkb_ctx *ctx) { return (uint32_t)ctx->sm_count * 2048u; }

} // namespace zkb

struct zkb_bases {
    zkb::MsmPlan plan;
    size_t n = 0;
};
