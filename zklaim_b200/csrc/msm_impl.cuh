// msm_impl.cuh -- shared declarations of the MSM translation units (see msm.cu for the design).
#pragma once
#include "common.cuh"

struct zkb_bases {
    int group = 0;
    int dev = 0;
    size_t n = 0;
    int c = 0, W = 0, F = 1, G = 0;
    void *d_pts = nullptr; // [F][n] affine
    size_t bytes = 0;
};

namespace zkb {

enum { SB_PLAIN = 0, SB_HIST, SB_START, SB_CURSOR, SB_BLOCKSUM, SB_ENT, SB_EBKT, SB_PARTS, SB_BUCKETS, SB_L1, SB_L2,
       SB_OUT, SB_SCAL, SB_PTS_STAGE };

constexpr int SCAN_T = 256, SCAN_E = 8, SCAN_BLK = SCAN_T * SCAN_E;

// msm.cu: scalar recoding + counting sort of the (point, window) pairs; fills ent / ebkt / start
int msm_sort_pairs(zkb_ctx *ctx, const zkb_bases *b, size_t offset, const Fr *d_scalars, size_t n);

// msm_acc_g1.cu / msm_g2.cu: bucket accumulation over fixed-length runs
template <class F>
int launch_accumulate(zkb_ctx *ctx, const Affine<F> *pts, const uint32_t *ent, const uint32_t *ebkt,
                      const uint32_t *start, uint32_t B, uint32_t S, XYZZ<F> *buckets, XYZZ<F> *parts, size_t runs);

// msm_g1.cu / msm_g2.cu: fix-up, bucket reduction, final combine; base precomputation
template <class F> int msm_finish(zkb_ctx *ctx, const zkb_bases *b, size_t n, uint32_t S, uint64_t *out_host);
template <class F> int bases_build(zkb_ctx *ctx, zkb_bases *b, const uint64_t *affine_host);

inline uint32_t msm_run_length(const zkb_ctx *ctx, size_t Emax)
{
    // enough runs to fill the machine, bounded scratch
    uint32_t S = 16;
    while (Emax / S > (1u << 21)) S <<= 1;
    while (S > 4 && Emax / S < (size_t)ctx->sm_count * 1024) S >>= 1;
    return S;
}

} // namespace zkb
