// msm_acc_impl.cuh -- bucket accumulation kernel (see msm.cu for the design); included by the
// per-group translation units so that G1 gets the fully inlined multiplier.
#pragma once
#include "msm_impl.cuh"

namespace zkb {
// ------------------------------------------------------------------------------------
// bucket accumulation: fixed-length runs of the sorted pair list
// ------------------------------------------------------------------------------------
// The sorted list holds point entries only; which bucket a position belongs to follows from start[]: a thread finds
// the bucket of its first position by binary search and then walks the boundaries.
// resident blocks per SM asked of the compiler for G1 (106 registers, no spills at 4).  Measured: 5 blocks (102
// registers, 84 B of spill loads) 6.19 ms against 6.21 ms per batch-64 launch -- noise; 6 blocks (85 registers) 6.37 ms.
#ifndef ZKB_ACC_MIN_BLOCKS_G1
#define ZKB_ACC_MIN_BLOCKS_G1 4
#endif
template <class F>
__global__ void __launch_bounds__(128, sizeof(F) == 32 ? ZKB_ACC_MIN_BLOCKS_G1 : 1)
k_accumulate(const Affine<F> *__restrict__ pts, const uint32_t *__restrict__ ent, const uint32_t *__restrict__ start,
             uint32_t B, uint32_t rmax, XYZZ<F> *__restrict__ buckets, XYZZ<F> *__restrict__ parts)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t E = start[B];
    const uint32_t S = msm_run_length<F>(E, rmax);
    const uint64_t lo64 = (uint64_t)t * S;
    if (lo64 >= E) return;
    const uint32_t lo = (uint32_t)lo64;
    const uint32_t hi = (lo64 + S < E) ? lo + S : E;

    // cur = the last bucket with start[cur] <= lo (it is non-empty: start[cur + 1] > lo)
    uint32_t cur = 0;
    {
        uint32_t a = 0, b = B; // invariant: start[a] <= lo < start[b]
        while (b - a > 1) {
            const uint32_t mid = a + ((b - a) >> 1);
            if (start[mid] <= lo)
                a = mid;
            else
                b = mid;
        }
        cur = a;
    }
    uint32_t cur_begin = start[cur], cur_end = start[cur + 1];
    XYZZ<F> acc = XYZZ<F>::inf();
    uint32_t seg_lo = lo;
    for (uint32_t e = lo; e < hi; e++) {
        if (e == cur_end) {
            // flush the finished segment [seg_lo, e) of bucket `cur`
            if (cur_begin >= lo) // began inside this run (and ended inside: it is complete)
                st_vec(buckets + cur, acc);
            else
                st_vec(parts + 2 * (size_t)t, acc); // reached back over the run's head
            acc = XYZZ<F>::inf();
            do { // next non-empty bucket
                cur++;
                cur_begin = cur_end;
                cur_end = start[cur + 1];
            } while (cur_end == cur_begin);
            seg_lo = e;
        }
        const uint32_t id = ent[e];
        const Affine<F> p = ld_vec_ro(pts + (id & 0x7fffffffu));
        acc.madd(p, (id >> 31) != 0);
    }
    // last segment [seg_lo, hi)
    const bool begins_inside = cur_begin >= lo;
    const bool ends_inside = cur_end <= hi;
    if (begins_inside && ends_inside)
        st_vec(buckets + cur, acc);
    else if (seg_lo == lo)
        st_vec(parts + 2 * (size_t)t, acc); // touches the head of the run (maybe the whole run)
    else
        st_vec(parts + 2 * (size_t)t + 1, acc); // touches only the tail
}

template <class F>
int launch_accumulate(zkb_ctx *ctx, const Affine<F> *pts, const uint32_t *ent, const uint32_t *start, uint32_t B,
                      uint32_t rmax, XYZZ<F> *buckets, XYZZ<F> *parts)
{
    k_accumulate<F><<<ceil_div(rmax, 128), 128, 0, ctx->stream>>>(pts, ent, start, B, rmax, buckets, parts);
    ZKB_LAUNCH_CHECK(ctx);
    return ZKB_OK;
}

} // namespace zkb
