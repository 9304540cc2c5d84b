// msm_acc_impl.cuh -- bucket accumulation kernel (see msm.cu for the design); included by the
// per-group translation units so that G1 gets the fully inlined multiplier.
#pragma once
#include "msm_impl.cuh"

namespace zkb {
// ------------------------------------------------------------------------------------
// bucket accumulation: fixed-length runs of the sorted pair list
// ------------------------------------------------------------------------------------
template <class F>
__global__ void __launch_bounds__(128)
k_accumulate(const Affine<F> *__restrict__ pts, const uint32_t *__restrict__ ent, const uint32_t *__restrict__ ebkt,
             const uint32_t *__restrict__ start, uint32_t B, uint32_t rmax, XYZZ<F> *__restrict__ buckets,
             XYZZ<F> *__restrict__ parts)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t E = start[B];
    const uint32_t S = msm_run_length(E, rmax);
    const uint64_t lo64 = (uint64_t)t * S;
    if (lo64 >= E) return;
    const uint32_t lo = (uint32_t)lo64;
    const uint32_t hi = (lo64 + S < E) ? lo + S : E;

    XYZZ<F> acc = XYZZ<F>::inf();
    uint32_t cur = ebkt[lo];
    uint32_t seg_lo = lo;
    for (uint32_t e = lo; e < hi; e++) {
        const uint32_t b = ebkt[e];
        if (b != cur) {
            // flush the finished segment [seg_lo, e) of bucket `cur`
            if (start[cur] >= lo) // began inside this run (and ended inside: it is complete)
                st_vec(buckets + cur, acc);
            else
                st_vec(parts + 2 * (size_t)t, acc); // reached back over the run's head
            acc = XYZZ<F>::inf();
            cur = b;
            seg_lo = e;
        }
        const uint32_t id = ent[e];
        const Affine<F> p = ld_vec_ro(pts + (id & 0x7fffffffu));
        acc.madd(p, (id >> 31) != 0);
    }
    // last segment [seg_lo, hi)
    const bool begins_inside = start[cur] >= lo;
    const bool ends_inside = start[cur + 1] <= hi;
    if (begins_inside && ends_inside)
        st_vec(buckets + cur, acc);
    else if (seg_lo == lo)
        st_vec(parts + 2 * (size_t)t, acc); // touches the head of the run (maybe the whole run)
    else
        st_vec(parts + 2 * (size_t)t + 1, acc); // touches only the tail
}

template <class F>
int launch_accumulate(zkb_ctx *ctx, const Affine<F> *pts, const uint32_t *ent, const uint32_t *ebkt,
                      const uint32_t *start, uint32_t B, uint32_t rmax, XYZZ<F> *buckets, XYZZ<F> *parts)
{
    k_accumulate<F><<<ceil_div(rmax, 128), 128, 0, ctx->stream>>>(pts, ent, ebkt, start, B, rmax, buckets, parts);
    ZKB_LAUNCH_CHECK(ctx);
    return ZKB_OK;
}

} // namespace zkb
