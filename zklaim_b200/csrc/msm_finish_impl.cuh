// msm_finish_impl.cuh -- fix-up of boundary partials, weighted bucket reduction, final combine and
// base precomputation (see msm.cu for the design).  Latency-bound: built with ZKB_NOINLINE_MUL.
#pragma once
#include "msm_impl.cuh"
#include <algorithm>

namespace zkb {
// word-striped shared-memory staging (conflict-free for any struct size)
template <class PT> __device__ __forceinline__ void sm_put(uint32_t *sm, int T, int t, const PT &v)
{
    const uint32_t *w = reinterpret_cast<const uint32_t *>(&v);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(PT) / 4); i++) sm[i * T + t] = w[i];
}
template <class PT> __device__ __forceinline__ PT sm_get(const uint32_t *sm, int T, int t)
{
    PT v;
    uint32_t *w = reinterpret_cast<uint32_t *>(&v);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(PT) / 4); i++) w[i] = sm[i * T + t];
    return v;
}

// merges the boundary partials of every bucket that spans more than one run; empty buckets -> infinity.
// One lane per bucket for short spans, the whole warp for medium spans; a bucket that spans very many runs
// (the "scalar == 1" bucket of a Groth16 witness holds thousands of pairs) is queued for k_fixup_wide, where a
// whole thread block reduces it.
constexpr int FIXUP_WIDE_T = 256;  // threads of a k_fixup_wide block
constexpr int FIXUP_CHUNK = 1024;  // runs per block there: a bucket of the "scalar == 1" kind spans 10^5 runs at 2^22 constraints
constexpr int FIXUP_T = 128, FIXUP_WARP = 8, FIXUP_BLOCK = 64;
template <class F>
__global__ void __launch_bounds__(FIXUP_T)
k_fixup(const uint32_t *__restrict__ start, uint32_t B, uint32_t rmax, XYZZ<F> *__restrict__ buckets,
        const XYZZ<F> *__restrict__ parts, uint32_t *__restrict__ wide_list, uint32_t max_chunks)
{
    extern __shared__ uint32_t sm[];
    const uint32_t S = msm_run_length<F>(start[B], rmax);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t *wsm = sm + (size_t)wid * 32 * (sizeof(XYZZ<F>) / 4);
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t s = 0, e = 0;
    if (b < B) {
        s = start[b];
        e = start[b + 1];
    }
    uint32_t t0 = 0, t1 = 0;
    bool multi = false;
    if (b < B) {
        if (s == e) {
            st_vec(buckets + b, XYZZ<F>::inf());
        } else {
            t0 = s / S;
            t1 = (e - 1) / S;
            multi = t1 > t0;
        }
    }
    // medium spans: the warp takes them one after another when only a few of its lanes have one (isolated heavy
    // buckets), each lane keeps its own when most lanes do (uniformly filled windows); long spans go to the queue
    // (neighbouring heavy buckets -- the small digits of window 0 -- would otherwise serialise inside one warp)
    const bool block_wide = multi && (t1 - t0) >= FIXUP_BLOCK;
    const bool medium = multi && !block_wide && (t1 - t0) >= FIXUP_WARP;
    const int n_medium = __popc(__ballot_sync(0xffffffffu, medium));
    const bool wide = medium && n_medium <= 4;
    // (queueing the 16 .. 63-run spans of a crowded warp as well was measured: MSM 2^16 0.82 -> 2.4 ms; not done)
    if (block_wide) { // queue layout: [0] chunk count, [1] bucket count, chunks (b, index), bucket records (b, first chunk, chunks)
        const uint32_t n = (t1 - t0 + FIXUP_CHUNK) / FIXUP_CHUNK;
        const uint32_t first = atomicAdd(wide_list, n), rec = atomicAdd(wide_list + 1, 1u);
        uint32_t *chunks = wide_list + 2, *recs = wide_list + 2 + 2 * (size_t)max_chunks;
        recs[3 * rec] = b;
        recs[3 * rec + 1] = first;
        recs[3 * rec + 2] = n;
        for (uint32_t i = 0; i < n; i++) {
            chunks[2 * (first + i)] = b;
            chunks[2 * (first + i) + 1] = i;
        }
    }
    if (multi && !wide && !block_wide) {
        XYZZ<F> acc = ld_vec(parts + 2 * (size_t)t0 + ((s == t0 * S) ? 0 : 1));
        for (uint32_t t = t0 + 1; t <= t1; t++) acc.add(ld_vec(parts + 2 * (size_t)t));
        st_vec(buckets + b, acc);
    }
    unsigned todo = __ballot_sync(0xffffffffu, wide);
    while (todo) {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        const uint32_t ws = __shfl_sync(0xffffffffu, s, src);
        const uint32_t wt0 = __shfl_sync(0xffffffffu, t0, src), wt1 = __shfl_sync(0xffffffffu, t1, src);
        const uint32_t wb = __shfl_sync(0xffffffffu, b, src);
        // partial k of the span: k = 0 is the head partial of run t0, k >= 1 the head partial of run t0 + k
        XYZZ<F> acc = XYZZ<F>::inf();
        for (uint32_t t = wt0 + lane; t <= wt1; t += 32) {
            const size_t idx = (t == wt0) ? 2 * (size_t)t + ((ws == t * S) ? 0 : 1) : 2 * (size_t)t;
            acc.add(ld_vec(parts + idx));
        }
        sm_put(wsm, 32, lane, acc);
        __syncwarp();
        for (int d = 16; d > 0; d >>= 1) {
            if (lane < d) {
                acc.add(sm_get<XYZZ<F>>(wsm, 32, lane + d));
                sm_put(wsm, 32, lane, acc);
            }
            __syncwarp();
        }
        if (lane == 0) st_vec(buckets + wb, acc);
        __syncwarp();
    }
}

// the queued buckets: a block per chunk of FIXUP_CHUNK runs (grid-stride over the chunk queue) ...
template <class F>
__global__ void __launch_bounds__(FIXUP_WIDE_T)
k_fixup_wide(const uint32_t *__restrict__ start, uint32_t B, uint32_t rmax, const XYZZ<F> *__restrict__ parts,
             const uint32_t *__restrict__ wide_list, XYZZ<F> *__restrict__ chunk_sums)
{
    extern __shared__ uint32_t sm[];
    const uint32_t S = msm_run_length<F>(start[B], rmax);
    const int T = blockDim.x, t = threadIdx.x;
    const uint32_t count = wide_list[0];
    const uint32_t *chunks = wide_list + 2;
    for (uint32_t i = blockIdx.x; i < count; i += gridDim.x) {
        const uint32_t b = chunks[2 * i], ch = chunks[2 * i + 1];
        const uint32_t s = start[b], e = start[b + 1];
        const uint32_t t0 = s / S, t1 = (e - 1) / S;
        const uint32_t lo = t0 + ch * FIXUP_CHUNK;
        const uint32_t hi = (t1 - lo >= (uint32_t)FIXUP_CHUNK) ? lo + FIXUP_CHUNK - 1 : t1;
        XYZZ<F> acc = XYZZ<F>::inf();
        for (uint32_t r = lo + t; r <= hi; r += T) {
            const size_t idx = (r == t0) ? 2 * (size_t)r + ((s == r * S) ? 0 : 1) : 2 * (size_t)r;
            acc.add(ld_vec(parts + idx));
        }
        sm_put(sm, T, t, acc);
        __syncthreads();
        for (int d = T >> 1; d > 0; d >>= 1) {
            if (t < d) {
                acc.add(sm_get<XYZZ<F>>(sm, T, t + d));
                sm_put(sm, T, t, acc);
            }
            __syncthreads();
        }
        if (t == 0) st_vec(chunk_sums + i, acc);
        __syncthreads();
    }
}
// ... then a block per queued bucket adds up its chunk sums
template <class F>
__global__ void __launch_bounds__(128)
k_fixup_wide_sum(XYZZ<F> *__restrict__ buckets, const uint32_t *__restrict__ wide_list, uint32_t max_chunks,
                 const XYZZ<F> *__restrict__ chunk_sums)
{
    extern __shared__ uint32_t sm[];
    const int T = blockDim.x, t = threadIdx.x;
    const uint32_t count = wide_list[1];
    const uint32_t *recs = wide_list + 2 + 2 * (size_t)max_chunks;
    for (uint32_t i = blockIdx.x; i < count; i += gridDim.x) {
        const uint32_t b = recs[3 * i], first = recs[3 * i + 1], n = recs[3 * i + 2];
        if (n == 1) {
            if (t == 0) st_vec(buckets + b, ld_vec(chunk_sums + first));
            continue;
        }
        XYZZ<F> acc = XYZZ<F>::inf();
        for (uint32_t k = t; k < n; k += T) acc.add(ld_vec(chunk_sums + first + k));
        sm_put(sm, T, t, acc);
        __syncthreads();
        for (int d = T >> 1; d > 0; d >>= 1) {
            if (t < d) {
                acc.add(sm_get<XYZZ<F>>(sm, T, t + d));
                sm_put(sm, T, t, acc);
            }
            __syncthreads();
        }
        if (t == 0) st_vec(buckets + b, acc);
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------
// weighted bucket sum  R = sum_j (j+1) * E_j  via block-parallel suffix scans
// ------------------------------------------------------------------------------------
// inclusive suffix scan over the block: v_t <- sum_{u >= t} v_u
template <class PT> __device__ __forceinline__ void block_suffix_scan(uint32_t *sm, int T, int t, PT &v)
{
    sm_put(sm, T, t, v);
    __syncthreads();
    for (int d = 1; d < T; d <<= 1) {
        const bool has = t + d < T;
        PT o;
        if (has) o = sm_get<PT>(sm, T, t + d);
        __syncthreads();
        if (has) {
            v.add(o);
            sm_put(sm, T, t, v);
        }
        __syncthreads();
    }
}
// sum over the block (T a power of two); valid in thread 0
template <class PT> __device__ __forceinline__ void block_sum(uint32_t *sm, int T, int t, PT &v)
{
    sm_put(sm, T, t, v);
    __syncthreads();
    for (int d = T >> 1; d > 0; d >>= 1) {
        if (t < d) {
            v.add(sm_get<PT>(sm, T, t + d));
            sm_put(sm, T, t, v);
        }
        __syncthreads();
    }
}

// level 1: block handles T*L consecutive buckets of one set; out[set][blk] = (R_blk, S_blk).
// set = blockIdx.y = proof * G + g; its buckets start at proof * Bp + bucket_off + g * nb.
template <class F>
__global__ void __launch_bounds__(128)
k_reduce_l1(const XYZZ<F> *__restrict__ buckets, uint32_t Bp, uint32_t bucket_off, uint32_t G, uint32_t nb, int L,
            int logL, XYZZ<F> *__restrict__ out)
{
    extern __shared__ uint32_t sm[];
    const int T = blockDim.x, t = threadIdx.x;
    const uint32_t set = blockIdx.y;
    const size_t base = (size_t)(set / G) * Bp + bucket_off + (size_t)(set % G) * nb;
    const XYZZ<F> *E = buckets + base + ((size_t)blockIdx.x * T + t) * L;
    XYZZ<F> acc = XYZZ<F>::inf(), tot = XYZZ<F>::inf();
    for (int u = L - 1; u >= 0; u--) {
        acc.add(ld_vec(E + u));
        tot.add(acc);
    }
    XYZZ<F> incl = acc;
    block_suffix_scan(sm, T, t, incl);
    // Above_t = incl_{t+1}
    XYZZ<F> above = XYZZ<F>::inf();
    if (t + 1 < T) above = sm_get<XYZZ<F>>(sm, T, t + 1);
    __syncthreads();
    for (int i = 0; i < logL; i++) above = above.dbl();
    tot.add(above);
    block_sum(sm, T, t, tot);
    if (t == 0) {
        XYZZ<F> *o = out + 2 * ((size_t)set * gridDim.x + blockIdx.x);
        st_vec(o, tot);
        st_vec(o + 1, incl);
    }
}

// level 2: one block per set, thread b holds (R_b, S_b) of level-1 block b (NB a power of two <= blockDim)
//   R_set = sum_b R_b + CH * sum_{b>=1} suffix_b(S),   CH = 2^logCH buckets per level-1 block
template <class F>
__global__ void __launch_bounds__(128)
k_reduce_l2(const XYZZ<F> *__restrict__ l1, int NB, int logCH, XYZZ<F> *__restrict__ out)
{
    extern __shared__ uint32_t sm[];
    const int T = blockDim.x, t = threadIdx.x;
    const uint32_t set = blockIdx.x;
    XYZZ<F> R = XYZZ<F>::inf(), S = XYZZ<F>::inf();
    if (t < NB) {
        R = ld_vec(l1 + 2 * ((size_t)set * NB + t));
        S = ld_vec(l1 + 2 * ((size_t)set * NB + t) + 1);
    }
    block_suffix_scan(sm, T, t, S);
    XYZZ<F> v = (t >= 1 && t < NB) ? S : XYZZ<F>::inf();
    block_sum(sm, T, t, v);
    __syncthreads();
    block_sum(sm, T, t, R);
    if (t == 0) {
        for (int i = 0; i < logCH; i++) v = v.dbl();
        R.add(v);
        st_vec(out + set, R);
    }
}

// ---- the two levels above for ALL jobs of a plan in one launch each -------------------------------------------------
// A single proof leaves the device nearly empty and the reduction of a job is a chain of ~45 dependent point
// additions (0.23 ms in G1): the jobs of a plan (A and B1 queries) are independent, so their chains run side by side
// instead of one after the other.  blockIdx.z = job; blocks outside a job's geometry leave at once.
struct RedGeo {
    uint32_t bucket_off, G, nb, sets;
    int T, L, logL, NB, logCH, c;
    uint32_t l1_off, l2_off; // first (R, S) pair / first set result of this job
};
struct RedGeoTab {
    RedGeo g[MSM_MAX_JOBS];
};
template <class F>
__global__ void __launch_bounds__(128)
k_reduce_l1_jobs(const XYZZ<F> *__restrict__ buckets, uint32_t Bp, RedGeoTab tab, XYZZ<F> *__restrict__ out)
{
    extern __shared__ uint32_t sm[];
    const RedGeo &J = tab.g[blockIdx.z];
    if (blockIdx.x >= (uint32_t)J.NB || blockIdx.y >= J.sets) return;
    const int T = blockDim.x, t = threadIdx.x;
    const uint32_t set = blockIdx.y;
    const size_t base = (size_t)(set / J.G) * Bp + J.bucket_off + (size_t)(set % J.G) * J.nb;
    XYZZ<F> acc = XYZZ<F>::inf(), tot = XYZZ<F>::inf();
    if (t < J.T) {
        const XYZZ<F> *E = buckets + base + ((size_t)blockIdx.x * J.T + t) * J.L;
        for (int u = J.L - 1; u >= 0; u--) {
            acc.add(ld_vec(E + u));
            tot.add(acc);
        }
    }
    XYZZ<F> incl = acc;
    block_suffix_scan(sm, T, t, incl);
    XYZZ<F> above = XYZZ<F>::inf();
    if (t + 1 < T) above = sm_get<XYZZ<F>>(sm, T, t + 1);
    __syncthreads();
    for (int i = 0; i < J.logL; i++) above = above.dbl();
    tot.add(above);
    block_sum(sm, T, t, tot);
    if (t == 0) {
        XYZZ<F> *o = out + 2 * ((size_t)J.l1_off + (size_t)set * J.NB + blockIdx.x);
        st_vec(o, tot);
        st_vec(o + 1, incl);
    }
}
template <class F>
__global__ void __launch_bounds__(128)
k_reduce_l2_jobs(const XYZZ<F> *__restrict__ l1, RedGeoTab tab, XYZZ<F> *__restrict__ out)
{
    extern __shared__ uint32_t sm[];
    const RedGeo &J = tab.g[blockIdx.y];
    if (blockIdx.x >= J.sets) return;
    const int T = blockDim.x, t = threadIdx.x;
    const uint32_t set = blockIdx.x;
    XYZZ<F> R = XYZZ<F>::inf(), S = XYZZ<F>::inf();
    if (t < J.NB) {
        R = ld_vec(l1 + 2 * ((size_t)J.l1_off + (size_t)set * J.NB + t));
        S = ld_vec(l1 + 2 * ((size_t)J.l1_off + (size_t)set * J.NB + t) + 1);
    }
    block_suffix_scan(sm, T, t, S);
    XYZZ<F> v = (t >= 1 && t < J.NB) ? S : XYZZ<F>::inf();
    block_sum(sm, T, t, v);
    __syncthreads();
    block_sum(sm, T, t, R);
    if (t == 0) {
        for (int i = 0; i < J.logCH; i++) v = v.dbl();
        R.add(v);
        st_vec(out + J.l2_off + set, R);
    }
}
template <class F>
__global__ void k_final_jobs(const XYZZ<F> *__restrict__ sets, int batch, RedGeoTab tab, int njobs, XYZZ<F> *__restrict__ res)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x, job = blockIdx.y;
    if (p >= batch) return;
    const RedGeo &J = tab.g[job];
    const XYZZ<F> *s = sets + J.l2_off + (size_t)p * J.G;
    XYZZ<F> r = ld_vec(s + (J.G - 1));
    for (int g = (int)J.G - 2; g >= 0; g--) {
        for (int i = 0; i < J.c; i++) r = r.dbl();
        r.add(ld_vec(s + g));
    }
    st_vec(res + (size_t)p * njobs + job, r);
}

// ---- the same weighted sum for large bucket sets, without a scan over the buckets -------------------------------
// View the nb = H2 * W2 buckets of a set as a grid, b = hi * W2 + lo:
//     sum_b (b + 1) E_b  =  W2 * sum_hi hi * Row_hi  +  sum_lo (lo + 1) * Col_lo,   Row_hi = sum_lo E, Col_lo = sum_hi E.
// The row and column sums are plain sums: two additions per bucket with every lane busy and no shared memory
// (k_reduce_rowcol: one thread per segment of a row or of a column); the two weighted sums that remain have H2 and
// W2 terms only (k_reduce_rowcol_finish, one block per set).  The running-sum form (k_reduce_l1 / l2) costs about
// 3.3 additions per bucket once its block-level scans are counted.
template <class F>
__global__ void __launch_bounds__(128)
k_reduce_rowcol(const XYZZ<F> *__restrict__ buckets, uint32_t Bp, uint32_t bucket_off, uint32_t G, uint32_t nb, int lw, int lh,
                int lseg, XYZZ<F> *__restrict__ part)
{
    const uint32_t set = blockIdx.y;
    const size_t base = (size_t)(set / G) * Bp + bucket_off + (size_t)(set % G) * nb;
    const uint32_t W2 = 1u << lw, H2 = 1u << lh;
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t nrow = H2 << lseg, ncol = W2 << lseg;
    if (t >= nrow + ncol) return;
    XYZZ<F> acc = XYZZ<F>::inf();
    if (t < nrow) { // a segment of a row: consecutive buckets
        const uint32_t len = W2 >> lseg;
        const XYZZ<F> *E = buckets + base + (size_t)(t >> lseg) * W2 + (size_t)(t & ((1u << lseg) - 1u)) * len;
        for (uint32_t u = 0; u < len; u++) acc.add(ld_vec(E + u));
    } else { // a segment of a column: consecutive threads read consecutive buckets
        const uint32_t u0 = t - nrow, col = u0 & (W2 - 1u), seg = u0 >> lw, len = H2 >> lseg;
        const XYZZ<F> *E = buckets + base + (size_t)seg * len * W2 + col;
        for (uint32_t u = 0; u < len; u++) acc.add(ld_vec(E + (size_t)u * W2));
    }
    st_vec(part + (size_t)set * (nrow + ncol) + t, acc);
}

// one block of 2 * H2 threads per set (H2 >= W2): the first half reduces the rows, the second half the columns, side
// by side;  R_set = W2 * sum_{hi >= 1} suffix_hi(Row) + sum_lo suffix_lo(Col)
template <class F>
__global__ void __launch_bounds__(512)
k_reduce_rowcol_finish(const XYZZ<F> *__restrict__ part, int lw, int lh, int lseg, XYZZ<F> *__restrict__ out)
{
    extern __shared__ uint32_t sm_all[];
    const int T = blockDim.x >> 1, half = threadIdx.x >= T ? 1 : 0, t = threadIdx.x - half * T;
    uint32_t *sm = sm_all + (size_t)half * T * (sizeof(XYZZ<F>) / 4);
    const uint32_t set = blockIdx.x, W2 = 1u << lw, H2 = 1u << lh, SEG = 1u << lseg;
    const XYZZ<F> *p = part + (size_t)set * ((H2 + W2) << lseg);
    XYZZ<F> v = XYZZ<F>::inf();
    if (!half) { // row t: its SEG segment sums are consecutive
        for (uint32_t s = 0; s < SEG; s++) v.add(ld_vec(p + (size_t)t * SEG + s));
    } else if ((uint32_t)t < W2) {
        for (uint32_t s = 0; s < SEG; s++) v.add(ld_vec(p + (size_t)(H2 << lseg) + (size_t)s * W2 + t));
    }
    block_suffix_scan(sm, T, t, v);
    if (!half && t == 0) v = XYZZ<F>::inf(); // rows carry the weight hi, columns lo + 1
    block_sum(sm, T, t, v);
    if (!half && t == 0) {
        for (int i = 0; i < lw; i++) v = v.dbl();
        sm_put(sm, T, 1, v);
    }
    __syncthreads();
    if (half && t == 0) {
        v.add(sm_get<XYZZ<F>>(sm_all, T, 1));
        st_vec(out + set, v);
    }
}

// Horner over the G bucket sets of every proof: res[p * njobs + job] = sum_g 2^(g*c) R_{p,g}
template <class F>
__global__ void k_final(const XYZZ<F> *__restrict__ sets, int batch, int G, int c, int njobs, int job,
                        XYZZ<F> *__restrict__ res)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= batch) return;
    const XYZZ<F> *s = sets + (size_t)p * G;
    XYZZ<F> r = ld_vec(s + (G - 1));
    for (int g = G - 2; g >= 0; g--) {
        for (int i = 0; i < c; i++) r = r.dbl();
        r.add(ld_vec(s + g));
    }
    st_vec(res + (size_t)p * njobs + job, r);
}

template <class F> __global__ void k_to_jacobian(const XYZZ<F> *__restrict__ in, Jacobian<F> *__restrict__ out, size_t n)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    st_vec(out + i, ld_vec(in + i).to_jacobian());
}

// ------------------------------------------------------------------------------------
// base precomputation: copy q = 2^(q*G*c) * P, affine-normalised
// ------------------------------------------------------------------------------------
template <class F>
__global__ void __launch_bounds__(128)
k_next_copy(const Affine<F> *__restrict__ prev, Affine<F> *__restrict__ next, size_t n, int doublings)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Affine<F> p = ld_vec(prev + i);
    XYZZ<F> x = XYZZ<F>::from_affine(p);
    for (int k = 0; k < doublings; k++) x = x.dbl();
    Affine<F> r;
    if (x.is_inf()) {
        r.x = F::zero();
        r.y = F::zero();
    } else {
        // x/ZZ, y/ZZZ with one inversion: inv = 1/(ZZ*ZZZ)
        const F inv = (x.ZZ * x.ZZZ).inverse();
        r.x = x.X * (inv * x.ZZZ);
        r.y = x.Y * (inv * x.ZZ);
    }
    st_vec(next + i, r);
}

// ---- measurement support: n distinct bases k_i * G made on the device (zkb_bases_synthetic) -------------------------
// k_i = 252 bits of a counter-based hash of (seed, i); double-and-add as curve_utils.tcc:13-34, affine-normalised.
__device__ __forceinline__ uint64_t synth_mix(uint64_t x)
{
    x += 0x9e3779b97f4a7c15ull;
    x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
    x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
    return x ^ (x >> 31);
}
template <class F>
__global__ void __launch_bounds__(128)
k_synth_bases(Affine<F> gen, uint64_t seed, size_t n, Affine<F> *__restrict__ out)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t k[8];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const uint64_t v = synth_mix(seed * 0x100000001b3ull + 4 * i + j);
        k[2 * j] = (uint32_t)v;
        k[2 * j + 1] = (uint32_t)(v >> 32);
    }
    k[7] &= 0x0fffffffu;
    k[0] |= 1u; // never zero
    const XYZZ<F> x = scalar_mul<F>(k, XYZZ<F>::from_affine(gen));
    Affine<F> r;
    const F inv = (x.ZZ * x.ZZZ).inverse();
    r.x = x.X * (inv * x.ZZZ);
    r.y = x.Y * (inv * x.ZZ);
    st_vec(out + i, r);
}
template <class F> int plan_synth_job(zkb_ctx *ctx, MsmPlan *plan, int job, const Affine<F> &gen, uint64_t seed)
{
    const MsmJob &J = plan->tab.j[job];
    const size_t n = J.n_total;
    Affine<F> *p = (Affine<F> *)plan->d_pts + J.pts_off;
    k_synth_bases<F><<<ceil_div(n, 128), 128, 0, ctx->stream>>>(gen, seed, n, p);
    ZKB_LAUNCH_CHECK(ctx);
    for (int q = 1; q < J.F; q++) {
        k_next_copy<F><<<ceil_div(n, 128), 128, 0, ctx->stream>>>(p + (size_t)(q - 1) * n, p + (size_t)q * n, n, J.G * J.c);
        ZKB_LAUNCH_CHECK(ctx);
    }
    return ZKB_OK;
}

template <class F>
int msm_finish(zkb_ctx *ctx, const MsmPlan *plan, const JobTab &tab, int batch, uint32_t rmax, XYZZ<F> *d_res)
{
    typedef XYZZ<F> PT;
    cudaStream_t st = ctx->stream;
    const int gs = msm_slot(ctx, plan->group);
    const size_t B = (size_t)tab.Bp * batch;
    uint32_t *start = ctx->scratch[gs + SB_START].as<uint32_t>();
    PT *parts = ctx->scratch[gs + SB_PARTS].as<PT>();
    PT *buckets = ctx->scratch[gs + SB_BUCKETS].as<PT>();
    // queue for the block-wide fix-up: a bucket enters it when it spans >= FIXUP_BLOCK runs (so it owns at least
    // FIXUP_BLOCK - 1 of the rmax runs), as ceil(span / FIXUP_CHUNK) chunks
    const uint32_t max_wide = rmax / (FIXUP_BLOCK / 2) + 16, max_chunks = max_wide + rmax / FIXUP_CHUNK + 16;
    ZKB_CUDA(ctx, ctx->scratch[gs + SB_WIDE].ensure((2 + 2 * (size_t)max_chunks + 3 * (size_t)max_wide) * 4));
    ZKB_CUDA(ctx, ctx->scratch[gs + SB_WIDESUM].ensure((size_t)max_chunks * sizeof(PT)));
    uint32_t *wide = ctx->scratch[gs + SB_WIDE].as<uint32_t>();
    PT *chunk_sums = ctx->scratch[gs + SB_WIDESUM].as<PT>();
    ZKB_CUDA(ctx, cudaMemsetAsync(wide, 0, 8, st));
    k_fixup<F><<<ceil_div(B, FIXUP_T), FIXUP_T, (FIXUP_T / 32) * 32 * sizeof(PT), st>>>(start, (uint32_t)B, rmax, buckets, parts, wide, max_chunks);
    ZKB_LAUNCH_CHECK(ctx);
    {
        const uint32_t attr_bit = plan->group == ZKB_G2 ? 4u : 2u; // function attributes are per device = per context
        if (!(ctx->func_attr_done & attr_bit)) {
            ZKB_CUDA(ctx, cudaFuncSetAttribute(k_fixup_wide<F>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(FIXUP_WIDE_T * sizeof(PT))));
            ctx->func_attr_done |= attr_bit;
        }
        k_fixup_wide<F><<<(unsigned)ctx->sm_count * 16, FIXUP_WIDE_T, FIXUP_WIDE_T * sizeof(PT), st>>>(start, (uint32_t)B, rmax, parts, wide, chunk_sums);
        ZKB_LAUNCH_CHECK(ctx);
        k_fixup_wide_sum<F><<<(unsigned)ctx->sm_count * 8, 128, 128 * sizeof(PT), st>>>(buckets, wide, max_chunks, chunk_sums);
        ZKB_LAUNCH_CHECK(ctx);
    }
    // several jobs, none with enough buckets for the row / column form: one launch per level for all of them
    bool together = tab.njobs > 1;
    for (uint32_t q = 0; q < tab.njobs && together; q++) {
        const uint32_t nb = 1u << (tab.j[q].c - 1);
        const size_t sets = (size_t)batch * tab.j[q].G;
        if (sets > 65535 || (nb >= 1024 && nb <= 65536 && sets * nb >= (size_t)ctx->sm_count * 2048)) together = false;
    }
    if (together) {
        RedGeoTab gt = {};
        size_t l1_total = 0, l2_total = 0, max_sets = 0;
        int max_nb_blocks = 0;
        for (uint32_t q = 0; q < tab.njobs; q++) {
            const MsmJob &J = tab.j[q];
            RedGeo &g = gt.g[q];
            g.bucket_off = J.bucket_off;
            g.G = J.G;
            g.c = J.c;
            g.nb = 1u << (J.c - 1);
            g.sets = (uint32_t)((size_t)batch * J.G);
            g.T = (int)(g.nb < 128 ? g.nb : 128);
            const size_t all_buckets = (size_t)g.nb * J.G * batch;
            int L = 1;
            while (L < 8 && L < (int)(g.nb / g.T) && all_buckets / (2 * L) >= (size_t)ctx->sm_count * 512) L <<= 1;
            while (g.nb / ((uint32_t)g.T * L) > 128) L <<= 1;
            g.L = L;
            g.NB = (int)(g.nb / ((uint32_t)g.T * L));
            if (g.NB > 128) return fail(ctx, ZKB_ERR_ARG, "msm: window too wide for the bucket reduction");
            while ((1 << g.logL) < L) g.logL++;
            while ((1u << g.logCH) < (uint32_t)g.T * L) g.logCH++;
            g.l1_off = (uint32_t)l1_total;
            g.l2_off = (uint32_t)l2_total;
            l1_total += (size_t)g.sets * g.NB;
            l2_total += g.sets;
            max_sets = std::max(max_sets, (size_t)g.sets);
            max_nb_blocks = std::max(max_nb_blocks, g.NB);
        }
        ZKB_CUDA(ctx, ctx->scratch[gs + SB_L1].ensure(2 * l1_total * sizeof(PT)));
        ZKB_CUDA(ctx, ctx->scratch[gs + SB_L2].ensure(l2_total * sizeof(PT)));
        PT *l1 = ctx->scratch[gs + SB_L1].as<PT>(), *l2 = ctx->scratch[gs + SB_L2].as<PT>();
        k_reduce_l1_jobs<F><<<dim3((unsigned)max_nb_blocks, (unsigned)max_sets, tab.njobs), 128, 128 * sizeof(PT), st>>>(buckets, tab.Bp, gt, l1);
        ZKB_LAUNCH_CHECK(ctx);
        k_reduce_l2_jobs<F><<<dim3((unsigned)max_sets, tab.njobs), 128, 128 * sizeof(PT), st>>>(l1, gt, l2);
        ZKB_LAUNCH_CHECK(ctx);
        k_final_jobs<F><<<dim3(ceil_div((size_t)batch, 32), tab.njobs), 32, 0, st>>>(l2, batch, gt, (int)tab.njobs, d_res);
        ZKB_LAUNCH_CHECK(ctx);
        return ZKB_OK;
    }
    for (uint32_t q = 0; q < tab.njobs; q++) {
        const MsmJob &J = tab.j[q];
        const uint32_t nb = 1u << (J.c - 1);
        const size_t sets = (size_t)batch * J.G;
        if (sets > 65535) return fail(ctx, ZKB_ERR_ARG, "msm: batch * bucket sets exceeds 65535");
        ZKB_CUDA(ctx, ctx->scratch[gs + SB_L2].ensure(sets * sizeof(PT)));
        PT *l2 = ctx->scratch[gs + SB_L2].as<PT>();
        if (nb >= 1024 && nb <= 65536 && sets * nb >= (size_t)ctx->sm_count * 2048) {
            // plenty of buckets (a batch): row / column sums over an H2 x W2 grid, segments per line so that about
            // 512 threads per SM exist; few buckets (a single proof) are latency-bound and keep the scan form below
            const int lw = (J.c - 1) / 2, lh = (J.c - 1) - lw;
            int lseg = 0;
            while (lseg < 4 && lseg < lw && (sets << (lh + 1 + lseg)) < (size_t)ctx->sm_count * 512) lseg++;
            const size_t per_set = ((size_t)1 << (lh + lseg)) + ((size_t)1 << (lw + lseg));
            ZKB_CUDA(ctx, ctx->scratch[gs + SB_L1].ensure(sets * per_set * sizeof(PT)));
            PT *part = ctx->scratch[gs + SB_L1].as<PT>();
            k_reduce_rowcol<F><<<dim3(ceil_div(per_set, 128), (unsigned)sets), 128, 0, st>>>(buckets, tab.Bp, J.bucket_off, J.G, nb, lw, lh, lseg, part);
            ZKB_LAUNCH_CHECK(ctx);
            const int T2 = 1 << lh; // lh >= lw
            const uint32_t attr_bit = plan->group == ZKB_G2 ? 32u : 16u;
            if (!(ctx->func_attr_done & attr_bit)) {
                ZKB_CUDA(ctx, cudaFuncSetAttribute(k_reduce_rowcol_finish<F>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(512 * sizeof(PT))));
                ctx->func_attr_done |= attr_bit;
            }
            k_reduce_rowcol_finish<F><<<(unsigned)sets, 2 * T2, (size_t)2 * T2 * sizeof(PT), st>>>(part, lw, lh, lseg, l2);
            ZKB_LAUNCH_CHECK(ctx);
        } else {
        // level-1 reduction geometry: nb = NB * T * L
        // buckets per thread: enough threads to fill the device when the batch is small (the serial chain is
        // 2 L + 2 log2 T additions deep), fewer scan steps per bucket when there is plenty of work
        const int T = (int)(nb < 128 ? nb : 128);
        const size_t all_buckets = (size_t)nb * J.G * batch;
        int L = 1;
        while (L < 8 && L < (int)(nb / T) && all_buckets / (2 * L) >= (size_t)ctx->sm_count * 512) L <<= 1;
        while (nb / ((uint32_t)T * L) > 128) L <<= 1; // level 2 handles at most 128 level-1 blocks per set
        const int NB = (int)(nb / ((uint32_t)T * L));
        if (NB > 128) return fail(ctx, ZKB_ERR_ARG, "msm: window too wide for the bucket reduction");
        int T2 = 1;
        while (T2 < NB) T2 <<= 1;
        ZKB_CUDA(ctx, ctx->scratch[gs + SB_L1].ensure(2 * sets * NB * sizeof(PT)));
        PT *l1 = ctx->scratch[gs + SB_L1].as<PT>();
        int logL = 0;
        while ((1 << logL) < L) logL++;
        k_reduce_l1<F><<<dim3(NB, (unsigned)sets), T, (size_t)T * sizeof(PT), st>>>(buckets, tab.Bp, J.bucket_off, J.G, nb, L,
                                                                                   logL, l1);
        ZKB_LAUNCH_CHECK(ctx);
        int logCH = 0;
        while ((1u << logCH) < (uint32_t)T * L) logCH++;
        k_reduce_l2<F><<<(unsigned)sets, T2, (size_t)T2 * sizeof(PT), st>>>(l1, NB, logCH, l2);
        ZKB_LAUNCH_CHECK(ctx);
        }
        k_final<F><<<ceil_div((size_t)batch, 32), 32, 0, st>>>(l2, batch, J.G, J.c, (int)tab.njobs, (int)q, d_res);
        ZKB_LAUNCH_CHECK(ctx);
    }
    return ZKB_OK;
}

template <class F> int results_to_host(zkb_ctx *ctx, const XYZZ<F> *d_res, size_t count, uint64_t *out_jacobian)
{
    cudaStream_t st = ctx->stream;
    const size_t bytes = count * sizeof(Jacobian<F>);
    ZKB_CUDA(ctx, ctx->scratch[SB_STAGE_O].ensure(bytes));
    if (ensure_pinned(ctx, bytes)) return ZKB_ERR_CUDA;
    Jacobian<F> *d_out = ctx->scratch[SB_STAGE_O].as<Jacobian<F>>();
    k_to_jacobian<F><<<ceil_div(count, 64), 64, 0, st>>>(d_res, d_out, count);
    ZKB_LAUNCH_CHECK(ctx);
    ZKB_CUDA(ctx, cudaEventRecord(ctx->ev1, st));
    ZKB_CUDA(ctx, cudaMemcpyAsync(ctx->pinned, d_out, bytes, cudaMemcpyDeviceToHost, st));
    ZKB_CUDA(ctx, cudaStreamSynchronize(st));
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->last_ms = ms;
    prof_collect(ctx);
    memcpy(out_jacobian, ctx->pinned, bytes);
    return ZKB_OK;
}

// installs the bases of one job: copy 0 from `affine` (host memory, or device memory when `from_device`), the other
// copies derived on the device (asynchronous on the stream)
template <class F> int plan_upload_job(zkb_ctx *ctx, MsmPlan *plan, int job, const uint64_t *affine, bool from_device)
{
    const MsmJob &J = plan->tab.j[job];
    const size_t n = J.n_total;
    Affine<F> *p = (Affine<F> *)plan->d_pts + J.pts_off;
    ZKB_CUDA(ctx, cudaMemcpyAsync(p, affine, n * sizeof(Affine<F>), from_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice,
                                  ctx->stream));
    for (int q = 1; q < J.F; q++) {
        k_next_copy<F><<<ceil_div(n, 128), 128, 0, ctx->stream>>>(p + (size_t)(q - 1) * n, p + (size_t)q * n, n, J.G * J.c);
        ZKB_LAUNCH_CHECK(ctx);
    }
    return ZKB_OK;
}

} // namespace zkb
