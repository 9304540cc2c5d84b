// msm.cu -- Pippenger bucket multi-scalar multiplication over alt_bn128 G1 / G2 for sm_100a.
//
// Replaces libff::multi_exp / multi_exp_with_mixed_addition with multi_exp_method_BDLO12
// (libff/algebra/scalar_multiplication/multiexp.tcc:165-282, :402-496) and libsnark's
// kc_multi_exp_with_mixed_addition (knowledge_commitment/kc_multiexp.tcc:21-89) as called by
// the Groth16 prover (r1cs_gg_ppzksnark.tcc:442-485).  The reference walks every window
// serially, adds full-Jacobian points into 2^c buckets and chains c doublings between
// windows.  This is a different algorithm with the same result (a group element):
//
//   0. an MSM *plan* holds up to four queries ("jobs") of one group that are always evaluated
//      together (for a proving key: A-query, B-query in G1, H- and L-query), and a run
//      evaluates the plan for a whole batch of proofs at once: one pair list, one sort, one
//      accumulation launch for all jobs of all proofs;
//   1. scalars leave Montgomery form once (fp.tcc:228-238) and are cut into W = ceil(255/c)
//      SIGNED c-bit digits, so a window needs 2^(c-1) buckets; zero scalars produce no pair,
//      scalars equal to one produce a single pair (the reference special-cases both,
//      multiexp.tcc:466-481);
//   2. the (item, window) pairs are counting-sorted by bucket: histogram (L2 atomics),
//      exclusive scan, scatter;
//   3. bucket accumulation is load-balanced by position, not by bucket: every thread owns a
//      fixed-length run of the sorted pair list, accumulates it with XYZZ mixed additions and
//      emits complete buckets directly and at most two boundary partials, which a fix-up pass
//      merges (warp-cooperatively for buckets that span many runs).  Skewed digit
//      distributions (the 0/1-heavy Groth16 witnesses) cost nothing extra;
//   4. bases are stored together with their 2^(j*G*c) multiples (computed once at upload):
//      windows j*G+g of all j then share bucket set g and no doublings remain at MSM time
//      when G == 1 -- HBM capacity (180 GB) is traded for the serial doubling chain;
//   5. the weighted bucket sum  sum_j (j+1) B_j  is a two-level block-parallel suffix scan
//      in shared memory, followed by a Horner pass over the G bucket sets.
//
// Layout in HBM: bases AoS affine (64 B G1 / 128 B G2 per point, 16-byte vector loads; a
// gathered point is exactly two / four full 32-byte sectors), copies stored copy-major
// [F][n]; buckets and partials XYZZ AoS (128 B / 256 B); the sorted pair list is one u32 array (point index | sign)
// plus the bucket boundaries start[B + 1].
#include "msm_impl.cuh"
#include <string.h>

using namespace zkb;

namespace {

// ------------------------------------------------------------------------------------
// scalar recoding
// ------------------------------------------------------------------------------------
__global__ void k_to_plain(const Fr *__restrict__ s, Fr *__restrict__ plain, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    st_vec(plain + i, ld_vec_ro(s + i).from_mont());
}

__device__ __forceinline__ uint32_t extract_bits(const uint32_t *__restrict__ k, int bit, int c)
{
    const int w = bit >> 5, off = bit & 31;
    if (w >= 8) return 0;
    uint32_t v = k[w] >> off;
    if (off + c > 32 && w + 1 < 8) v |= k[w + 1] << (32 - off);
    return v & ((1u << c) - 1u);
}

// calls fn(window, signed digit) for every non-zero signed digit of the plain scalar k
template <class Fn> __device__ __forceinline__ void for_each_digit(const uint32_t *__restrict__ k, int c, int W, Fn fn)
{
    uint32_t carry = 0;
    const uint32_t half = 1u << (c - 1);
    for (int w = 0; w < W; w++) {
        uint32_t raw = extract_bits(k, w * c, c) + carry;
        int d;
        if (raw > half) {
            d = (int)raw - (int)(1u << c);
            carry = 1;
        } else {
            d = (int)raw;
            carry = 0;
        }
        if (d != 0) fn(w, d);
    }
}

// item i of proof p: finds its job, loads its plain scalar; returns false when the scalar is zero
__device__ __forceinline__ bool load_item(const JobTab &tab, const uint32_t *__restrict__ sidx,
                                          const uint32_t *__restrict__ plain, size_t zstride, uint32_t i, uint32_t p,
                                          int &job, uint32_t &li, uint32_t *k)
{
    job = 0;
#pragma unroll
    for (int q = 1; q < MSM_MAX_JOBS; q++)
        if (q < (int)tab.njobs && i >= tab.j[q].item_begin) job = q;
    li = i - tab.j[job].item_begin;
    const uint32_t si = sidx ? sidx[i] : tab.j[job].scal_first + li;
    const uint4 *src = reinterpret_cast<const uint4 *>(plain + (size_t)p * zstride + 8 * (size_t)si);
    const uint4 lo = __ldg(src), hi = __ldg(src + 1);
    k[0] = lo.x; k[1] = lo.y; k[2] = lo.z; k[3] = lo.w; k[4] = hi.x; k[5] = hi.y; k[6] = hi.z; k[7] = hi.w;
    return (k[0] | k[1] | k[2] | k[3] | k[4] | k[5] | k[6] | k[7]) != 0;
}

__global__ void k_count(JobTab tab, const uint32_t *__restrict__ sidx, const uint32_t *__restrict__ plain,
                        size_t zstride, uint32_t *__restrict__ hist)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t p = blockIdx.y;
    if (i >= tab.n_items) return;
    int job;
    uint32_t li, k[8];
    if (!load_item(tab, sidx, plain, zstride, i, p, job, li, k)) return;
    const MsmJob J = tab.j[job];
    const uint32_t nb = 1u << (J.c - 1);
    uint32_t *h = hist + (size_t)p * tab.Bp + J.bucket_off;
    for_each_digit(k, J.c, J.W, [&](int w, int d) {
        const uint32_t mag = (uint32_t)(d < 0 ? -d : d);
        atomicAdd(&h[(uint32_t)(w % J.G) * nb + (mag - 1)], 1u);
    });
}

__global__ void k_scatter(JobTab tab, const uint32_t *__restrict__ sidx, const uint32_t *__restrict__ plain,
                          size_t zstride, uint32_t *__restrict__ cursor, uint32_t *__restrict__ ent)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t p = blockIdx.y;
    if (i >= tab.n_items) return;
    int job;
    uint32_t li, k[8];
    if (!load_item(tab, sidx, plain, zstride, i, p, job, li, k)) return;
    const MsmJob J = tab.j[job];
    const uint32_t nb = 1u << (J.c - 1);
    const uint32_t b0 = p * tab.Bp + J.bucket_off;
    const uint32_t pt0 = J.pts_off + J.base_first + li;
    for_each_digit(k, J.c, J.W, [&](int w, int d) {
        const uint32_t mag = (uint32_t)(d < 0 ? -d : d);
        const uint32_t b = b0 + (uint32_t)(w % J.G) * nb + (mag - 1);
        const uint32_t pos = atomicAdd(&cursor[b], 1u);
        const uint32_t q = (uint32_t)(w / J.G);
        ent[pos] = (pt0 + q * J.n_total) | (d < 0 ? 0x80000000u : 0u);
    });
}

// ---- block-local form of the two kernels above (one proof's bucket space fits in shared memory) -------------
// Block (j, p) owns the items [j * chunk, (j + 1) * chunk) of proof p.  It counts its pairs in a shared-memory
// histogram of the proof's Bp buckets and writes the counts to hist[(p * NB + j) * Bp + b] (coalesced);
// k_bucket_totals / the scan / k_block_cursors turn them into one write cursor per (block, bucket), which the
// scatter pass loads back into shared memory: no global atomic is left in either pass.  Scalars equal to one --
// nearly half of a Groth16 witness, all landing in ONE bucket -- are counted once per warp (ballot) instead of lane
// by lane.  NB is chosen so that the proofs in flight at any time write a region of the sorted list that fits in L2
// (4-byte scattered stores that miss L2 cost a DRAM read-modify-write each).
constexpr int SORT_T = 256;
__device__ __forceinline__ bool is_plain_one(const uint32_t *k)
{
    return k[0] == 1u && (k[1] | k[2] | k[3] | k[4] | k[5] | k[6] | k[7]) == 0u;
}
// `SCATTER` = false: count; true: place.  sh: Bp words (counts, or cursors).
template <bool SCATTER>
__global__ void __launch_bounds__(SORT_T)
k_sort_local(JobTab tab, const uint32_t *__restrict__ sidx, const uint32_t *__restrict__ plain, size_t zstride, uint32_t chunk,
             uint32_t *__restrict__ hist, uint32_t *__restrict__ ent)
{
    extern __shared__ uint32_t sh[];
    const uint32_t p = blockIdx.y, j = blockIdx.x, NB = gridDim.x, Bp = tab.Bp;
    uint32_t *g = hist + ((size_t)p * NB + j) * Bp;
    for (uint32_t b = threadIdx.x; b < Bp; b += SORT_T) sh[b] = SCATTER ? g[b] : 0u;
    __syncthreads();
    const uint32_t lo = j * chunk, hi = min(lo + chunk, tab.n_items);
    const uint32_t lane = threadIdx.x & 31;
    for (uint32_t i0 = lo; i0 < hi; i0 += SORT_T) { // chunk is a multiple of SORT_T: whole warps stay together
        const uint32_t i = i0 + threadIdx.x;
        int job = 0;
        uint32_t li = 0, k[8];
        const bool live = i < hi && load_item(tab, sidx, plain, zstride, i, p, job, li, k);
        const MsmJob J = tab.j[job];
        const uint32_t nb = 1u << (J.c - 1);
        const uint32_t pt0 = J.pts_off + J.base_first + li;
        const bool one = live && is_plain_one(k);
        // scalar == 1: a single pair in bucket (window 0, digit 1) of the item's job; aggregated over the lanes that
        // share the job
        unsigned rem = __ballot_sync(0xffffffffu, one);
        while (rem) { // one iteration per job present among the warp's ones (usually a single one)
            const int leader = __ffs(rem) - 1;
            const int lead_job = __shfl_sync(0xffffffffu, job, leader);
            const unsigned grp = __ballot_sync(0xffffffffu, one && job == lead_job);
            uint32_t base = 0;
            if ((int)lane == leader) base = atomicAdd(&sh[J.bucket_off], (uint32_t)__popc(grp));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (SCATTER && one && job == lead_job) {
                const uint32_t pos = base + (uint32_t)__popc(grp & ((1u << lane) - 1u));
                ent[pos] = pt0;
            }
            rem &= ~grp;
        }
        if (live && !one) {
            for_each_digit(k, J.c, J.W, [&](int w, int d) {
                const uint32_t mag = (uint32_t)(d < 0 ? -d : d);
                const uint32_t b = J.bucket_off + (uint32_t)(w % J.G) * nb + (mag - 1);
                const uint32_t pos = atomicAdd(&sh[b], 1u);
                if (SCATTER) {
                    const uint32_t q = (uint32_t)(w / J.G);
                    ent[pos] = (pt0 + q * J.n_total) | (d < 0 ? 0x80000000u : 0u);
                }
            });
        }
    }
    if (!SCATTER) {
        __syncthreads();
        for (uint32_t b = threadIdx.x; b < Bp; b += SORT_T) g[b] = sh[b];
    }
}

// tot[p * Bp + b] = sum over the NB blocks of proof p of their count for bucket b
__global__ void k_bucket_totals(const uint32_t *__restrict__ hist, uint32_t Bp, uint32_t NB, uint32_t *__restrict__ tot)
{
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x, p = blockIdx.y;
    if (b >= Bp) return;
    const uint32_t *h = hist + (size_t)p * NB * Bp + b;
    uint32_t s = 0;
    for (uint32_t j = 0; j < NB; j++) s += h[(size_t)j * Bp];
    tot[(size_t)p * Bp + b] = s;
}
// hist[p][j][b] <- start[p * Bp + b] + the counts of the blocks j' < j: the first position block j writes in bucket b.
// Eight loads are issued before the first store, so the column walk is not one dependent memory round trip per block.
__global__ void k_block_cursors(uint32_t *hist, uint32_t Bp, uint32_t NB, const uint32_t *__restrict__ start)
{
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x, p = blockIdx.y;
    if (b >= Bp) return;
    uint32_t *h = hist + (size_t)p * NB * Bp + b;
    uint32_t run = start[(size_t)p * Bp + b];
    for (uint32_t j0 = 0; j0 < NB; j0 += 8) {
        uint32_t c[8];
#pragma unroll
        for (int u = 0; u < 8; u++) c[u] = j0 + u < NB ? h[(size_t)(j0 + u) * Bp] : 0u;
#pragma unroll
        for (int u = 0; u < 8; u++) {
            if (j0 + u < NB) h[(size_t)(j0 + u) * Bp] = run;
            run += c[u];
        }
    }
}

// ------------------------------------------------------------------------------------
// exclusive scan of u32 (three-phase; 2048 elements per block)
// ------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t block_excl_scan_256(uint32_t v, uint32_t *sm, uint32_t &total)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t x = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t y = __shfl_up_sync(0xffffffffu, x, d);
        if (lane >= d) x += y;
    }
    if (lane == 31) sm[wid] = x;
    __syncthreads();
    if (wid == 0) {
        uint32_t s = lane < (SCAN_T / 32) ? sm[lane] : 0;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, s, d);
            if (lane >= d) s += y;
        }
        if (lane < (SCAN_T / 32)) sm[lane] = s;
    }
    __syncthreads();
    const uint32_t base = wid ? sm[wid - 1] : 0;
    total = sm[SCAN_T / 32 - 1];
    __syncthreads();
    return base + x - v;
}

__global__ void k_scan_sums(const uint32_t *__restrict__ in, size_t n, uint32_t *__restrict__ block_sums)
{
    __shared__ uint32_t sm[32];
    size_t base = (size_t)blockIdx.x * SCAN_BLK + (size_t)threadIdx.x * SCAN_E;
    uint32_t s = 0;
#pragma unroll
    for (int j = 0; j < SCAN_E; j++)
        if (base + j < n) s += in[base + j];
    uint32_t total;
    block_excl_scan_256(s, sm, total);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}

// single block: exclusive scan of the block sums (sequential chunks of 256)
__global__ void k_scan_block_sums(uint32_t *__restrict__ block_sums, unsigned nblocks)
{
    __shared__ uint32_t sm[32];
    uint32_t carry = 0;
    for (unsigned base = 0; base < nblocks; base += SCAN_T) {
        unsigned i = base + threadIdx.x;
        uint32_t v = i < nblocks ? block_sums[i] : 0;
        uint32_t total;
        uint32_t e = block_excl_scan_256(v, sm, total);
        if (i < nblocks) block_sums[i] = e + carry;
        carry += total;
    }
}

// start[i] = exclusive prefix of in[0..n), start[n] = total; `cursor` (optional, may alias in) receives a second copy
__global__ void k_scan_final(const uint32_t *in, size_t n, const uint32_t *__restrict__ block_sums,
                             uint32_t *cursor, uint32_t *__restrict__ start)
{
    __shared__ uint32_t sm[32];
    size_t base = (size_t)blockIdx.x * SCAN_BLK + (size_t)threadIdx.x * SCAN_E;
    uint32_t v[SCAN_E];
    uint32_t s = 0;
#pragma unroll
    for (int j = 0; j < SCAN_E; j++) {
        v[j] = (base + j < n) ? in[base + j] : 0;
        s += v[j];
    }
    uint32_t total;
    uint32_t e = block_excl_scan_256(s, sm, total) + block_sums[blockIdx.x];
#pragma unroll
    for (int j = 0; j < SCAN_E; j++) {
        if (base + j < n) {
            if (cursor) cursor[base + j] = e;
            start[base + j] = e;
        }
        e += v[j];
        if (base + j + 1 == n) start[n] = e;
    }
}

} // namespace

namespace zkb {

int msm_choose_c(size_t n_eff, int group, int precompute_factor)
{
    // cost in base-field multiplications: XYZZ mixed add 10 (G1) / 30 (G2), full add 14 / 42
    const double madd = group == ZKB_G2 ? 30 : 10, add = group == ZKB_G2 ? 42 : 14, dbl = group == ZKB_G2 ? 27 : 9;
    int best = 4;
    double best_cost = 1e300;
    for (int c = 4; c <= MSM_MAX_C; c++) {
        const int W = (255 + c - 1) / c;
        int F = precompute_factor <= 0 ? W : precompute_factor;
        if (F > W) F = W;
        const int G = (W + F - 1) / F;
        // the bucket reduction runs at about half the arithmetic efficiency of the accumulation (scan overhead,
        // short dependent chains): its additions are charged twice
        const double cost = (double)n_eff * W * madd + 2.0 * G * (double)(1u << c) * add + (double)(G - 1) * c * dbl * 64;
        if (cost < best_cost) {
            best_cost = cost;
            best = c;
        }
    }
    return best;
}

int plan_add_job(zkb_ctx *ctx, MsmPlan *plan, size_t n_total, int c, int precompute_factor)
{
    JobTab &t = plan->tab;
    if (t.njobs >= MSM_MAX_JOBS) return fail(ctx, ZKB_ERR_ARG, "msm plan: too many jobs");
    if (c < 2 || c > MSM_MAX_C) return fail(ctx, ZKB_ERR_ARG, "msm plan: window bits must be 2..20");
    MsmJob &J = t.j[t.njobs];
    const int W = (255 + c - 1) / c;
    int F = precompute_factor <= 0 ? W : precompute_factor;
    if (F > W) F = W;
    const int G = (W + F - 1) / F;
    F = (W + G - 1) / G;
    const double pts = (double)plan->pts_count + (double)n_total * F;
    if (pts >= 2147483648.0 || (double)t.n_items + (double)n_total >= 4294967295.0)
        return fail(ctx, ZKB_ERR_ARG, "msm plan: n * precomputed copies exceeds 2^31 points");
    J.item_begin = t.n_items;
    J.n = (uint32_t)n_total;
    J.base_first = 0;
    J.n_total = (uint32_t)n_total;
    J.pts_off = (uint32_t)plan->pts_count;
    J.scal_first = t.n_items;
    J.bucket_off = t.Bp;
    J.c = (uint16_t)c;
    J.W = (uint16_t)W;
    J.G = (uint16_t)G;
    J.F = (uint16_t)F;
    t.n_items += (uint32_t)n_total;
    t.Bp += (uint32_t)G << (c - 1);
    t.njobs++;
    plan->pts_count += n_total * (size_t)F;
    return ZKB_OK;
}

int plan_alloc(zkb_ctx *ctx, MsmPlan *plan)
{
    const size_t psz = plan->group == ZKB_G2 ? sizeof(Affine<Fq2>) : sizeof(Affine<Fq>);
    plan->dev = ctx->dev;
    plan->bytes = plan->pts_count * psz;
    ZKB_CUDA(ctx, cudaMalloc(&plan->d_pts, plan->bytes ? plan->bytes : 16));
    return ZKB_OK;
}

void plan_free(MsmPlan *plan)
{
    if (plan->d_pts) cudaFree(plan->d_pts);
    if (plan->d_sidx) cudaFree(plan->d_sidx);
    plan->d_pts = nullptr;
    plan->d_sidx = nullptr;
}

// shared-memory budget of the block-local sort: a proof's bucket space must fit (4 bytes per bucket)
constexpr uint32_t SORT_LOCAL_MAX_BP = 49152;

int msm_sort_pairs(zkb_ctx *ctx, const MsmPlan *plan, const JobTab &tab, const uint32_t *d_plain, size_t zstride,
                   int batch)
{
    cudaStream_t st = ctx->stream;
    const int gs = msm_slot(ctx, plan->group);
    size_t Emax = 0;
    for (uint32_t q = 0; q < tab.njobs; q++) Emax += (size_t)tab.j[q].n * tab.j[q].W;
    const size_t e_proof = Emax;
    Emax *= (size_t)batch;
    const size_t B = (size_t)tab.Bp * batch;
    if (Emax >= (1ull << 32) || B >= (1ull << 32)) return fail(ctx, ZKB_ERR_ARG, "msm: more than 2^32 pairs in one run");
    // block-local sort: NB blocks per proof.  About three blocks per SM are resident; the proofs in flight
    // (resident / NB) should write no more than ~48 MB of the sorted list at a time; every block wants >= 1024 items.
    const bool local = tab.Bp <= SORT_LOCAL_MAX_BP;
    uint32_t NB = 1, chunk = 0;
    if (local) {
        const uint32_t resident = (uint32_t)ctx->sm_count * 3;
        uint32_t want = (resident + (uint32_t)batch - 1) / (uint32_t)batch;              // fill the device
        const size_t inflight_max = ((size_t)48 << 20) / (e_proof * 4 + 1) + 1;          // proofs whose output fits in L2
        if ((size_t)resident / want > inflight_max) want = (uint32_t)((resident + inflight_max - 1) / inflight_max);
        const uint32_t most = (tab.n_items + 1023) / 1024;
        NB = want < most ? want : most;
        if (NB < 1) NB = 1;
        if (NB > 1024) NB = 1024;
        chunk = ((tab.n_items + NB - 1) / NB + SORT_T - 1) / SORT_T * SORT_T;
        NB = (tab.n_items + chunk - 1) / chunk;
        if (B * NB >= (1ull << 32)) return fail(ctx, ZKB_ERR_ARG, "msm: bucket space too large");
    }
    const unsigned scan_blocks = ceil_div(B, SCAN_BLK);
    ZKB_CUDA(ctx, ctx->scratch[gs + SB_HIST].ensure(B * 4));
    ZKB_CUDA(ctx, ctx->scratch[gs + SB_START].ensure((B + 1) * 4));
    ZKB_CUDA(ctx, ctx->scratch[gs + SB_CURSOR].ensure(B * NB * 4));
    ZKB_CUDA(ctx, ctx->scratch[gs + SB_BLOCKSUM].ensure((size_t)scan_blocks * 4 + 16));
    ZKB_CUDA(ctx, ctx->scratch[gs + SB_ENT].ensure(Emax * 4 + 16));
    uint32_t *tot = ctx->scratch[gs + SB_HIST].as<uint32_t>();      // pairs per bucket
    uint32_t *start = ctx->scratch[gs + SB_START].as<uint32_t>();
    uint32_t *cursor = ctx->scratch[gs + SB_CURSOR].as<uint32_t>(); // local: counts, then cursors, per (proof, block, bucket)
    uint32_t *bsum = ctx->scratch[gs + SB_BLOCKSUM].as<uint32_t>();
    uint32_t *ent = ctx->scratch[gs + SB_ENT].as<uint32_t>();

    const size_t smem = (size_t)tab.Bp * 4;
    const dim3 bgrid(ceil_div(tab.Bp, 256), (unsigned)batch);
    if (local) {
        const uint32_t attr_bit = 8u;
        if (!(ctx->func_attr_done & attr_bit)) {
            ZKB_CUDA(ctx, cudaFuncSetAttribute(k_sort_local<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(SORT_LOCAL_MAX_BP * 4)));
            ZKB_CUDA(ctx, cudaFuncSetAttribute(k_sort_local<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(SORT_LOCAL_MAX_BP * 4)));
            ctx->func_attr_done |= attr_bit;
        }
        k_sort_local<false><<<dim3(NB, (unsigned)batch), SORT_T, smem, st>>>(tab, plan->d_sidx, d_plain, zstride, chunk, cursor, nullptr);
        ZKB_LAUNCH_CHECK(ctx);
        k_bucket_totals<<<bgrid, 256, 0, st>>>(cursor, tab.Bp, NB, tot);
        ZKB_LAUNCH_CHECK(ctx);
    } else {
        ZKB_CUDA(ctx, cudaMemsetAsync(tot, 0, B * 4, st));
        const dim3 grid(ceil_div(tab.n_items, 128), (unsigned)batch);
        k_count<<<grid, 128, 0, st>>>(tab, plan->d_sidx, d_plain, zstride, tot);
        ZKB_LAUNCH_CHECK(ctx);
    }
    k_scan_sums<<<scan_blocks, SCAN_T, 0, st>>>(tot, B, bsum);
    ZKB_LAUNCH_CHECK(ctx);
    k_scan_block_sums<<<1, SCAN_T, 0, st>>>(bsum, scan_blocks);
    ZKB_LAUNCH_CHECK(ctx);
    k_scan_final<<<scan_blocks, SCAN_T, 0, st>>>(tot, B, bsum, local ? nullptr : cursor, start);
    ZKB_LAUNCH_CHECK(ctx);
    if (local) {
        k_block_cursors<<<bgrid, 256, 0, st>>>(cursor, tab.Bp, NB, start);
        ZKB_LAUNCH_CHECK(ctx);
        k_sort_local<true><<<dim3(NB, (unsigned)batch), SORT_T, smem, st>>>(tab, plan->d_sidx, d_plain, zstride, chunk, cursor, ent);
    } else {
        const dim3 grid(ceil_div(tab.n_items, 128), (unsigned)batch);
        k_scatter<<<grid, 128, 0, st>>>(tab, plan->d_sidx, d_plain, zstride, cursor, ent);
    }
    ZKB_LAUNCH_CHECK(ctx);
    return ZKB_OK;
}

template <class F>
int msm_run(zkb_ctx *ctx, const MsmPlan *plan, const JobTab &tab, const uint32_t *d_plain, size_t zstride, int batch,
            XYZZ<F> *d_res)
{
    typedef XYZZ<F> PT;
    const int gs = msm_slot(ctx, plan->group);
    if (batch <= 0 || batch > 65535) return fail(ctx, ZKB_ERR_ARG, "msm: batch must be 1..65535");
    const size_t B = (size_t)tab.Bp * batch;
    const uint32_t rmax = msm_max_runs(ctx);
    int rc = msm_sort_pairs(ctx, plan, tab, d_plain, zstride, batch);
    if (rc) return rc;
    ZKB_CUDA(ctx, ctx->scratch[gs + SB_PARTS].ensure(2 * ((size_t)rmax + 1) * sizeof(PT)));
    ZKB_CUDA(ctx, ctx->scratch[gs + SB_BUCKETS].ensure(B * sizeof(PT)));
    {
        const int g = plan->group == ZKB_G2 ? 1 : 0;
        const bool prof = ctx->prof && !ctx->prof_pending[g] && (g == 1 || ctx->lane_bias == 0); // one sample per synchronisation
        if (prof) ZKB_CUDA(ctx, cudaEventRecord(ctx->pe[g][0], ctx->stream));
        rc = launch_accumulate<F>(ctx, (const Affine<F> *)plan->d_pts, ctx->scratch[gs + SB_ENT].as<uint32_t>(),
                                  ctx->scratch[gs + SB_START].as<uint32_t>(),
                                  (uint32_t)B, rmax, ctx->scratch[gs + SB_BUCKETS].as<PT>(), ctx->scratch[gs + SB_PARTS].as<PT>());
        if (rc) return rc;
        if (prof) {
            ZKB_CUDA(ctx, cudaEventRecord(ctx->pe[g][1], ctx->stream));
            ZKB_CUDA(ctx, cudaMemcpyAsync(ctx->prof_pairs_host + g, ctx->scratch[gs + SB_START].as<uint32_t>() + B, 4,
                                          cudaMemcpyDeviceToHost, ctx->stream));
            ctx->prof_pending[g] = true;
        }
    }
    return msm_finish<F>(ctx, plan, tab, batch, rmax, d_res);
}

int msm_to_plain(zkb_ctx *ctx, const void *d_mont, uint32_t *d_plain, size_t n)
{
    if (!n) return ZKB_OK;
    k_to_plain<<<ceil_div(n, 128), 128, 0, ctx->stream>>>((const Fr *)d_mont, (Fr *)d_plain, n);
    ZKB_LAUNCH_CHECK(ctx);
    return ZKB_OK;
}

template int msm_run<Fq>(zkb_ctx *, const MsmPlan *, const JobTab &, const uint32_t *, size_t, int, XYZZ<Fq> *);
template int msm_run<Fq2>(zkb_ctx *, const MsmPlan *, const JobTab &, const uint32_t *, size_t, int, XYZZ<Fq2> *);

} // namespace zkb

extern "C" {

int zkb_bases_upload(zkb_ctx *ctx, int group, const uint64_t *affine, size_t n, int window_bits, int precompute_factor,
                     zkb_bases **out)
{
    if (!ctx || !affine || !out || n == 0 || (group != ZKB_G1 && group != ZKB_G2))
        return fail(ctx, ZKB_ERR_ARG, "zkb_bases_upload: bad argument");
    if (window_bits < 0 || window_bits > MSM_MAX_C || window_bits == 1 || precompute_factor < 0)
        return fail(ctx, ZKB_ERR_ARG, "zkb_bases_upload: window_bits must be 0 or 2..20");
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    zkb_bases *b = new zkb_bases();
    b->plan.group = group;
    b->n = n;
    const int c = window_bits ? window_bits : msm_choose_c(n, group, precompute_factor);
    int rc = plan_add_job(ctx, &b->plan, n, c, precompute_factor);
    if (!rc) rc = plan_alloc(ctx, &b->plan);
    if (!rc) rc = group == ZKB_G1 ? plan_upload_job<Fq>(ctx, &b->plan, 0, affine) : plan_upload_job<Fq2>(ctx, &b->plan, 0, affine);
    if (!rc && cudaStreamSynchronize(ctx->stream) != cudaSuccess) rc = fail(ctx, ZKB_ERR_CUDA, "zkb_bases_upload: precomputation failed");
    if (rc) {
        plan_free(&b->plan);
        delete b;
        return rc;
    }
    *out = b;
    return ZKB_OK;
}

int zkb_bases_synthetic(zkb_ctx *ctx, size_t n, uint64_t seed, int window_bits, int precompute_factor, zkb_bases **out)
{
    if (!ctx || !out || n == 0) return fail(ctx, ZKB_ERR_ARG, "zkb_bases_synthetic: bad argument");
    if (window_bits < 0 || window_bits > MSM_MAX_C || window_bits == 1 || precompute_factor < 0)
        return fail(ctx, ZKB_ERR_ARG, "zkb_bases_synthetic: window_bits must be 0 or 2..20");
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    zkb_bases *b = new zkb_bases();
    b->plan.group = ZKB_G1;
    b->n = n;
    const int c = window_bits ? window_bits : msm_choose_c(n, ZKB_G1, precompute_factor);
    Affine<Fq> gen; // G1 generator (1, 2), alt_bn128_init.cpp:148-150
    gen.x = Fq::one();
    gen.y = Fq::one() + Fq::one();
    int rc = plan_add_job(ctx, &b->plan, n, c, precompute_factor);
    if (!rc) rc = plan_alloc(ctx, &b->plan);
    if (!rc) rc = plan_synth_job<Fq>(ctx, &b->plan, 0, gen, seed);
    if (!rc && cudaStreamSynchronize(ctx->stream) != cudaSuccess) rc = fail(ctx, ZKB_ERR_CUDA, "zkb_bases_synthetic: generation failed");
    if (rc) {
        plan_free(&b->plan);
        delete b;
        return rc;
    }
    *out = b;
    return ZKB_OK;
}

void zkb_bases_free(zkb_bases *b)
{
    if (!b) return;
    plan_free(&b->plan);
    delete b;
}

size_t zkb_bases_device_bytes(const zkb_bases *b) { return b ? b->plan.bytes : 0; }

int zkb_msm_device(zkb_ctx *ctx, const zkb_bases *bases, size_t offset, const void *d_scalars, size_t n, uint64_t *out)
{
    if (!ctx || !bases || !out || (!d_scalars && n)) return fail(ctx, ZKB_ERR_ARG, "zkb_msm: bad argument");
    if (offset + n > bases->n) return fail(ctx, ZKB_ERR_ARG, "zkb_msm: range exceeds the uploaded bases");
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    const bool g1 = bases->plan.group == ZKB_G1;
    if (n == 0) {
        // empty sum = the reference's zero (0, 1, 0)  (alt_bn128_init.cpp:145-147)
        memset(out, 0, (g1 ? 12 : 24) * 8);
        const uint64_t one_q[4] = {0xd35d438dc58f0d9dull, 0x0a78eb28f5c70b3dull, 0x666ea36f7879462cull,
                                   0x0e0a77c19a07df2full};
        memcpy(out + (g1 ? 4 : 8), one_q, 32);
        return ZKB_OK;
    }
    ZKB_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    ZKB_CUDA(ctx, ctx->scratch[SB_PLAIN].ensure(n * 32));
    uint32_t *plain = ctx->scratch[SB_PLAIN].as<uint32_t>();
    int rc = msm_to_plain(ctx, d_scalars, plain, n);
    if (rc) return rc;
    JobTab tab = bases->plan.tab;
    tab.j[0].n = (uint32_t)n;
    tab.j[0].base_first = (uint32_t)offset;
    tab.j[0].scal_first = 0;
    tab.n_items = (uint32_t)n;
    const int gs = msm_slot(ctx, bases->plan.group);
    ZKB_CUDA(ctx, ctx->scratch[gs + SB_RES].ensure(g1 ? sizeof(G1X) : sizeof(G2X)));
    if (g1) {
        rc = msm_run<Fq>(ctx, &bases->plan, tab, plain, 0, 1, ctx->scratch[gs + SB_RES].as<G1X>());
        if (!rc) rc = results_to_host<Fq>(ctx, ctx->scratch[gs + SB_RES].as<G1X>(), 1, out);
    } else {
        rc = msm_run<Fq2>(ctx, &bases->plan, tab, plain, 0, 1, ctx->scratch[gs + SB_RES].as<G2X>());
        if (!rc) rc = results_to_host<Fq2>(ctx, ctx->scratch[gs + SB_RES].as<G2X>(), 1, out);
    }
    return rc;
}

int zkb_msm(zkb_ctx *ctx, const zkb_bases *bases, size_t offset, const uint64_t *scalars, size_t n, uint64_t *out)
{
    if (!ctx || !bases || !out || (!scalars && n)) return fail(ctx, ZKB_ERR_ARG, "zkb_msm: bad argument");
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    if (n) {
        ZKB_CUDA(ctx, ctx->scratch[SB_SCAL].ensure(n * 32));
        ZKB_CUDA(ctx, cudaMemcpyAsync(ctx->scratch[SB_SCAL].p, scalars, n * 32, cudaMemcpyHostToDevice, ctx->stream));
    }
    return zkb_msm_device(ctx, bases, offset, ctx->scratch[SB_SCAL].p, n, out);
}

} // extern "C"
