// msm.cu -- Pippenger bucket multi-scalar multiplication over alt_bn128 G1 / G2 for sm_100a.
//
// Replaces libff::multi_exp / multi_exp_with_mixed_addition with multi_exp_method_BDLO12
// (libff/algebra/scalar_multiplication/multiexp.tcc:165-282, :402-496) and libsnark's
// kc_multi_exp_with_mixed_addition (knowledge_commitment/kc_multiexp.tcc:21-89) as called by
// the Groth16 prover (r1cs_gg_ppzksnark.tcc:442-485).  The reference walks every window
// serially, adds full-Jacobian points into 2^c buckets and chains c doublings between
// windows.  This is a different algorithm with the same result (a group element):
//
//   1. scalars leave Montgomery form once (fp.tcc:228-238) and are cut into W = ceil(255/c)
//      SIGNED c-bit digits, so a window needs 2^(c-1) buckets;
//   2. the (point, window) pairs are counting-sorted by bucket: histogram (L2 atomics),
//      exclusive scan, scatter;
//   3. bucket accumulation is load-balanced by position, not by bucket: every thread owns a
//      fixed-length run of the sorted pair list, accumulates it with XYZZ mixed additions and
//      emits complete buckets directly and at most two boundary partials, which a fix-up pass
//      merges.  Skewed digit distributions (the 0/1-heavy Groth16 witnesses) cost nothing extra;
//   4. bases may be stored together with their 2^(j*G*c) multiples (computed once at upload):
//      windows j*G+g of all j then share bucket set g and no doublings remain at MSM time
//      when G == 1 -- HBM capacity (180 GB) is traded for the serial doubling chain;
//   5. the weighted bucket sum  sum_j (j+1) B_j  is a two-level block-parallel suffix scan
//      in shared memory, followed by a Horner pass over the G bucket sets.
//
// Layout in HBM: bases AoS affine (64 B G1 / 128 B G2 per point, 16-byte vector loads; a
// gathered point is exactly two / four full 32-byte sectors), copies stored copy-major
// [F][n]; buckets and partials XYZZ AoS (128 B / 256 B); the pair list is two u32 arrays.
#include "msm_impl.cuh"

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

__global__ void k_count(const uint32_t *__restrict__ plain, size_t n, int c, int W, int G, uint32_t *__restrict__ hist)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t nb = 1u << (c - 1);
    uint32_t k[8];
    const uint4 *p = reinterpret_cast<const uint4 *>(plain + 8 * i);
    uint4 lo = __ldg(p), hi = __ldg(p + 1);
    k[0] = lo.x; k[1] = lo.y; k[2] = lo.z; k[3] = lo.w; k[4] = hi.x; k[5] = hi.y; k[6] = hi.z; k[7] = hi.w;
    if ((k[0] | k[1] | k[2] | k[3] | k[4] | k[5] | k[6] | k[7]) == 0) return;
    for_each_digit(k, c, W, [&](int w, int d) {
        const uint32_t mag = (uint32_t)(d < 0 ? -d : d);
        atomicAdd(&hist[(uint32_t)(w % G) * nb + (mag - 1)], 1u);
    });
}

__global__ void k_scatter(const uint32_t *__restrict__ plain, size_t n, size_t n_total, size_t offset, int c, int W, int G,
                          uint32_t *__restrict__ cursor, uint32_t *__restrict__ ent, uint32_t *__restrict__ ebkt)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t nb = 1u << (c - 1);
    uint32_t k[8];
    const uint4 *p = reinterpret_cast<const uint4 *>(plain + 8 * i);
    uint4 lo = __ldg(p), hi = __ldg(p + 1);
    k[0] = lo.x; k[1] = lo.y; k[2] = lo.z; k[3] = lo.w; k[4] = hi.x; k[5] = hi.y; k[6] = hi.z; k[7] = hi.w;
    if ((k[0] | k[1] | k[2] | k[3] | k[4] | k[5] | k[6] | k[7]) == 0) return;
    for_each_digit(k, c, W, [&](int w, int d) {
        const uint32_t mag = (uint32_t)(d < 0 ? -d : d);
        const uint32_t b = (uint32_t)(w % G) * nb + (mag - 1);
        const uint32_t pos = atomicAdd(&cursor[b], 1u);
        const uint32_t q = (uint32_t)(w / G);
        ent[pos] = (uint32_t)(q * n_total + offset + i) | (d < 0 ? 0x80000000u : 0u);
        ebkt[pos] = b;
    });
}

// ------------------------------------------------------------------------------------
// exclusive scan of u32 (three-phase; 2048 elements per block)
// ------------------------------------------------------------------------------------

__device__ __forceinline__ uint32_t block_excl_scan_256(uint32_t v, uint32_t *sm, uint32_t &total)
{
    // warp scan + cross-warp
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

// single block: exclusive scan of up to 256*? block sums (sequential chunks of 256)
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

// out[i] = exclusive prefix; out[n] = total.  `out2` (optional) receives a second copy of out[0..n).
__global__ void k_scan_final(const uint32_t *__restrict__ in, size_t n, const uint32_t *__restrict__ block_sums,
                             uint32_t *__restrict__ out, uint32_t *__restrict__ out2)
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
            out[base + j] = e;
            if (out2) out2[base + j] = e;
        }
        e += v[j];
        if (base + j + 1 == n) out[n] = e;
    }
}


} // namespace

namespace zkb {

int msm_sort_pairs(zkb_ctx *ctx, const zkb_bases *b, size_t offset, const Fr *d_scalars, size_t n)
{
    cudaStream_t st = ctx->stream;
    const int c = b->c, W = b->W, G = b->G;
    const uint32_t nb = 1u << (c - 1);
    const uint32_t B = nb * (uint32_t)G;
    const size_t Emax = n * (size_t)W;
    if (Emax >= (1ull << 32)) return fail(ctx, ZKB_ERR_ARG, "msm: n * windows exceeds 2^32 pairs");
    const unsigned scan_blocks = ceil_div((size_t)B, SCAN_BLK);
    ZKB_CUDA(ctx, ctx->scratch[SB_PLAIN].ensure(n * 32));
    ZKB_CUDA(ctx, ctx->scratch[SB_HIST].ensure((size_t)B * 4));
    ZKB_CUDA(ctx, ctx->scratch[SB_START].ensure(((size_t)B + 1) * 4));
    ZKB_CUDA(ctx, ctx->scratch[SB_CURSOR].ensure((size_t)B * 4));
    ZKB_CUDA(ctx, ctx->scratch[SB_BLOCKSUM].ensure((size_t)scan_blocks * 4 + 16));
    ZKB_CUDA(ctx, ctx->scratch[SB_ENT].ensure(Emax * 4 + 16));
    ZKB_CUDA(ctx, ctx->scratch[SB_EBKT].ensure(Emax * 4 + 16));
    uint32_t *plain = ctx->scratch[SB_PLAIN].as<uint32_t>();
    uint32_t *hist = ctx->scratch[SB_HIST].as<uint32_t>();
    uint32_t *start = ctx->scratch[SB_START].as<uint32_t>();
    uint32_t *cursor = ctx->scratch[SB_CURSOR].as<uint32_t>();
    uint32_t *bsum = ctx->scratch[SB_BLOCKSUM].as<uint32_t>();
    uint32_t *ent = ctx->scratch[SB_ENT].as<uint32_t>();
    uint32_t *ebkt = ctx->scratch[SB_EBKT].as<uint32_t>();

    ZKB_CUDA(ctx, cudaMemsetAsync(hist, 0, (size_t)B * 4, st));
    k_to_plain<<<ceil_div(n, 128), 128, 0, st>>>(d_scalars, (Fr *)plain, n);
    ZKB_LAUNCH_CHECK(ctx);
    k_count<<<ceil_div(n, 128), 128, 0, st>>>(plain, n, c, W, G, hist);
    ZKB_LAUNCH_CHECK(ctx);
    k_scan_sums<<<scan_blocks, SCAN_T, 0, st>>>(hist, B, bsum);
    ZKB_LAUNCH_CHECK(ctx);
    k_scan_block_sums<<<1, SCAN_T, 0, st>>>(bsum, scan_blocks);
    ZKB_LAUNCH_CHECK(ctx);
    k_scan_final<<<scan_blocks, SCAN_T, 0, st>>>(hist, B, bsum, start, cursor);
    ZKB_LAUNCH_CHECK(ctx);
    k_scatter<<<ceil_div(n, 128), 128, 0, st>>>(plain, n, b->n, offset, c, W, G, cursor, ent, ebkt);
    ZKB_LAUNCH_CHECK(ctx);
    return ZKB_OK;
}

template <class F>
static int msm_run(zkb_ctx *ctx, const zkb_bases *b, size_t offset, const Fr *d_scalars, size_t n, uint64_t *out_host)
{
    typedef XYZZ<F> PT;
    const uint32_t nb = 1u << (b->c - 1);
    const uint32_t B = nb * (uint32_t)b->G;
    const size_t Emax = n * (size_t)b->W;
    const uint32_t S = msm_run_length(ctx, Emax);
    const size_t runs = (Emax + S - 1) / S;
    ZKB_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    int rc = msm_sort_pairs(ctx, b, offset, d_scalars, n);
    if (rc) return rc;
    ZKB_CUDA(ctx, ctx->scratch[SB_PARTS].ensure(2 * runs * sizeof(PT)));
    ZKB_CUDA(ctx, ctx->scratch[SB_BUCKETS].ensure((size_t)B * sizeof(PT)));
    rc = launch_accumulate<F>(ctx, (const Affine<F> *)b->d_pts, ctx->scratch[SB_ENT].as<uint32_t>(),
                              ctx->scratch[SB_EBKT].as<uint32_t>(), ctx->scratch[SB_START].as<uint32_t>(), B, S,
                              ctx->scratch[SB_BUCKETS].as<PT>(), ctx->scratch[SB_PARTS].as<PT>(), runs);
    if (rc) return rc;
    return msm_finish<F>(ctx, b, n, S, out_host);
}

} // namespace zkb

namespace {
int ilog2_floor(size_t x)
{
    int l = 0;
    while (x >>= 1) l++;
    return l;
}


} // namespace

extern "C" {

int zkb_bases_upload(zkb_ctx *ctx, int group, const uint64_t *affine, size_t n, int window_bits, int precompute_factor,
                     zkb_bases **out)
{
    if (!ctx || !affine || !out || n == 0 || (group != ZKB_G1 && group != ZKB_G2))
        return fail(ctx, ZKB_ERR_ARG, "zkb_bases_upload: bad argument");
    if (window_bits < 0 || window_bits > 16 || precompute_factor < 0)
        return fail(ctx, ZKB_ERR_ARG, "zkb_bases_upload: window_bits must be 0..16");
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    zkb_bases *b = new zkb_bases();
    b->group = group;
    b->dev = ctx->dev;
    b->n = n;
    int c = window_bits;
    if (c == 0) {
        const int lg = ilog2_floor(n);
        c = lg - (precompute_factor == 1 ? 4 : 1);
        if (group == ZKB_G2) c -= 1;
        if (c > 16) c = 16;
        if (c < 4) c = 4;
    }
    if (c < 2) c = 2;
    b->c = c;
    b->W = (255 + c - 1) / c;
    int F = precompute_factor == 0 ? b->W : precompute_factor;
    if (F > b->W) F = b->W;
    b->G = (b->W + F - 1) / F;
    b->F = (b->W + b->G - 1) / b->G;
    if ((double)n * b->F >= 2147483648.0) {
        delete b;
        return fail(ctx, ZKB_ERR_ARG, "zkb_bases_upload: n * precompute_factor exceeds 2^31");
    }
    int rc = group == ZKB_G1 ? bases_build<Fq>(ctx, b, affine) : bases_build<Fq2>(ctx, b, affine);
    if (rc) {
        if (b->d_pts) cudaFree(b->d_pts);
        delete b;
        return rc;
    }
    *out = b;
    return ZKB_OK;
}

void zkb_bases_free(zkb_bases *b)
{
    if (!b) return;
    if (b->d_pts) cudaFree(b->d_pts);
    delete b;
}

size_t zkb_bases_device_bytes(const zkb_bases *b) { return b ? b->bytes : 0; }

int zkb_msm_device(zkb_ctx *ctx, const zkb_bases *bases, size_t offset, const void *d_scalars, size_t n, uint64_t *out)
{
    if (!ctx || !bases || !out || (!d_scalars && n)) return fail(ctx, ZKB_ERR_ARG, "zkb_msm: bad argument");
    if (offset + n > bases->n) return fail(ctx, ZKB_ERR_ARG, "zkb_msm: range exceeds the uploaded bases");
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    if (n == 0) {
        // empty sum = the reference's zero (0, 1, 0)  (alt_bn128_init.cpp:145-147)
        const size_t limbs = bases->group == ZKB_G1 ? 12 : 24;
        memset(out, 0, limbs * 8);
        const uint64_t one_q[4] = {0xd35d438dc58f0d9dull, 0x0a78eb28f5c70b3dull, 0x666ea36f7879462cull,
                                   0x0e0a77c19a07df2full};
        memcpy(out + (bases->group == ZKB_G1 ? 4 : 8), one_q, 32);
        return ZKB_OK;
    }
    if (bases->group == ZKB_G1) return msm_run<Fq>(ctx, bases, offset, (const Fr *)d_scalars, n, out);
    return msm_run<Fq2>(ctx, bases, offset, (const Fr *)d_scalars, n, out);
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
