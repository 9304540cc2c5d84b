// ntt.cu -- number-theoretic transforms over alt_bn128 Fr and libfqfft's evaluation domains for sm_100a.
//
// Replaces, on the prover path, libfqfft's basic_radix2_domain and step_radix2_domain
// (libfqfft/evaluation_domain/domains/basic_radix2_domain.tcc:41-112, basic_radix2_domain_aux.tcc:45-79,
// :172-180, step_radix2_domain.tcc:21-248) and the domain factory
// (get_evaluation_domain.tcc:33-52).  The reference runs a serial in-place radix-2 DIT with an
// explicit bit-reversal.  Field arithmetic is exact, so any schedule that evaluates the same
// DFT gives bit-identical vectors; the schedule here is chosen for the GPU:
//
//   * a transform of size 2^L is cut into passes of at most 8 butterfly levels; a thread block
//     stages C columns x 2^k elements in shared memory (word-planar, XOR-swizzled), runs the
//     k levels there and writes back in place, so one pass costs one HBM read and one write
//     of the vector, in 256-byte rows;
//   * the witness map needs iFFT followed by cosetFFT: the inverse runs decimation-in-
//     frequency (natural in, bit-reversed out), the coset scaling is applied in bit-reversed
//     order (folded into the last pass), and the forward transform runs decimation-in-time
//     (bit-reversed in, natural out), so no permutation pass exists on that path;
//   * all transforms are batched over vectors (the A, B, C polynomials of a whole batch of
//     proofs go through one launch per pass).
//
// Twiddles w^i, i < n/2 are tabulated once per domain in HBM (L2-resident for the sizes of
// the zklaim circuit).
#include "ntt.cuh"
#include "ntt_tile.cuh"
#include <mutex>
#include <stdlib.h>
#include <string.h>

using namespace zkb;

namespace {

constexpr int NTT_T = 256;     // threads per block
constexpr int NTT_EMAX = 2048; // elements staged per block (64 KB of shared memory)
#ifndef NTT_TFAST
#define NTT_TFAST 1 // 1: lanes walk the t index in the staging loops (coalesced rows, but shared-memory stride C); 0: always the column index
#endif

__device__ __forceinline__ uint32_t brev(uint32_t x, int bits) { return bits ? (__brev(x) >> (32 - bits)) : 0u; }

// Word-planar staging with an XOR swizzle of the low five index bits by the next five: the transposing load / store
// phase (lanes walk t, index stride C = 2 .. 8 words) would otherwise hit 32 / C banks only (ncu: 16.4 M conflicts
// in 20 M shared-memory wavefronts of a 2^20 pass, profiles/r01_k_ntt_pass_ncu.txt); butterfly levels whose warp
// touches 32 consecutive aligned indices stay conflict-free (the swizzle is a permutation inside such a group).
#ifndef NTT_SWIZZLE
#define NTT_SWIZZLE 1
#endif
__device__ __forceinline__ int smw_pos(int idx) { return NTT_SWIZZLE ? (idx ^ ((idx >> 5) & 31)) : idx; }
__device__ __forceinline__ void smw_put(uint32_t *sm, int E, int idx, const Fr &v)
{
    const int p = smw_pos(idx);
#pragma unroll
    for (int i = 0; i < 8; i++) sm[i * E + p] = v.v[i];
}
__device__ __forceinline__ Fr smw_get(const uint32_t *sm, int E, int idx)
{
    const int p = smw_pos(idx);
    Fr v;
#pragma unroll
    for (int i = 0; i < 8; i++) v.v[i] = sm[i * E + p];
    return v;
}

// One pass: butterfly levels on index bits [lobits, lobits + k) of every vector, in place.
// Element index = hi << (lobits + k) | t << lobits | lo; column = hi << lobits | lo.
// post (optional): result multiplied by post[bitrev_L(index)] on the way out.
template <bool DIT>
__global__ void __launch_bounds__(NTT_T)
k_ntt_pass(Fr *__restrict__ data, size_t vstride, int L, int lobits, int k, int logC, const Fr *__restrict__ tw,
           const Fr *__restrict__ post)
{
    extern __shared__ uint32_t sm[];
    const int C = 1 << logC, E = C << k;
    Fr *a = data + (size_t)blockIdx.y * vstride;
    const uint32_t col0 = blockIdx.x << logC;
    const uint32_t lomask = (1u << lobits) - 1u;
    const bool tfast = NTT_TFAST && lobits < logC;

    for (int e = threadIdx.x; e < E; e += NTT_T) {
        const int col = tfast ? (e >> k) : (e & (C - 1));
        const int t = tfast ? (e & ((1 << k) - 1)) : (e >> logC);
        const uint32_t gcol = col0 + col;
        const size_t idx = ((size_t)(gcol >> lobits) << (lobits + k)) | ((size_t)t << lobits) | (gcol & lomask);
        smw_put(sm, E, (t << logC) + col, ld_vec(a + idx));
    }
    __syncthreads();

    for (int s = 0; s < k; s++) {
        const int b = DIT ? s : (k - 1 - s);
        const int sh = L - 1 - lobits - b;
        // (measured and rejected: skipping the unit twiddles of the level on index bit 0 -- uniform per block and level,
        //  bit-exact -- made the kernel slower, 8.4 -> 10.9 ms per witness map at 2^22: the conditional defeats the
        //  two-butterfly software pipelining)
        // two butterflies per iteration, loads first: two independent Montgomery products in flight per thread
        for (int j = threadIdx.x; j < (E >> 1); j += 2 * NTT_T) {
            int i0[2], i1[2];
            Fr w[2], u[2], v[2];
            const bool two = j + NTT_T < (E >> 1);
#pragma unroll
            for (int q = 0; q < 2; q++) {
                const int jj = (q == 0 || two) ? j + q * NTT_T : j;
                const int col = jj & (C - 1);
                const int tt = jj >> logC;
                const int low = tt & ((1 << b) - 1);
                const int t0 = ((tt >> b) << (b + 1)) | low;
                i0[q] = (t0 << logC) + col;
                i1[q] = i0[q] + (1 << (b + logC));
                const uint32_t lo = (col0 + col) & lomask;
                const uint32_t ex = (((uint32_t)low << lobits) | lo) << sh;
                w[q] = ld_vec_ro(tw + ex);
                u[q] = smw_get(sm, E, i0[q]);
                v[q] = smw_get(sm, E, i1[q]);
            }
            if (DIT) {
#pragma unroll
                for (int q = 0; q < 2; q++) v[q] = v[q] * w[q];
#pragma unroll
                for (int q = 0; q < 2; q++) {
                    if (q == 1 && !two) break;
                    smw_put(sm, E, i0[q], u[q] + v[q]);
                    smw_put(sm, E, i1[q], u[q] - v[q]);
                }
            } else {
                Fr d[2];
#pragma unroll
                for (int q = 0; q < 2; q++) d[q] = (u[q] - v[q]) * w[q];
#pragma unroll
                for (int q = 0; q < 2; q++) {
                    if (q == 1 && !two) break;
                    smw_put(sm, E, i0[q], u[q] + v[q]);
                    smw_put(sm, E, i1[q], d[q]);
                }
            }
        }
        __syncthreads();
    }

    for (int e = threadIdx.x; e < E; e += NTT_T) {
        const int col = tfast ? (e >> k) : (e & (C - 1));
        const int t = tfast ? (e & ((1 << k) - 1)) : (e >> logC);
        const uint32_t gcol = col0 + col;
        const size_t idx = ((size_t)(gcol >> lobits) << (lobits + k)) | ((size_t)t << lobits) | (gcol & lomask);
        Fr v = smw_get(sm, E, (t << logC) + col);
        if (post) v = v * ld_vec_ro(post + brev((uint32_t)idx, L));
        st_vec(a + idx, v);
    }
}

// The register-radix pass (ntt_tile.cuh): eight elements per thread, two or three butterfly levels per round in
// registers, one shared-memory exchange per round.  Two blocks of 256 threads per SM (64 KB of shared memory each).
template <bool DIT>
__global__ void __launch_bounds__(256, 2)
k_ntt_r8(Fr *__restrict__ data, size_t vstride, const NttPass P, const Fr *__restrict__ tw, const Fr *__restrict__ post)
{
    extern __shared__ uint32_t sm[];
    Fr *a = data + (size_t)blockIdx.y * vstride;
    const int tid = threadIdx.x, T = blockDim.x;
    ntt_tile_load(P, tid, T, sm, a, blockIdx.x);
    __syncthreads();
    for (int rd = 0; rd < P.nrounds; rd++) {
        ntt_tile_round_any<DIT>(P, rd, tid, sm, blockIdx.x, tw);
        __syncthreads();
    }
    ntt_tile_store(P, tid, T, sm, a, blockIdx.x, post);
}

// the same pass with the fused steps of the witness map's last transform (NttFuse): a separate kernel, so that the plain
// passes carry neither the extra parameters nor the extra code
template <bool DIT>
__global__ void __launch_bounds__(256, 2)
k_ntt_r8_h(Fr *__restrict__ data, size_t vstride, const NttPass P, const Fr *__restrict__ tw, NttFuse f, size_t hstride)
{
    extern __shared__ uint32_t sm[];
    const size_t voff = (size_t)blockIdx.y * vstride;
    Fr *a = data + voff;
    if (f.B) { // B and C live at fixed distances from the A vector (the caller passes them for vector 0)
        f.B += voff;
        f.C += voff;
    }
    if (f.h_mont) {
        f.h_mont += (size_t)blockIdx.y * hstride;
        if (f.h_plain) f.h_plain += (size_t)blockIdx.y * hstride;
    }
    const int tid = threadIdx.x, T = blockDim.x;
    ntt_tile_load(P, tid, T, sm, a, blockIdx.x, &f);
    __syncthreads();
    for (int rd = 0; rd < P.nrounds; rd++) {
        ntt_tile_round_any<DIT>(P, rd, tid, sm, blockIdx.x, tw);
        __syncthreads();
    }
    ntt_tile_store(P, tid, T, sm, a, blockIdx.x, nullptr, &f);
}

// out[bitrev(i)] = in[i]
__global__ void k_bitrev(const Fr *__restrict__ in, Fr *__restrict__ out, size_t vstride, int L)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >> L) return;
    const size_t v = blockIdx.y;
    st_vec(out + v * vstride + brev((uint32_t)i, L), ld_vec(in + v * vstride + i));
}

// a[i] *= tab[i] (tab may be NULL) * cst (when use_cst)
__global__ void k_scale(Fr *__restrict__ a, size_t vstride, size_t n, const Fr *__restrict__ tab, Fr cst, int use_cst)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fr *p = a + (size_t)blockIdx.y * vstride + i;
    Fr v = ld_vec(p);
    if (tab) v = v * ld_vec_ro(tab + i);
    if (use_cst) v = v * cst;
    st_vec(p, v);
}

// out[i] = scale * base^i, i < n
__global__ void k_powers(Fr *__restrict__ out, size_t n, Fr base, Fr scale)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fr acc = scale, b = base;
    for (size_t e = i; e; e >>= 1) {
        if (e & 1) acc = acc * b;
        b = b.sqr();
    }
    st_vec(out + i, acc);
}

// basic domain, last step of the witness map: position p of `a` holds m * g^i * h_i for i = bitrev(p)
// (unscaled DIF output).  h_i = a[p] * cos_inv[i] -> natural order, Montgomery and plain.
__global__ void k_h_finish(const Fr *__restrict__ a, size_t vstride, int L, const Fr *__restrict__ cos_inv,
                           Fr *__restrict__ h_mont, Fr *__restrict__ h_plain, size_t hstride)
{
    const size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >> L) return;
    const size_t v = blockIdx.y;
    const uint32_t i = brev((uint32_t)p, L);
    const Fr h = ld_vec(a + v * vstride + p) * ld_vec_ro(cos_inv + i);
    st_vec(h_mont + v * hstride + i, h);
    if (h_plain) st_vec(h_plain + v * hstride + i, h.from_mont());
}

} // namespace

namespace {

// ---- step domain kernels (step_radix2_domain.tcc:39-146) ---------------------------------
// The two power-of-two transforms of a step-domain FFT run decimation-in-time (bit-reversed in, natural out) and
// those of an inverse run decimation-in-frequency (natural in, bit-reversed out); the permutation is folded into
// the writes of the front end / the reads of the back end below, so no separate bit-reversal or copy pass exists.
// FFT front end: out = [c (big, bit-reversed) | e (small, bit-reversed)]; in_tab (optional): a[i] *= in_tab[i] first
__global__ void k_step_fft_pre(const Fr *__restrict__ a, Fr *__restrict__ out, size_t vstride, int lb, int ls,
                               const Fr *__restrict__ om_fwd, const Fr *__restrict__ in_tab)
{
    const size_t big = (size_t)1 << lb, small = (size_t)1 << ls;
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= small) return;
    const Fr *x = a + (size_t)blockIdx.y * vstride;
    Fr *o = out + (size_t)blockIdx.y * vstride;
    Fr lo = ld_vec(x + i), hi = ld_vec(x + big + i);
    if (in_tab) {
        lo = lo * ld_vec_ro(in_tab + i);
        hi = hi * ld_vec_ro(in_tab + big + i);
    }
    st_vec(o + brev((uint32_t)i, lb), lo + hi);
    Fr e = (lo - hi) * ld_vec_ro(om_fwd + i);
    for (size_t k = i + small; k < big; k += small) {
        Fr v = ld_vec(x + k);
        if (in_tab) v = v * ld_vec_ro(in_tab + k);
        st_vec(o + brev((uint32_t)k, lb), v);
        e = e + v * ld_vec_ro(om_fwd + k);
    }
    st_vec(o + big + brev((uint32_t)i, ls), e);
}

// iFFT back end: u = [NTT_big^-1 (unscaled, bit-reversed) | NTT_small^-1 (unscaled, bit-reversed)] -> coefficients in
// natural order, times out_tab[i] when given, to out (stride ostride) and, when plain != NULL, their plain form
__global__ void k_step_ifft_post(const Fr *__restrict__ u, size_t vstride, Fr *__restrict__ out, Fr *__restrict__ plain,
                                 size_t ostride, int lb, int ls, const Fr *__restrict__ om_fwd, const Fr *__restrict__ om_inv,
                                 Fr big_inv, Fr small_inv, Fr half, const Fr *__restrict__ out_tab)
{
    const size_t big = (size_t)1 << lb, small = (size_t)1 << ls;
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= small) return;
    const Fr *x = u + (size_t)blockIdx.y * vstride;
    Fr *o = out + (size_t)blockIdx.y * ostride;
    Fr *pl = plain ? plain + (size_t)blockIdx.y * ostride : nullptr;
    auto emit = [&](size_t idx, Fr v) {
        if (out_tab) v = v * ld_vec_ro(out_tab + idx);
        st_vec(o + idx, v);
        if (pl) st_vec(pl + idx, v.from_mont());
    };
    const Fr u0 = ld_vec(x + brev((uint32_t)i, lb)) * big_inv;
    Fr u1 = ld_vec(x + big + brev((uint32_t)i, ls)) * small_inv;
    for (size_t k = i + small; k < big; k += small) {
        const Fr v = ld_vec(x + brev((uint32_t)k, lb)) * big_inv;
        emit(k, v);
        u1 = u1 - v * ld_vec_ro(om_fwd + k);
    }
    u1 = u1 * ld_vec_ro(om_inv + i);
    emit(i, (u0 + u1) * half);
    emit(big + i, (u0 - u1) * half);
}

// 1 / Z(g x_i) on the coset (step_radix2_domain.tcc:222-246)
__global__ void k_step_zinv(Fr *__restrict__ tab, size_t big, size_t small, Fr c0, Fr c1, Fr om2s, Fr z1inv)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= big + small) return;
    if (i >= big) {
        st_vec(tab + i, z1inv);
        return;
    }
    Fr elt = Fr::one(), b = om2s;
    for (size_t e = i; e; e >>= 1) {
        if (e & 1) elt = elt * b;
        b = b.sqr();
    }
    st_vec(tab + i, (c0 * elt - c1).inverse());
}

// ---- host-side field helpers (portable path of fp.cuh) -------------------------------------
Fr fr_from_u64(uint64_t x)
{
    Fr r = Fr::zero();
    r.v[0] = (uint32_t)x;
    r.v[1] = (uint32_t)(x >> 32);
    return r.to_mont();
}
Fr fr_pow(Fr b, uint64_t e)
{
    Fr acc = Fr::one();
    for (; e; e >>= 1) {
        if (e & 1) acc = acc * b;
        b = b.sqr();
    }
    return acc;
}
// Fr::root_of_unity (alt_bn128_init.cpp:60), Montgomery limbs; order 2^28
Fr fr_root_of_unity_2_28()
{
    static const uint32_t w[8] = {0x80d13d9cu, 0x636e7355u, 0x2445ffd6u, 0xa22bf374u,
                                  0x1eb203d8u, 0x56452ac0u, 0x2963f9e7u, 0x1860ef94u};
    Fr r;
    for (int i = 0; i < 8; i++) r.v[i] = w[i];
    return r;
}
// libff::get_root_of_unity<Fr>(2^logn)  (field_utils.tcc:44-60)
Fr fr_root_of_unity(int logn)
{
    Fr w = fr_root_of_unity_2_28();
    for (int i = 28; i > logn; i--) w = w.sqr();
    return w;
}
int ceil_log2(size_t n) // libff::log2 (libff/common/utils.cpp)
{
    int r = ((n & (n - 1)) == 0 ? 0 : 1);
    while (n > 1) {
        n >>= 1;
        r++;
    }
    return r;
}

int powers(zkb_ctx *ctx, Fr *out, size_t n, const Fr &base, const Fr &scale)
{
    if (!n) return ZKB_OK;
    k_powers<<<ceil_div(n, 128), 128, 0, ctx->stream>>>(out, n, base, scale);
    ZKB_LAUNCH_CHECK(ctx);
    return ZKB_OK;
}

int r2_build(zkb_ctx *ctx, R2Tables *t, int logn, size_t *bytes)
{
    t->logn = logn;
    const size_t half = logn ? ((size_t)1 << (logn - 1)) : 1;
    ZKB_CUDA(ctx, cudaMalloc(&t->fwd, half * sizeof(Fr)));
    ZKB_CUDA(ctx, cudaMalloc(&t->inv, half * sizeof(Fr)));
    *bytes += 2 * half * sizeof(Fr);
    const Fr w = fr_root_of_unity(logn);
    int rc = powers(ctx, t->fwd, half, w, Fr::one());
    if (!rc) rc = powers(ctx, t->inv, half, w.inverse(), Fr::one());
    return rc;
}

// all butterfly levels of size 2^T.logn on nvec vectors, in place.
// dit = false: natural in, bit-reversed out; dit = true: bit-reversed in, natural out.
// fuse (optional): the H transform of the witness map -- the first pass computes (A B - C) zinv while loading, the
// last pass writes h = value * cos_inv in natural order to fuse->h_mont / h_plain (see NttFuse); needs r8 passes throughout
int r2_levels(zkb_ctx *ctx, const R2Tables &T, bool inverse, bool dit, Fr *a, size_t vstride, int nvec, const Fr *post,
              const NttFuse *fuse = nullptr, size_t hstride = 0)
{
    const int L = T.logn;
    if (L == 0) return ZKB_OK;
    if (!(ctx->func_attr_done & 1u)) { // per device = per context
        ZKB_CUDA(ctx, cudaFuncSetAttribute(k_ntt_pass<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, NTT_EMAX * 32));
        ZKB_CUDA(ctx, cudaFuncSetAttribute(k_ntt_pass<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, NTT_EMAX * 32));
        ZKB_CUDA(ctx, cudaFuncSetAttribute(k_ntt_r8<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, NTT_EMAX * 32));
        ZKB_CUDA(ctx, cudaFuncSetAttribute(k_ntt_r8<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, NTT_EMAX * 32));
        ZKB_CUDA(ctx, cudaFuncSetAttribute(k_ntt_r8_h<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, NTT_EMAX * 32));
        ZKB_CUDA(ctx, cudaFuncSetAttribute(k_ntt_r8_h<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, NTT_EMAX * 32));
        ZKB_CUDA(ctx, cudaFuncSetAttribute(k_ntt_r8_h<false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        ZKB_CUDA(ctx, cudaFuncSetAttribute(k_ntt_r8_h<true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        ZKB_CUDA(ctx, cudaFuncSetAttribute(k_ntt_r8<false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        ZKB_CUDA(ctx, cudaFuncSetAttribute(k_ntt_r8<true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        ctx->func_attr_done |= 1u;
    }
    static const int tile_env = [] { // development knob: ZKB_NTT_TILE = 10 or 11
        const char *e = getenv("ZKB_NTT_TILE");
        const int v = e ? atoi(e) : 0;
        return v == 10 || v == 11 ? v : 0;
    }();
    // 1 024-element tiles (128 threads, four blocks per SM: the load / compute / store phases of neighbouring blocks
    // interleave better) unless 2 048-element tiles save a whole pass (a pass takes at most tile_log - 2 levels).
    // Measured (witness map, ms, tile 10 / 11): 2^16 0.287 / 0.309, 2^17 0.416 / 0.441, 2^18 0.609 / 0.584 (3 passes
    // against 2), 2^19 1.03 / 1.10, 2^20 1.97 / 2.08, 2^21 3.87 / 3.98, 2^22 7.85 / 8.03.
    const int tile_log = tile_env ? tile_env : (L >= 18 && (L + 8) / 9 < (L + 7) / 8) ? 11 : 10;
    NttPlan pl;
    ntt_plan_transform(&pl, L, dit, tile_log);
    if (fuse)
        for (int p = 0; p < pl.npass; p++)
            if (!pl.r8[p]) return fail(ctx, ZKB_ERR_ARG, "r2_levels: fused H transform needs a domain of at least 256 points");
    const Fr *tw = inverse ? T.inv : T.fwd;
    for (int p = 0; p < pl.npass; p++) {
        const NttPass &P = pl.pass[p];
        const unsigned blocks = 1u << (L - P.k - P.logC);
        const size_t smem = ((size_t)32 << P.k) << P.logC;
        const Fr *pp = p == pl.npass - 1 ? post : nullptr;
        if (pl.r8[p]) {
            const unsigned threads = 1u << (P.k + P.logC - 3);
            const bool first = p == 0, last = p == pl.npass - 1;
            if (fuse && (first || last)) {
                NttFuse f;
                memset(&f, 0, sizeof(f));
                if (first) {
                    f.B = fuse->B;
                    f.C = fuse->C;
                    f.zinv = fuse->zinv;
                }
                if (last) {
                    f.h_mont = fuse->h_mont;
                    f.h_plain = fuse->h_plain;
                    f.cos_inv = fuse->cos_inv;
                }
                if (dit)
                    k_ntt_r8_h<true><<<dim3(blocks, nvec), threads, smem, ctx->stream>>>(a, vstride, P, tw, f, hstride);
                else
                    k_ntt_r8_h<false><<<dim3(blocks, nvec), threads, smem, ctx->stream>>>(a, vstride, P, tw, f, hstride);
            } else if (dit)
                k_ntt_r8<true><<<dim3(blocks, nvec), threads, smem, ctx->stream>>>(a, vstride, P, tw, pp);
            else
                k_ntt_r8<false><<<dim3(blocks, nvec), threads, smem, ctx->stream>>>(a, vstride, P, tw, pp);
        } else if (dit) { // transforms below 256 points
            k_ntt_pass<true><<<dim3(blocks, nvec), NTT_T, smem, ctx->stream>>>(a, vstride, L, P.lobits, P.k, P.logC, tw, pp);
        } else {
            k_ntt_pass<false><<<dim3(blocks, nvec), NTT_T, smem, ctx->stream>>>(a, vstride, L, P.lobits, P.k, P.logC, tw, pp);
        }
        ZKB_LAUNCH_CHECK(ctx);
    }
    return ZKB_OK;
}

// natural-order DFT of size 2^T.logn (unscaled), via tmp
int r2_natural(zkb_ctx *ctx, const R2Tables &T, bool inverse, Fr *a, size_t vstride, int nvec, Fr *tmp)
{
    const size_t n = (size_t)1 << T.logn;
    if (T.logn == 0) return ZKB_OK;
    int rc = r2_levels(ctx, T, inverse, false, a, vstride, nvec, nullptr);
    if (rc) return rc;
    k_bitrev<<<dim3(ceil_div(n, 128), nvec), 128, 0, ctx->stream>>>(a, tmp, vstride, T.logn);
    ZKB_LAUNCH_CHECK(ctx);
    ZKB_CUDA(ctx, cudaMemcpy2DAsync(a, vstride * sizeof(Fr), tmp, vstride * sizeof(Fr), n * sizeof(Fr), nvec,
                                    cudaMemcpyDeviceToDevice, ctx->stream));
    return ZKB_OK;
}

int scale(zkb_ctx *ctx, Fr *a, size_t vstride, size_t n, int nvec, const Fr *tab, const Fr *cst)
{
    k_scale<<<dim3(ceil_div(n, 128), nvec), 128, 0, ctx->stream>>>(a, vstride, n, tab, cst ? *cst : Fr::one(), cst ? 1 : 0);
    ZKB_LAUNCH_CHECK(ctx);
    return ZKB_OK;
}

} // namespace

namespace zkb {

Fr host_fr_from_u64(uint64_t x) { return fr_from_u64(x); }
Fr host_fr_pow(Fr b, uint64_t e) { return fr_pow(b, e); }
Fr host_fr_root_of_unity(int logn) { return fr_root_of_unity(logn); }
int host_ceil_log2(size_t n) { return ceil_log2(n); }

int domain_select_host(size_t min_size, size_t *m_out, int *kind_out)
{
    if (min_size <= 1) return ZKB_ERR_DOMAIN;
    auto is_pow2 = [](size_t x) { return x && !(x & (x - 1)); };
    auto basic_ok = [&](size_t m) { return m > 1 && is_pow2(m) && ceil_log2(m) <= 28; };
    auto step_ok = [&](size_t m) {
        if (m <= 1) return false;
        const int lg = ceil_log2(m);
        const size_t big = (size_t)1 << (lg - 1), small = m - big;
        return is_pow2(small) && lg <= 28;
    };
    const size_t big = (size_t)1 << (ceil_log2(min_size) - 1);
    const size_t small = min_size - big;
    const size_t rounded_small = (size_t)1 << ceil_log2(small);
    // order of get_evaluation_domain.tcc:41-49; extended_radix2 (log m == 29) and the geometric / arithmetic
    // sequence domains are not supported: sizes that would reach them report ZKB_ERR_DOMAIN
    const size_t cand[2] = {min_size, big + rounded_small};
    for (int t = 0; t < 2; t++) {
        if (basic_ok(cand[t])) {
            *m_out = cand[t];
            *kind_out = ZKB_DOMAIN_BASIC_RADIX2;
            return ZKB_OK;
        }
        if (ceil_log2(cand[t]) == 29) return ZKB_ERR_DOMAIN;
        if (step_ok(cand[t])) {
            *m_out = cand[t];
            *kind_out = ZKB_DOMAIN_STEP_RADIX2;
            return ZKB_OK;
        }
    }
    return ZKB_ERR_DOMAIN;
}

void domain_free(Domain *d)
{
    Fr *ptrs[] = {d->big.fwd, d->big.inv, d->small.fwd, d->small.inv, d->cos_fwd, d->cos_inv, d->zinv_tab, d->om_fwd, d->om_inv};
    for (Fr *p : ptrs)
        if (p) cudaFree(p);
    *d = Domain();
}

int domain_build(zkb_ctx *ctx, size_t m, int kind, Domain *d)
{
    *d = Domain();
    d->kind = kind;
    d->dev = ctx->dev;
    d->m = m;
    const Fr g = fr_from_u64(5); // Fr::multiplicative_generator (alt_bn128_init.cpp:59)
    const Fr ginv = g.inverse();
    int rc;
    ZKB_CUDA(ctx, cudaMalloc(&d->cos_fwd, m * sizeof(Fr)));
    ZKB_CUDA(ctx, cudaMalloc(&d->cos_inv, m * sizeof(Fr)));
    d->bytes += 2 * m * sizeof(Fr);
    if (kind == ZKB_DOMAIN_BASIC_RADIX2) {
        d->big_m = m;
        if ((rc = r2_build(ctx, &d->big, ceil_log2(m), &d->bytes))) return rc;
        const Fr minv = fr_from_u64(m).inverse();
        if ((rc = powers(ctx, d->cos_fwd, m, g, minv))) return rc;
        if ((rc = powers(ctx, d->cos_inv, m, ginv, minv))) return rc;
        d->zinv = (fr_pow(g, m) - Fr::one()).inverse();
    } else if (kind == ZKB_DOMAIN_STEP_RADIX2) {
        const int lg = ceil_log2(m);
        d->big_m = (size_t)1 << (lg - 1);
        d->small_m = m - d->big_m;
        if ((rc = r2_build(ctx, &d->big, lg - 1, &d->bytes))) return rc;
        if ((rc = r2_build(ctx, &d->small, ceil_log2(d->small_m), &d->bytes))) return rc;
        if ((rc = powers(ctx, d->cos_fwd, m, g, Fr::one()))) return rc;
        if ((rc = powers(ctx, d->cos_inv, m, ginv, Fr::one()))) return rc;
        const Fr omega = fr_root_of_unity(lg);
        ZKB_CUDA(ctx, cudaMalloc(&d->om_fwd, d->big_m * sizeof(Fr)));
        ZKB_CUDA(ctx, cudaMalloc(&d->om_inv, d->small_m * sizeof(Fr)));
        ZKB_CUDA(ctx, cudaMalloc(&d->zinv_tab, m * sizeof(Fr)));
        d->bytes += (d->big_m + d->small_m + m) * sizeof(Fr);
        if ((rc = powers(ctx, d->om_fwd, d->big_m, omega, Fr::one()))) return rc;
        if ((rc = powers(ctx, d->om_inv, d->small_m, omega.inverse(), Fr::one()))) return rc;
        // step_radix2_domain.tcc:222-246
        const Fr Z0 = fr_pow(g, d->big_m) - Fr::one();
        const Fr c0 = fr_pow(g, d->small_m) * Z0;
        const Fr c1 = fr_pow(omega, d->small_m) * Z0;
        const Fr om2s = fr_pow(omega, 2 * d->small_m);
        const Fr go = g * omega;
        const Fr Z1 = (fr_pow(go, d->big_m) - Fr::one()) * (fr_pow(go, d->small_m) - fr_pow(omega, d->small_m));
        k_step_zinv<<<ceil_div(m, 128), 128, 0, ctx->stream>>>(d->zinv_tab, d->big_m, d->small_m, c0, c1, om2s, Z1.inverse());
        ZKB_LAUNCH_CHECK(ctx);
    } else {
        return fail(ctx, ZKB_ERR_DOMAIN, "unsupported evaluation domain kind");
    }
    ZKB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return ZKB_OK;
}

// ---- step domain (step_radix2_domain.tcc:39-146) ------------------------------------------------------------
// FFT of src (natural order, optionally pre-multiplied by in_tab) -> dst (natural order); src is left intact
static int step_fft(zkb_ctx *ctx, const Domain *d, const Fr *src, Fr *dst, size_t vstride, int nvec, const Fr *in_tab)
{
    const int lb = d->big.logn, ls = d->small.logn;
    k_step_fft_pre<<<dim3(ceil_div(d->small_m, 128), nvec), 128, 0, ctx->stream>>>(src, dst, vstride, lb, ls, d->om_fwd, in_tab);
    ZKB_LAUNCH_CHECK(ctx);
    int rc = r2_levels(ctx, d->big, false, true, dst, vstride, nvec, nullptr);
    if (!rc) rc = r2_levels(ctx, d->small, false, true, dst + d->big_m, vstride, nvec, nullptr);
    return rc;
}

// iFFT of a (natural order; destroyed) -> out (natural order, stride ostride), times out_tab[i] when given;
// plain != NULL also receives the plain-integer form
static int step_ifft(zkb_ctx *ctx, const Domain *d, Fr *a, size_t vstride, int nvec, Fr *out, Fr *plain, size_t ostride,
                     const Fr *out_tab)
{
    int rc = r2_levels(ctx, d->big, true, false, a, vstride, nvec, nullptr);
    if (!rc) rc = r2_levels(ctx, d->small, true, false, a + d->big_m, vstride, nvec, nullptr);
    if (rc) return rc;
    const Fr big_inv = fr_from_u64(d->big_m).inverse(), small_inv = fr_from_u64(d->small_m).inverse();
    const Fr half = fr_from_u64(2).inverse();
    k_step_ifft_post<<<dim3(ceil_div(d->small_m, 128), nvec), 128, 0, ctx->stream>>>(
        a, vstride, out, plain, ostride, d->big.logn, d->small.logn, d->om_fwd, d->om_inv, big_inv, small_inv, half, out_tab);
    ZKB_LAUNCH_CHECK(ctx);
    return ZKB_OK;
}

static int copy_back(zkb_ctx *ctx, Fr *a, const Fr *tmp, size_t vstride, size_t m, int nvec)
{
    ZKB_CUDA(ctx, cudaMemcpy2DAsync(a, vstride * sizeof(Fr), tmp, vstride * sizeof(Fr), m * sizeof(Fr), nvec,
                                    cudaMemcpyDeviceToDevice, ctx->stream));
    return ZKB_OK;
}

int domain_op(zkb_ctx *ctx, const Domain *d, int op, Fr *a, size_t vstride, int nvec, Fr *tmp)
{
    const bool basic = d->kind == ZKB_DOMAIN_BASIC_RADIX2;
    const Fr m_scal = fr_from_u64(d->m);       // undoes the 1/m folded into the basic coset tables
    const Fr minv = m_scal.inverse();
    int rc = ZKB_OK;
    switch (op) {
    case ZKB_NTT_COSET_FFT:
    case ZKB_NTT_FFT:
        if (basic) {
            if (op == ZKB_NTT_COSET_FFT && (rc = scale(ctx, a, vstride, d->m, nvec, d->cos_fwd, &m_scal))) return rc;
            return r2_natural(ctx, d->big, false, a, vstride, nvec, tmp);
        }
        if ((rc = step_fft(ctx, d, a, tmp, vstride, nvec, op == ZKB_NTT_COSET_FFT ? d->cos_fwd : nullptr))) return rc;
        return copy_back(ctx, a, tmp, vstride, d->m, nvec);
    case ZKB_NTT_IFFT:
    case ZKB_NTT_ICOSET_FFT:
        if (basic) {
            if ((rc = r2_natural(ctx, d->big, true, a, vstride, nvec, tmp))) return rc;
            // 1/m, and g^-i for the coset variant (cos_inv already carries the 1/m)
            return op == ZKB_NTT_IFFT ? scale(ctx, a, vstride, d->m, nvec, nullptr, &minv)
                                      : scale(ctx, a, vstride, d->m, nvec, d->cos_inv, nullptr);
        }
        if ((rc = step_ifft(ctx, d, a, vstride, nvec, tmp, nullptr, vstride, op == ZKB_NTT_IFFT ? nullptr : d->cos_inv))) return rc;
        return copy_back(ctx, a, tmp, vstride, d->m, nvec);
    case ZKB_NTT_DIV_Z_ON_COSET:
        return basic ? scale(ctx, a, vstride, d->m, nvec, nullptr, &d->zinv) : scale(ctx, a, vstride, d->m, nvec, d->zinv_tab, nullptr);
    default:
        return fail(ctx, ZKB_ERR_ARG, "zkb_domain_op: unknown operation");
    }
}

int domain_to_coset(zkb_ctx *ctx, const Domain *d, Fr *a, size_t vstride, int nvec, Fr *tmp)
{
    if (d->kind == ZKB_DOMAIN_BASIC_RADIX2) {
        // DIF inverse (bit-reversed out), times g^i / m in bit-reversed order, DIT forward (natural out)
        int rc = r2_levels(ctx, d->big, true, false, a, vstride, nvec, d->cos_fwd);
        if (!rc) rc = r2_levels(ctx, d->big, false, true, a, vstride, nvec, nullptr);
        return rc;
    }
    // coefficients times g^i land in tmp, the forward transform brings them back to a: no copy, no permutation pass
    int rc = step_ifft(ctx, d, a, vstride, nvec, tmp, nullptr, vstride, d->cos_fwd);
    if (!rc) rc = step_fft(ctx, d, tmp, a, vstride, nvec, nullptr);
    return rc;
}

// (A B - C) / Z on the coset and the inverse coset transform in one go, for the basic domain: the pointwise step rides on
// the loads of the first pass, the g^-i / m scaling, the un-permutation and the plain-integer copy on the stores of the last
// pass (r1cs_to_qap.tcc:270-311).  abc = [A (nvec vectors) | B | C], vector v of matrix k at abc + (k * nvec + v) * m.
// Returns ZKB_ERR_DOMAIN when the domain does not qualify (step domain, fewer than 256 points): callers use the two-step form.
int domain_h_fused(zkb_ctx *ctx, const Domain *d, Fr *abc, int nvec, Fr *h_mont, Fr *h_plain, size_t hstride)
{
    if (d->kind != ZKB_DOMAIN_BASIC_RADIX2 || d->big.logn < 8) return ZKB_ERR_DOMAIN;
    NttFuse f;
    memset(&f, 0, sizeof(f));
    f.B = abc + (size_t)nvec * d->m;
    f.C = abc + (size_t)2 * nvec * d->m;
    f.zinv = d->zinv;
    f.h_mont = h_mont;
    f.h_plain = h_plain;
    f.cos_inv = d->cos_inv;
    return r2_levels(ctx, d->big, true, false, abc, d->m, nvec, nullptr, &f, hstride);
}

int domain_h_from_coset(zkb_ctx *ctx, const Domain *d, Fr *a, size_t vstride, int nvec, Fr *tmp, Fr *h_mont, Fr *h_plain,
                        size_t hstride)
{
    // 1 / Z on the coset is the caller's: a constant for the basic domain (it commutes with the transform), the
    // table Domain::zinv_tab for the step domain -- both belong in the caller's pointwise pass
    (void)tmp;
    if (d->kind == ZKB_DOMAIN_BASIC_RADIX2) {
        const dim3 grid(ceil_div(d->m, 128), nvec);
        int rc = r2_levels(ctx, d->big, true, false, a, vstride, nvec, nullptr);
        if (rc) return rc;
        k_h_finish<<<grid, 128, 0, ctx->stream>>>(a, vstride, d->big.logn, d->cos_inv, h_mont, h_plain, hstride);
        ZKB_LAUNCH_CHECK(ctx);
        return ZKB_OK;
    }
    return step_ifft(ctx, d, a, vstride, nvec, h_mont, h_plain, hstride, d->cos_inv);
}

} // namespace zkb

extern "C" {

int zkb_domain_select(size_t min_size, size_t *m_out, int *kind_out)
{
    if (!m_out || !kind_out) return ZKB_ERR_ARG;
    return domain_select_host(min_size, m_out, kind_out);
}

int zkb_domain_op(zkb_ctx *ctx, size_t min_size, int op, uint64_t *data)
{
    if (!ctx || !data) return fail(ctx, ZKB_ERR_ARG, "zkb_domain_op: bad argument");
    size_t m;
    int kind;
    if (domain_select_host(min_size, &m, &kind)) return fail(ctx, ZKB_ERR_DOMAIN, "zkb_domain_op: no supported evaluation domain");
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    Domain d;
    int rc = domain_build(ctx, m, kind, &d);
    if (!rc) {
        rc = ZKB_ERR_CUDA;
        if (ctx->scratch[SB_STAGE_A].ensure(m * sizeof(Fr)) == cudaSuccess && ctx->scratch[SB_STAGE_B].ensure(m * sizeof(Fr)) == cudaSuccess) {
            Fr *a = ctx->scratch[SB_STAGE_A].as<Fr>(), *tmp = ctx->scratch[SB_STAGE_B].as<Fr>();
            cudaMemcpyAsync(a, data, m * sizeof(Fr), cudaMemcpyHostToDevice, ctx->stream);
            cudaEventRecord(ctx->ev0, ctx->stream);
            rc = domain_op(ctx, &d, op, a, m, 1, tmp);
            cudaEventRecord(ctx->ev1, ctx->stream);
            if (!rc) {
                cudaMemcpyAsync(data, a, m * sizeof(Fr), cudaMemcpyDeviceToHost, ctx->stream);
                if (cudaStreamSynchronize(ctx->stream) != cudaSuccess)
                    rc = fail(ctx, ZKB_ERR_CUDA, std::string("zkb_domain_op: ") + cudaGetErrorString(cudaGetLastError()));
                float ms = 0;
                cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
                ctx->last_ms = ms;
            }
        } else
            fail(ctx, ZKB_ERR_CUDA, "zkb_domain_op: out of device memory");
    }
    domain_free(&d);
    return rc;
}

} // extern "C"

// ---- an entry point that allocates nothing (no counterpart in the reference) ----
// Callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it) get
// the natural-order radix-2 FFT / inverse FFT over Fr -- what r2_natural + the 1/n scaling do for zkb_domain_op on a
// basic_radix2_domain -- with every byte of scratch in ONE file-scope device array: the working copy of the vector
// (r2_natural's `tmp`, which zkb_domain_op takes from the context's grow-only cudaMalloc'ed scratch) and the twiddle
// table w^i, i < n/2 (which zkb_domain_op cudaMallocs per call in r2_build).  Every GPU that loads libzkb.so holds the
// array until exit: ZKB_FFT_NOALLOC_MAX_LOG2 = 14 -> 2^14 x 32 B (vector) + 2^13 x 32 B (twiddles) = 768 KiB.
// The array is shared by every call on a device and keeps the last call's contents; nothing in it has to start at zero
// (both regions are fully rewritten by each call before they are read), and calls that overlap on the GPU race on it.
#define ZKB_FFT_NOALLOC_MAX_LOG2 14
static __device__ __align__(256) unsigned char g_fft_noalloc_scratch[((size_t)32 << ZKB_FFT_NOALLOC_MAX_LOG2) + ((size_t)32 << (ZKB_FFT_NOALLOC_MAX_LOG2 - 1))];

namespace {
// In host code `g_fft_noalloc_scratch` is a host-side stand-in for the symbol, NOT a device address: the address is asked
// for once per device, with that device current, and cached here.
constexpr int NOALLOC_MAX_DEV = 64;
void *g_noalloc_addr[NOALLOC_MAX_DEV] = {nullptr};
std::mutex g_noalloc_mu;

int noalloc_scratch(zkb_ctx *ctx, unsigned char **out)
{
    if (ctx->dev < 0 || ctx->dev >= NOALLOC_MAX_DEV) return fail(ctx, ZKB_ERR_ARG, "zkb_fft_noalloc: device ordinal out of range");
    std::lock_guard<std::mutex> g(g_noalloc_mu);
    if (!g_noalloc_addr[ctx->dev]) {
        void *p = nullptr;
        ZKB_CUDA(ctx, cudaGetSymbolAddress(&p, g_fft_noalloc_scratch)); // ctx->dev is current (set by the caller below)
        if (!p) return fail(ctx, ZKB_ERR_CUDA, "zkb_fft_noalloc: cudaGetSymbolAddress returned NULL");
        g_noalloc_addr[ctx->dev] = p;
    }
    *out = (unsigned char *)g_noalloc_addr[ctx->dev];
    return ZKB_OK;
}
} // namespace

extern "C" {

int zkb_fft_noalloc_max_log2(void) { return ZKB_FFT_NOALLOC_MAX_LOG2; }

int zkb_fft_noalloc(zkb_ctx *ctx, int log2n, int inverse, const void *d_in, void *d_out)
{
    if (!ctx || !d_in || !d_out) return fail(ctx, ZKB_ERR_ARG, "zkb_fft_noalloc: bad argument");
    if (log2n < 1 || log2n > ZKB_FFT_NOALLOC_MAX_LOG2)
        return fail(ctx, ZKB_ERR_ARG, "zkb_fft_noalloc: log2n must be 1 .. " + std::to_string(ZKB_FFT_NOALLOC_MAX_LOG2) +
                                          " (the static scratch array holds 2^" + std::to_string(ZKB_FFT_NOALLOC_MAX_LOG2) +
                                          " elements); larger transforms: zkb_domain_op");
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    unsigned char *base = nullptr;
    int rc = noalloc_scratch(ctx, &base);
    if (rc) return rc;
    const size_t n = (size_t)1 << log2n;
    Fr *work = reinterpret_cast<Fr *>(base);                                                  // n elements
    Fr *tw = reinterpret_cast<Fr *>(base + ((size_t)32 << ZKB_FFT_NOALLOC_MAX_LOG2));          // n / 2 elements
    cudaStream_t st = ctx->stream;
    // twiddles of this size and direction (rebuilt every call: the array may hold another call's table)
    const Fr w = fr_root_of_unity(log2n);
    if ((rc = powers(ctx, tw, n / 2, inverse ? w.inverse() : w, Fr::one()))) return rc;
    ZKB_CUDA(ctx, cudaMemcpyAsync(work, d_in, n * sizeof(Fr), cudaMemcpyDeviceToDevice, st));
    R2Tables T;
    T.logn = log2n;
    T.fwd = tw;
    T.inv = tw;
    // decimation in frequency in place (natural in, bit-reversed out), then the permutation straight into d_out
    if ((rc = r2_levels(ctx, T, inverse != 0, false, work, n, 1, nullptr))) return rc;
    k_bitrev<<<dim3(ceil_div(n, 128), 1), 128, 0, st>>>(work, reinterpret_cast<Fr *>(d_out), n, log2n);
    ZKB_LAUNCH_CHECK(ctx);
    if (inverse) { // iFFT = 1/n times the inverse transform (basic_radix2_domain.tcc:52-62)
        const Fr ninv = fr_from_u64(n).inverse();
        if ((rc = scale(ctx, reinterpret_cast<Fr *>(d_out), n, n, 1, nullptr, &ninv))) return rc;
    }
    ZKB_CUDA(ctx, cudaStreamSynchronize(st)); // the shared array is free again when the call returns
    return ZKB_OK;
}

} // extern "C"
