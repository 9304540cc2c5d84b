// ntt_tile.cuh -- one pass of a radix-2 number-theoretic transform over Fr as a thread block executes it:
// up to nine butterfly levels on a tile of 2048 elements staged in shared memory, three (or two) levels at a
// time in REGISTERS (radix-8 / radix-4 groups of eight elements per thread), one shared-memory exchange
// between the rounds.  Replaces the butterfly loops of libfqfft's _basic_serial_radix2_FFT
// (libfqfft/evaluation_domain/domains/basic_radix2_domain_aux.tcc:45-79); arithmetic is exact, so any
// schedule of the same butterflies gives bit-identical vectors.
//
// The phases (load, round 0 .. n-1, store) are plain functions of (thread id, tile): the kernel in ntt.cu calls
// them with __syncthreads() in between; tests/host/ntt_host.cpp compiles this very file for the CPU and runs the
// threads of a phase one after the other against the reference's transforms.
//
// Tile geometry: a pass handles index bits [lobits, lobits + k) of every vector.  A block owns C = 2^logC
// "columns" (all other index bits fixed) of 2^k elements: tile index = t << logC | col, t = the k pass bits.
// Round r of a pass covers tile bits [p, p + R), p = logC + b0[r]; a thread owns the 8 elements that differ in
// those bits (R = 3) or two such groups of 4 (R = 2).  Which tile bits the bits of the thread id stand for is
// chosen on the host per round (NttPass::F) so that the 32 lanes of a warp always hit 32 different banks of the
// word-planar, XOR-swizzled staging area.
#pragma once
#include "fp.cuh"
#include <stddef.h>

namespace zkb {

constexpr int NTT_TILE_LOG_MAX = 11;      // elements per tile: at most 2048 = 64 KB of shared memory
constexpr int NTT_MAX_ROUNDS = 4;

struct NttPass {
    int L;        // transform size 2^L
    int lobits;   // the pass works on index bits [lobits, lobits + k)
    int k, logC;  // tile = 2^k x 2^logC elements
    int nrounds;
    int unit;     // round whose twiddles start at w^0 for every thread (tile bit 0 of the lowest pass), or -1
    unsigned char b0[NTT_MAX_ROUNDS]; // first pass bit of the round
    unsigned char R[NTT_MAX_ROUNDS];  // levels of the round: 2 or 3
    unsigned char F[NTT_MAX_ROUNDS][12]; // thread-id bit i -> tile-index bit; for R == 2, F[r][nbits] = the bit that tells the two groups apart
};

// word-planar staging with an XOR swizzle of the low five index bits by the next five
ZKB_HD int ntt_pos(int idx) { return idx ^ ((idx >> 5) & 31); }
ZKB_HD void ntt_sm_put(uint32_t *sm, int E, int idx, const Fr &v)
{
    const int p = ntt_pos(idx);
#pragma unroll
    for (int i = 0; i < 8; i++) sm[i * E + p] = v.v[i];
}
ZKB_HD Fr ntt_sm_get(const uint32_t *sm, int E, int idx)
{
    const int p = ntt_pos(idx);
    Fr v;
#pragma unroll
    for (int i = 0; i < 8; i++) v.v[i] = sm[i * E + p];
    return v;
}

ZKB_HD Fr ntt_ld(const Fr *p)
{
#if defined(__CUDA_ARCH__)
    Fr r;
    const uint4 *s = reinterpret_cast<const uint4 *>(p);
    uint4 *d = reinterpret_cast<uint4 *>(&r);
    d[0] = s[0];
    d[1] = s[1];
    return r;
#else
    return *p;
#endif
}
ZKB_HD Fr ntt_ld_ro(const Fr *p)
{
#if defined(__CUDA_ARCH__)
    Fr r;
    const uint4 *s = reinterpret_cast<const uint4 *>(p);
    uint4 *d = reinterpret_cast<uint4 *>(&r);
    d[0] = __ldg(s);
    d[1] = __ldg(s + 1);
    return r;
#else
    return *p;
#endif
}
ZKB_HD void ntt_st(Fr *p, const Fr &v)
{
#if defined(__CUDA_ARCH__)
    uint4 *d = reinterpret_cast<uint4 *>(p);
    const uint4 *s = reinterpret_cast<const uint4 *>(&v);
    d[0] = s[0];
    d[1] = s[1];
#else
    *p = v;
#endif
}

ZKB_HD uint32_t ntt_brev(uint32_t x, int bits)
{
    uint32_t r = 0;
#if defined(__CUDA_ARCH__)
    r = bits ? (__brev(x) >> (32 - bits)) : 0u;
#else
    for (int i = 0; i < bits; i++) r |= ((x >> i) & 1u) << (bits - 1 - i);
#endif
    return r;
}

// global index of tile element (t, col) of block `blk`
ZKB_HD size_t ntt_gidx(const NttPass &P, uint32_t blk, int t, int col)
{
    const uint32_t gcol = (blk << P.logC) + (uint32_t)col;
    const uint32_t lomask = (1u << P.lobits) - 1u;
    return ((size_t)(gcol >> P.lobits) << (P.lobits + P.k)) | ((size_t)t << P.lobits) | (gcol & lomask);
}

// ---- load / store phases: lanes walk consecutive global addresses ------------------------------------------
// (Measured and rejected in round 2: letting the first round load and the last round store straight from / to global
//  memory, without the staging here -- the lanes of a round then touch 32 different 128-byte lines per instruction:
//  witness map at 2^22 8.13 -> 8.38 ms.  Also rejected: dropping the UNIT round variant to shrink the kernel from
//  8 560 to 6 496 SASS instructions (137 -> 104 KB; the profile shows instruction-fetch stalls): the seven extra
//  products by w^0 per thread cost more than the fetches, 7.83 -> 8.05 ms.)
// Optional fused steps of r1cs_to_qap_witness_map (r1cs_to_qap.tcc:270-311), used by its last transform only:
struct NttFuse {
    const Fr *B, *C; // load: the staged value becomes (a * B[i] - C[i]) * zinv   (H on the coset, divided by Z)
    Fr zinv;
    Fr *h_mont, *h_plain; // store: instead of the in-place store, h[bitrev_L(i)] = value * cos_inv[bitrev_L(i)], natural
    const Fr *cos_inv;    // order, Montgomery form to h_mont and (optional) plain integers to h_plain
};

// lobits < logC: the 2^logC columns of a tile do not form one contiguous run, the t index does (tfast)
ZKB_HD void ntt_tile_load(const NttPass &P, int tid, int T, uint32_t *sm, const Fr *a, uint32_t blk, const NttFuse *f = nullptr)
{
    const int E = 1 << (P.k + P.logC), C = 1 << P.logC;
    const bool tfast = P.lobits < P.logC;
    for (int e = tid; e < E; e += T) {
        const int col = tfast ? (e >> P.k) : (e & (C - 1));
        const int t = tfast ? (e & ((1 << P.k) - 1)) : (e >> P.logC);
        const size_t g = ntt_gidx(P, blk, t, col);
        Fr v = ntt_ld(a + g);
        if (f && f->B) v = Fr::mont_mul_inline(Fr::mont_mul_inline(v, ntt_ld(f->B + g)) - ntt_ld(f->C + g), f->zinv);
        ntt_sm_put(sm, E, (t << P.logC) + col, v);
    }
}
// post (optional): result multiplied by post[bitrev_L(index)] on the way out
ZKB_HD void ntt_tile_store(const NttPass &P, int tid, int T, const uint32_t *sm, Fr *a, uint32_t blk, const Fr *post,
                           const NttFuse *f = nullptr)
{
    const int E = 1 << (P.k + P.logC), C = 1 << P.logC;
    const bool tfast = P.lobits < P.logC;
    for (int e = tid; e < E; e += T) {
        const int col = tfast ? (e >> P.k) : (e & (C - 1));
        const int t = tfast ? (e & ((1 << P.k) - 1)) : (e >> P.logC);
        const size_t g = ntt_gidx(P, blk, t, col);
        Fr v = ntt_sm_get(sm, E, (t << P.logC) + col);
        if (f && f->h_mont) {
            const uint32_t i = ntt_brev((uint32_t)g, P.L);
            v = Fr::mont_mul_inline(v, ntt_ld_ro(f->cos_inv + i));
            ntt_st(f->h_mont + i, v);
            if (f->h_plain) ntt_st(f->h_plain + i, v.from_mont());
            continue;
        }
        if (post) v = Fr::mont_mul_inline(v, ntt_ld_ro(post + ntt_brev((uint32_t)g, P.L)));
        ntt_st(a + g, v);
    }
}

// ---- butterflies ----------------------------------------------------------------------------------------
// decimation in time: (u, v) <- (u + w v, u - w v);  decimation in frequency: (u, v) <- (u + v, (u - v) w)
template <bool DIT, bool MUL> ZKB_HD void ntt_bf(Fr &u, Fr &v, const Fr &w)
{
    if (DIT) {
        if (MUL) v = Fr::mont_mul_inline(v, w);
        const Fr d = u - v;
        u = u + v;
        v = d;
    } else {
        const Fr d = u - v;
        u = u + v;
        v = MUL ? Fr::mont_mul_inline(d, w) : d;
    }
}

// One round.  Level q of the round works on pass bit b0 + q; the twiddle of a butterfly whose lower element has
// global index i is w^((i mod 2^gb) << (L - 1 - gb)), gb = lobits + b0 + q (the table holds w^e, e < 2^(L-1)).
// UNIT: every thread's twiddle base is 0 (lowest bits of the whole transform): the w^0 products are skipped.
template <bool DIT, int R, bool UNIT>
ZKB_HD void ntt_tile_round(const NttPass &P, int rd, int tid, uint32_t *sm, uint32_t blk, const Fr *tw)
{
    const int logC = P.logC, E = 1 << (P.k + logC), b0 = P.b0[rd], p = logC + b0;
    const int nbits = P.k + logC - 3;
    int idx0 = 0;
    for (int i = 0; i < nbits; i++) idx0 |= ((tid >> i) & 1) << P.F[rd][i];
    const int hb = R == 2 ? P.F[rd][nbits] : 0;
    constexpr int G = R == 3 ? 1 : 2; // groups per thread
    constexpr int N = 1 << R;         // elements per group
    Fr x[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const int idx = idx0 | ((j >> R) << hb) | ((j & (N - 1)) << p);
        x[j] = ntt_sm_get(sm, E, idx);
    }
    // twiddle exponent base of each group: the index bits below the round's bits
    const uint32_t lomask = (1u << P.lobits) - 1u;
    const int s = b0 + P.lobits;              // position of the round's lowest bit in the global index
    const int sh0 = P.L - 1 - s;              // twiddle stride shift of level 0
    uint32_t base[G];
#pragma unroll
    for (int g = 0; g < G; g++) {
        const int idg = idx0 | (g << hb);
        const uint32_t t0 = (uint32_t)(idg >> logC), col = (uint32_t)(idg & ((1 << logC) - 1));
        base[g] = UNIT ? 0u : (((t0 & ((1u << b0) - 1u)) << P.lobits) | (((blk << logC) + col) & lomask));
    }
#pragma unroll
    for (int lv = 0; lv < R; lv++) {
        const int q = DIT ? lv : (R - 1 - lv);
        // level q: pairs (j, j + 2^q) inside each group; the twiddle depends on the low q bits of j
#pragma unroll
        for (int g = 0; g < G; g++)
#pragma unroll
            for (int j = 0; j < N; j++) {
                if (j & (1 << q)) continue;
                const int jl = j & ((1 << q) - 1);
                if (UNIT && jl == 0) {
                    ntt_bf<DIT, false>(x[g * N + j], x[g * N + j + (1 << q)], x[0]);
                } else {
                    const Fr w = ntt_ld_ro(tw + (((size_t)base[g] | ((size_t)jl << s)) << (sh0 - q)));
                    ntt_bf<DIT, true>(x[g * N + j], x[g * N + j + (1 << q)], w);
                }
            }
    }
#pragma unroll
    for (int j = 0; j < 8; j++) {
        const int idx = idx0 | ((j >> R) << hb) | ((j & (N - 1)) << p);
        ntt_sm_put(sm, E, idx, x[j]);
    }
}

template <bool DIT> ZKB_HD void ntt_tile_round_any(const NttPass &P, int rd, int tid, uint32_t *sm, uint32_t blk, const Fr *tw)
{
    if (P.R[rd] == 3) {
        if (rd == P.unit)
            ntt_tile_round<DIT, 3, true>(P, rd, tid, sm, blk, tw);
        else
            ntt_tile_round<DIT, 3, false>(P, rd, tid, sm, blk, tw);
    } else {
        ntt_tile_round<DIT, 2, false>(P, rd, tid, sm, blk, tw);
    }
}

// ---- host side: the round plan of a pass ---------------------------------------------------------------------
// Returns false when the pass cannot be expressed (k < 2 or a tile below 256 elements): callers fall back.
inline bool ntt_plan_pass(NttPass *P, int L, int lobits, int k, int logC, bool dit, int tile_log = NTT_TILE_LOG_MAX)
{
    P->L = L;
    P->lobits = lobits;
    P->k = k;
    P->logC = logC;
    P->unit = -1;
    const int nidx = k + logC;
    if (k < 2 || k > 9 || nidx < 8 || nidx > tile_log || tile_log > NTT_TILE_LOG_MAX) return false;
    // k = 3 a + 2 b
    int n3 = k / 3, n2 = 0;
    if (k % 3 == 1) {
        n3 -= 1;
        n2 = 2;
    } else if (k % 3 == 2) {
        n2 = 1;
    }
    P->nrounds = n3 + n2;
    // rounds in the order of bit position (ascending), threes first so that the round on bit 0 is a radix-8 one;
    // DIT executes them in ascending, DIF in descending order of position
    int sizes[NTT_MAX_ROUNDS], n = 0;
    for (int i = 0; i < n3; i++) sizes[n++] = 3;
    for (int i = 0; i < n2; i++) sizes[n++] = 2;
    int pos = 0;
    for (int i = 0; i < n; i++) {
        const int slot = dit ? i : (n - 1 - i);
        P->b0[slot] = (unsigned char)pos;
        P->R[slot] = (unsigned char)sizes[i];
        if (pos == 0 && lobits == 0 && sizes[i] == 3) P->unit = slot;
        pos += sizes[i];
    }
    for (int rd = 0; rd < n; rd++) {
        const int R = P->R[rd], p = logC + P->b0[rd];
        bool used[16] = {false};
        for (int q = p; q < p + R; q++) used[q] = true;
        int F[16], nf = 0;
        // lanes: the free bits among the five bank bits, then for every group bit q below 5 the bit 5 + q that the
        // swizzle folds onto it
        for (int u = 0; u < 5 && u < nidx; u++)
            if (!used[u]) {
                F[nf++] = u;
                used[u] = true;
            }
        for (int q = p; q < p + R && q < 5; q++)
            if (5 + q < nidx && !used[5 + q]) {
                F[nf++] = 5 + q;
                used[5 + q] = true;
            }
        for (int u = 0; u < nidx; u++)
            if (!used[u]) {
                F[nf++] = u;
                used[u] = true;
            }
        if (nf != nidx - R) return false;
        for (int i = 0; i < nf; i++) P->F[rd][i] = (unsigned char)F[i];
        // R == 2: F[nidx - 3] (the highest free bit) separates the thread's two groups
    }
    return true;
}

// the passes of a transform of size 2^L: pass i works on k[i] index bits starting at lobits[i]; r8[i] says whether
// the register-radix kernel can run it (otherwise the one-level-per-exchange kernel of ntt.cu does)
struct NttPlan {
    int npass;
    NttPass pass[4];
    bool r8[4];
};
// tile_log: log2 of the elements a block stages (11: 256 threads, two blocks per SM; 10: 128 threads, four blocks per SM);
// a pass takes at most tile_log - 2 levels, so that its rows in HBM are at least four elements = 128 bytes long
inline void ntt_plan_transform(NttPlan *pl, int L, bool dit, int tile_log = NTT_TILE_LOG_MAX)
{
    const int kmax = tile_log - 2;
    const int n = (L + kmax - 1) / kmax;
    pl->npass = n;
    int done = 0;
    for (int i = 0; i < n; i++) {
        const int k = L / n + (i < L % n ? 1 : 0);
        const int lobits = dit ? done : (L - done - k);
        int logC = 0;
        while ((2 << logC << k) <= (1 << tile_log) && logC + 1 <= L - k) logC++;
        pl->r8[i] = ntt_plan_pass(&pl->pass[i], L, lobits, k, logC, dit, tile_log);
        done += k;
    }
}

} // namespace zkb
