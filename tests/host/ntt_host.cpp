// CPU build of zklaim_b200/csrc/ntt_tile.cuh: the phases of the register-radix NTT pass (tile load, rounds,
// tile store) run here with the threads of a phase executed one after the other -- the barrier semantics of the
// kernel -- so the GPU-less test-suite checks the index / twiddle / round-plan logic the GPU executes.
#include "../../zklaim_b200/csrc/ntt_tile.cuh"
#include <cstring>
#include <vector>
using namespace zkb;

static Fr fr_pow(Fr b, uint64_t e)
{
    Fr acc = Fr::one();
    for (; e; e >>= 1) {
        if (e & 1) acc = acc * b;
        b = b.sqr();
    }
    return acc;
}

// all butterfly levels of a 2^L transform on `a` (in place) with root w (Montgomery limbs); dit = 0: natural in,
// bit-reversed out; dit = 1: bit-reversed in, natural out.  post (optional, 2^L elements) as in the kernel.
// Returns the number of passes the radix-8 path could NOT run (those are left to the caller: error).
extern "C" int hostntt_levels(int L, int dit, uint32_t *data, const uint32_t *w_limbs, const uint32_t *post_limbs, int tile_log)
{
    const size_t n = (size_t)1 << L;
    Fr w;
    memcpy(w.v, w_limbs, 32);
    std::vector<Fr> tw(n / 2 ? n / 2 : 1);
    Fr acc = Fr::one();
    for (size_t i = 0; i < n / 2; i++) {
        tw[i] = acc;
        acc = acc * w;
    }
    Fr *a = reinterpret_cast<Fr *>(data);
    const Fr *post = reinterpret_cast<const Fr *>(post_limbs);
    NttPlan pl;
    ntt_plan_transform(&pl, L, dit != 0, tile_log);
    int bad = 0;
    for (int p = 0; p < pl.npass; p++) {
        if (!pl.r8[p]) {
            bad++;
            continue;
        }
        const NttPass &P = pl.pass[p];
        const int E = 1 << (P.k + P.logC), T = E / 8;
        const uint32_t blocks = 1u << (L - P.k - P.logC);
        std::vector<uint32_t> sm((size_t)8 * E);
        const bool last = p == pl.npass - 1;
        for (uint32_t blk = 0; blk < blocks; blk++) {
            for (int t = 0; t < T; t++) ntt_tile_load(P, t, T, sm.data(), a, blk);
            for (int rd = 0; rd < P.nrounds; rd++)
                for (int t = 0; t < T; t++) {
                    if (dit)
                        ntt_tile_round_any<true>(P, rd, t, sm.data(), blk, tw.data());
                    else
                        ntt_tile_round_any<false>(P, rd, t, sm.data(), blk, tw.data());
                }
            for (int t = 0; t < T; t++) ntt_tile_store(P, t, T, sm.data(), a, blk, last ? post : nullptr);
        }
    }
    return bad;
}

// the lanes of a warp must hit 32 different banks in every shared-memory access of a round: returns the worst
// number of lanes that share a bank over all rounds / element slots of all passes of a 2^L transform
extern "C" int hostntt_worst_bank_conflict(int L, int dit, int tile_log)
{
    NttPlan pl;
    ntt_plan_transform(&pl, L, dit != 0, tile_log);
    int worst = 1;
    for (int p = 0; p < pl.npass; p++) {
        if (!pl.r8[p]) continue;
        const NttPass &P = pl.pass[p];
        const int nbits = P.k + P.logC - 3, T = 1 << nbits;
        for (int rd = 0; rd < P.nrounds; rd++) {
            const int R = P.R[rd], pp = P.logC + P.b0[rd], hb = R == 2 ? P.F[rd][nbits] : 0, N = 1 << R;
            for (int wbase = 0; wbase < T; wbase += 32)
                for (int j = 0; j < 8; j++) {
                    int cnt[32] = {0};
                    for (int lane = 0; lane < 32 && wbase + lane < T; lane++) {
                        int idx0 = 0;
                        for (int i = 0; i < nbits; i++) idx0 |= (((wbase + lane) >> i) & 1) << P.F[rd][i];
                        const int idx = idx0 | ((j >> R) << hb) | ((j & (N - 1)) << pp);
                        const int c = ++cnt[ntt_pos(idx) & 31];
                        if (c > worst) worst = c;
                    }
                }
        }
    }
    return worst;
}
