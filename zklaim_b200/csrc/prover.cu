// prover.cu -- the Groth16 prover of libsnark's r1cs_gg_ppzksnark on one B200, batched over proofs.
//
// Replaces r1cs_gg_ppzksnark_prover (libsnark/zk_proof_systems/ppzksnark/r1cs_gg_ppzksnark/
// r1cs_gg_ppzksnark.tcc:390-505) and r1cs_to_qap_witness_map with d1 = d2 = d3 = 0
// (libsnark/reductions/r1cs_to_qap/r1cs_to_qap.tcc:205-334).  A proving key is uploaded once
// (zkb_key_upload); every later call works on a batch of witnesses with one launch per stage
// for the whole batch:
//
//   Z (one row per proof) = [ 1 | w (nv) | r | s | -rs | H coefficients (m + 1) ]   Montgomery Fr
//   P = the same row as plain integers (MSM digits are cut from it)
//
//   1. sparse matrix-vector products  aA = A z, aB = B z, aC = C z  (z = [1 | w], CSR rows =
//      constraints, r1cs_to_qap.tcc:228-238, :277-283) plus the satisfiability check;
//   2. iFFT + cosetFFT of the 3 * batch vectors (ntt.cu), H = (A B - C) / Z on the coset,
//      icosetFFT -> H coefficients written straight into Z and P;
//   3. three MSM runs (msm.cu).  The fixed key elements ride along as extra bases so that no variable-time
//      work is left for them:
//        run G1ab: job A  = A_query + {alpha_g1 (scalar 1), delta_g1 (scalar r)}      = g1_A  (tcc:488)
//                  job B1 = B_query.h + {beta_g1 (1), delta_g1 (s)}                     = g1_B  (tcc:491)
//        run G2:   job B2 = B_query.g + {beta_g2 (1), delta_g2 (s)}                     = g2_B  (tcc:492)
//        run G1c:  job C  = H_query + L_query + {delta_g1 (-rs)}                        = Ht + Lt - rs delta
//      G1ab and G2 need only the witness: they run on two side streams, concurrently with stages 1-2 and
//      with G1c (which needs H) on the main stream; so do the two scalar multiplications s g1_A, r g1_B;
//   4. assembly: g1_C = job C + s g1_A + r g1_B  (tcc:495).
//
// zkb_key_tune() re-plans the window widths of the witness-driven jobs from the scalar statistics of a sample
// witness (the reference special-cases 0 / 1 scalars for the same reason, multiexp.tcc:466-481).
#define ZKB_NOINLINE_MUL 1
#include "msm_impl.cuh"
#include "ntt.cuh"
#include "glv.cuh"
#include <string.h>
#include <algorithm>

using namespace zkb;

struct zkb_key {
    int dev = 0;
    size_t nc = 0, nv = 0, ni = 0, m = 0;
    int kind = 0;
    Domain dom;
    uint32_t *rowptr[3] = {nullptr, nullptr, nullptr};
    uint32_t *rowperm[3] = {nullptr, nullptr, nullptr};
    uint32_t nlong[3] = {0, 0, 0};
    uint32_t ngiant[3] = {0, 0, 0};       // rows with >= SPMV_GIANT terms: cut into segments, one block each
    uint32_t *giant_seg = nullptr;        // per segment: mat, row, e_lo, e_hi
    uint32_t *giant_row = nullptr;        // per giant row: mat, row, first segment, segment count
    uint32_t n_giant_seg = 0, n_giant_row = 0;
    std::vector<uint32_t> h_giant_seg, h_giant_row; // staging while the three matrices are uploaded
    uint32_t *col[3] = {nullptr, nullptr, nullptr};
    Fr *coeff[3] = {nullptr, nullptr, nullptr};
    size_t nnz[3] = {0, 0, 0};
    MsmPlan g1ab, g1c, g2;              // see the header: jobs {A, B1}, {C}, {B2}
    std::vector<uint32_t> h_sidx[3];    // host copies of the scalar maps (g1ab, g1c, g2), for zkb_key_tune
    size_t nnzB = 0;
    bool cs_only = false;                // uploaded without queries: zkb_witness_map only
    int shard_rank = 0, shard_world = 1; // which contiguous range of every query this key holds
    size_t zlen = 0; // elements per row of Z
    size_t bytes = 0;
    // offsets into a row of Z
    size_t z_w() const { return 1; }
    size_t z_r() const { return nv + 1; }
    size_t z_s() const { return nv + 2; }
    size_t z_nrs() const { return nv + 3; }
    size_t z_h() const { return nv + 4; }
};

namespace {

enum { SP_Z = SB_PROVER0, SP_P, SP_ABC, SP_TMP, SP_RS, SP_FLAGS, SP_PROOF, SP_W, SP_RES_AB, SP_RES_C, SP_RES2, SP_SMUL, SP_GIANT };

struct CsrDev {
    const uint32_t *rowptr[3];
    const uint32_t *col[3];
    const Fr *coeff[3];
    const uint32_t *perm[3]; // rows by decreasing length
    uint32_t nlong[3];       // how many of them have >= SPMV_LONG terms
    uint32_t ngiant[3];      // the first ngiant of those are handled by k_spmv_giant
};

__device__ __forceinline__ bool is_mont_one(const Fr &a)
{
    uint32_t o = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) o |= a.v[i] ^ FrP::ONE(i);
    return o == 0;
}

// abc[(mat * batch + p) * m + row] = <row of matrix mat, z_p>; rows >= nc: the input-consistency
// constraints of r1cs_to_qap.tcc:228-231 for A, zero for B and C.
// Rows are visited in order of decreasing length (csr.perm, built at key upload) so that the threads of a
// warp see rows of equal length; the csr.nlong[mat] longest rows (>= SPMV_LONG terms: the packing
// constraints of the SHA-256 / comparison gadgets) get a whole warp each.
constexpr int SPMV_LONG = 48;
__device__ __forceinline__ Fr spmv_term(const CsrDev &csr, int mat, const Fr *__restrict__ z, uint32_t e)
{
    const Fr c = ld_vec_ro(csr.coeff[mat] + e);
    const Fr x = ld_vec(z + csr.col[mat][e]);
    return is_mont_one(c) ? x : c * x;
}
__global__ void __launch_bounds__(128)
k_spmv(CsrDev csr, uint32_t long_blocks, const Fr *__restrict__ Z, size_t zlen, size_t nc, size_t ni, size_t m, int batch,
       Fr *__restrict__ abc, int mat_base)
{
    const int mat = blockIdx.y + mat_base, p = blockIdx.z;
    const Fr *z = Z + (size_t)p * zlen;
    Fr *out = abc + ((size_t)mat * batch + p) * m;
    const uint32_t nlong = csr.nlong[mat];
    if (blockIdx.x < long_blocks) {
        const uint32_t w = blockIdx.x * 4 + (threadIdx.x >> 5) + csr.ngiant[mat];
        const int lane = threadIdx.x & 31;
        if (w >= nlong) return;
        const uint32_t row = csr.perm[mat][w];
        const uint32_t lo = csr.rowptr[mat][row], hi = csr.rowptr[mat][row + 1];
        Fr acc = Fr::zero();
        for (uint32_t e = lo + lane; e < hi; e += 32) acc = acc + spmv_term(csr, mat, z, e);
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            Fr o;
#pragma unroll
            for (int i = 0; i < 8; i++) o.v[i] = __shfl_down_sync(0xffffffffu, acc.v[i], d);
            acc = acc + o;
        }
        if (lane == 0) st_vec(out + row, acc);
        return;
    }
    const size_t s = (size_t)(blockIdx.x - long_blocks) * blockDim.x + threadIdx.x + nlong;
    if (s >= m) return;
    Fr acc = Fr::zero();
    size_t row = s;
    if (s < nc) {
        row = csr.perm[mat][s];
        const uint32_t lo = csr.rowptr[mat][row], hi = csr.rowptr[mat][row + 1];
        for (uint32_t e = lo; e < hi; e++) acc = acc + spmv_term(csr, mat, z, e);
    } else if (mat == 0 && s - nc <= ni) {
        acc = ld_vec(z + (s - nc));
    }
    st_vec(out + row, acc);
}

// Rows with tens of thousands of terms (libsnark's example circuits end with one constraint over ALL variables;
// r1cs_examples.tcc:70-90) are cut into segments of SPMV_GIANT_SEG terms: one block per (segment, proof) writes a
// partial sum, k_spmv_giant_finish adds the partials of a row.
constexpr int SPMV_GIANT = 8192, SPMV_GIANT_SEG = 4096;
__global__ void __launch_bounds__(128)
k_spmv_giant(CsrDev csr, const uint32_t *__restrict__ seg, const Fr *__restrict__ Z, size_t zlen, uint32_t nseg,
             Fr *__restrict__ partial)
{
    __shared__ Fr sh[4];
    const uint32_t g = blockIdx.x, p = blockIdx.y;
    const int mat = (int)seg[4 * g], lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const uint32_t lo = seg[4 * g + 2], hi = seg[4 * g + 3];
    const Fr *z = Z + (size_t)p * zlen;
    Fr acc = Fr::zero();
    for (uint32_t e = lo + threadIdx.x; e < hi; e += blockDim.x) acc = acc + spmv_term(csr, mat, z, e);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        Fr o;
#pragma unroll
        for (int i = 0; i < 8; i++) o.v[i] = __shfl_down_sync(0xffffffffu, acc.v[i], d);
        acc = acc + o;
    }
    if (lane == 0) sh[wid] = acc;
    __syncthreads();
    if (threadIdx.x == 0) st_vec(partial + (size_t)p * nseg + g, sh[0] + sh[1] + (sh[2] + sh[3]));
}
__global__ void __launch_bounds__(32)
k_spmv_giant_finish(const uint32_t *__restrict__ rows, const Fr *__restrict__ partial, uint32_t nseg, size_t m, int batch,
                    Fr *__restrict__ abc)
{
    const uint32_t g = blockIdx.x, p = blockIdx.y;
    const uint32_t mat = rows[4 * g], row = rows[4 * g + 1], s0 = rows[4 * g + 2], cnt = rows[4 * g + 3];
    Fr acc = Fr::zero();
    for (uint32_t k = threadIdx.x; k < cnt; k += 32) acc = acc + ld_vec(partial + (size_t)p * nseg + s0 + k);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        Fr o;
#pragma unroll
        for (int i = 0; i < 8; i++) o.v[i] = __shfl_down_sync(0xffffffffu, acc.v[i], d);
        acc = acc + o;
    }
    if (threadIdx.x == 0) st_vec(abc + ((size_t)mat * batch + p) * m + row, acc);
}

// r1cs_constraint_system::is_satisfied (r1cs.tcc): <a,z> * <b,z> == <c,z> for every constraint
__global__ void k_sat_check(const Fr *__restrict__ abc, size_t nc, size_t m, int batch, uint32_t *__restrict__ flags)
{
    const size_t row = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= nc) return;
    const int p = blockIdx.y;
    const Fr a = ld_vec(abc + ((size_t)0 * batch + p) * m + row);
    const Fr b = ld_vec(abc + ((size_t)1 * batch + p) * m + row);
    const Fr c = ld_vec(abc + ((size_t)2 * batch + p) * m + row);
    if (a * b != c) atomicOr(flags + p, 1u);
}

// A <- (A * B - C) / Z   on the coset (r1cs_to_qap.tcc:270-303); 1 / Z is the constant zinv (basic domain) or
// zinv_tab[i] (step domain, step_radix2_domain.tcc:222-246)
__global__ void k_h_pointwise(Fr *__restrict__ abc, size_t m, int batch, Fr zinv, const Fr *__restrict__ zinv_tab)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const int p = blockIdx.y;
    Fr *a = abc + ((size_t)0 * batch + p) * m + i;
    const Fr b = ld_vec(abc + ((size_t)1 * batch + p) * m + i);
    const Fr c = ld_vec(abc + ((size_t)2 * batch + p) * m + i);
    const Fr h = (ld_vec(a) * b - c) * (zinv_tab ? ld_vec_ro(zinv_tab + i) : zinv);
    st_vec(a, h);
}

// fixed entries of a row of Z: 1, r, s, -rs, and the top coefficient H[m] = 0 (d1 = d2 = d3 = 0)
__global__ void k_z_fill(Fr *__restrict__ Z, size_t zlen, size_t nv, size_t m, const Fr *__restrict__ rs, int batch)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= batch) return;
    Fr *z = Z + (size_t)p * zlen;
    const Fr r = ld_vec(rs + 2 * p), s = ld_vec(rs + 2 * p + 1);
    st_vec(z, Fr::one());
    st_vec(z + nv + 1, r);
    st_vec(z + nv + 2, s);
    st_vec(z + nv + 3, (r * s).neg());
    st_vec(z + nv + 4 + m, Fr::zero());
}

// P[0 .. n) = plain form of Z[0 .. n) for every proof
__global__ void k_z_plain(const Fr *__restrict__ Z, Fr *__restrict__ P, size_t zlen, size_t n)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const size_t o = (size_t)blockIdx.y * zlen + i;
    st_vec(P + o, ld_vec(Z + o).from_mont());
}

// ---- the two variable-base scalar multiplications of r1cs_gg_ppzksnark.tcc:495 ------------------------------
// sm[2p] = s * g1_A, sm[2p + 1] = r * g1_B.  The scalar is split with the curve's endomorphism (glv.cuh: k B =
// k1 B + |k2| (-phi(B)), halves below 2^128), which leaves a chain of 128 doublings; that chain is serial and a warp instruction costs the
// same for one active lane as for 32, so the parallelism available is INSIDE a point operation: the independent
// field products of an XYZZ doubling (9 products, depth 3) and addition (14 products, depth 4) are spread over four
// lanes that run one uniform instruction stream -- a tiny register machine whose registers live in shared memory.
// An instruction = (op, dst, a, b) per lane and step; a step is either "every lane multiplies" or "every lane
// adds / subtracts".  4-bit fixed windows; table entries are registers 32 + 4 d .. 32 + 4 d + 3.
namespace vm {
enum { MUL = 0, ADD = 1, SUB = 2, SUB2 = 3 /* a - b - b */, TRIPLE = 4 /* 3 a */, NOP = 5 };
constexpr uint32_t ins(int op, int d, int a, int b) { return (uint32_t)op << 24 | (uint32_t)b << 16 | (uint32_t)a << 8 | (uint32_t)d; }
constexpr int X = 0, Y = 1, ZZ = 2, ZZZ = 3, T0 = 4; // accumulator, temporaries T0 .. T0 + 11
constexpr int Q = 0x40;                              // operand point: Q + 0 .. Q + 3 (resolved at run time)
constexpr int DUMP = 28;                             // + lane: target of idle lanes
constexpr uint32_t IDLE = ins(NOP, DUMP, X, X);
// dbl-2008-s-1 (a = 0), see XYZZ::dbl in ec.cuh
constexpr int U = T0, V = T0 + 1, W = T0 + 2, S = T0 + 3, XX = T0 + 4, M = T0 + 5, MM = T0 + 6, D = T0 + 7, YA = T0 + 8, YB = T0 + 9;
__constant__ uint32_t DBL[8][4] = {
    {ins(ADD, U, Y, Y), IDLE, IDLE, IDLE},
    {ins(MUL, V, U, U), ins(MUL, XX, X, X), IDLE, IDLE},
    {ins(TRIPLE, M, XX, XX), IDLE, IDLE, IDLE},
    {ins(MUL, W, U, V), ins(MUL, S, X, V), ins(MUL, MM, M, M), ins(MUL, ZZ, V, ZZ)},
    {ins(SUB2, X, MM, S), IDLE, IDLE, IDLE},
    {ins(SUB, D, S, X), IDLE, IDLE, IDLE},
    {ins(MUL, YA, M, D), ins(MUL, YB, W, Y), ins(MUL, ZZZ, W, ZZZ), IDLE},
    {ins(SUB, Y, YA, YB), IDLE, IDLE, IDLE},
};
constexpr uint32_t DBL_MUL_STEPS = 0b01001010;
// add-2008-s, see XYZZ::add in ec.cuh; part A ends with P = U2 - U1 (P == 0: doubling / inverse case, handled outside)
constexpr int U1 = T0, U2 = T0 + 1, S1 = T0 + 2, S2 = T0 + 3, P = T0 + 4, R = T0 + 5, PP = T0 + 6, RR = T0 + 7, Z12 = T0 + 8, Z123 = T0 + 9,
              PPP = T0 + 10, QQ = T0 + 11;
__constant__ uint32_t ADDA[2][4] = {
    {ins(MUL, U1, X, Q + 2), ins(MUL, U2, Q + 0, ZZ), ins(MUL, S1, Y, Q + 3), ins(MUL, S2, Q + 1, ZZZ)},
    {ins(SUB, P, U2, U1), ins(SUB, R, S2, S1), IDLE, IDLE},
};
constexpr uint32_t ADDA_MUL_STEPS = 0b01;
__constant__ uint32_t ADDB[7][4] = {
    {ins(MUL, PP, P, P), ins(MUL, RR, R, R), ins(MUL, Z12, ZZ, Q + 2), ins(MUL, Z123, ZZZ, Q + 3)},
    {ins(MUL, PPP, P, PP), ins(MUL, QQ, U1, PP), ins(MUL, ZZ, Z12, PP), IDLE},
    {ins(SUB, X, RR, PPP), IDLE, IDLE, IDLE},
    {ins(SUB2, X, X, QQ), IDLE, IDLE, IDLE},
    {ins(SUB, U2, QQ, X), IDLE, IDLE, IDLE},
    {ins(MUL, S2, R, U2), ins(MUL, S1, S1, PPP), ins(MUL, ZZZ, Z123, PPP), IDLE},
    {ins(SUB, Y, S2, S1), IDLE, IDLE, IDLE},
};
constexpr uint32_t ADDB_MUL_STEPS = 0b0100011;

__device__ __forceinline__ void run(const uint32_t (*prog)[4], int nsteps, uint32_t mul_steps, Fq *Rg, int qbase, int lane)
{
    for (int s = 0; s < nsteps; s++) {
        const uint32_t in = prog[s][lane];
        int d = in & 0xff, a = (in >> 8) & 0xff, b = (in >> 16) & 0xff;
        const int op = in >> 24;
        if (a >= Q) a = qbase + (a - Q);
        if (b >= Q) b = qbase + (b - Q);
        if (op == NOP) d = DUMP + lane;
        const Fq x = Rg[a], y = Rg[b];
        Fq t;
        if ((mul_steps >> s) & 1u) {
            t = Fq::mont_mul_inline(x, y); // the one product of this machine: expanded in place (the out-of-line multiplier of
                                           // this translation unit passes its operands through local memory)
        } else {
            const Fq sum = x + y, dif = x - y;
            t = op == ADD ? sum : op == SUB ? dif : op == SUB2 ? dif - y : sum + x;
        }
        __syncwarp(0xfu); // every operand has been read (steps update registers in place)
        Rg[d] = t;
        __syncwarp(0xfu);
    }
}
} // namespace vm

__global__ void __launch_bounds__(64)
k_smul(const G1X *__restrict__ rab, const Fr *__restrict__ Pl, size_t zlen, size_t off_r, size_t off_s, G1X *__restrict__ sm)
{
    // registers 0 .. 31: the machine's; 32 + 4 d: T[d] = d B; 96 + 4 d: U[d] = -phi(T[d]) = (beta x, -y) (glv.cuh)
    __shared__ Fq regs[2][160];
    __shared__ uint32_t halves[2][2][5];
    const int p = blockIdx.x, which = threadIdx.x >> 5, lane = threadIdx.x & 31; // 0: s * g1_A, 1: r * g1_B
    if (lane >= 4) return;
    Fq *Rg = regs[which];
    const Fr k = ld_vec(Pl + (size_t)p * zlen + (which == 0 ? off_s : off_r)); // plain integers
    if (lane == 0) glv::decompose(k.v, halves[which][0], halves[which][1]);   // k = k1 - |k2| lambda, both below 2^128
    const G1X *src = rab + 2 * (size_t)p + which;
    G1X *dst = sm + 2 * (size_t)p + which;
    const Fq *srcw = reinterpret_cast<const Fq *>(src);
    // lane c owns coordinate c of every point copy
    auto copy = [&](int to, int from) {
        const Fq v = Rg[from + lane];
        __syncwarp(0xfu);
        Rg[to + lane] = v;
        __syncwarp(0xfu);
    };
    Rg[32 + 4 + lane] = ld_vec(srcw + lane); // T[1] = base
    __syncwarp(0xfu);
    const bool base_inf = Rg[32 + 4 + vm::ZZ].is_zero();
    bool fallback = false;
    if (!base_inf) {
        for (int d = 2; d < 16; d++) { // T[d] = 2 T[d / 2] (+ T[1])
            copy(vm::X, 32 + 4 * (d >> 1));
            vm::run(vm::DBL, 8, vm::DBL_MUL_STEPS, Rg, 0, lane);
            if (d & 1) {
                vm::run(vm::ADDA, 2, vm::ADDA_MUL_STEPS, Rg, 32 + 4, lane);
                fallback |= Rg[vm::P].is_zero();
                vm::run(vm::ADDB, 7, vm::ADDB_MUL_STEPS, Rg, 32 + 4, lane);
            }
            copy(32 + 4 * d, vm::X);
        }
        {
            Fq beta;
            const uint32_t bm[8] = ZKB_GLV_BETA_MONT;
#pragma unroll
            for (int i = 0; i < 8; i++) beta.v[i] = bm[i];
            for (int e = 1 + lane; e < 16; e += 4) { // U[e]: one product per entry, entries dealt over the four lanes
                Rg[96 + 4 * e + vm::X] = Fq::mont_mul_inline(Rg[32 + 4 * e + vm::X], beta);
                Rg[96 + 4 * e + vm::Y] = Rg[32 + 4 * e + vm::Y].neg();
                Rg[96 + 4 * e + vm::ZZ] = Rg[32 + 4 * e + vm::ZZ];
                Rg[96 + 4 * e + vm::ZZZ] = Rg[32 + 4 * e + vm::ZZZ];
            }
            __syncwarp(0xfu);
        }
        const uint32_t *k1 = halves[which][0], *k2 = halves[which][1];
        bool acc_inf = true;
        auto add_entry = [&](int entry) {
            if (acc_inf) {
                copy(vm::X, entry);
                acc_inf = false;
                return;
            }
            vm::run(vm::ADDA, 2, vm::ADDA_MUL_STEPS, Rg, entry, lane);
            fallback |= Rg[vm::P].is_zero(); // the accumulator never equals +- a table entry for halves below 2^128 (the lattice of
                                             // glv.cuh has no vector that short); guarded all the same
            vm::run(vm::ADDB, 7, vm::ADDB_MUL_STEPS, Rg, entry, lane);
        };
        for (int w = 32; w >= 0; w--) { // 33 windows of 4 bits: one past the 128 bits of the halves
            if (!acc_inf)
                for (int i = 0; i < 4; i++) vm::run(vm::DBL, 8, vm::DBL_MUL_STEPS, Rg, 0, lane);
            const int d1 = (int)((k1[w >> 3] >> ((w & 7) * 4)) & 15u), d2 = (int)((k2[w >> 3] >> ((w & 7) * 4)) & 15u);
            if (d1) add_entry(32 + 4 * d1);
            if (d2) add_entry(96 + 4 * d2);
        }
        if (acc_inf) { // k == 0
            Rg[lane] = Fq::zero();
            __syncwarp(0xfu);
        }
    }
    if (base_inf || fallback) {
        if (lane == 0) st_vec(dst, base_inf ? G1X::inf() : scalar_mul<Fq>(k.v, ld_vec(src)));
        return;
    }
    st_vec(reinterpret_cast<Fq *>(dst) + lane, Rg[lane]);
}

// g1_C = job C + s * g1_A + r * g1_B (r1cs_gg_ppzksnark.tcc:495); proof = (g1_A, g2_B, g1_C) as Jacobian limbs
__global__ void __launch_bounds__(64)
k_assemble(const G1X *__restrict__ rab, const G1X *__restrict__ rc, const G1X *__restrict__ sm, const G2X *__restrict__ r2,
           int batch, zkb_proof *__restrict__ out)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= batch) return;
    G1X c = ld_vec(rc + p);
    c.add(ld_vec(sm + 2 * (size_t)p));
    c.add(ld_vec(sm + 2 * (size_t)p + 1));
    st_vec(reinterpret_cast<G1J *>(out[p].A), ld_vec(rab + 2 * (size_t)p).to_jacobian());
    st_vec(reinterpret_cast<G1J *>(out[p].C), c.to_jacobian());
    st_vec(reinterpret_cast<G2J *>(out[p].B), ld_vec(r2 + p).to_jacobian());
}

// a shard's share of the four sums of one proof, as Jacobian limbs (zkb_prove_partial)
__global__ void k_export_partial(const G1X *__restrict__ rab, const G1X *__restrict__ rc, const G2X *__restrict__ r2, int batch,
                                 zkb_partial *__restrict__ out)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= batch) return;
    st_vec(reinterpret_cast<G1J *>(out[p].A), ld_vec(rab + 2 * (size_t)p).to_jacobian());
    st_vec(reinterpret_cast<G1J *>(out[p].B1), ld_vec(rab + 2 * (size_t)p + 1).to_jacobian());
    st_vec(reinterpret_cast<G1J *>(out[p].C), ld_vec(rc + p).to_jacobian());
    st_vec(reinterpret_cast<G2J *>(out[p].B2), ld_vec(r2 + p).to_jacobian());
}

// sums the shares of n shards (multiexp.tcc:433-438) into the layout k_smul / k_assemble expect; rs: r, s Montgomery -> plain
__global__ void k_combine_partials(const zkb_partial *__restrict__ parts, int n, const Fr *__restrict__ rs, G1X *__restrict__ rab,
                                   G1X *__restrict__ rc, G2X *__restrict__ r2, Fr *__restrict__ rs_plain)
{
    const int t = threadIdx.x >> 5; // one warp per component: divergent code stays in separate warps
    if (threadIdx.x & 31) return;
    if (t < 3) {
        G1X acc = G1X::inf();
        for (int i = 0; i < n; i++) {
            const uint64_t *src = t == 0 ? parts[i].A : t == 1 ? parts[i].B1 : parts[i].C;
            acc.add(G1X::from_jacobian(ld_vec(reinterpret_cast<const G1J *>(src))));
        }
        st_vec(t == 2 ? rc : rab + t, acc);
    } else if (t == 3) {
        G2X acc = G2X::inf();
        for (int i = 0; i < n; i++) acc.add(G2X::from_jacobian(ld_vec(reinterpret_cast<const G2J *>(parts[i].B2))));
        st_vec(r2, acc);
    } else {
        st_vec(rs_plain, ld_vec(rs).from_mont());
        st_vec(rs_plain + 1, ld_vec(rs + 1).from_mont());
    }
}

} // namespace

namespace {

int upload_csr(zkb_ctx *ctx, zkb_key *key, int mat, const zkb_csr &h)
{
    const size_t nc = key->nc;
    if (!h.row_ptr || (h.row_ptr[nc] && (!h.col || !h.coeff))) return fail(ctx, ZKB_ERR_ARG, "zkb_key_upload: CSR pointers missing");
    const size_t nnz = h.row_ptr[nc];
    if (nnz >= (1ull << 32)) return fail(ctx, ZKB_ERR_ARG, "zkb_key_upload: more than 2^32 matrix entries");
    std::vector<uint32_t> rp(nc + 1);
    for (size_t i = 0; i <= nc; i++) {
        if (h.row_ptr[i] > nnz || (i && h.row_ptr[i] < h.row_ptr[i - 1])) return fail(ctx, ZKB_ERR_ARG, "zkb_key_upload: CSR row_ptr not monotone");
        rp[i] = (uint32_t)h.row_ptr[i];
    }
    for (size_t e = 0; e < nnz; e++)
        if (h.col[e] > key->nv) return fail(ctx, ZKB_ERR_ARG, "zkb_key_upload: CSR column out of range");
    key->nnz[mat] = nnz;
    ZKB_CUDA(ctx, cudaMalloc(&key->rowptr[mat], (nc + 1) * 4));
    ZKB_CUDA(ctx, cudaMalloc(&key->col[mat], (nnz + 1) * 4));
    ZKB_CUDA(ctx, cudaMalloc(&key->coeff[mat], (nnz + 1) * sizeof(Fr)));
    key->bytes += (nc + 1) * 4 + (nnz + 1) * 36;
    ZKB_CUDA(ctx, cudaMemcpy(key->rowptr[mat], rp.data(), (nc + 1) * 4, cudaMemcpyHostToDevice));
    // rows by decreasing length (counting sort on the length, stable)
    {
        uint32_t maxlen = 0;
        for (size_t i = 0; i < nc; i++) maxlen = std::max(maxlen, rp[i + 1] - rp[i]);
        std::vector<uint32_t> cnt((size_t)maxlen + 2, 0), perm(nc + 1);
        for (size_t i = 0; i < nc; i++) cnt[maxlen - (rp[i + 1] - rp[i]) + 1]++;
        for (size_t l = 1; l < cnt.size(); l++) cnt[l] += cnt[l - 1];
        key->nlong[mat] = 0;
        for (size_t i = 0; i < nc; i++) {
            const uint32_t len = rp[i + 1] - rp[i];
            perm[cnt[maxlen - len]++] = (uint32_t)i;
            if (len >= (uint32_t)SPMV_LONG) key->nlong[mat]++;
        }
        key->ngiant[mat] = 0;
        for (size_t k = 0; k < nc && rp[perm[k] + 1] - rp[perm[k]] >= (uint32_t)SPMV_GIANT; k++) {
            const uint32_t row = perm[k], lo = rp[row], hi = rp[row + 1];
            const uint32_t first = (uint32_t)(key->h_giant_seg.size() / 4);
            for (uint32_t e = lo; e < hi; e += SPMV_GIANT_SEG) {
                const uint32_t v[4] = {(uint32_t)mat, row, e, std::min(hi, e + (uint32_t)SPMV_GIANT_SEG)};
                key->h_giant_seg.insert(key->h_giant_seg.end(), v, v + 4);
            }
            const uint32_t r[4] = {(uint32_t)mat, row, first, (uint32_t)(key->h_giant_seg.size() / 4) - first};
            key->h_giant_row.insert(key->h_giant_row.end(), r, r + 4);
            key->ngiant[mat]++;
        }
        ZKB_CUDA(ctx, cudaMalloc(&key->rowperm[mat], (nc + 1) * 4));
        ZKB_CUDA(ctx, cudaMemcpy(key->rowperm[mat], perm.data(), nc * 4, cudaMemcpyHostToDevice));
        key->bytes += (nc + 1) * 4;
    }
    if (nnz) {
        ZKB_CUDA(ctx, cudaMemcpy(key->col[mat], h.col, nnz * 4, cudaMemcpyHostToDevice));
        ZKB_CUDA(ctx, cudaMemcpy(key->coeff[mat], h.coeff, nnz * sizeof(Fr), cudaMemcpyHostToDevice));
    }
    return ZKB_OK;
}

// appends `n` points of `words` u64 each to a host staging vector
void append_pts(std::vector<uint64_t> &v, const uint64_t *src, size_t n, int words) { v.insert(v.end(), src, src + n * words); }

void key_release(zkb_key *key)
{
    if (!key) return;
    cudaSetDevice(key->dev);
    for (int k = 0; k < 3; k++) {
        if (key->rowptr[k]) cudaFree(key->rowptr[k]);
        if (key->rowperm[k]) cudaFree(key->rowperm[k]);
        if (k == 0 && key->giant_seg) cudaFree(key->giant_seg);
        if (k == 0 && key->giant_row) cudaFree(key->giant_row);
        if (key->col[k]) cudaFree(key->col[k]);
        if (key->coeff[k]) cudaFree(key->coeff[k]);
    }
    plan_free(&key->g1ab);
    plan_free(&key->g1c);
    plan_free(&key->g2);
    domain_free(&key->dom);
    delete key;
}

int key_build(zkb_ctx *ctx, const zkb_key_desc *d, zkb_key *key, int rank, int world)
{
    key->dev = ctx->dev;
    key->nc = d->num_constraints;
    key->nv = d->num_variables;
    key->ni = d->num_inputs;
    const size_t nv = key->nv, ni = key->ni;
    if (ni > nv) return fail(ctx, ZKB_ERR_ARG, "zkb_key_upload: num_inputs > num_variables");
    // r1cs_to_qap.tcc:218: domain for num_constraints + num_inputs + 1 points
    if (domain_select_host(key->nc + ni + 1, &key->m, &key->kind))
        return fail(ctx, ZKB_ERR_DOMAIN, "zkb_key_upload: no supported evaluation domain for this constraint system");
    const size_t m = key->m;
    key->cs_only = d->A_query == nullptr;
    if (!key->cs_only && d->H_size != m - 1) return fail(ctx, ZKB_ERR_ARG, "zkb_key_upload: H_query size must be domain size - 1");
    if (nv + m + 8 >= (1ull << 31)) return fail(ctx, ZKB_ERR_ARG, "zkb_key_upload: key too large");
    for (size_t k = 0; d->B_index && k < d->B_nnz; k++)
        if (d->B_index[k] > nv) return fail(ctx, ZKB_ERR_ARG, "zkb_key_upload: B_query index out of range");
    int rc;
    if ((rc = domain_build(ctx, m, key->kind, &key->dom))) return rc;
    key->bytes += key->dom.bytes;
    const zkb_csr *mats[3] = {&d->A, &d->B, &d->C};
    for (int k = 0; k < 3; k++)
        if ((rc = upload_csr(ctx, key, k, *mats[k]))) return rc;
    key->n_giant_seg = (uint32_t)(key->h_giant_seg.size() / 4);
    key->n_giant_row = (uint32_t)(key->h_giant_row.size() / 4);
    if (key->n_giant_row) {
        ZKB_CUDA(ctx, cudaMalloc(&key->giant_seg, key->h_giant_seg.size() * 4));
        ZKB_CUDA(ctx, cudaMalloc(&key->giant_row, key->h_giant_row.size() * 4));
        ZKB_CUDA(ctx, cudaMemcpy(key->giant_seg, key->h_giant_seg.data(), key->h_giant_seg.size() * 4, cudaMemcpyHostToDevice));
        ZKB_CUDA(ctx, cudaMemcpy(key->giant_row, key->h_giant_row.data(), key->h_giant_row.size() * 4, cudaMemcpyHostToDevice));
        key->bytes += (key->h_giant_seg.size() + key->h_giant_row.size()) * 4;
    }
    key->zlen = (nv + 4 + m + 1 + 3) & ~(size_t)3;
    key->nnzB = d->B_nnz;
    if (key->cs_only) { // the constraint system and its domain are all zkb_witness_map needs
        ZKB_CUDA(ctx, cudaDeviceSynchronize());
        return ZKB_OK;
    }

    // ---- MSM plans: {A, B1}, {C} in G1 and {B2} in G2 ----
    // A shard (rank of world) keeps the contiguous range [lo, hi) of every query -- libff's own chunking,
    // multiexp.tcc:417-431 -- and rank 0 also keeps the fixed key elements; world == 1 is the whole key.
    auto range = [&](size_t total, size_t &lo, size_t &hi) {
        if (total < (size_t)world || world == 1) {
            lo = 0;
            hi = rank == 0 ? total : 0;
            return;
        }
        const size_t one = total / world;
        lo = rank * one;
        hi = rank == world - 1 ? total : lo + one;
    };
    const size_t nL_all = nv - ni;
    size_t loA, hiA, loB, hiB, loH, hiH, loL, hiL;
    range(nv + 1, loA, hiA);
    range(d->B_nnz, loB, hiB);
    range(m - 1, loH, hiH);
    range(nL_all, loL, hiL);
    const bool fixed = rank == 0; // alpha, beta, delta ride along on rank 0 only
    key->shard_rank = rank;
    key->shard_world = world;
    static const uint64_t inf_pt[16] = {0};
    std::vector<uint64_t> stA, stB, stC, stB2;
    std::vector<uint32_t> &sab = key->h_sidx[0], &sc = key->h_sidx[1], &s2 = key->h_sidx[2];
    // job A
    append_pts(stA, d->A_query + 8 * loA, hiA - loA, 8);
    for (size_t i = loA; i < hiA; i++) sab.push_back((uint32_t)i);
    if (fixed) {
        append_pts(stA, d->alpha_g1, 1, 8);
        append_pts(stA, d->delta_g1, 1, 8);
        sab.push_back(0);
        sab.push_back((uint32_t)key->z_r());
    }
    if (stA.empty()) { // an empty range still needs one (neutral) item: the point at infinity
        append_pts(stA, inf_pt, 1, 8);
        sab.push_back(0);
    }
    const size_t nA = stA.size() / 8;
    // job B1 / B2
    if (hiB > loB) {
        append_pts(stB, d->B_g1 + 8 * loB, hiB - loB, 8);
        append_pts(stB2, d->B_g2 + 16 * loB, hiB - loB, 16);
    }
    for (size_t k = loB; k < hiB; k++) {
        sab.push_back(d->B_index[k]);
        s2.push_back(d->B_index[k]);
    }
    if (fixed) {
        append_pts(stB, d->beta_g1, 1, 8);
        append_pts(stB, d->delta_g1, 1, 8);
        append_pts(stB2, d->beta_g2, 1, 16);
        append_pts(stB2, d->delta_g2, 1, 16);
        for (std::vector<uint32_t> *v : {&sab, &s2}) {
            v->push_back(0);
            v->push_back((uint32_t)key->z_s());
        }
    }
    if (stB.empty()) {
        append_pts(stB, inf_pt, 1, 8);
        append_pts(stB2, inf_pt, 1, 16);
        sab.push_back(0);
        s2.push_back(0);
    }
    const size_t nB = stB.size() / 8;
    // job C
    append_pts(stC, d->H_query + 8 * loH, hiH - loH, 8);
    for (size_t j = loH; j < hiH; j++) sc.push_back((uint32_t)(key->z_h() + j));
    if (hiL > loL) append_pts(stC, d->L_query + 8 * loL, hiL - loL, 8);
    for (size_t j = loL; j < hiL; j++) sc.push_back((uint32_t)(1 + ni + j));
    if (fixed) {
        append_pts(stC, d->delta_g1, 1, 8);
        sc.push_back((uint32_t)key->z_nrs());
    }
    if (stC.empty()) {
        append_pts(stC, inf_pt, 1, 8);
        sc.push_back(0);
    }
    const size_t nC = stC.size() / 8;

    key->g1ab.group = ZKB_G1;
    key->g1c.group = ZKB_G1;
    key->g2.group = ZKB_G2;
    // witness-driven queries are mostly 0 / 1 scalars in practice (r1cs_gg_ppzksnark.tcc:442-485 uses the
    // mixed-addition variants for them); H is dense.  zkb_key_tune refines the guess from a real witness.
    if ((rc = plan_add_job(ctx, &key->g1ab, nA, msm_choose_c(nA / 4 + 1, ZKB_G1, 0), 0))) return rc;
    if ((rc = plan_add_job(ctx, &key->g1ab, nB, msm_choose_c(nB / 4 + 1, ZKB_G1, 0), 0))) return rc;
    if ((rc = plan_add_job(ctx, &key->g1c, nC, msm_choose_c((hiH - loH) + (hiL - loL) / 4 + 1, ZKB_G1, 0), 0))) return rc;
    if ((rc = plan_add_job(ctx, &key->g2, nB, msm_choose_c(nB / 4 + 1, ZKB_G2, 0), 0))) return rc;
    if ((rc = plan_alloc(ctx, &key->g1ab))) return rc;
    if ((rc = plan_alloc(ctx, &key->g1c))) return rc;
    if ((rc = plan_alloc(ctx, &key->g2))) return rc;
    if ((rc = plan_upload_job<Fq>(ctx, &key->g1ab, 0, stA.data()))) return rc;
    if ((rc = plan_upload_job<Fq>(ctx, &key->g1ab, 1, stB.data()))) return rc;
    if ((rc = plan_upload_job<Fq>(ctx, &key->g1c, 0, stC.data()))) return rc;
    if ((rc = plan_upload_job<Fq2>(ctx, &key->g2, 0, stB2.data()))) return rc;
    ZKB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));

    MsmPlan *plans[3] = {&key->g1ab, &key->g1c, &key->g2};
    for (int k = 0; k < 3; k++) {
        const std::vector<uint32_t> &v = key->h_sidx[k];
        ZKB_CUDA(ctx, cudaMalloc(&plans[k]->d_sidx, v.size() * 4 + 4));
        ZKB_CUDA(ctx, cudaMemcpy(plans[k]->d_sidx, v.data(), v.size() * 4, cudaMemcpyHostToDevice));
        key->bytes += plans[k]->bytes + v.size() * 4;
    }
    // the synchronous copies above come from pageable memory and may return before their DMA has landed; the
    // context's streams are non-blocking, so make the key complete before anyone can enqueue work that reads it
    ZKB_CUDA(ctx, cudaDeviceSynchronize());
    return ZKB_OK;
}

// rebuilds `plan` with the window widths `c[job]`: copy 0 of every job is taken from the old plan on the device
template <class F> int replan(zkb_ctx *ctx, MsmPlan *plan, const int *c)
{
    MsmPlan np;
    np.group = plan->group;
    int rc;
    for (uint32_t q = 0; q < plan->tab.njobs; q++)
        if ((rc = plan_add_job(ctx, &np, plan->tab.j[q].n_total, c[q], 0))) return rc;
    if ((rc = plan_alloc(ctx, &np))) return rc;
    for (uint32_t q = 0; q < plan->tab.njobs; q++) {
        const uint64_t *src = (const uint64_t *)((const Affine<F> *)plan->d_pts + plan->tab.j[q].pts_off);
        if ((rc = plan_upload_job<F>(ctx, &np, (int)q, src, true))) {
            plan_free(&np);
            return rc;
        }
    }
    ZKB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    np.d_sidx = plan->d_sidx;
    plan->d_sidx = nullptr;
    plan_free(plan);
    *plan = np;
    return ZKB_OK;
}

size_t pairs_per_proof(const MsmPlan &pl)
{
    size_t e = 0;
    for (uint32_t q = 0; q < pl.tab.njobs; q++) e += (size_t)pl.tab.j[q].n * pl.tab.j[q].W;
    return e;
}

// proofs per chunk: bounded by the 2^32 pair-list positions of one MSM run and a scratch-memory budget
int chunk_size(const zkb_key *key, size_t count)
{
    const size_t eab = pairs_per_proof(key->g1ab), ec = pairs_per_proof(key->g1c), e2 = pairs_per_proof(key->g2);
    const size_t e1 = eab > ec ? eab : ec;
    const size_t b1 = key->g1ab.tab.Bp > key->g1c.tab.Bp ? key->g1ab.tab.Bp : key->g1c.tab.Bp;
    const size_t per_proof = (eab + ec) * 8 + (eab + ec) / 4 * 16 + ((size_t)key->g1ab.tab.Bp + key->g1c.tab.Bp) * 132 + e2 * 8 +
                             e2 / 4 * 32 + (size_t)key->g2.tab.Bp * 260 + key->m * 32 * 6 + key->zlen * 64;
    size_t c = 64;
    const size_t budget = (size_t)32 << 30;
    if (c > budget / per_proof) c = budget / per_proof;
    if (c > ((1ull << 31) / (e1 + 1))) c = (1ull << 31) / (e1 + 1);
    if (c > ((1ull << 31) / (b1 + 1))) c = (1ull << 31) / (b1 + 1);
    if (c > count) c = count;
    return c ? (int)c : 1;
}

// abc[(mat * batch + p) * m + row] = <row of matrix mat, z_p> for the matrices mat_lo .. mat_lo + nmat - 1 (stage 1)
int spmv_stage(zkb_ctx *ctx, const zkb_key *key, int batch, Fr *abc, int mat_lo, int nmat)
{
    cudaStream_t st = ctx->stream;
    const size_t m = key->m;
    Fr *Z = ctx->scratch[SP_Z].as<Fr>();
    CsrDev csr;
    for (int k = 0; k < 3; k++) {
        csr.rowptr[k] = key->rowptr[k];
        csr.col[k] = key->col[k];
        csr.coeff[k] = key->coeff[k];
        csr.perm[k] = key->rowperm[k];
        csr.nlong[k] = key->nlong[k];
        csr.ngiant[k] = key->ngiant[k];
    }
    if (key->n_giant_row) { // the few giant rows of all three matrices (cheap: done whatever the subset)
        ZKB_CUDA(ctx, ctx->scratch[SP_GIANT].ensure((size_t)batch * key->n_giant_seg * sizeof(Fr)));
        Fr *partial = ctx->scratch[SP_GIANT].as<Fr>();
        k_spmv_giant<<<dim3(key->n_giant_seg, batch), 128, 0, st>>>(csr, key->giant_seg, Z, key->zlen, key->n_giant_seg, partial);
        ZKB_LAUNCH_CHECK(ctx);
        k_spmv_giant_finish<<<dim3(key->n_giant_row, batch), 32, 0, st>>>(key->giant_row, partial, key->n_giant_seg, m, batch, abc);
        ZKB_LAUNCH_CHECK(ctx);
    }
    uint32_t nl = 0;
    for (int k = mat_lo; k < mat_lo + nmat; k++) nl = std::max(nl, key->nlong[k]);
    const uint32_t long_blocks = (nl + 3) / 4;
    k_spmv<<<dim3(long_blocks + ceil_div(m, 128), nmat, batch), 128, 0, st>>>(csr, long_blocks, Z, key->zlen, key->nc, key->ni, m, batch, abc, mat_lo);
    ZKB_LAUNCH_CHECK(ctx);
    return ZKB_OK;
}

// Z rows for `batch` proofs are in place (w, r, s uploaded); runs stages 1-2.  Leaves H in Z and P.
int witness_map_chunk(zkb_ctx *ctx, const zkb_key *key, int batch, uint32_t *d_flags)
{
    cudaStream_t st = ctx->stream;
    const size_t m = key->m;
    Fr *Z = ctx->scratch[SP_Z].as<Fr>(), *P = ctx->scratch[SP_P].as<Fr>();
    ZKB_CUDA(ctx, ctx->scratch[SP_ABC].ensure((size_t)3 * batch * m * sizeof(Fr)));
    Fr *abc = ctx->scratch[SP_ABC].as<Fr>();
    Fr *tmp = nullptr;
    if (key->kind != ZKB_DOMAIN_BASIC_RADIX2) {
        ZKB_CUDA(ctx, ctx->scratch[SP_TMP].ensure((size_t)3 * batch * m * sizeof(Fr)));
        tmp = ctx->scratch[SP_TMP].as<Fr>();
    }
    int rc = spmv_stage(ctx, key, batch, abc, 0, 3);
    if (rc) return rc;
    if (key->nc) {
        k_sat_check<<<dim3(ceil_div(key->nc, 128), batch), 128, 0, st>>>(abc, key->nc, m, batch, d_flags);
        ZKB_LAUNCH_CHECK(ctx);
    }
    rc = domain_to_coset(ctx, &key->dom, abc, m, 3 * batch, tmp);
    if (rc) return rc;
    rc = domain_h_fused(ctx, &key->dom, abc, batch, Z + key->z_h(), P + key->z_h(), key->zlen);
    if (rc != ZKB_ERR_DOMAIN) return rc;
    const bool basic = key->kind == ZKB_DOMAIN_BASIC_RADIX2;
    k_h_pointwise<<<dim3(ceil_div(m, 128), batch), 128, 0, st>>>(abc, m, batch, key->dom.zinv, basic ? nullptr : key->dom.zinv_tab);
    ZKB_LAUNCH_CHECK(ctx);
    return domain_h_from_coset(ctx, &key->dom, abc, m, batch, tmp, Z + key->z_h(), P + key->z_h(), key->zlen);
}

// stages the fixed entries of Z / P for a chunk whose witnesses are already at Z[p][1 .. nv]
int stage_rows(zkb_ctx *ctx, const zkb_key *key, int batch, const Fr *d_rs, uint32_t *d_flags)
{
    cudaStream_t st = ctx->stream;
    Fr *Z = ctx->scratch[SP_Z].as<Fr>(), *P = ctx->scratch[SP_P].as<Fr>();
    ZKB_CUDA(ctx, cudaMemsetAsync(d_flags, 0, (size_t)batch * 4, st));
    k_z_fill<<<ceil_div((size_t)batch, 64), 64, 0, st>>>(Z, key->zlen, key->nv, key->m, d_rs, batch);
    ZKB_LAUNCH_CHECK(ctx);
    k_z_plain<<<dim3(ceil_div(key->nv + 4, 128), batch), 128, 0, st>>>(Z, P, key->zlen, key->nv + 4);
    ZKB_LAUNCH_CHECK(ctx);
    return ZKB_OK;
}

// the witness-only MSM runs and the two scalar multiplications, on the side streams (forked after stage_rows)
int msm_side_lanes(zkb_ctx *ctx, const zkb_key *key, int batch, bool partial)
{
    const uint32_t *P = ctx->scratch[SP_P].as<uint32_t>();
    ZKB_CUDA(ctx, ctx->scratch[SP_RES_AB].ensure((size_t)2 * batch * sizeof(G1X)));
    ZKB_CUDA(ctx, ctx->scratch[SP_SMUL].ensure((size_t)2 * batch * sizeof(G1X)));
    ZKB_CUDA(ctx, ctx->scratch[SP_RES2].ensure((size_t)batch * sizeof(G2X)));
    G1X *rab = ctx->scratch[SP_RES_AB].as<G1X>(), *sm = ctx->scratch[SP_SMUL].as<G1X>();
    G2X *r2 = ctx->scratch[SP_RES2].as<G2X>();
    ZKB_CUDA(ctx, cudaEventRecord(ctx->ev_fork, ctx->main_stream));
    {
        LaneScope lane(ctx, ctx->side_stream[0], LANE_G1_SIDE);
        ZKB_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_fork, 0));
        int rc = msm_run<Fq>(ctx, &key->g1ab, key->g1ab.tab, P, key->zlen * 8, batch, rab);
        if (rc) return rc;
        if (!partial) {
            k_smul<<<batch, 64, 0, ctx->stream>>>(rab, ctx->scratch[SP_P].as<Fr>(), key->zlen, key->z_r(), key->z_s(), sm);
            ZKB_LAUNCH_CHECK(ctx);
        }
        ZKB_CUDA(ctx, cudaEventRecord(ctx->ev_join[0], ctx->stream));
    }
    {
        LaneScope lane(ctx, ctx->side_stream[1], 0);
        ZKB_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_fork, 0));
        int rc = msm_run<Fq2>(ctx, &key->g2, key->g2.tab, P, key->zlen * 8, batch, r2);
        if (rc) return rc;
        ZKB_CUDA(ctx, cudaEventRecord(ctx->ev_join[1], ctx->stream));
    }
    return ZKB_OK;
}

// job C (needs H) on the main stream, then the join with the side lanes and the assembly
int msm_main_lane_and_assemble(zkb_ctx *ctx, const zkb_key *key, int batch, zkb_proof *d_out, zkb_partial *d_part)
{
    const uint32_t *P = ctx->scratch[SP_P].as<uint32_t>();
    ZKB_CUDA(ctx, ctx->scratch[SP_RES_C].ensure((size_t)batch * sizeof(G1X)));
    G1X *rc_ = ctx->scratch[SP_RES_C].as<G1X>();
    int rc = msm_run<Fq>(ctx, &key->g1c, key->g1c.tab, P, key->zlen * 8, batch, rc_);
    if (rc) return rc;
    ZKB_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_join[0], 0));
    ZKB_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_join[1], 0));
    if (d_part)
        k_export_partial<<<ceil_div((size_t)batch, 64), 64, 0, ctx->stream>>>(ctx->scratch[SP_RES_AB].as<G1X>(), rc_,
                                                                            ctx->scratch[SP_RES2].as<G2X>(), batch, d_part);
    else
        k_assemble<<<ceil_div((size_t)batch, 64), 64, 0, ctx->stream>>>(ctx->scratch[SP_RES_AB].as<G1X>(), rc_, ctx->scratch[SP_SMUL].as<G1X>(),
                                                                      ctx->scratch[SP_RES2].as<G2X>(), batch, d_out);
    ZKB_LAUNCH_CHECK(ctx);
    return ZKB_OK;
}

} // namespace

namespace {

// shared body of zkb_witness_map / zkb_prove*: `dev_in` says whether w, r, s are device pointers
int prove_impl(zkb_ctx *ctx, const zkb_key *key, size_t count, const uint64_t *w, const uint64_t *r, const uint64_t *s,
               bool dev_in, zkb_proof *out, uint64_t *H_out, bool map_only, zkb_partial *part_out = nullptr)
{
    if (!ctx || !key || !w || (!map_only && (!r || !s || (!out && !part_out)))) return fail(ctx, ZKB_ERR_ARG, "zkb_prove: bad argument");
    if (key->dev != ctx->dev) return fail(ctx, ZKB_ERR_ARG, "zkb_prove: key lives on another device");
    if (!map_only && key->cs_only) return fail(ctx, ZKB_ERR_ARG, "zkb_prove: this key was uploaded without queries (witness map only)");
    if (!map_only && !part_out && key->shard_world != 1)
        return fail(ctx, ZKB_ERR_ARG, "zkb_prove: this key holds one shard of the queries: use zkb_prove_partial + zkb_prove_combine");
    const bool partial = part_out != nullptr;
    static_assert(sizeof(zkb_partial) >= sizeof(zkb_proof), "the proof staging buffer is sized for partials");
    if (count == 0) return ZKB_OK;
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    cudaStream_t st = ctx->stream;
    const size_t nv = key->nv, m = key->m, zlen = key->zlen;
    const int chunk = chunk_size(key, count);
    const cudaMemcpyKind kin = dev_in ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    double total_ms = 0;
    bool unsat = false;
    ctx->batch_status.assign(count, 0);
    for (size_t base = 0; base < count; base += chunk) {
        const int batch = (int)((count - base < (size_t)chunk) ? count - base : chunk);
        ZKB_CUDA(ctx, ctx->scratch[SP_Z].ensure((size_t)batch * zlen * sizeof(Fr)));
        ZKB_CUDA(ctx, ctx->scratch[SP_P].ensure((size_t)batch * zlen * sizeof(Fr)));
        ZKB_CUDA(ctx, ctx->scratch[SP_RS].ensure((size_t)batch * 2 * sizeof(Fr)));
        ZKB_CUDA(ctx, ctx->scratch[SP_FLAGS].ensure((size_t)batch * 4));
        ZKB_CUDA(ctx, ctx->scratch[SP_PROOF].ensure((size_t)batch * sizeof(zkb_partial)));
        Fr *Z = ctx->scratch[SP_Z].as<Fr>();
        Fr *rs = ctx->scratch[SP_RS].as<Fr>();
        uint32_t *flags = ctx->scratch[SP_FLAGS].as<uint32_t>();
        zkb_proof *d_out = ctx->scratch[SP_PROOF].as<zkb_proof>();
        if (nv && dev_in)
            ZKB_CUDA(ctx, cudaMemcpy2DAsync(Z + 1, zlen * sizeof(Fr), w + base * nv * 4, nv * sizeof(Fr), nv * sizeof(Fr), batch, kin, st));
        if (nv && !dev_in && upload_rows(ctx, Z + 1, zlen * sizeof(Fr), w + base * nv * 4, nv * sizeof(Fr), nv * sizeof(Fr), batch, st))
            return ZKB_ERR_CUDA;
        if (map_only) {
            ZKB_CUDA(ctx, cudaMemsetAsync(rs, 0, (size_t)batch * 2 * sizeof(Fr), st));
        } else {
            ZKB_CUDA(ctx, cudaMemcpy2DAsync(rs, 2 * sizeof(Fr), r + base * 4, sizeof(Fr), sizeof(Fr), batch, kin, st));
            ZKB_CUDA(ctx, cudaMemcpy2DAsync(rs + 1, 2 * sizeof(Fr), s + base * 4, sizeof(Fr), sizeof(Fr), batch, kin, st));
        }
        ZKB_CUDA(ctx, cudaEventRecord(ctx->ev0, st));
        int rc = stage_rows(ctx, key, batch, rs, flags);
        if (!rc && !map_only) rc = msm_side_lanes(ctx, key, batch, partial);
        if (!rc) rc = witness_map_chunk(ctx, key, batch, flags);
        if (!rc && !map_only) rc = msm_main_lane_and_assemble(ctx, key, batch, d_out, partial ? (zkb_partial *)d_out : nullptr);
        if (rc) {
            cudaDeviceSynchronize(); // never leave a side stream running behind an error return
            return rc;
        }
        ZKB_CUDA(ctx, cudaEventRecord(ctx->ev1, st));
        const size_t out_sz = partial ? sizeof(zkb_partial) : sizeof(zkb_proof);
        const size_t pin_bytes = (size_t)batch * (out_sz + 4);
        if (ensure_pinned(ctx, pin_bytes)) return ZKB_ERR_CUDA;
        uint32_t *h_flags = (uint32_t *)((char *)ctx->pinned + (size_t)batch * out_sz);
        if (!map_only) ZKB_CUDA(ctx, cudaMemcpyAsync(ctx->pinned, d_out, (size_t)batch * out_sz, cudaMemcpyDeviceToHost, st));
        ZKB_CUDA(ctx, cudaMemcpyAsync(h_flags, flags, (size_t)batch * 4, cudaMemcpyDeviceToHost, st));
        if (H_out)
            ZKB_CUDA(ctx, cudaMemcpy2DAsync(H_out + base * (m + 1) * 4, (m + 1) * sizeof(Fr), Z + key->z_h(), zlen * sizeof(Fr),
                                            (m + 1) * sizeof(Fr), batch, cudaMemcpyDeviceToHost, st));
        ZKB_CUDA(ctx, cudaStreamSynchronize(st));
        float ms = 0;
        cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
        total_ms += ms;
        prof_collect(ctx);
        if (!map_only && partial) memcpy(part_out + base, ctx->pinned, (size_t)batch * sizeof(zkb_partial));
        if (!map_only && !partial) memcpy(out + base, ctx->pinned, (size_t)batch * sizeof(zkb_proof));
        for (int p = 0; p < batch; p++) {
            if (!h_flags[p]) continue;
            unsat = true;
            ctx->batch_status[base + p] = ZKB_ERR_UNSAT;
            // what the arithmetic produced for an unsatisfying assignment is not a proof: hand back zeros
            if (!map_only && partial) memset(part_out + base + p, 0, sizeof(zkb_partial));
            if (!map_only && !partial) memset(out + base + p, 0, sizeof(zkb_proof));
        }
    }
    ctx->last_ms = total_ms;
    if (unsat) return fail(ctx, ZKB_ERR_UNSAT, "the witness does not satisfy the constraint system");
    return ZKB_OK;
}

} // namespace

extern "C" {

int zkb_key_upload(zkb_ctx *ctx, const zkb_key_desc *desc, zkb_key **out) { return zkb_key_upload_shard(ctx, desc, 0, 1, out); }

int zkb_key_upload_shard(zkb_ctx *ctx, const zkb_key_desc *desc, int rank, int world, zkb_key **out)
{
    if (!ctx || !desc || !out) return fail(ctx, ZKB_ERR_ARG, "zkb_key_upload: bad argument");
    if (world < 1 || rank < 0 || rank >= world) return fail(ctx, ZKB_ERR_ARG, "zkb_key_upload_shard: bad rank / world");
    const bool no_queries = !desc->alpha_g1 && !desc->beta_g1 && !desc->delta_g1 && !desc->beta_g2 && !desc->delta_g2 &&
                            !desc->A_query && !desc->H_query && !desc->L_query && !desc->B_g1 && !desc->B_g2;
    if (!no_queries &&
        (!desc->alpha_g1 || !desc->beta_g1 || !desc->delta_g1 || !desc->beta_g2 || !desc->delta_g2 || !desc->A_query ||
         !desc->H_query || (desc->B_nnz && (!desc->B_index || !desc->B_g1 || !desc->B_g2)) ||
         (desc->num_variables > desc->num_inputs && !desc->L_query)))
        return fail(ctx, ZKB_ERR_ARG, "zkb_key_upload: missing key arrays");
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    zkb_key *key = new zkb_key();
    const int rc = key_build(ctx, desc, key, rank, world);
    if (rc) {
        key_release(key);
        return rc;
    }
    *out = key;
    return ZKB_OK;
}

int zkb_key_tune(zkb_ctx *ctx, zkb_key *key, const uint64_t *w)
{
    if (!ctx || !key || !w) return fail(ctx, ZKB_ERR_ARG, "zkb_key_tune: bad argument");
    if (key->dev != ctx->dev) return fail(ctx, ZKB_ERR_ARG, "zkb_key_tune: key lives on another device");
    if (key->cs_only) return ZKB_OK;
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    ZKB_CUDA(ctx, cudaDeviceSynchronize());
    static const uint64_t one_r[4] = {0xac96341c4ffffffbull, 0x36fc76959f60cd29ull, 0x666ea36f7879462eull, 0x0e0a77c19a07df2full};
    const size_t nv = key->nv;
    // is the scalar at index si of a row of Z neither 0 nor 1?  (0 = the constant, 1..nv = w, above = r, s, -rs, H: dense)
    auto dense = [&](uint32_t si) -> bool {
        if (si == 0) return false;
        if (si > nv) return true;
        const uint64_t *x = w + 4 * (size_t)(si - 1);
        if ((x[0] | x[1] | x[2] | x[3]) == 0) return false;
        return memcmp(x, one_r, 32) != 0;
    };
    MsmPlan *plans[3] = {&key->g1ab, &key->g1c, &key->g2};
    for (int k = 0; k < 3; k++) {
        MsmPlan *pl = plans[k];
        int c[MSM_MAX_JOBS];
        bool changed = false;
        for (uint32_t q = 0; q < pl->tab.njobs; q++) {
            const MsmJob &J = pl->tab.j[q];
            size_t nd = 0;
            for (uint32_t i = 0; i < J.n_total; i++) nd += dense(key->h_sidx[k][J.item_begin + i]) ? 1 : 0;
            c[q] = msm_choose_c(nd + 1, pl->group, 0);
            if (c[q] < 7) c[q] = 7; // at most 37 precomputed copies of the bases
            changed |= c[q] != J.c;
        }
        if (!changed) continue;
        key->bytes -= pl->bytes;
        const int rc = pl->group == ZKB_G1 ? replan<Fq>(ctx, pl, c) : replan<Fq2>(ctx, pl, c);
        key->bytes += pl->bytes;
        if (rc) return rc;
    }
    return ZKB_OK;
}

int zkb_key_windows(const zkb_key *key, int out[4])
{
    if (!key || !out) return ZKB_ERR_ARG;
    if (key->cs_only) {
        out[0] = out[1] = out[2] = out[3] = 0;
        return ZKB_OK;
    }
    out[0] = key->g1ab.tab.j[0].c;
    out[1] = key->g1ab.tab.j[1].c;
    out[2] = key->g1c.tab.j[0].c;
    out[3] = key->g2.tab.j[0].c;
    return ZKB_OK;
}

void zkb_key_free(zkb_key *key) { key_release(key); }

size_t zkb_key_device_bytes(const zkb_key *key) { return key ? key->bytes : 0; }

int zkb_key_num_variables(const zkb_key *key, size_t *nv_out)
{
    if (!key || !nv_out) return ZKB_ERR_ARG;
    *nv_out = key->nv;
    return ZKB_OK;
}

int zkb_key_domain(const zkb_key *key, size_t *m_out, int *kind_out)
{
    if (!key || !m_out || !kind_out) return ZKB_ERR_ARG;
    *m_out = key->m;
    *kind_out = key->kind;
    return ZKB_OK;
}

int zkb_witness_map(zkb_ctx *ctx, const zkb_key *key, const uint64_t *w, uint64_t *H_out)
{
    if (!H_out) return fail(ctx, ZKB_ERR_ARG, "zkb_witness_map: H_out is NULL");
    return prove_impl(ctx, key, 1, w, nullptr, nullptr, false, nullptr, H_out, true);
}

int zkb_prove(zkb_ctx *ctx, const zkb_key *key, const uint64_t *w, const uint64_t r[4], const uint64_t s[4], zkb_proof *out,
              uint64_t *H_out)
{
    return prove_impl(ctx, key, 1, w, r, s, false, out, H_out, false);
}

int zkb_prove_partial(zkb_ctx *ctx, const zkb_key *key, const uint64_t *w, const uint64_t r[4], const uint64_t s[4],
                      zkb_partial *out, uint64_t *H_out)
{
    if (!out) return fail(ctx, ZKB_ERR_ARG, "zkb_prove_partial: out is NULL");
    return prove_impl(ctx, key, 1, w, r, s, false, nullptr, H_out, false, out);
}

int zkb_prove_combine(zkb_ctx *ctx, const zkb_partial *parts, size_t n, const uint64_t r[4], const uint64_t s[4], zkb_proof *out)
{
    if (!ctx || !parts || !n || !r || !s || !out || n > 4096) return fail(ctx, ZKB_ERR_ARG, "zkb_prove_combine: bad argument");
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    cudaStream_t st = ctx->stream;
    ZKB_CUDA(ctx, ctx->scratch[SP_W].ensure(n * sizeof(zkb_partial) + 4 * sizeof(Fr)));
    ZKB_CUDA(ctx, ctx->scratch[SP_RES_AB].ensure(2 * sizeof(G1X)));
    ZKB_CUDA(ctx, ctx->scratch[SP_RES_C].ensure(sizeof(G1X)));
    ZKB_CUDA(ctx, ctx->scratch[SP_RES2].ensure(sizeof(G2X)));
    ZKB_CUDA(ctx, ctx->scratch[SP_SMUL].ensure(2 * sizeof(G1X)));
    ZKB_CUDA(ctx, ctx->scratch[SP_PROOF].ensure(sizeof(zkb_partial)));
    char *stage = ctx->scratch[SP_W].as<char>();
    zkb_partial *d_parts = (zkb_partial *)stage;
    Fr *d_rs = (Fr *)(stage + n * sizeof(zkb_partial)), *d_rs_plain = d_rs + 2;
    ZKB_CUDA(ctx, cudaMemcpyAsync(d_parts, parts, n * sizeof(zkb_partial), cudaMemcpyHostToDevice, st));
    ZKB_CUDA(ctx, cudaMemcpyAsync(d_rs, r, sizeof(Fr), cudaMemcpyHostToDevice, st));
    ZKB_CUDA(ctx, cudaMemcpyAsync(d_rs + 1, s, sizeof(Fr), cudaMemcpyHostToDevice, st));
    G1X *rab = ctx->scratch[SP_RES_AB].as<G1X>(), *rc = ctx->scratch[SP_RES_C].as<G1X>(), *sm = ctx->scratch[SP_SMUL].as<G1X>();
    G2X *r2 = ctx->scratch[SP_RES2].as<G2X>();
    zkb_proof *d_out = ctx->scratch[SP_PROOF].as<zkb_proof>();
    k_combine_partials<<<1, 160, 0, st>>>(d_parts, (int)n, d_rs, rab, rc, r2, d_rs_plain);
    ZKB_LAUNCH_CHECK(ctx);
    k_smul<<<1, 64, 0, st>>>(rab, d_rs_plain, 0, 0, 1, sm);
    ZKB_LAUNCH_CHECK(ctx);
    k_assemble<<<1, 64, 0, st>>>(rab, rc, sm, r2, 1, d_out);
    ZKB_LAUNCH_CHECK(ctx);
    if (ensure_pinned(ctx, sizeof(zkb_proof))) return ZKB_ERR_CUDA;
    ZKB_CUDA(ctx, cudaMemcpyAsync(ctx->pinned, d_out, sizeof(zkb_proof), cudaMemcpyDeviceToHost, st));
    ZKB_CUDA(ctx, cudaStreamSynchronize(st));
    memcpy(out, ctx->pinned, sizeof(zkb_proof));
    return ZKB_OK;
}

// ---- one proof over several GPUs with the witness map split by polynomial (optional row) ------
int zkb_prove_split_begin(zkb_ctx *ctx, const zkb_key *key, const uint64_t *w, const uint64_t r[4], const uint64_t s[4], int which,
                          int check, void *d_part)
{
    if (!ctx || !key || !w || !r || !s || which < -1 || which > 2 || (which >= 0 && !d_part))
        return fail(ctx, ZKB_ERR_ARG, "zkb_prove_split_begin: bad argument");
    if (key->dev != ctx->dev) return fail(ctx, ZKB_ERR_ARG, "zkb_prove_split_begin: key lives on another device");
    if (key->cs_only) return fail(ctx, ZKB_ERR_ARG, "zkb_prove_split_begin: this key has no queries");
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    cudaStream_t st = ctx->stream;
    const size_t nv = key->nv, m = key->m, zlen = key->zlen;
    ZKB_CUDA(ctx, ctx->scratch[SP_Z].ensure(zlen * sizeof(Fr)));
    ZKB_CUDA(ctx, ctx->scratch[SP_P].ensure(zlen * sizeof(Fr)));
    ZKB_CUDA(ctx, ctx->scratch[SP_RS].ensure(2 * sizeof(Fr)));
    ZKB_CUDA(ctx, ctx->scratch[SP_FLAGS].ensure(4));
    ZKB_CUDA(ctx, ctx->scratch[SP_PROOF].ensure(sizeof(zkb_partial)));
    Fr *Z = ctx->scratch[SP_Z].as<Fr>();
    Fr *rs = ctx->scratch[SP_RS].as<Fr>();
    uint32_t *flags = ctx->scratch[SP_FLAGS].as<uint32_t>();
    // w: host memory, or device memory of this GPU (a witness gathered over NVLink, sharding.sharded_prove_split)
    if (nv && upload_rows(ctx, Z + 1, nv * sizeof(Fr), w, nv * sizeof(Fr), nv * sizeof(Fr), 1, st)) return ZKB_ERR_CUDA;
    ZKB_CUDA(ctx, cudaMemcpyAsync(rs, r, sizeof(Fr), cudaMemcpyHostToDevice, st));
    ZKB_CUDA(ctx, cudaMemcpyAsync(rs + 1, s, sizeof(Fr), cudaMemcpyHostToDevice, st));
    ZKB_CUDA(ctx, cudaEventRecord(ctx->ev0, st));
    int rc = stage_rows(ctx, key, 1, rs, flags);
    if (!rc) rc = msm_side_lanes(ctx, key, 1, true);
    if (!rc && (check || which >= 0)) {
        ZKB_CUDA(ctx, ctx->scratch[SP_ABC].ensure((size_t)3 * m * sizeof(Fr)));
        Fr *abc = ctx->scratch[SP_ABC].as<Fr>();
        Fr *tmp = nullptr;
        if (key->kind != ZKB_DOMAIN_BASIC_RADIX2) {
            ZKB_CUDA(ctx, ctx->scratch[SP_TMP].ensure((size_t)3 * m * sizeof(Fr)));
            tmp = ctx->scratch[SP_TMP].as<Fr>();
        }
        rc = check ? spmv_stage(ctx, key, 1, abc, 0, 3) : spmv_stage(ctx, key, 1, abc, which, 1);
        if (!rc && check && key->nc) {
            k_sat_check<<<dim3(ceil_div(key->nc, 128), 1), 128, 0, st>>>(abc, key->nc, m, 1, flags);
            ZKB_LAUNCH_CHECK(ctx);
        }
        if (!rc && which >= 0) {
            rc = domain_to_coset(ctx, &key->dom, abc + (size_t)which * m, m, 1, tmp);
            if (!rc) ZKB_CUDA(ctx, cudaMemcpyAsync(d_part, abc + (size_t)which * m, m * sizeof(Fr), cudaMemcpyDeviceToDevice, st));
        }
    }
    if (rc) {
        cudaDeviceSynchronize();
        return rc;
    }
    ZKB_CUDA(ctx, cudaEventRecord(ctx->ev1, st));
    ZKB_CUDA(ctx, cudaStreamSynchronize(st)); // d_part is complete; the side lanes may still be running
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->last_ms = ms;
    return ZKB_OK;
}

int zkb_prove_split_chain(zkb_ctx *ctx, const zkb_key *key, int which, void *d_part)
{
    if (!ctx || !key || which < 0 || which > 2 || !d_part) return fail(ctx, ZKB_ERR_ARG, "zkb_prove_split_chain: bad argument");
    if (key->dev != ctx->dev) return fail(ctx, ZKB_ERR_ARG, "zkb_prove_split_chain: key lives on another device");
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    cudaStream_t st = ctx->stream;
    const size_t m = key->m;
    if (ctx->scratch[SP_Z].cap < key->zlen * sizeof(Fr)) return fail(ctx, ZKB_ERR_ARG, "zkb_prove_split_chain: call zkb_prove_split_begin first");
    ZKB_CUDA(ctx, ctx->scratch[SP_ABC].ensure((size_t)3 * m * sizeof(Fr)));
    Fr *abc = ctx->scratch[SP_ABC].as<Fr>();
    Fr *tmp = nullptr;
    if (key->kind != ZKB_DOMAIN_BASIC_RADIX2) {
        ZKB_CUDA(ctx, ctx->scratch[SP_TMP].ensure((size_t)3 * m * sizeof(Fr)));
        tmp = ctx->scratch[SP_TMP].as<Fr>();
    }
    ZKB_CUDA(ctx, cudaEventRecord(ctx->ev2, st));
    int rc = spmv_stage(ctx, key, 1, abc, which, 1);
    if (!rc) rc = domain_to_coset(ctx, &key->dom, abc + (size_t)which * m, m, 1, tmp);
    if (rc) {
        cudaDeviceSynchronize();
        return rc;
    }
    ZKB_CUDA(ctx, cudaMemcpyAsync(d_part, abc + (size_t)which * m, m * sizeof(Fr), cudaMemcpyDeviceToDevice, st));
    ZKB_CUDA(ctx, cudaEventRecord(ctx->ev3, st));
    ZKB_CUDA(ctx, cudaStreamSynchronize(st));
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev2, ctx->ev3);
    ctx->last_ms += ms;
    return ZKB_OK;
}

int zkb_prove_split_combine(zkb_ctx *ctx, const zkb_key *key, void *d_abc, void *d_H_plain)
{
    if (!ctx || !key || !d_abc || !d_H_plain) return fail(ctx, ZKB_ERR_ARG, "zkb_prove_split_combine: bad argument");
    if (key->dev != ctx->dev) return fail(ctx, ZKB_ERR_ARG, "zkb_prove_split_combine: key lives on another device");
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    cudaStream_t st = ctx->stream;
    const size_t m = key->m;
    Fr *abc = (Fr *)d_abc;
    Fr *tmp = nullptr;
    if (key->kind != ZKB_DOMAIN_BASIC_RADIX2) {
        ZKB_CUDA(ctx, ctx->scratch[SP_TMP].ensure((size_t)3 * m * sizeof(Fr)));
        tmp = ctx->scratch[SP_TMP].as<Fr>();
    }
    ZKB_CUDA(ctx, ctx->scratch[SP_W].ensure((m + 1) * sizeof(Fr))); // the Montgomery form of H is not needed downstream
    ZKB_CUDA(ctx, cudaEventRecord(ctx->ev2, st));
    const bool basic = key->kind == ZKB_DOMAIN_BASIC_RADIX2;
    k_h_pointwise<<<dim3(ceil_div(m, 128), 1), 128, 0, st>>>(abc, m, 1, key->dom.zinv, basic ? nullptr : key->dom.zinv_tab);
    ZKB_LAUNCH_CHECK(ctx);
    int rc = domain_h_from_coset(ctx, &key->dom, abc, m, 1, tmp, ctx->scratch[SP_W].as<Fr>(), (Fr *)d_H_plain, m);
    if (rc) return rc;
    ZKB_CUDA(ctx, cudaEventRecord(ctx->ev3, st));
    ZKB_CUDA(ctx, cudaStreamSynchronize(st));
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev2, ctx->ev3);
    ctx->last_ms += ms;
    return ZKB_OK;
}

int zkb_prove_split_end(zkb_ctx *ctx, const zkb_key *key, const void *d_H_plain, zkb_partial *out)
{
    if (!ctx || !key || !d_H_plain || !out) return fail(ctx, ZKB_ERR_ARG, "zkb_prove_split_end: bad argument");
    if (key->dev != ctx->dev) return fail(ctx, ZKB_ERR_ARG, "zkb_prove_split_end: key lives on another device");
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    cudaStream_t st = ctx->stream;
    Fr *P = ctx->scratch[SP_P].as<Fr>();
    uint32_t *flags = ctx->scratch[SP_FLAGS].as<uint32_t>();
    zkb_proof *d_out = ctx->scratch[SP_PROOF].as<zkb_proof>();
    ZKB_CUDA(ctx, cudaEventRecord(ctx->ev2, st));
    ZKB_CUDA(ctx, cudaMemcpyAsync(P + key->z_h(), d_H_plain, key->m * sizeof(Fr), cudaMemcpyDeviceToDevice, st));
    int rc = msm_main_lane_and_assemble(ctx, key, 1, d_out, (zkb_partial *)d_out);
    if (rc) {
        cudaDeviceSynchronize();
        return rc;
    }
    ZKB_CUDA(ctx, cudaEventRecord(ctx->ev3, st));
    if (ensure_pinned(ctx, sizeof(zkb_partial) + 4)) return ZKB_ERR_CUDA;
    uint32_t *h_flags = (uint32_t *)((char *)ctx->pinned + sizeof(zkb_partial));
    ZKB_CUDA(ctx, cudaMemcpyAsync(ctx->pinned, d_out, sizeof(zkb_partial), cudaMemcpyDeviceToHost, st));
    ZKB_CUDA(ctx, cudaMemcpyAsync(h_flags, flags, 4, cudaMemcpyDeviceToHost, st));
    ZKB_CUDA(ctx, cudaStreamSynchronize(st));
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev2, ctx->ev3);
    ctx->last_ms += ms;
    prof_collect(ctx);
    memcpy(out, ctx->pinned, sizeof(zkb_partial));
    if (*h_flags) {
        memset(out, 0, sizeof(zkb_partial));
        return fail(ctx, ZKB_ERR_UNSAT, "the witness does not satisfy the constraint system");
    }
    return ZKB_OK;
}

int zkb_prove_batch(zkb_ctx *ctx, const zkb_key *key, size_t count, const uint64_t *w, const uint64_t *r, const uint64_t *s,
                    zkb_proof *out)
{
    return prove_impl(ctx, key, count, w, r, s, false, out, nullptr, false);
}

int zkb_prove_batch_status(const zkb_ctx *ctx, int32_t *status, size_t n)
{
    if (!ctx || (!status && n)) return ZKB_ERR_ARG;
    for (size_t i = 0; i < n; i++) status[i] = i < ctx->batch_status.size() ? ctx->batch_status[i] : 0;
    return ZKB_OK;
}

int zkb_prove_batch_submit(zkb_ctx *ctx, const zkb_key *key, size_t count, const uint64_t *w, const uint64_t *r, const uint64_t *s)
{
    if (!ctx || !key || !w || !r || !s || !count) return fail(ctx, ZKB_ERR_ARG, "zkb_prove_batch_submit: bad argument");
    if (key->dev != ctx->dev) return fail(ctx, ZKB_ERR_ARG, "zkb_prove_batch_submit: key lives on another device");
    if (ctx->in_pending >= 2) return fail(ctx, ZKB_ERR_ARG, "zkb_prove_batch_submit: two batches are already in flight, collect one first");
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    const int slot = (ctx->in_head + ctx->in_pending) & 1;
    const size_t wb = count * key->nv * sizeof(Fr), sb = count * sizeof(Fr);
    ZKB_CUDA(ctx, ctx->in_w[slot].ensure(wb ? wb : 32));
    ZKB_CUDA(ctx, ctx->in_rs[slot].ensure(2 * sb));
    if (wb) ZKB_CUDA(ctx, cudaMemcpyAsync(ctx->in_w[slot].p, w, wb, cudaMemcpyHostToDevice, ctx->copy_stream));
    ZKB_CUDA(ctx, cudaMemcpyAsync(ctx->in_rs[slot].p, r, sb, cudaMemcpyHostToDevice, ctx->copy_stream));
    ZKB_CUDA(ctx, cudaMemcpyAsync((char *)ctx->in_rs[slot].p + sb, s, sb, cudaMemcpyHostToDevice, ctx->copy_stream));
    ZKB_CUDA(ctx, cudaEventRecord(ctx->ev_in[slot], ctx->copy_stream));
    ctx->in_key[slot] = key;
    ctx->in_count[slot] = count;
    ctx->in_pending++;
    return ZKB_OK;
}

int zkb_prove_batch_collect(zkb_ctx *ctx, zkb_proof *out)
{
    if (!ctx || !out) return fail(ctx, ZKB_ERR_ARG, "zkb_prove_batch_collect: bad argument");
    if (!ctx->in_pending) return fail(ctx, ZKB_ERR_ARG, "zkb_prove_batch_collect: nothing submitted");
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    const int slot = ctx->in_head;
    const size_t count = ctx->in_count[slot];
    ctx->in_head ^= 1; // the slot is released whatever happens below: prove_impl returns with the device idle
    ctx->in_pending--;
    ZKB_CUDA(ctx, cudaStreamWaitEvent(ctx->main_stream, ctx->ev_in[slot], 0));
    const uint64_t *dw = (const uint64_t *)ctx->in_w[slot].p, *dr = (const uint64_t *)ctx->in_rs[slot].p;
    return prove_impl(ctx, ctx->in_key[slot], count, dw, dr, dr + count * 4, true, out, nullptr, false);
}

int zkb_prove_batch_device(zkb_ctx *ctx, const zkb_key *key, size_t count, const void *d_w, const void *d_r, const void *d_s,
                           zkb_proof *out)
{
    return prove_impl(ctx, key, count, (const uint64_t *)d_w, (const uint64_t *)d_r, (const uint64_t *)d_s, true, out, nullptr,
                      false);
}

} // extern "C"
