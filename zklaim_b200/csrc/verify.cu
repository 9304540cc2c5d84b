// verify.cu -- batch Groth16 verification on the GPU.
//
// Replaces, for a batch of proofs under one verification key, the reference's
//   r1cs_gg_ppzksnark_verifier_strong_IC / _online_verifier_weak_IC
//       (libsnark/zk_proof_systems/ppzksnark/r1cs_gg_ppzksnark/r1cs_gg_ppzksnark.tcc:523-622)
//   accumulation_vector::accumulate_chunk (libsnark/common/data_structures/accumulation_vector.tcc:75-118)
//   alt_bn128_pp::miller_loop / double_miller_loop / final_exponentiation (alt_bn128_pairing.cpp:109-500)
// reached from libsnark_verify (zklaim/libsnark_wrapper.cpp:252-276).  Per proof the same decision as the
// reference, from the same values: QAP = FE(ML(A, B) * conj(ML(acc, gamma) * ML(C, delta))) compared with
// vk.alpha_g1_beta_g2 limb for limb, AND the three proof points lying on their curves.
//
// Parallel shape: the batch is the parallel axis.  Every Miller loop has the same control flow (the loop
// count is a constant), so the 3 n Miller loops of n proofs run in lock step, and the n final exponentiations
// likewise: one lane each for large batches, 6 or 18 cooperating lanes each (pairing_coop.cuh) while the batch is
// too small to fill the device.  The public-input sums input_j * gamma_ABC[j + 1] are one lane per
// (proof, input, 32-bit chunk) over a per-key table.  Latency-bound (dependent 254-bit products): built with the
// out-of-line multiplier.
#define ZKB_NOINLINE_MUL 1
#include "common.cuh"
#include "pairing_coop.cuh"

#include <mutex>
#include <stdlib.h>

using namespace zkb;

namespace {
enum { SV_CONST = 58, SV_IN, SV_T, SV_F, SV_OUT, SV_TAB }; // scratch slots of this file (prover.cu ends at 57)

// affine limbs at the ABI: all-zero = infinity (zkb.h conventions)
template <class F> __device__ void affine_in(const Affine<F> &a, F &x, F &y)
{
    if (a.is_inf()) {
        x = F::zero();
        y = F::one();
    } else {
        x = a.x;
        y = a.y;
    }
}

// Public-input sums with a per-key table: input_j * P_j = sum_c chunk_c * (2^(32 c) P_j), eight 32-bit chunks.
// The 8 multiples of every gamma_ABC[j + 1] are made once per verification key (k_vf_tab; zkb_verify_batch keeps them
// while the key stays the same), after which a lane's chain is 32 doublings instead of 254.
constexpr uint32_t VF_CHUNKS = 8;
__global__ void k_vf_tab(const G1Affine *gamma_abc, G1Affine *tab, uint32_t ni)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ni * VF_CHUNKS) return;
    const uint32_t j = t / VF_CHUNKS, c = t % VF_CHUNKS;
    G1X p = G1X::from_affine(ld_vec(gamma_abc + j + 1));
    for (uint32_t k = 0; k < 32 * c; k++) p = p.dbl();
    G1Affine a;
    if (p.is_inf()) {
        a.x = Fq::zero();
        a.y = Fq::zero();
    } else {
        jac_to_affine<Fq>(p.to_jacobian(), a.x, a.y);
    }
    st_vec(tab + t, a);
}
// lane t = (proof i, input j, chunk c): T[t] = chunk_c(input_ij) * tab[j][c]
__global__ void k_vf_inputs_tab(const Fr *primary, const G1Affine *tab, G1X *T, uint32_t n, uint32_t ni)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n * ni * VF_CHUNKS) return;
    const uint32_t c = t % VF_CHUNKS, ij = t / VF_CHUNKS, j = ij % ni;
    const Fr k = ld_vec(primary + ij).from_mont();
    uint32_t kc[8] = {k.v[c], 0, 0, 0, 0, 0, 0, 0};
    st_vec(T + t, scalar_mul<Fq>(kc, G1X::from_affine(ld_vec(tab + j * VF_CHUNKS + c))));
}

struct ProofJ { // zkb_proof_out layout
    G1J A;
    G2J B;
    G1J C;
};

// lane t = which * n + i: Miller loop `which` of proof i (0: (A, B), 1: (acc, gamma), 2: (C, delta)); also the
// well-formedness bits of the points this lane touches
__global__ void k_vf_miller(const ProofJ *proofs, const G1X *T, const G1Affine *gamma_abc, const G2Affine *gamma_g2,
                            const G2Affine *delta_g2, const PairingConsts *Kp, Fq12 *F, uint32_t *wf, uint32_t n,
                            uint32_t ni /* terms of acc per proof */)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= 3 * n) return;
    const uint32_t which = t / n, i = t % n;
    const PairingConsts &K = *Kp;
    Fq px, py;
    Fq2 qx, qy;
    bool ok = true;
    if (which == 0) {
        const G1J A = ld_vec(&proofs[i].A);
        const G2J B = ld_vec(&proofs[i].B);
        Fq three = Fq::one();
        three = three + three.dbl();
        ok = jacobian_well_formed<Fq>(A.X, A.Y, A.Z, three) && jacobian_well_formed<Fq2>(B.X, B.Y, B.Z, K.twist_b);
        jac_to_affine<Fq>(A, px, py);
        jac_to_affine<Fq2>(B, qx, qy);
    } else if (which == 1) {
        G1X acc = G1X::from_affine(ld_vec(gamma_abc));
        for (uint32_t j = 0; j < ni; j++) acc.add(ld_vec(T + (size_t)i * ni + j));
        jac_to_affine<Fq>(acc.to_jacobian(), px, py);
        affine_in<Fq2>(ld_vec(gamma_g2), qx, qy);
    } else {
        const G1J C = ld_vec(&proofs[i].C);
        Fq three = Fq::one();
        three = three + three.dbl();
        ok = jacobian_well_formed<Fq>(C.X, C.Y, C.Z, three);
        jac_to_affine<Fq>(C, px, py);
        affine_in<Fq2>(ld_vec(delta_g2), qx, qy);
    }
    wf[t] = ok ? 1u : 0u;
    st_vec(F + t, miller_loop(px, py, qx, qy, K));
}

// lane i: the decision for proof i
__global__ void k_vf_final(const Fq12 *F, const uint32_t *wf, const Fq12 *alpha_beta, const PairingConsts *Kp, int32_t *accept, uint32_t n)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const bool eq = groth16_pairing_check(ld_vec(F + i), ld_vec(F + n + i), ld_vec(F + 2 * n + i), ld_vec(alpha_beta), *Kp);
    const bool ok = wf[i] && wf[2 * n + i] && eq;
    accept[i] = ok ? 1 : 0;
}

// lane i: alt_bn128_ate_miller_loop / alt_bn128_reduced_pairing of one affine pair
__global__ void k_pairing(const G1Affine *P, const G2Affine *Q, const PairingConsts *Kp, int reduced, Fq12 *out, uint32_t n)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fq px, py;
    Fq2 qx, qy;
    affine_in<Fq>(ld_vec(P + i), px, py);
    affine_in<Fq2>(ld_vec(Q + i), qx, qy);
    Fq12 f = miller_loop(px, py, qx, qy, *Kp);
    if (reduced) f = final_exponentiation(f, *Kp);
    st_vec(out + i, f);
}

// ---- NL = 6 or 18 lanes per Miller loop / final exponentiation (pairing_coop.cuh) ---------------------------------

// group t = which * n + i as in k_vf_miller; lanes 0 and 1 prepare the affine G1 / G2 inputs, then the group runs
// the loop together
template <int NL> __global__ void __launch_bounds__(32) k_vf_miller_coop(const ProofJ *proofs, const G1X *T, const G1Affine *gamma_abc,
                                                       const G2Affine *gamma_g2, const G2Affine *delta_g2, const PairingConsts *Kp,
                                                       Fq12 *F, uint32_t *wf, uint32_t n, uint32_t nt /* terms of acc per proof */)
{
    constexpr int COOP_GROUP = coop_group(NL), GROUPS = 32 / COOP_GROUP, COOP_LANES = 6;
    __shared__ Fq2 S[GROUPS][COOP_MILLER_REGS];
    __shared__ G1X part[GROUPS][COOP_LANES];
    __shared__ Fq2 Qin[GROUPS][2];
    __shared__ Fq Pin[GROUPS][2];
    __shared__ uint32_t ok_bits[GROUPS][2];
    const int grp = threadIdx.x / COOP_GROUP, lane = threadIdx.x % COOP_GROUP;
    uint32_t t = blockIdx.x * GROUPS + grp;
    const bool live = t < 3 * n;
    if (!live) t = 3 * n - 1; // a whole warp walks the same phases: spare groups redo the last unit and drop the result
    const uint32_t which = t / n, i = t % n;
    const PairingConsts &K = *Kp;
    if (which == 1 && lane < COOP_LANES) { // the nt terms of acc, strided over the lanes
        G1X s = G1X::inf();
        for (uint32_t k = lane; k < nt; k += COOP_LANES) s.add(ld_vec(T + (size_t)i * nt + k));
        part[grp][lane] = s;
    }
    __syncwarp();
    if (lane == 0) { // the G1 side
        Fq three = Fq::one();
        three = three + three.dbl();
        bool ok = true;
        if (which == 1) {
            G1X acc = G1X::from_affine(ld_vec(gamma_abc));
            for (int k = 0; k < COOP_LANES; k++) acc.add(part[grp][k]);
            jac_to_affine<Fq>(acc.to_jacobian(), Pin[grp][0], Pin[grp][1]);
        } else {
            const G1J A = ld_vec(which == 0 ? &proofs[i].A : &proofs[i].C);
            ok = jacobian_well_formed<Fq>(A.X, A.Y, A.Z, three);
            jac_to_affine<Fq>(A, Pin[grp][0], Pin[grp][1]);
        }
        ok_bits[grp][0] = ok;
    }
    if (lane == 1) { // the G2 side
        bool ok = true;
        if (which == 0) {
            const G2J B = ld_vec(&proofs[i].B);
            ok = jacobian_well_formed<Fq2>(B.X, B.Y, B.Z, K.twist_b);
            jac_to_affine<Fq2>(B, Qin[grp][0], Qin[grp][1]);
        } else {
            affine_in<Fq2>(ld_vec(which == 1 ? gamma_g2 : delta_g2), Qin[grp][0], Qin[grp][1]);
        }
        ok_bits[grp][1] = ok;
    }
    __syncwarp();
    const Fq px = Pin[grp][0], py = Pin[grp][1];
    const Fq2 qx = Qin[grp][0], qy = Qin[grp][1];
    coop_miller_loop<NL>(S[grp], px, py, qx, qy, K);
    if (live && lane < COOP_LANES) st_vec((Fq2 *)(F + t) + lane, S[grp][lane]);
    if (live && lane == 0) wf[t] = ok_bits[grp][0] & ok_bits[grp][1];
}

// group i: the decision for proof i
template <int NL> __global__ void __launch_bounds__(32) k_vf_final_coop(const Fq12 *F, const uint32_t *wf, const Fq12 *alpha_beta, const PairingConsts *Kp,
                                                      int32_t *accept, uint32_t n)
{
    constexpr int COOP_GROUP = coop_group(NL), GROUPS = 32 / COOP_GROUP, COOP_LANES = 6;
    __shared__ Fq2 S[GROUPS][COOP_FE_REGS];
    const int grp = threadIdx.x / COOP_GROUP, lane = threadIdx.x % COOP_GROUP;
    uint32_t i = blockIdx.x * GROUPS + grp;
    const bool live = i < n;
    if (!live) i = n - 1;
    Fq2 *s = S[grp];
    // s[0..5] = ML(A, B) * conj(ML(acc, gamma) * ML(C, delta))
    if (lane < COOP_LANES) {
        s[6 + lane] = ld_vec((const Fq2 *)(F + n + i) + lane);
        s[12 + lane] = ld_vec((const Fq2 *)(F + 2 * n + i) + lane);
        s[lane] = ld_vec((const Fq2 *)(F + i) + lane);
    }
    __syncwarp();
    coop_mul12<NL>(s + 6, s + 6, s + 12, s + 60);
    coop_conj12<NL>(s + 6, s + 6);
    coop_mul12<NL>(s, s, s + 6, s + 60);
    coop_final_exponentiation<NL>(s, *Kp);
    if (live && lane == 0) {
        bool eq = true;
        const Fq2 *ab = (const Fq2 *)alpha_beta;
        for (int k = 0; k < 6; k++) eq = eq && s[k] == ld_vec(ab + k);
        accept[i] = (wf[i] && wf[2 * n + i] && eq) ? 1 : 0;
    }
}

// group i: alt_bn128_ate_miller_loop / alt_bn128_reduced_pairing of one affine pair
template <int NL> __global__ void __launch_bounds__(32) k_pairing_coop(const G1Affine *P, const G2Affine *Q, const PairingConsts *Kp, int reduced, Fq12 *out,
                                                     uint32_t n)
{
    constexpr int COOP_GROUP = coop_group(NL), GROUPS = 32 / COOP_GROUP, COOP_LANES = 6;
    __shared__ Fq2 S[GROUPS][COOP_FE_REGS];
    const int grp = threadIdx.x / COOP_GROUP, lane = threadIdx.x % COOP_GROUP;
    uint32_t i = blockIdx.x * GROUPS + grp;
    const bool live = i < n;
    if (!live) i = n - 1;
    Fq px, py;
    Fq2 qx, qy;
    affine_in<Fq>(ld_vec(P + i), px, py);
    affine_in<Fq2>(ld_vec(Q + i), qx, qy);
    coop_miller_loop<NL>(S[grp], px, py, qx, qy, *Kp);
    if (reduced) coop_final_exponentiation<NL>(S[grp], *Kp);
    if (live && lane < COOP_LANES) st_vec((Fq2 *)(out + i) + lane, S[grp][lane]);
}

const PairingConsts &host_consts()
{
    static PairingConsts K;
    static std::once_flag once;
    std::call_once(once, [] { pairing_consts_init(K); });
    return K;
}

int consts_on_device(zkb_ctx *ctx, const PairingConsts **out)
{
    DevBuf &b = ctx->scratch[SV_CONST];
    if (!b.p) {
        ZKB_CUDA(ctx, b.ensure(sizeof(PairingConsts)));
        // the source is a process-lifetime static: a stream-ordered copy from it is safe
        ZKB_CUDA(ctx, cudaMemcpyAsync(b.p, &host_consts(), sizeof(PairingConsts), cudaMemcpyHostToDevice, ctx->stream));
    }
    *out = b.as<PairingConsts>();
    return ZKB_OK;
}

constexpr int LANES = 32; // one warp per block: the lanes of a warp run identical instruction streams
// Lanes per pairing by batch size: 18 while a warp per pairing still leaves the device under-filled, 6 up to the size
// where one lane per pairing has filled it anyway.  ZKB_VERIFY_COOP = 0 / 6 / 18 forces a form (measurement switch).
int coop_lanes(size_t n)
{
    static const int forced = [] {
        const char *e = getenv("ZKB_VERIFY_COOP");
        return e ? atoi(e) : -1;
    }();
    if (forced == 0) return 1;
    if (forced == 6 || forced == 18) return forced;
    return n <= 128 ? 18 : (n <= 1024 ? 6 : 1);
}
} // namespace

extern "C" {

int zkb_pairing(zkb_ctx *ctx, const uint64_t *g1_affine, const uint64_t *g2_affine, size_t n, int reduced, uint64_t *out)
{
    if (!ctx || !g1_affine || !g2_affine || !out) return fail(ctx, ZKB_ERR_ARG, "zkb_pairing: bad argument");
    if (n == 0) return ZKB_OK;
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    const PairingConsts *K;
    int rc = consts_on_device(ctx, &K);
    if (rc) return rc;
    const size_t in_bytes = n * (sizeof(G1Affine) + sizeof(G2Affine));
    ZKB_CUDA(ctx, ctx->scratch[SV_IN].ensure(in_bytes));
    ZKB_CUDA(ctx, ctx->scratch[SV_F].ensure(n * sizeof(Fq12)));
    if (ensure_pinned(ctx, n * sizeof(Fq12))) return ZKB_ERR_CUDA;
    G2Affine *dQ = ctx->scratch[SV_IN].as<G2Affine>();
    G1Affine *dP = (G1Affine *)(dQ + n);
    ZKB_CUDA(ctx, cudaMemcpyAsync(dQ, g2_affine, n * sizeof(G2Affine), cudaMemcpyHostToDevice, ctx->stream));
    ZKB_CUDA(ctx, cudaMemcpyAsync(dP, g1_affine, n * sizeof(G1Affine), cudaMemcpyHostToDevice, ctx->stream));
    ZKB_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    const int lanes = coop_lanes(n);
    if (lanes == 18)
        k_pairing_coop<18><<<(unsigned)n, 32, 0, ctx->stream>>>(dP, dQ, K, reduced, ctx->scratch[SV_F].as<Fq12>(), (uint32_t)n);
    else if (lanes == 6)
        k_pairing_coop<6><<<ceil_div(n, 4), 32, 0, ctx->stream>>>(dP, dQ, K, reduced, ctx->scratch[SV_F].as<Fq12>(), (uint32_t)n);
    else
        k_pairing<<<ceil_div(n, LANES), LANES, 0, ctx->stream>>>(dP, dQ, K, reduced, ctx->scratch[SV_F].as<Fq12>(), (uint32_t)n);
    ZKB_LAUNCH_CHECK(ctx);
    ZKB_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    ZKB_CUDA(ctx, cudaMemcpyAsync(ctx->pinned, ctx->scratch[SV_F].p, n * sizeof(Fq12), cudaMemcpyDeviceToHost, ctx->stream));
    ZKB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->last_ms = ms;
    memcpy(out, ctx->pinned, n * sizeof(Fq12));
    return ZKB_OK;
}

int zkb_verify_batch(zkb_ctx *ctx, const zkb_vk_desc *vk, const uint64_t *primary, const uint64_t *proofs, size_t n, int32_t *accept)
{
    if (!ctx || !vk || !proofs || !accept || !vk->alpha_g1_beta_g2 || !vk->gamma_g2 || !vk->delta_g2 || !vk->gamma_ABC_g1 ||
        (!primary && vk->num_inputs))
        return fail(ctx, ZKB_ERR_ARG, "zkb_verify_batch: bad argument");
    if (n == 0) return ZKB_OK;
    if (n > (1u << 24) || vk->num_inputs > (1u << 24)) return fail(ctx, ZKB_ERR_ARG, "zkb_verify_batch: batch too large");
    // the input-sum kernels index their lanes with 32 bits
    if ((uint64_t)n * (vk->num_inputs ? vk->num_inputs : 1) * VF_CHUNKS >= ((uint64_t)1 << 32) || 3 * (uint64_t)n >= ((uint64_t)1 << 31))
        return fail(ctx, ZKB_ERR_ARG, "zkb_verify_batch: proofs x inputs exceeds 2^32 / 8 lanes: verify in smaller batches");
    ZKB_CUDA(ctx, cudaSetDevice(ctx->dev));
    const PairingConsts *K;
    int rc = consts_on_device(ctx, &K);
    if (rc) return rc;
    const uint32_t ni = (uint32_t)vk->num_inputs, nn = (uint32_t)n;
    // one staging buffer: alpha_beta | gamma_g2 | delta_g2 | gamma_ABC | proofs | primary   (all multiples of 16 bytes)
    const size_t o_ab = 0, o_g = o_ab + sizeof(Fq12), o_d = o_g + sizeof(G2Affine), o_abc = o_d + sizeof(G2Affine);
    const size_t o_pr = o_abc + (size_t)(ni + 1) * sizeof(G1Affine), o_in = o_pr + n * sizeof(ProofJ);
    const size_t total = o_in + n * (size_t)ni * sizeof(Fr);
    static_assert(sizeof(ProofJ) == 48 * 8, "zkb_proof_out layout");
    ZKB_CUDA(ctx, ctx->scratch[SV_IN].ensure(total));
    ZKB_CUDA(ctx, ctx->scratch[SV_T].ensure((size_t)nn * (ni ? ni : 1) * VF_CHUNKS * sizeof(G1X)));
    ZKB_CUDA(ctx, ctx->scratch[SV_F].ensure(3 * n * sizeof(Fq12)));
    ZKB_CUDA(ctx, ctx->scratch[SV_OUT].ensure(4 * n * sizeof(uint32_t)));
    if (ensure_pinned(ctx, n * sizeof(int32_t))) return ZKB_ERR_CUDA;
    char *base = ctx->scratch[SV_IN].as<char>();
    cudaStream_t st = ctx->stream;
    ZKB_CUDA(ctx, cudaMemcpyAsync(base + o_ab, vk->alpha_g1_beta_g2, sizeof(Fq12), cudaMemcpyHostToDevice, st));
    ZKB_CUDA(ctx, cudaMemcpyAsync(base + o_g, vk->gamma_g2, sizeof(G2Affine), cudaMemcpyHostToDevice, st));
    ZKB_CUDA(ctx, cudaMemcpyAsync(base + o_d, vk->delta_g2, sizeof(G2Affine), cudaMemcpyHostToDevice, st));
    ZKB_CUDA(ctx, cudaMemcpyAsync(base + o_abc, vk->gamma_ABC_g1, (size_t)(ni + 1) * sizeof(G1Affine), cudaMemcpyHostToDevice, st));
    ZKB_CUDA(ctx, cudaMemcpyAsync(base + o_pr, proofs, n * sizeof(ProofJ), cudaMemcpyHostToDevice, st));
    if (ni) ZKB_CUDA(ctx, cudaMemcpyAsync(base + o_in, primary, n * (size_t)ni * sizeof(Fr), cudaMemcpyHostToDevice, st));
    ZKB_CUDA(ctx, cudaEventRecord(ctx->ev0, st));
    const G1Affine *d_abc = (const G1Affine *)(base + o_abc);
    G1X *dT = ctx->scratch[SV_T].as<G1X>();
    Fq12 *dF = ctx->scratch[SV_F].as<Fq12>();
    uint32_t *d_wf = ctx->scratch[SV_OUT].as<uint32_t>();
    int32_t *d_acc = (int32_t *)(d_wf + 3 * n);
    const int lanes = coop_lanes(n);
    uint32_t nt = ni; // terms of acc per proof
    if (ni) {
        // window table of the key: rebuilt only when gamma_ABC_g1 changes
        const size_t abc_words = (size_t)(ni + 1) * 8;
        const bool same = ctx->vf_abc_host.size() == abc_words && !memcmp(ctx->vf_abc_host.data(), vk->gamma_ABC_g1, abc_words * 8);
        if (!same) ctx->vf_abc_host.clear(); // before the buffer may move: a failed call must not leave a stale "same key" behind
        ZKB_CUDA(ctx, ctx->scratch[SV_TAB].ensure((size_t)ni * VF_CHUNKS * sizeof(G1Affine)));
        G1Affine *d_tab = ctx->scratch[SV_TAB].as<G1Affine>();
        if (!same) {
            ctx->vf_abc_host.clear();
            k_vf_tab<<<ceil_div((size_t)ni * VF_CHUNKS, LANES), LANES, 0, st>>>(d_abc, d_tab, ni);
            ZKB_LAUNCH_CHECK(ctx);
        }
        nt = ni * VF_CHUNKS;
        k_vf_inputs_tab<<<ceil_div((size_t)nn * nt, LANES), LANES, 0, st>>>((const Fr *)(base + o_in), d_tab, dT, nn, ni);
        ZKB_LAUNCH_CHECK(ctx);
        if (!same) ctx->vf_abc_host.assign(vk->gamma_ABC_g1, vk->gamma_ABC_g1 + abc_words);
    }
    // small batches: 18 or 6 lanes per Miller loop / final exponentiation (the device is empty, latency is what counts);
    // large batches: one lane each (fewer instructions in total once every warp slot is busy)
    if (lanes == 18) {
        k_vf_miller_coop<18><<<3 * nn, 32, 0, st>>>((const ProofJ *)(base + o_pr), dT, d_abc, (const G2Affine *)(base + o_g),
                                                   (const G2Affine *)(base + o_d), K, dF, d_wf, nn, nt);
        ZKB_LAUNCH_CHECK(ctx);
        k_vf_final_coop<18><<<nn, 32, 0, st>>>(dF, d_wf, (const Fq12 *)(base + o_ab), K, d_acc, nn);
        ZKB_LAUNCH_CHECK(ctx);
    } else if (lanes == 6) {
        k_vf_miller_coop<6><<<ceil_div(3 * n, 4), 32, 0, st>>>((const ProofJ *)(base + o_pr), dT, d_abc, (const G2Affine *)(base + o_g),
                                                             (const G2Affine *)(base + o_d), K, dF, d_wf, nn, nt);
        ZKB_LAUNCH_CHECK(ctx);
        k_vf_final_coop<6><<<ceil_div(n, 4), 32, 0, st>>>(dF, d_wf, (const Fq12 *)(base + o_ab), K, d_acc, nn);
        ZKB_LAUNCH_CHECK(ctx);
    } else {
        k_vf_miller<<<ceil_div(3 * n, LANES), LANES, 0, st>>>((const ProofJ *)(base + o_pr), dT, d_abc, (const G2Affine *)(base + o_g),
                                                            (const G2Affine *)(base + o_d), K, dF, d_wf, nn, nt);
        ZKB_LAUNCH_CHECK(ctx);
        k_vf_final<<<ceil_div(n, LANES), LANES, 0, st>>>(dF, d_wf, (const Fq12 *)(base + o_ab), K, d_acc, nn);
        ZKB_LAUNCH_CHECK(ctx);
    }
    ZKB_CUDA(ctx, cudaEventRecord(ctx->ev1, st));
    ZKB_CUDA(ctx, cudaMemcpyAsync(ctx->pinned, d_acc, n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    ZKB_CUDA(ctx, cudaStreamSynchronize(st));
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->last_ms = ms;
    memcpy(accept, ctx->pinned, n * sizeof(int32_t));
    // untrusted input: the field arithmetic assumes canonical residues (< q for coordinates, < r for inputs).  An encoding
    // outside that range is not a proof the reference could have serialised: rejected here, whatever the kernels made of it.
    {
        static const uint64_t Q[4] = {0x3c208c16d87cfd47ull, 0x97816a916871ca8dull, 0xb85045b68181585dull, 0x30644e72e131a029ull};
        static const uint64_t R[4] = {0x43e1f593f0000001ull, 0x2833e84879b97091ull, 0xb85045b68181585dull, 0x30644e72e131a029ull};
        auto below = [](const uint64_t *x, const uint64_t *m) {
            for (int i = 3; i >= 0; i--)
                if (x[i] != m[i]) return x[i] < m[i];
            return false;
        };
        for (size_t i = 0; i < n; i++) {
            bool ok = true;
            for (int k = 0; k < 12 && ok; k++) ok = below(proofs + i * 48 + 4 * k, Q);
            for (size_t k = 0; k < vk->num_inputs && ok; k++) ok = below(primary + (i * vk->num_inputs + k) * 4, R);
            if (!ok) accept[i] = 0;
        }
    }
    return ZKB_OK;
}

} // extern "C"
