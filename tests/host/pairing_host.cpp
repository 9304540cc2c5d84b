// CPU build of zklaim_b200/csrc/pairing.cuh's portable path: lets the (GPU-less) test-suite check the
// tower arithmetic, the Miller loop and the final exponentiation against libff (oracle/_ref) before
// the same source runs as device code.
#include "../../zklaim_b200/csrc/pairing_coop.cuh"
#include <cstring>
using namespace zkb;

static const PairingConsts &consts()
{
    static PairingConsts K;
    static bool done = false;
    if (!done) {
        pairing_consts_init(K);
        done = true;
    }
    return K;
}
// p: affine G1 (16 words), q: affine G2 (32 words), out: Fq12 (96 words, libff's member order)
extern "C" void hostpair_pairing(const uint32_t *p, const uint32_t *q, int reduced, uint32_t *out)
{
    Fq px, py;
    Fq2 qx, qy;
    memcpy(&px, p, 32);
    memcpy(&py, p + 8, 32);
    memcpy(&qx, q, 64);
    memcpy(&qy, q + 16, 64);
    Fq12 f = miller_loop(px, py, qx, qy, consts());
    if (reduced) f = final_exponentiation(f, consts());
    memcpy(out, &f, sizeof f);
}
extern "C" void hostpair_fq12_op(int op, const uint32_t *a, const uint32_t *b, uint32_t *out)
{
    Fq12 x, y, z;
    memcpy(&x, a, sizeof x);
    if (b) memcpy(&y, b, sizeof y);
    switch (op) {
    case 0: z = mul12(x, y); break;
    case 1: z = sqr12(x); break;
    case 2: z = inv12(x); break;
    case 3: z = frobenius12(x, 1, consts()); break;
    case 4: z = frobenius12(x, 2, consts()); break;
    default: z = frobenius12(x, 3, consts()); break;
    }
    memcpy(out, &z, sizeof z);
}
extern "C" void hostpair_consts(uint32_t *out) { memcpy(out, &consts(), sizeof(PairingConsts)); }
extern "C" int hostpair_consts_size(void) { return (int)sizeof(PairingConsts); }

// The whole per-proof decision (r1cs_gg_ppzksnark_verifier_strong_IC, r1cs_gg_ppzksnark.tcc:523-622) composed on
// the host from the same pieces the kernels of verify.cu use.  Layouts as zkb_verify_batch.
extern "C" int hostpair_verify(const uint32_t *alpha_beta, const uint32_t *gamma_g2, const uint32_t *delta_g2, const uint32_t *gamma_abc,
                               uint32_t ni, const uint32_t *primary, const uint32_t *proof)
{
    const PairingConsts &K = consts();
    Fq12 ab;
    memcpy(&ab, alpha_beta, sizeof ab);
    G2Affine g, d;
    memcpy(&g, gamma_g2, sizeof g);
    memcpy(&d, delta_g2, sizeof d);
    G1J A, C;
    G2J B;
    memcpy(&A, proof, sizeof A);
    memcpy(&B, proof + 24, sizeof B);
    memcpy(&C, proof + 72, sizeof C);
    Fq three = Fq::one();
    three = three + three.dbl();
    const bool wf = jacobian_well_formed<Fq>(A.X, A.Y, A.Z, three) && jacobian_well_formed<Fq2>(B.X, B.Y, B.Z, K.twist_b) &&
                    jacobian_well_formed<Fq>(C.X, C.Y, C.Z, three);
    G1Affine first;
    memcpy(&first, gamma_abc, sizeof first);
    G1X acc = G1X::from_affine(first);
    for (uint32_t j = 0; j < ni; j++) {
        G1Affine b;
        Fr k;
        memcpy(&b, gamma_abc + 16 * (j + 1), sizeof b);
        memcpy(&k, primary + 8 * j, sizeof k);
        k = k.from_mont();
        acc.add(scalar_mul<Fq>(k.v, G1X::from_affine(b)));
    }
    Fq ax, ay, accx, accy, cx, cy;
    Fq2 bx, by;
    jac_to_affine<Fq>(A, ax, ay);
    jac_to_affine<Fq2>(B, bx, by);
    jac_to_affine<Fq>(acc.to_jacobian(), accx, accy);
    jac_to_affine<Fq>(C, cx, cy);
    const bool eq = groth16_pairing_check(miller_loop(ax, ay, bx, by, K), miller_loop(accx, accy, g.x, g.y, K),
                                          miller_loop(cx, cy, d.x, d.y, K), ab, K);
    return wf && eq;
}

// the cooperative forms (pairing_coop.cuh: 6 or 18 lanes per pairing), lanes emulated by a loop per phase
template <int NL> static void pairing_coop(const Fq &px, const Fq &py, const Fq2 &qx, const Fq2 &qy, int reduced, uint32_t *out)
{
    static Fq2 S[COOP_FE_REGS > COOP_MILLER_REGS ? COOP_FE_REGS : COOP_MILLER_REGS];
    coop_miller_loop<NL>(S, px, py, qx, qy, consts());
    if (reduced) coop_final_exponentiation<NL>(S, consts());
    memcpy(out, S, 6 * sizeof(Fq2));
}
extern "C" void hostpair_pairing_coop(const uint32_t *p, const uint32_t *q, int reduced, int lanes, uint32_t *out)
{
    Fq px, py;
    Fq2 qx, qy;
    memcpy(&px, p, 32);
    memcpy(&py, p + 8, 32);
    memcpy(&qx, q, 64);
    memcpy(&qy, q + 16, 64);
    if (lanes == 18)
        pairing_coop<18>(px, py, qx, qy, reduced, out);
    else
        pairing_coop<6>(px, py, qx, qy, reduced, out);
}
