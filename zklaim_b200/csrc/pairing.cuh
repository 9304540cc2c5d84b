// pairing.cuh -- the alt_bn128 optimal ate pairing for the verifier side of the path.
//
// Replaces (same values, bit for bit) libff's
//   Fp6_3over2_model   libff/algebra/fields/fp6_3over2.tcc:52-170     (Fq6  = Fq2[v]/(v^3 - xi), xi = 9 + u)
//   Fp12_2over3over2   libff/algebra/fields/fp12_2over3over2.tcc:52-360 (Fq12 = Fq6[w]/(w^2 - v))
//   alt_bn128_ate_precompute_G2 + alt_bn128_ate_miller_loop   alt_bn128_pairing.cpp:234-420
//   alt_bn128_final_exponentiation                            alt_bn128_pairing.cpp:109-240
// with the constants of alt_bn128_init.cpp:90-141, 268-272.  Every field result is canonical, so the
// unreduced Miller value and the reduced pairing are bit-identical to the reference's although the
// tower arithmetic below is organised differently:
//   - the Frobenius maps use ONE table gamma_i^k = xi^(k (q^i - 1) / 6), k = 0..5, derived at start-up
//     from xi alone (the reference keeps three tables of decimal literals);
//   - the G2 line coefficients are produced on the fly inside the Miller loop (no precomputed vector
//     of 102 coefficient triples per G2 point) -- one thread owns one Miller loop;
//   - the sparse line product is two sparse Fq6 products + one dense one (Karatsuba over w).
// Plain C++ when compiled by the host compiler (tests/host/pairing_host.cpp checks it against libff
// without a GPU); device code under nvcc.
#pragma once
#include "ec.cuh"
#include "fp.cuh"

#if defined(__CUDACC__)
#define ZKB_NI __host__ __device__ __noinline__
#else
#define ZKB_NI inline
#endif

namespace zkb {

// xi = 9 + u, the non-residue of both extensions (alt_bn128_init.cpp:102, 118)
ZKB_HD Fq2 mul_xi(const Fq2 &a)
{
    // (a0 + a1 u)(9 + u) = (9 a0 - a1) + (9 a1 + a0) u
    const Fq2 a8 = a.dbl().dbl().dbl();
    return Fq2{a8.c0 + a.c0 - a.c1, a8.c1 + a.c1 + a.c0};
}
ZKB_HD Fq2 conj(const Fq2 &a) { return Fq2{a.c0, a.c1.neg()}; }
// The Fq2 products of this file expand their two or three independent base-field products in ONE basic block
// (Fq::mont_mul_inline), so the instruction scheduler interleaves the carry chains: the kernels built from them
// run one warp per scheduler and are bound by dependent-issue latency, not by throughput.
ZKB_NI Fq2 mul_fq(const Fq2 &a, const Fq &k) { return Fq2{Fq::mont_mul_inline(a.c0, k), Fq::mont_mul_inline(a.c1, k)}; }
ZKB_NI Fq2 mul2(const Fq2 &a, const Fq2 &b)
{
    const Fq s0 = a.c0 + a.c1, s1 = b.c0 + b.c1;
    const Fq aa = Fq::mont_mul_inline(a.c0, b.c0);
    const Fq bb = Fq::mont_mul_inline(a.c1, b.c1);
    const Fq s = Fq::mont_mul_inline(s0, s1);
    return Fq2{aa - bb, s - aa - bb};
}
ZKB_NI Fq2 sqr2(const Fq2 &a)
{
    const Fq s = a.c0 + a.c1, d = a.c0 - a.c1;
    const Fq ab = Fq::mont_mul_inline(a.c0, a.c1);
    const Fq t = Fq::mont_mul_inline(s, d);
    return Fq2{t, ab.dbl()};
}

struct Fq6 {
    Fq2 c0, c1, c2;
    static ZKB_HD Fq6 zero() { return Fq6{Fq2::zero(), Fq2::zero(), Fq2::zero()}; }
    static ZKB_HD Fq6 one() { return Fq6{Fq2::one(), Fq2::zero(), Fq2::zero()}; }
    ZKB_HD bool operator==(const Fq6 &b) const { return c0 == b.c0 && c1 == b.c1 && c2 == b.c2; }
    friend ZKB_HD Fq6 operator+(const Fq6 &a, const Fq6 &b) { return Fq6{a.c0 + b.c0, a.c1 + b.c1, a.c2 + b.c2}; }
    friend ZKB_HD Fq6 operator-(const Fq6 &a, const Fq6 &b) { return Fq6{a.c0 - b.c0, a.c1 - b.c1, a.c2 - b.c2}; }
    ZKB_HD Fq6 neg() const { return Fq6{c0.neg(), c1.neg(), c2.neg()}; }
    ZKB_HD Fq6 dbl() const { return Fq6{c0.dbl(), c1.dbl(), c2.dbl()}; }
    // times v: (c0, c1, c2) -> (xi c2, c0, c1)
    ZKB_HD Fq6 mul_v() const { return Fq6{mul_xi(c2), c0, c1}; }
};

// dense product, 6 Fq2 products (value of fp6_3over2.tcc:73-95)
ZKB_NI Fq6 mul6(const Fq6 &a, const Fq6 &b)
{
    const Fq2 v0 = mul2(a.c0, b.c0), v1 = mul2(a.c1, b.c1), v2 = mul2(a.c2, b.c2);
    const Fq2 t12 = mul2(a.c1 + a.c2, b.c1 + b.c2) - v1 - v2;
    const Fq2 t01 = mul2(a.c0 + a.c1, b.c0 + b.c1) - v0 - v1;
    const Fq2 t02 = mul2(a.c0 + a.c2, b.c0 + b.c2) - v0 - v2;
    return Fq6{v0 + mul_xi(t12), t01 + mul_xi(v2), t02 + v1};
}
// a * (x + z v^2)
ZKB_NI Fq6 mul6_by_02(const Fq6 &a, const Fq2 &x, const Fq2 &z)
{
    const Fq2 v0 = mul2(a.c0, x), v2 = mul2(a.c2, z);
    const Fq2 t02 = mul2(a.c0 + a.c2, x + z) - v0 - v2; // a0 z + a2 x
    return Fq6{v0 + mul_xi(mul2(a.c1, z)), mul2(a.c1, x) + mul_xi(v2), t02};
}
// a * (y v)
ZKB_NI Fq6 mul6_by_1(const Fq6 &a, const Fq2 &y) { return Fq6{mul_xi(mul2(a.c2, y)), mul2(a.c0, y), mul2(a.c1, y)}; }
// (fp6_3over2.tcc:117-145: same value)
ZKB_NI Fq6 inv6(const Fq6 &a)
{
    const Fq2 t0 = sqr2(a.c0) - mul_xi(mul2(a.c1, a.c2));
    const Fq2 t1 = mul_xi(sqr2(a.c2)) - mul2(a.c0, a.c1);
    const Fq2 t2 = sqr2(a.c1) - mul2(a.c0, a.c2);
    const Fq2 d = (mul2(a.c0, t0) + mul_xi(mul2(a.c2, t1) + mul2(a.c1, t2))).inverse();
    return Fq6{mul2(t0, d), mul2(t1, d), mul2(t2, d)};
}

struct Fq12 {
    Fq6 c0, c1;
    static ZKB_HD Fq12 one() { return Fq12{Fq6::one(), Fq6::zero()}; }
    ZKB_HD bool operator==(const Fq12 &b) const { return c0 == b.c0 && c1 == b.c1; }
    // unitary_inverse (fp12_2over3over2.tcc:160-170): the inverse of an element of the cyclotomic subgroup
    ZKB_HD Fq12 conj() const { return Fq12{c0, c1.neg()}; }
};

ZKB_NI Fq12 mul12(const Fq12 &a, const Fq12 &b)
{
    const Fq6 t0 = mul6(a.c0, b.c0), t1 = mul6(a.c1, b.c1);
    const Fq6 t2 = mul6(a.c0 + a.c1, b.c0 + b.c1);
    return Fq12{t0 + t1.mul_v(), t2 - t0 - t1};
}
ZKB_NI Fq12 sqr12(const Fq12 &a)
{
    // (a0 + a1 w)^2 = (a0 + a1)(a0 + v a1) - ab - v ab  +  2 ab w,   ab = a0 a1
    const Fq6 ab = mul6(a.c0, a.c1);
    const Fq6 t = mul6(a.c0 + a.c1, a.c0 + a.c1.mul_v());
    return Fq12{t - ab - ab.mul_v(), ab.dbl()};
}
// a^2 for a in the cyclotomic subgroup (Granger-Scott; value of cyclotomic_squared, fp12_2over3over2.tcc:172-237):
// with z0..z5 = c0.c0, c1.c1, c1.c0, c0.c2, c0.c1, c1.c2 the pairs (z0, z1), (z2, z3), (z4, z5) are elements of
// Fq4 = Fq2[s]/(s^2 - xi); three Fq4 squarings (two Fq2 products each) instead of the twelve products of sqr12.
ZKB_HD void sqr4(const Fq2 &a, const Fq2 &b, Fq2 &even, Fq2 &odd)
{
    const Fq2 m = mul2(a, b);
    even = mul2(a + b, a + mul_xi(b)) - m - mul_xi(m);
    odd = m.dbl();
}
ZKB_NI Fq12 cyc_sqr12(const Fq12 &a)
{
    Fq2 t0, t1, t2, t3, t4, t5;
    sqr4(a.c0.c0, a.c1.c1, t0, t1);
    sqr4(a.c1.c0, a.c0.c2, t2, t3);
    sqr4(a.c0.c1, a.c1.c2, t4, t5);
    auto minus = [](const Fq2 &t, const Fq2 &z) { return t.dbl() + t - z.dbl(); }; // 3t - 2z
    auto plus = [](const Fq2 &t, const Fq2 &z) { return t.dbl() + t + z.dbl(); };  // 3t + 2z
    Fq12 r;
    r.c0.c0 = minus(t0, a.c0.c0);
    r.c1.c1 = plus(t1, a.c1.c1);
    r.c1.c0 = plus(mul_xi(t5), a.c1.c0);
    r.c0.c2 = minus(t4, a.c0.c2);
    r.c0.c1 = minus(t2, a.c0.c1);
    r.c1.c2 = plus(t3, a.c1.c2);
    return r;
}
ZKB_NI Fq12 inv12(const Fq12 &a)
{
    const Fq6 d = inv6(mul6(a.c0, a.c0) - mul6(a.c1, a.c1).mul_v());
    return Fq12{mul6(a.c0, d), mul6(a.c1, d).neg()};
}
// f * (e0 + eVV v^2 + eVW v w): the line value, value of mul_by_024 (fp12_2over3over2.tcc:239-300)
ZKB_NI Fq12 mul12_line(const Fq12 &f, const Fq2 &e0, const Fq2 &eVW, const Fq2 &eVV)
{
    const Fq6 t0 = mul6_by_02(f.c0, e0, eVV);
    const Fq6 t1 = mul6_by_1(f.c1, eVW);
    const Fq6 t2 = mul6(f.c0 + f.c1, Fq6{e0, eVW, eVV});
    return Fq12{t0 + t1.mul_v(), t2 - t0 - t1};
}

// ---------------------------------------------------------------------------------------
// constants derived from xi (host, once per context) and kept in device memory
// ---------------------------------------------------------------------------------------
struct PairingConsts {
    Fq2 g[4][6];  // g[i][k] = xi^(k (q^i - 1) / 6): Frobenius_coeffs of Fq12 (k = 1), Fq6 c1 (k = 2), Fq6 c2 (k = 4);
                  // g[1][2], g[1][3] = alt_bn128_twist_mul_by_q_X / _Y (alt_bn128_init.cpp:139-142)
    Fq2 twist_b;  // 3 / xi (alt_bn128_init.cpp:136)
    Fq two_inv;
};

// Frobenius_map(i), i = 1..3 (fp12_2over3over2.tcc:146-152 with fp6_3over2.tcc:107-114): in the basis
// 1, w, .., w^5 (c0 = (w^0, w^2, w^4), c1 = (w^1, w^3, w^5)) coefficient k becomes frob^i(z_k) * g[i][k].
ZKB_NI Fq12 frobenius12(const Fq12 &a, int i, const PairingConsts &K)
{
    const bool odd = i & 1;
    auto fr = [&](const Fq2 &z, int k) { return mul2(odd ? conj(z) : z, K.g[i][k]); };
    Fq12 r;
    r.c0.c0 = odd ? conj(a.c0.c0) : a.c0.c0;
    r.c0.c1 = fr(a.c0.c1, 2);
    r.c0.c2 = fr(a.c0.c2, 4);
    r.c1.c0 = fr(a.c1.c0, 1);
    r.c1.c1 = fr(a.c1.c1, 3);
    r.c1.c2 = fr(a.c1.c2, 5);
    return r;
}

// host only: g[1][1] = xi^((q - 1) / 6) by square-and-multiply, the rest by products and conjugation
// ((q^i - 1)/6 = (q - 1)/6 * (1 + q + .. + q^(i-1)) and x^q = conj(x) in Fq2).
inline void pairing_consts_init(PairingConsts &K)
{
    Fq nine = Fq::zero();
    const Fq one = Fq::one();
    for (int i = 0; i < 9; i++) nine = nine + one;
    const Fq2 xi{nine, one};
    uint32_t e[8];
    for (int i = 0; i < 8; i++) e[i] = FqP::P(i);
    e[0] -= 1;
    uint64_t rem = 0; // e = (q - 1) / 6
    for (int i = 7; i >= 0; i--) {
        const uint64_t cur = (rem << 32) | e[i];
        e[i] = (uint32_t)(cur / 6);
        rem = cur % 6;
    }
    Fq2 g1 = Fq2::one();
    for (int i = 255; i >= 0; i--) {
        g1 = g1.sqr();
        if ((e[i >> 5] >> (i & 31)) & 1) g1 = g1 * xi;
    }
    Fq2 gi = Fq2::one();
    for (int i = 0; i < 4; i++) {
        if (i) gi = g1 * conj(gi); // gamma_i = gamma_1 * gamma_(i-1)^q
        K.g[i][0] = Fq2::one();
        for (int k = 1; k < 6; k++) K.g[i][k] = K.g[i][k - 1] * gi;
    }
    const Fq three = one + one + one;
    K.twist_b = mul_fq(xi.inverse(), three);
    K.two_inv = (one + one).inverse();
}

// ---------------------------------------------------------------------------------------
// Miller loop (flipped ate, loop count 6z + 2 = 29793968203157093288, alt_bn128_init.cpp:268)
// ---------------------------------------------------------------------------------------
struct G2Proj {
    Fq2 X, Y, Z;
};
struct Line {
    Fq2 e0, eVW, eVV;
};

// tangent at R and R <- 2R, homogeneous projective coordinates on the twist (value of
// doubling_step_for_flipped_miller_loop, alt_bn128_pairing.cpp:247-272)
ZKB_NI Line line_double(G2Proj &R, const PairingConsts &K)
{
    const Fq2 A = mul_fq(mul2(R.X, R.Y), K.two_inv);
    const Fq2 B = sqr2(R.Y);
    const Fq2 C = sqr2(R.Z);
    const Fq2 E = mul2(K.twist_b, C.dbl() + C);
    const Fq2 F = E.dbl() + E;
    const Fq2 G = mul_fq(B + F, K.two_inv);
    const Fq2 H = sqr2(R.Y + R.Z) - (B + C);
    const Fq2 J = sqr2(R.X);
    const Fq2 E2 = sqr2(E);
    Line l;
    l.e0 = mul_xi(E - B);
    l.eVW = H.neg();
    l.eVV = J.dbl() + J;
    R.X = mul2(A, B - F);
    R.Y = sqr2(G) - (E2.dbl() + E2);
    R.Z = mul2(B, H);
    return l;
}
// chord through R and the affine point (x2, y2), R <- R + (x2, y2) (value of
// mixed_addition_step_for_flipped_miller_loop, alt_bn128_pairing.cpp:274-298)
ZKB_NI Line line_add(G2Proj &R, const Fq2 &x2, const Fq2 &y2)
{
    const Fq2 D = R.X - mul2(x2, R.Z);
    const Fq2 E = R.Y - mul2(y2, R.Z);
    const Fq2 F = sqr2(D);
    const Fq2 G = sqr2(E);
    const Fq2 H = mul2(D, F);
    const Fq2 I = mul2(R.X, F);
    const Fq2 J = H + mul2(R.Z, G) - I.dbl();
    Line l;
    l.e0 = mul_xi(mul2(E, x2) - mul2(D, y2));
    l.eVV = E.neg();
    l.eVW = D;
    R.X = mul2(D, J);
    R.Y = mul2(E, I - J) - mul2(H, R.Y);
    R.Z = mul2(R.Z, H);
    return l;
}

ZKB_HD Fq12 mul_line_at(const Fq12 &f, const Line &l, const Fq &px, const Fq &py)
{
    return mul12_line(f, l.e0, mul_fq(l.eVW, py), mul_fq(l.eVV, px));
}

// alt_bn128_ate_miller_loop(precompute_G1(P), precompute_G2(Q)) for affine P = (px, py), Q = (qx, qy)
// (alt_bn128_pairing.cpp:300-420).  Points at infinity arrive as the reference's to_affine_coordinates
// leaves them: (0, 1).
ZKB_NI Fq12 miller_loop(const Fq &px, const Fq &py, const Fq2 &qx, const Fq2 &qy, const PairingConsts &K)
{
    const uint64_t loop = 0x9d797039be763ba8ull; // bits 63..0 of 6z + 2; bit 64 (the leading one) is skipped
    G2Proj R{qx, qy, Fq2::one()};
    Fq12 f = Fq12::one();
    for (int i = 63; i >= 0; i--) {
        f = sqr12(f);
        f = mul_line_at(f, line_double(R, K), px, py);
        if ((loop >> i) & 1) f = mul_line_at(f, line_add(R, qx, qy), px, py);
    }
    // Q1 = pi(Q), Q2 = -pi^2(Q)  (alt_bn128_G2::mul_by_q, alt_bn128_g2.cpp:367-374)
    const Fq2 q1x = mul2(conj(qx), K.g[1][2]), q1y = mul2(conj(qy), K.g[1][3]);
    const Fq2 q2x = mul2(conj(q1x), K.g[1][2]), q2y = mul2(conj(q1y), K.g[1][3]).neg();
    f = mul_line_at(f, line_add(R, q1x, q1y), px, py);
    f = mul_line_at(f, line_add(R, q2x, q2y), px, py);
    return f;
}

// ---------------------------------------------------------------------------------------
// final exponentiation f^((q^12 - 1) / r * 2z(6z^2 + 3z + 1))  (alt_bn128_pairing.cpp:109-240)
// ---------------------------------------------------------------------------------------
// elt^(-z), z = 4965661367192848881 (alt_bn128_init.cpp:271); elt in the cyclotomic subgroup, where the
// cyclotomic squaring applies
ZKB_NI Fq12 exp_by_neg_z(const Fq12 &a)
{
    const uint64_t z = 0x44e992b44a6909f1ull;
    Fq12 r = a; // leading one at bit 62
    for (int i = 61; i >= 0; i--) {
        r = cyc_sqr12(r);
        if ((z >> i) & 1) r = mul12(r, a);
    }
    return r.conj();
}

ZKB_NI Fq12 final_exponentiation(const Fq12 &f, const PairingConsts &K)
{
    // easy part: f^((q^6 - 1)(q^2 + 1))
    const Fq12 c = mul12(f.conj(), inv12(f));
    const Fq12 m = mul12(frobenius12(c, 2, K), c);
    // hard part: m^(l0 + l1 q + l2 q^2 + l3 q^3) with l3 = 12z^3 + 6z^2 + 4z - 1, l2 = 12z^3 + 6z^2 + 6z,
    // l1 = 12z^3 + 6z^2 + 4z, l0 = 12z^3 + 12z^2 + 6z + 1 (Fuentes-Castaneda, Knapp, Rodriguez-Henriquez)
    const Fq12 mz = exp_by_neg_z(m);                 // m^-z
    const Fq12 m2z = cyc_sqr12(mz);                      // m^-2z
    const Fq12 m6z = mul12(cyc_sqr12(m2z), m2z);         // m^-6z
    const Fq12 m6z2 = exp_by_neg_z(m6z);             // m^(6z^2)
    const Fq12 m12z3 = exp_by_neg_z(cyc_sqr12(m6z2)).conj(); // m^(12z^3)
    const Fq12 a = mul12(mul12(m12z3, m6z2), m6z.conj()); // m^(12z^3 + 6z^2 + 6z)       = m^l2
    const Fq12 b = mul12(a, m2z);                         // m^(12z^3 + 6z^2 + 4z)       = m^l1
    const Fq12 l0 = mul12(mul12(a, m6z2), m);             // m^(12z^3 + 12z^2 + 6z + 1)  = m^l0
    const Fq12 l3 = mul12(m.conj(), b);                   // m^l3
    Fq12 r = mul12(frobenius12(b, 1, K), l0);
    r = mul12(frobenius12(a, 2, K), r);
    return mul12(frobenius12(l3, 3, K), r);
}

// Jacobian point on y^2 = x^3 + b: Y^2 = X^3 + b Z^6 (alt_bn128_g1.cpp:360-384, alt_bn128_g2.cpp:376-400);
// the zero point is well formed.
template <class F> ZKB_HD bool jacobian_well_formed(const F &X, const F &Y, const F &Z, const F &b)
{
    if (Z.is_zero()) return true;
    const F Z2 = Z.sqr();
    const F Z6 = (Z2.sqr()) * Z2;
    return Y.sqr() == X.sqr() * X + b * Z6;
}

// the reference's to_affine_coordinates (alt_bn128_g1.cpp:60-77): zero becomes (0, 1)
template <class F> ZKB_HD void jac_to_affine(const Jacobian<F> &j, F &x, F &y)
{
    if (j.Z.is_zero()) {
        x = F::zero();
        y = F::one();
        return;
    }
    const F zi = j.Z.inverse();
    const F zi2 = zi.sqr();
    x = j.X * zi2;
    y = j.Y * (zi2 * zi);
}

// the pairing check of r1cs_gg_ppzksnark_online_verifier_weak_IC (r1cs_gg_ppzksnark.tcc:556-570):
// FE(ML(A, B) * unitary_inverse(ML(acc, gamma) * ML(C, delta))) == vk.alpha_g1_beta_g2
ZKB_NI bool groth16_pairing_check(const Fq12 &f_ab, const Fq12 &f_acc_gamma, const Fq12 &f_c_delta, const Fq12 &alpha_beta,
                                  const PairingConsts &K)
{
    const Fq12 qap2 = mul12(f_acc_gamma, f_c_delta);
    return final_exponentiation(mul12(f_ab, qap2.conj()), K) == alpha_beta;
}

} // namespace zkb
