// pairing_coop.cuh -- the pairing of pairing.cuh spread over 6 or 18 cooperating lanes.
//
// At verification-sized batches (64 proofs) the device is almost empty and the time is one lane's chain of
// dependent 254-bit products (pairing.cuh: ~9 000 per Miller loop).  The tower arithmetic has 6-way parallelism
// at the Fq2 level -- an Fq6 Karatsuba product is six independent Fq2 products -- so a group of lanes
// keeps an Fq12 as six Fq2 "registers" in shared memory and every step is: each lane multiplies one (or a few)
// operand pairs, the group synchronises, each lane assembles one output coefficient.  Same values as
// pairing.cuh / libff (every result canonical); file:line citations of the reference are in pairing.cuh.
//
// Discipline that makes the host build (lanes emulated by a loop, tests/host/pairing_host.cpp) equivalent to
// the device: data crosses a phase boundary only through the arrays passed in, and no phase reads an array
// element that another lane writes in the same phase.
#pragma once
#include "pairing.cuh"

namespace zkb {

// A group is NL lanes working on one pairing: NL = 6 (one lane per Fq12 coefficient; a lane multiplies 3 operand
// pairs per Fq12 product; four groups share a warp) or NL = 18 (one operand pair per lane; one group per warp) -- the
// wide form is for the smallest batches, where warps are free and only the length of the dependent chain counts.
ZKB_HD constexpr int coop_group(int nl) { return nl <= 6 ? 8 : 32; } // lanes a group occupies in its warp

#if defined(__CUDA_ARCH__)
#define COOP_FOR_N(l, n)                                                                                               \
    {                                                                                                                  \
        const int l = threadIdx.x & (coop_group(NL) - 1);                                                              \
        if (l < (n)) {
#define COOP_END                                                                                                       \
        }                                                                                                              \
    }                                                                                                                  \
    __syncwarp();
#else
#define COOP_FOR_N(l, n)                                                                                               \
    {                                                                                                                  \
        for (int l = 0; l < (n); l++) {
#define COOP_END                                                                                                       \
        }                                                                                                              \
    }
#endif
#define COOP_FOR(l) COOP_FOR_N(l, NL) /* every lane of the group: the product phases */
#define COOP_FOR6(l) COOP_FOR_N(l, 6) /* the six coefficient lanes: everything else */

// An Fq12 as registers: index 0..5 = c0.c0, c0.c1, c0.c2, c1.c0, c1.c1, c1.c2 (memory order of struct Fq12).
//
// SIMT rule of this file: the lanes of a group share one instruction stream with the other groups of their warp, so
// a lane may differ from its neighbours only in WHICH operands it picks (selects on the lane number) -- never in
// which multiplication routine it calls.  Every phase is therefore "pick (x, y) by lane; one common mul2; store".

ZKB_HD Fq2 sel(bool c, const Fq2 &a, const Fq2 &b) { return c ? a : b; }
ZKB_HD Fq2 pick3(int c, const Fq2 &a0, const Fq2 &a1, const Fq2 &a2) { return c == 0 ? a0 : (c == 1 ? a1 : a2); }
// a / 2 (no multiplication: (a + (a odd ? p : 0)) >> 1; halving commutes with the Montgomery factor)
ZKB_HD Fq half(const Fq &a)
{
    const bool odd = a.v[0] & 1u;
    uint32_t t[8];
    uint64_t c = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const uint64_t s = (uint64_t)a.v[i] + (odd ? FqP::P(i) : 0u) + c;
        t[i] = (uint32_t)s;
        c = s >> 32;
    }
    Fq r; // a + p < 2^255: no carry out
#pragma unroll
    for (int i = 0; i < 8; i++) r.v[i] = (t[i] >> 1) | (i < 7 ? t[i + 1] << 31 : 0u);
    return r;
}
ZKB_HD Fq2 half(const Fq2 &a) { return Fq2{half(a.c0), half(a.c1)}; }
ZKB_HD Fq2 embed(const Fq &k) { return Fq2{k, Fq::zero()}; } // mul2(a, embed(k)) == mul_fq(a, k)

// Karatsuba over v of A = (A0, A1, A2) and B = (B0, B1, B2): operand pair j = 0, 1, 2 is (Aj, Bj); j = 3, 4, 5 are
// (A1 + A2, B1 + B2), (A0 + A1, B0 + B1), (A0 + A2, B0 + B2) -- the product phases below pick theirs by lane.
// coefficient c of that product from the six partial products p[0..5]:
//   c = 0: p0 + xi (p3 - p1 - p2);  c = 1: (p4 - p0 - p1) + xi p2;  c = 2: (p5 - p0 - p2) + p1
ZKB_HD Fq2 k6_coef(const Fq2 *p, int c)
{
    const Fq2 u = p[3 + c] - p[c == 0 ? 1 : 0] - p[c == 1 ? 1 : 2];
    const Fq2 w = p[c == 0 ? 0 : (c == 1 ? 2 : 1)];
    const Fq2 xt = mul_xi(sel(c == 0, u, w));
    return sel(c == 0, w, u) + sel(c == 2, w, xt);
}
// coefficient l of (t0 + v t1) + (t2 - t0 - t1) w from the partial products of t0, t1, t2 (18 values)
ZKB_HD Fq2 k12_coef(const Fq2 *P, int l)
{
    const bool hi = l >= 3;
    const int c = hi ? l - 3 : l;
    const Fq2 A = k6_coef(P, c);
    const Fq2 B = k6_coef(P + 6, hi ? c : (c + 2) % 3);
    const Fq2 C = k6_coef(P + 12, c);
    const Fq2 lo = A + sel(c == 0, mul_xi(B), B);
    return sel(hi, C - A - B, lo);
}

// out = a * b.  P: 18 scratch registers.  out may be a or b.
template <int NL> ZKB_NI void coop_mul12(Fq2 *out, const Fq2 *a, const Fq2 *b, Fq2 *P)
{
    COOP_FOR(l)
    for (int pl = l; pl < 18; pl += NL) { // product pl = 6 r + j: pair j of a0 b0 (r = 0), a1 b1 (1), (a0 + a1)(b0 + b1) (2)
        const int r = pl / 6, j = pl - 6 * r;
        const int i0 = j < 3 ? j : (j == 3 ? 1 : 0), i1 = j == 4 ? 1 : 2;
        auto ld = [&](const Fq2 *v, int i) { return r == 2 ? v[i] + v[3 + i] : v[3 * r + i]; };
        Fq2 x = ld(a, i0), y = ld(b, i0);
        if (j >= 3) {
            x = x + ld(a, i1);
            y = y + ld(b, i1);
        }
        P[pl] = mul2(x, y);
    }
    COOP_END
    COOP_FOR6(l)
    out[l] = k12_coef(P, l);
    COOP_END
}

// out = a^2.  P: 12 scratch registers.  out may be a.
template <int NL> ZKB_NI void coop_sqr12(Fq2 *out, const Fq2 *a, Fq2 *P)
{
    COOP_FOR(l)
    for (int pl = l; pl < 12; pl += NL) { // r = 0: ab = a0 a1;  r = 1: t = (a0 + a1)(a0 + v a1)
        const int r = pl / 6, jj = pl - 6 * r;
        const int i0 = jj < 3 ? jj : (jj == 3 ? 1 : 0), i1 = jj == 4 ? 1 : 2;
        auto X = [&](int i) { return r ? a[i] + a[3 + i] : a[i]; };
        auto Y = [&](int i) { return r ? (i == 0 ? a[0] + mul_xi(a[5]) : a[i] + a[2 + i]) : a[3 + i]; };
        Fq2 x = X(i0), y = Y(i0);
        if (jj >= 3) {
            x = x + X(i1);
            y = y + Y(i1);
        }
        P[pl] = mul2(x, y);
    }
    COOP_END
    COOP_FOR6(l)
    const bool hi = l >= 3;
    const int c = hi ? l - 3 : l;
    const Fq2 ab = k6_coef(P, c);
    const Fq2 abp = k6_coef(P, (c + 2) % 3); // coefficient c of v * ab is xi^[c == 0] * ab[c - 1]
    const Fq2 t = k6_coef(P + 6, c);
    out[l] = sel(hi, ab.dbl(), t - ab - sel(c == 0, mul_xi(abp), abp));
    COOP_END
}

// a <- a^2 for a in the cyclotomic subgroup (cyc_sqr12 of pairing.cuh): six products, one per lane.  P: 6 registers.
// Register l holds z0, z4, z3, z2, z1, z5; the Fq4 pairs are (f0, f4), (f3, f2), (f1, f5).
template <int NL> ZKB_NI void coop_cyc_sqr12(Fq2 *a, Fq2 *P)
{
    COOP_FOR6(l)
    const int k = l >> 1; // pair k: lane 2k makes m = x y, lane 2k + 1 makes n = (x + y)(x + xi y)
    const Fq2 x = a[k == 0 ? 0 : k == 1 ? 3 : 1], y = a[k == 0 ? 4 : k == 1 ? 2 : 5];
    const bool second = l & 1;
    P[l] = mul2(sel(second, x + y, x), sel(second, x + mul_xi(y), y));
    COOP_END
    COOP_FOR6(l)
    // register l takes: 0: 3 te(0) - 2z;  1: 3 te(1) - 2z;  2: 3 te(2) - 2z;  3: 3 xi to(2) + 2z;  4: 3 to(0) + 2z;  5: 3 to(1) + 2z
    const bool even = l < 3;
    const int k = even ? l : (l == 3 ? 2 : l - 4);
    const Fq2 m = P[2 * k], n = P[2 * k + 1], z = a[l];
    const Fq2 xm = mul_xi(m);
    const Fq2 t = sel(even, n - m - xm, sel(l == 3, xm.dbl(), m.dbl()));
    const Fq2 t3 = t.dbl() + t, z2 = z.dbl();
    a[l] = sel(even, t3 - z2, t3 + z2);
    COOP_END
}

// f = f * (e0 + eVV v^2 + eVW v w), the (already scaled) line value.  P: 14 scratch registers.
template <int NL> ZKB_NI void coop_mul12_line(Fq2 *f, const Fq2 *L /* e0, eVW, eVV */, Fq2 *P)
{
    COOP_FOR(l)
    const Fq2 x = L[0], y = L[1], z = L[2];
    for (int pl = l; pl < 14; pl += NL) {
        Fq2 p, q;
        if (pl < 6) { // t2 = (f.c0 + f.c1) * (x, y, z), dense: pair pl
            const int i0 = pl < 3 ? pl : (pl == 3 ? 1 : 0), i1 = pl == 4 ? 1 : 2;
            p = f[i0] + f[3 + i0];
            q = pick3(i0, x, y, z);
            if (pl >= 3) {
                p = p + f[i1] + f[3 + i1];
                q = q + pick3(i1, x, y, z);
            }
        } else if (pl < 12) { // t0 = f.c0 * (x, 0, z): q0 = a0 x, q1 = a2 z, q2 = (a0 + a2)(x + z), q3 = a1 z, q4 = a1 x;  r0 = b2 y
            const int k = pl - 6;
            p = k == 0 ? f[0] : k == 1 ? f[2] : k == 2 ? f[0] + f[2] : k == 5 ? f[5] : f[1];
            q = k == 0 ? x : k == 1 ? z : k == 2 ? x + z : k == 3 ? z : k == 4 ? x : y;
        } else { // t1 = f.c1 * (0, y, 0): r1 = b0 y, r2 = b1 y
            p = f[3 + pl - 12];
            q = y;
        }
        P[pl] = mul2(p, q); // P[0..5] t2's pairs, P[6..10] = q0..q4, P[11] = r0, P[12], P[13] = r1, r2
    }
    COOP_END
    COOP_FOR6(l)
    // t0 = (q0 + xi q3, q4 + xi q1, q2 - q0 - q1), t1 = (xi r0, r1, r2);  out = (t0 + v t1) + (t2 - t0 - t1) w
    const bool hi = l >= 3;
    const int c = hi ? l - 3 : l;
    const Fq2 t0c = pick3(c, P[6] + mul_xi(P[9]), P[10] + mul_xi(P[7]), P[8] - P[6] - P[7]);
    const Fq2 xr0 = mul_xi(P[11]);
    const Fq2 t1c = pick3(c, xr0, P[12], P[13]);
    const Fq2 vt1c = pick3(c, mul_xi(P[13]), xr0, P[12]); // coefficient c of v * t1
    f[l] = sel(hi, k6_coef(P, c) - t0c - t1c, t0c + vt1c);
    COOP_END
}

// Frobenius_map(i): register l holds the coefficient of w^k, k = (0, 2, 4, 1, 3, 5)[l]
template <int NL> ZKB_NI void coop_frobenius12(Fq2 *out, const Fq2 *a, int i, const PairingConsts &K)
{
    COOP_FOR6(l)
    const int k = l < 3 ? 2 * l : 2 * (l - 3) + 1;
    const Fq2 z = (i & 1) ? conj(a[l]) : a[l];
    out[l] = mul2(z, K.g[i][k]); // g[i][0] = 1
    COOP_END
}
// unitary inverse; out may be a
template <int NL> ZKB_NI void coop_conj12(Fq2 *out, const Fq2 *a)
{
    COOP_FOR6(l)
    const Fq2 v = a[l];
    out[l] = sel(l < 3, v, v.neg());
    COOP_END
}
template <int NL> ZKB_NI void coop_copy12(Fq2 *out, const Fq2 *a)
{
    COOP_FOR6(l)
    out[l] = a[l];
    COOP_END
}
// full inverse: one lane (an extended-field inversion is a serial chain ending in one Fq inversion)
template <int NL> ZKB_NI void coop_inv12(Fq2 *out, const Fq2 *a)
{
    COOP_FOR6(l)
    if (l == 0) {
        Fq12 t;
        t.c0 = Fq6{a[0], a[1], a[2]};
        t.c1 = Fq6{a[3], a[4], a[5]};
        t = inv12(t);
        out[0] = t.c0.c0; out[1] = t.c0.c1; out[2] = t.c0.c2;
        out[3] = t.c1.c0; out[4] = t.c1.c1; out[5] = t.c1.c2;
    }
    COOP_END
}

// tangent step: R <- 2R, L <- the line scaled by (px, py).  T: 8 scratch registers.
template <int NL> ZKB_NI void coop_line_double(Fq2 *R, Fq2 *L, const Fq &px, const Fq &py, Fq2 *T, const PairingConsts &K)
{
    COOP_FOR6(l)
    // T0 = X Y, T1 = B = Y^2, T2 = C = Z^2, T3 = (Y + Z)^2, T4 = J = X^2
    if (l < 5) {
        const Fq2 yz = R[1] + R[2];
        const Fq2 x = l == 0 ? R[0] : l == 1 ? R[1] : l == 2 ? R[2] : l == 3 ? yz : R[0];
        const Fq2 y = l == 0 ? R[1] : x;
        T[l] = mul2(x, y);
    }
    COOP_END
    COOP_FOR6(l)
    if (l == 0) T[5] = mul2(K.twist_b, T[2].dbl() + T[2]); // E = b' 3C
    COOP_END
    COOP_FOR6(l)
    // lane 0: X3 = X Y (B - F) / 2;  1: G^2, G = (B + F) / 2;  2: E^2;  3: Z3 = B H;  4: eVW = -H py;  5: eVV = 3J px
    const Fq2 B = T[1], E = T[5], F = E.dbl() + E, H = T[3] - (B + T[2]), G = half(B + F), J3 = T[4].dbl() + T[4];
    const Fq2 x = l == 0 ? T[0] : l == 1 ? G : l == 2 ? E : l == 3 ? B : l == 4 ? H.neg() : J3;
    const Fq2 y = l == 0 ? half(B - F) : l == 1 ? G : l == 2 ? E : l == 3 ? H : l == 4 ? embed(py) : embed(px);
    const Fq2 r = mul2(x, y);
    Fq2 *dst = l == 0 ? &R[0] : l == 1 ? &T[6] : l == 2 ? &T[7] : l == 3 ? &R[2] : l == 4 ? &L[1] : &L[2];
    *dst = r; // nobody reads R or L in this phase
    if (l == 5) L[0] = mul_xi(E - B);
    COOP_END
    COOP_FOR6(l)
    if (l == 0) R[1] = T[6] - (T[7].dbl() + T[7]); // Y3 = G^2 - 3 E^2
    COOP_END
}

// chord step: R <- R + (x2, y2), L <- the line scaled by (px, py).  T: 11 scratch registers.
template <int NL> ZKB_NI void coop_line_add(Fq2 *R, Fq2 *L, const Fq2 &x2, const Fq2 &y2, const Fq &px, const Fq &py, Fq2 *T)
{
    COOP_FOR6(l)
    if (l < 2) T[l] = R[l] - mul2(l == 0 ? x2 : y2, R[2]); // D = X1 - x2 Z1, E = Y1 - y2 Z1
    COOP_END
    COOP_FOR6(l)
    // T2 = F = D^2, T3 = G = E^2, T4 = E x2, T5 = D y2
    if (l < 4) {
        const Fq2 x = (l == 0 || l == 3) ? T[0] : T[1];
        const Fq2 y = l == 0 ? T[0] : l == 1 ? T[1] : l == 2 ? x2 : y2;
        T[2 + l] = mul2(x, y);
    }
    COOP_END
    COOP_FOR6(l)
    // lane 0: H = D F;  1: I = X1 F;  2: Z1 G;  3: eVW = D py;  4: eVV = -E px;  5: e0 = xi (E x2 - D y2)
    if (l < 5) {
        const Fq2 x = l == 0 ? T[0] : l == 1 ? R[0] : l == 2 ? R[2] : l == 3 ? T[0] : T[1].neg();
        const Fq2 y = l < 2 ? T[2] : l == 2 ? T[3] : l == 3 ? embed(py) : embed(px);
        const Fq2 r = mul2(x, y);
        Fq2 *dst = l == 0 ? &T[6] : l == 1 ? &T[7] : l == 2 ? &T[8] : l == 3 ? &L[1] : &L[2];
        *dst = r;
    } else {
        L[0] = mul_xi(T[4] - T[5]);
    }
    COOP_END
    COOP_FOR6(l)
    // lane 0: X3 = D J;  1: E (I - J);  2: H Y1;  3: Z3 = Z1 H     (J = H + Z1 G - 2 I)
    if (l < 4) {
        const Fq2 H = T[6], I = T[7], J = H + T[8] - I.dbl();
        const Fq2 x = l == 0 ? T[0] : l == 1 ? T[1] : l == 2 ? H : R[2];
        const Fq2 y = l == 0 ? J : l == 1 ? I - J : l == 2 ? R[1] : H;
        const Fq2 r = mul2(x, y);
        Fq2 *dst = l == 0 ? &R[0] : l == 1 ? &T[9] : l == 2 ? &T[10] : &R[2];
        *dst = r; // R[0] is read by nobody here; R[2] only by the lane that rewrites it; R[1] is not written
    }
    COOP_END
    COOP_FOR6(l)
    if (l == 0) R[1] = T[9] - T[10];
    COOP_END
}

// Registers a Miller loop needs: f 6 | R 3 | L 3 | Q1, Q2 4 | scratch 18
constexpr int COOP_MILLER_REGS = 34;
// f (S[0..5]) = alt_bn128_ate_miller_loop of affine P = (px, py), Q = (qx, qy); infinity as (0, 1)
template <int NL> ZKB_NI void coop_miller_loop(Fq2 *S, const Fq &px, const Fq &py, const Fq2 &qx, const Fq2 &qy, const PairingConsts &K)
{
    Fq2 *f = S, *R = S + 6, *L = S + 9, *Qf = S + 12, *T = S + 16;
    COOP_FOR6(l)
    f[l] = l == 0 ? Fq2::one() : Fq2::zero();
    if (l == 0) { R[0] = qx; R[1] = qy; R[2] = Fq2::one(); }
    // Q1 = pi(Q), Q2 = -pi^2(Q)
    if (l == 1) {
        const Fq2 q1x = mul2(conj(qx), K.g[1][2]);
        Qf[0] = q1x;
        Qf[2] = mul2(conj(q1x), K.g[1][2]);
    }
    if (l == 2) {
        const Fq2 q1y = mul2(conj(qy), K.g[1][3]);
        Qf[1] = q1y;
        Qf[3] = mul2(conj(q1y), K.g[1][3]).neg();
    }
    COOP_END
    const uint64_t loop = 0x9d797039be763ba8ull;
    for (int i = 63; i >= 0; i--) {
        coop_sqr12<NL>(f, f, T);
        coop_line_double<NL>(R, L, px, py, T, K);
        coop_mul12_line<NL>(f, L, T);
        if ((loop >> i) & 1) {
            coop_line_add<NL>(R, L, qx, qy, px, py, T);
            coop_mul12_line<NL>(f, L, T);
        }
    }
    for (int k = 0; k < 2; k++) {
        Fq2 x2, y2; // read before the step's first phase: Qf is never written again
        x2 = Qf[2 * k];
        y2 = Qf[2 * k + 1];
        coop_line_add<NL>(R, L, x2, y2, px, py, T);
        coop_mul12_line<NL>(f, L, T);
    }
}

// elt^(-z) for elt in the cyclotomic subgroup.  out != a.
template <int NL> ZKB_NI void coop_exp_by_neg_z(Fq2 *out, const Fq2 *a, Fq2 *T)
{
    const uint64_t z = 0x44e992b44a6909f1ull;
    coop_copy12<NL>(out, a);
    for (int i = 61; i >= 0; i--) {
        coop_cyc_sqr12<NL>(out, T);
        if ((z >> i) & 1) coop_mul12<NL>(out, out, a, T);
    }
    coop_conj12<NL>(out, out);
}

// Registers the final exponentiation needs: 10 values of 6 + scratch 18
constexpr int COOP_FE_REGS = 78;
// S[0..5] = final_exponentiation(S[0..5])
template <int NL> ZKB_NI void coop_final_exponentiation(Fq2 *S, const PairingConsts &K)
{
    Fq2 *f = S, *m = S + 6, *mz = S + 12, *m2z = S + 18, *m6z = S + 24, *m6z2 = S + 30, *a = S + 36, *b = S + 42, *t = S + 48,
        *u = S + 54, *T = S + 60;
    // easy part: m = c^(q^2) c, c = conj(f) / f
    coop_inv12<NL>(t, f);
    coop_conj12<NL>(u, f);
    coop_mul12<NL>(t, u, t, T);
    coop_frobenius12<NL>(u, t, 2, K);
    coop_mul12<NL>(m, u, t, T);
    // hard part (see final_exponentiation in pairing.cuh)
    coop_exp_by_neg_z<NL>(mz, m, T);          // m^-z
    coop_copy12<NL>(m2z, mz);
    coop_cyc_sqr12<NL>(m2z, T);               // m^-2z
    coop_copy12<NL>(t, m2z);
    coop_cyc_sqr12<NL>(t, T);
    coop_mul12<NL>(m6z, t, m2z, T);           // m^-6z
    coop_exp_by_neg_z<NL>(m6z2, m6z, T);      // m^(6z^2)
    coop_copy12<NL>(t, m6z2);
    coop_cyc_sqr12<NL>(t, T);
    coop_exp_by_neg_z<NL>(u, t, T);
    coop_conj12<NL>(u, u);                    // m^(12z^3)
    coop_mul12<NL>(a, u, m6z2, T);
    coop_conj12<NL>(t, m6z);
    coop_mul12<NL>(a, a, t, T);               // a = m^(12z^3 + 6z^2 + 6z)
    coop_mul12<NL>(b, a, m2z, T);             // b = m^(12z^3 + 6z^2 + 4z)
    coop_mul12<NL>(t, a, m6z2, T);
    coop_mul12<NL>(t, t, m, T);               // l0
    coop_frobenius12<NL>(u, b, 1, K);
    coop_mul12<NL>(f, u, t, T);               // b^q l0
    coop_frobenius12<NL>(u, a, 2, K);
    coop_mul12<NL>(f, u, f, T);
    coop_conj12<NL>(t, m);
    coop_mul12<NL>(t, t, b, T);               // l3
    coop_frobenius12<NL>(u, t, 3, K);
    coop_mul12<NL>(f, u, f, T);
}

} // namespace zkb
