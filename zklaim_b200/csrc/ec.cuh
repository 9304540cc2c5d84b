// ec.cuh -- alt_bn128 G1 / G2 group arithmetic (y^2 = x^3 + b, a = 0) for sm_100a.
//
// Replaces libff's alt_bn128_G1 / alt_bn128_G2 group law on the prover path
// (alt_bn128_g1.cpp:139-195 operator+, :256-323 mixed_add, :325-358 dbl, :60-77 to_affine;
// alt_bn128_g2.cpp:149-205, :266-333, :335-365).  The reference keeps Jacobian (X,Y,Z) and
// adds with add-2007-bl / madd-2007-bl / dbl-2009-l.  Here bucket accumulators use the
// extended-Jacobian XYZZ form (x = X/ZZ, y = Y/ZZZ, ZZ^3 = ZZZ^2), whose mixed addition is
// 8M+2S against 7M+4S(+reductions) and needs no inversions; results are converted to the
// reference's Jacobian layout at the C-ABI.  Group elements, not coordinates, are what the
// reference serialises (affine-normalised on write, alt_bn128_g1.cpp:404-418), so any
// representative gives identical proof bytes.
//
// All formulas are complete in the sense the reference's are: infinity operands, P + P
// (falls to doubling, alt_bn128_g1.cpp:175-179 / :297-301) and P + (-P) are handled.
//
// Everything is a template over the coordinate field F (Fq for G1, Fq2 for G2).
#pragma once
#include "fp.cuh"

namespace zkb {

// Affine point; the point at infinity is encoded as x = y = 0 (not on the curve: b != 0).
template <class F> struct Affine {
    F x, y;
    ZKB_HD bool is_inf() const { return x.is_zero() && y.is_zero(); }
};

// Reference layout (alt_bn128_g1.hpp:35): Jacobian X, Y, Z; infinity <=> Z == 0.
template <class F> struct Jacobian {
    F X, Y, Z;
};

template <class F> struct XYZZ {
    F X, Y, ZZ, ZZZ;

    static ZKB_HD XYZZ inf()
    {
        XYZZ r;
        r.X = F::zero();
        r.Y = F::zero();
        r.ZZ = F::zero();
        r.ZZZ = F::zero();
        return r;
    }
    ZKB_HD bool is_inf() const { return ZZ.is_zero(); }

    static ZKB_HD XYZZ from_affine(const Affine<F> &p)
    {
        if (p.is_inf()) return inf();
        XYZZ r;
        r.X = p.x;
        r.Y = p.y;
        r.ZZ = F::one();
        r.ZZZ = F::one();
        return r;
    }

    // Jacobian representative with Z = ZZZ:  X' = X*ZZ^2, Y' = Y*ZZ^3  (ZZZ^2 = ZZ^3).
    ZKB_HD Jacobian<F> to_jacobian() const
    {
        Jacobian<F> j;
        if (is_inf()) {
            j.X = F::zero();
            j.Y = F::one();
            j.Z = F::zero();
            return j;
        }
        const F zz2 = ZZ.sqr();
        j.X = X * zz2;
        j.Y = Y * (zz2 * ZZ);
        j.Z = ZZZ;
        return j;
    }

    static ZKB_HD XYZZ from_jacobian(const Jacobian<F> &j)
    {
        if (j.Z.is_zero()) return inf();
        XYZZ r;
        r.X = j.X;
        r.Y = j.Y;
        r.ZZ = j.Z.sqr();
        r.ZZZ = r.ZZ * j.Z;
        return r;
    }

    ZKB_HD XYZZ neg() const
    {
        XYZZ r = *this;
        r.Y = Y.neg();
        return r;
    }

    // dbl-2008-s-1 (a = 0): 6M + 4S... U=2Y, V=U^2, W=U*V, S=X*V, M=3X^2
    ZKB_HD XYZZ dbl() const
    {
        if (is_inf()) return *this;
        const F U = Y.dbl();
        const F V = U.sqr();
        const F W = U * V;
        const F S = X * V;
        const F X2 = X.sqr();
        const F M = X2.dbl() + X2;
        XYZZ r;
        r.X = M.sqr() - S.dbl();
        r.Y = M * (S - r.X) - W * Y;
        r.ZZ = V * ZZ;
        r.ZZZ = W * ZZZ;
        return r;
    }

    // this += (x2, y2) affine, optionally negated.  madd-2008-s (8M + 2S); mdbl-2008-s-1 when equal.
    ZKB_HD void madd(const Affine<F> &p, bool negate = false)
    {
        if (p.is_inf()) return;
        const F y2 = negate ? p.y.neg() : p.y;
        if (is_inf()) {
            X = p.x;
            Y = y2;
            ZZ = F::one();
            ZZZ = F::one();
            return;
        }
        const F U2 = p.x * ZZ;
        const F S2 = y2 * ZZZ;
        const F P = U2 - X;
        const F R = S2 - Y;
        if (P.is_zero()) {
            if (R.is_zero()) {
                // doubling of the affine point
                const F U = y2.dbl();
                const F V = U.sqr();
                const F W = U * V;
                const F S = p.x * V;
                const F X2 = p.x.sqr();
                const F M = X2.dbl() + X2;
                X = M.sqr() - S.dbl();
                Y = M * (S - X) - W * y2;
                ZZ = V;
                ZZZ = W;
            } else {
                *this = inf();
            }
            return;
        }
        const F PP = P.sqr();
        const F PPP = P * PP;
        const F Q = X * PP;
        const F X3 = R.sqr() - PPP - Q.dbl();
        Y = R * (Q - X3) - Y * PPP;
        X = X3;
        ZZ = ZZ * PP;
        ZZZ = ZZZ * PPP;
    }

    // this += o.  add-2008-s (12M + 2S)
    ZKB_HD void add(const XYZZ &o)
    {
        if (o.is_inf()) return;
        if (is_inf()) {
            *this = o;
            return;
        }
        const F U1 = X * o.ZZ;
        const F U2 = o.X * ZZ;
        const F S1 = Y * o.ZZZ;
        const F S2 = o.Y * ZZZ;
        const F P = U2 - U1;
        const F R = S2 - S1;
        if (P.is_zero()) {
            if (R.is_zero())
                *this = dbl();
            else
                *this = inf();
            return;
        }
        const F PP = P.sqr();
        const F PPP = P * PP;
        const F Q = U1 * PP;
        const F X3 = R.sqr() - PPP - Q.dbl();
        Y = R * (Q - X3) - S1 * PPP;
        X = X3;
        ZZ = ZZ * o.ZZ * PP;
        ZZZ = ZZZ * o.ZZZ * PPP;
    }
};

typedef Affine<Fq> G1Affine;
typedef Affine<Fq2> G2Affine;
typedef XYZZ<Fq> G1X;
typedef XYZZ<Fq2> G2X;
typedef Jacobian<Fq> G1J;
typedef Jacobian<Fq2> G2J;

// 256-bit plain-integer scalar (already out of Montgomery form) times a point,
// MSB-first double-and-add as in curve_utils.tcc:13-34.
template <class F> ZKB_HD XYZZ<F> scalar_mul(const uint32_t *k, const XYZZ<F> &p)
{
    XYZZ<F> acc = XYZZ<F>::inf();
    bool started = false;
    for (int i = 255; i >= 0; i--) {
        if (started) acc = acc.dbl();
        if ((k[i >> 5] >> (i & 31)) & 1) {
            started = true;
            acc.add(p);
        }
    }
    return acc;
}

} // namespace zkb
