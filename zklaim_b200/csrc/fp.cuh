// fp.cuh -- 254-bit Montgomery prime-field arithmetic on 8 x 32-bit limbs for sm_100a.
//
// Replaces libff's Fp_model<4, modulus> (libff/algebra/fields/fp.hpp:39-126; mul_reduce
// fp.tcc:23-187, += :310-420, -= :423-510, negation :575-591, squared :594-639) for the two
// alt_bn128 fields Fr and Fq (constants: alt_bn128_init.cpp:40-62, 66-88).  Same value
// domain as the reference: Montgomery form with R = 2^256, every result canonical (< p),
// so a device element is bit-identical to the reference's 4 x u64 `mont_repr`.
//
// The multiplier is a word-serial (CIOS) Montgomery product arranged as two carry chains
// per row: products of the even-indexed limbs of `a` tile columns (0,1)(2,3)(4,5)(6,7) and
// those of the odd-indexed limbs tile columns (1,2)(3,4)(5,6)(7,8), so that each row of
// a*b[i] and of m*p is two mad.lo.cc/madc.hi.cc chains without cross-chain carry fix-ups.
// The 32-bit right shift after each row is a role swap of the two accumulators; the one
// word that changes alignment is folded into the head of the next row's odd chain.
// 8 x (32 IMAD + 1 IMUL) integer multiply-adds per product; no tensor cores (integer work).
//
// Every chain is one asm block (carry flag never crosses a block), so the file also
// compiles as plain C++ (ZKB_PORTABLE path below) -- used by the CPU-side unit test of
// the algorithm and by the host-side helpers of the library.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define ZKB_HD __host__ __device__ __forceinline__
#else
#define ZKB_HD inline
#endif

#if defined(__CUDACC__) && defined(ZKB_NOINLINE_MUL)
#define ZKB_MUL_ATTR __host__ __device__ __noinline__
#elif defined(__CUDACC__)
#define ZKB_MUL_ATTR __host__ __device__ __forceinline__
#else
#define ZKB_MUL_ATTR inline
#endif

#if defined(__CUDA_ARCH__)
#define ZKB_ASM 1
#else
#define ZKB_ASM 0
#endif

namespace zkb {

// ---------------------------------------------------------------------------------------
// Field parameters (little-endian 32-bit words).  alt_bn128_init.cpp:40-62 (Fr), :66-88 (Fq)
// ---------------------------------------------------------------------------------------
struct FqP {
    // q = 21888242871839275222246405745257275088696311157297823662689037894645226208583
    static ZKB_HD constexpr uint32_t P(int i)
    {
        constexpr uint32_t t[8] = {0xd87cfd47u, 0x3c208c16u, 0x6871ca8du, 0x97816a91u, 0x8181585du, 0xb85045b6u, 0xe131a029u, 0x30644e72u};
        return t[i];
    }
    // R mod q (Montgomery one)
    static ZKB_HD constexpr uint32_t ONE(int i)
    {
        constexpr uint32_t t[8] = {0xc58f0d9du, 0xd35d438du, 0xf5c70b3du, 0x0a78eb28u, 0x7879462cu, 0x666ea36fu, 0x9a07df2fu, 0x0e0a77c1u};
        return t[i];
    }
    // R^2 mod q
    static ZKB_HD constexpr uint32_t R2(int i)
    {
        constexpr uint32_t t[8] = {0x538afa89u, 0xf32cfc5bu, 0xd44501fbu, 0xb5e71911u, 0x0a417ff6u, 0x47ab1effu, 0xcab8351fu, 0x06d89f71u};
        return t[i];
    }
    static constexpr uint32_t INV = 0xe4866389u; // -q^-1 mod 2^32  (low half of 0x87d20782e4866389)
};
struct FrP {
    // r = 21888242871839275222246405745257275088548364400416034343698204186575808495617
    static ZKB_HD constexpr uint32_t P(int i)
    {
        constexpr uint32_t t[8] = {0xf0000001u, 0x43e1f593u, 0x79b97091u, 0x2833e848u, 0x8181585du, 0xb85045b6u, 0xe131a029u, 0x30644e72u};
        return t[i];
    }
    static ZKB_HD constexpr uint32_t ONE(int i)
    {
        constexpr uint32_t t[8] = {0x4ffffffbu, 0xac96341cu, 0x9f60cd29u, 0x36fc7695u, 0x7879462eu, 0x666ea36fu, 0x9a07df2fu, 0x0e0a77c1u};
        return t[i];
    }
    static ZKB_HD constexpr uint32_t R2(int i)
    {
        constexpr uint32_t t[8] = {0xae216da7u, 0x1bb8e645u, 0xe35c59e3u, 0x53fe3ab1u, 0x53bb8085u, 0x8c49833du, 0x7f4e44a5u, 0x0216d0b1u};
        return t[i];
    }
    static constexpr uint32_t INV = 0xefffffffu; // low half of 0xc2e1f593efffffff
};

// ---------------------------------------------------------------------------------------
// Carry-chain building blocks.  x: accumulator aligned on column 0; y: aligned on column 1.
// ---------------------------------------------------------------------------------------

// x[0..7] += {a0,a2,a4,a6} * b  (column-aligned pairs); carry out of column 7 is added to `top`.
ZKB_HD void chain_even(uint32_t *x, uint32_t &top, uint32_t a0, uint32_t a2, uint32_t a4, uint32_t a6, uint32_t b)
{
#if ZKB_ASM
    asm("mad.lo.cc.u32  %0, %9,  %13, %0;\n\t"
        "madc.hi.cc.u32 %1, %9,  %13, %1;\n\t"
        "madc.lo.cc.u32 %2, %10, %13, %2;\n\t"
        "madc.hi.cc.u32 %3, %10, %13, %3;\n\t"
        "madc.lo.cc.u32 %4, %11, %13, %4;\n\t"
        "madc.hi.cc.u32 %5, %11, %13, %5;\n\t"
        "madc.lo.cc.u32 %6, %12, %13, %6;\n\t"
        "madc.hi.cc.u32 %7, %12, %13, %7;\n\t"
        "addc.u32       %8, %8, 0;"
        : "+r"(x[0]), "+r"(x[1]), "+r"(x[2]), "+r"(x[3]), "+r"(x[4]), "+r"(x[5]), "+r"(x[6]), "+r"(x[7]), "+r"(top)
        : "r"(a0), "r"(a2), "r"(a4), "r"(a6), "r"(b));
#else
    const uint32_t a[4] = {a0, a2, a4, a6};
    uint64_t c = 0;
    for (int j = 0; j < 4; j++) {
        const uint64_t prod = (uint64_t)a[j] * b;
        uint64_t lo = (uint64_t)x[2 * j] + (uint32_t)prod + c;
        x[2 * j] = (uint32_t)lo;
        uint64_t hi = (uint64_t)x[2 * j + 1] + (prod >> 32) + (lo >> 32);
        x[2 * j + 1] = (uint32_t)hi;
        c = hi >> 32;
    }
    top += (uint32_t)c;
#endif
}

// First row: x = {a0,a2,a4,a6} * b, y = {a1,a3,a5,a7} * b  (no carries at all).
ZKB_HD void row_first(uint32_t *x, uint32_t *y, const uint32_t *a, uint32_t b)
{
#if ZKB_ASM
#pragma unroll
    for (int j = 0; j < 8; j += 2) {
        asm("mul.lo.u32 %0, %2, %3;\n\tmul.hi.u32 %1, %2, %3;" : "=r"(x[j]), "=r"(x[j + 1]) : "r"(a[j]), "r"(b));
        asm("mul.lo.u32 %0, %2, %3;\n\tmul.hi.u32 %1, %2, %3;" : "=r"(y[j]), "=r"(y[j + 1]) : "r"(a[j + 1]), "r"(b));
    }
#else
    for (int j = 0; j < 8; j += 2) {
        uint64_t p0 = (uint64_t)a[j] * b, p1 = (uint64_t)a[j + 1] * b;
        x[j] = (uint32_t)p0;
        x[j + 1] = (uint32_t)(p0 >> 32);
        y[j] = (uint32_t)p1;
        y[j + 1] = (uint32_t)(p1 >> 32);
    }
#endif
}

// Shift-and-accumulate for the odd chain.  On entry `xo` is the previous row's column-0
// accumulator (xo[0] == 0 after reduction) and `xn0` is word 0 of the previous row's column-1
// accumulator, which becomes the new column-0 accumulator.  Performs
//     xn0 += xo[1]                       (the one word whose alignment changes)
//     yn[k] = xo[k+2] + {a1,a3,a5,a7}*b  (k = 0..7, xo[8], xo[9] read as 0), carry-in from above.
ZKB_HD void chain_odd_shift(uint32_t *yn, uint32_t &xn0, const uint32_t *xo, uint32_t a1, uint32_t a3, uint32_t a5,
                            uint32_t a7, uint32_t b)
{
#if ZKB_ASM
    asm("add.cc.u32     %8, %8, %9;\n\t"
        "madc.lo.cc.u32 %0, %16, %20, %10;\n\t"
        "madc.hi.cc.u32 %1, %16, %20, %11;\n\t"
        "madc.lo.cc.u32 %2, %17, %20, %12;\n\t"
        "madc.hi.cc.u32 %3, %17, %20, %13;\n\t"
        "madc.lo.cc.u32 %4, %18, %20, %14;\n\t"
        "madc.hi.cc.u32 %5, %18, %20, %15;\n\t"
        "madc.lo.cc.u32 %6, %19, %20, 0;\n\t"
        "madc.hi.u32    %7, %19, %20, 0;"
        : "=&r"(yn[0]), "=&r"(yn[1]), "=&r"(yn[2]), "=&r"(yn[3]), "=&r"(yn[4]), "=&r"(yn[5]), "=&r"(yn[6]), "=&r"(yn[7]),
          "+r"(xn0)
        : "r"(xo[1]), "r"(xo[2]), "r"(xo[3]), "r"(xo[4]), "r"(xo[5]), "r"(xo[6]), "r"(xo[7]), "r"(a1), "r"(a3),
          "r"(a5), "r"(a7), "r"(b));
#else
    const uint32_t a[4] = {a1, a3, a5, a7};
    uint64_t s = (uint64_t)xn0 + xo[1];
    xn0 = (uint32_t)s;
    uint64_t c = s >> 32;
    for (int j = 0; j < 4; j++) {
        const uint64_t prod = (uint64_t)a[j] * b;
        const uint32_t in0 = (2 * j + 2 < 8) ? xo[2 * j + 2] : 0u;
        const uint32_t in1 = (2 * j + 3 < 8) ? xo[2 * j + 3] : 0u;
        uint64_t lo = (uint64_t)in0 + (uint32_t)prod + c;
        yn[2 * j] = (uint32_t)lo;
        uint64_t hi = (uint64_t)in1 + (prod >> 32) + (lo >> 32);
        yn[2 * j + 1] = (uint32_t)hi;
        c = hi >> 32;
    }
#endif
}

// y[0..7] += {p1,p3,p5,p7} * m  (y aligned on column 1; no carry out of column 8 by range).
ZKB_HD void chain_odd(uint32_t *y, uint32_t p1, uint32_t p3, uint32_t p5, uint32_t p7, uint32_t m)
{
#if ZKB_ASM
    asm("mad.lo.cc.u32  %0, %8,  %12, %0;\n\t"
        "madc.hi.cc.u32 %1, %8,  %12, %1;\n\t"
        "madc.lo.cc.u32 %2, %9,  %12, %2;\n\t"
        "madc.hi.cc.u32 %3, %9,  %12, %3;\n\t"
        "madc.lo.cc.u32 %4, %10, %12, %4;\n\t"
        "madc.hi.cc.u32 %5, %10, %12, %5;\n\t"
        "madc.lo.cc.u32 %6, %11, %12, %6;\n\t"
        "madc.hi.u32    %7, %11, %12, %7;"
        : "+r"(y[0]), "+r"(y[1]), "+r"(y[2]), "+r"(y[3]), "+r"(y[4]), "+r"(y[5]), "+r"(y[6]), "+r"(y[7])
        : "r"(p1), "r"(p3), "r"(p5), "r"(p7), "r"(m));
#else
    const uint32_t p[4] = {p1, p3, p5, p7};
    uint64_t c = 0;
    for (int j = 0; j < 4; j++) {
        const uint64_t prod = (uint64_t)p[j] * m;
        uint64_t lo = (uint64_t)y[2 * j] + (uint32_t)prod + c;
        y[2 * j] = (uint32_t)lo;
        uint64_t hi = (uint64_t)y[2 * j + 1] + (prod >> 32) + (lo >> 32);
        y[2 * j + 1] = (uint32_t)hi;
        c = hi >> 32;
    }
#endif
}

// r = a + b over 8 words, returns carry out.
ZKB_HD uint32_t add8(uint32_t *r, const uint32_t *a, const uint32_t *b)
{
    uint32_t c;
#if ZKB_ASM
    asm("add.cc.u32  %0, %9,  %17;\n\t"
        "addc.cc.u32 %1, %10, %18;\n\t"
        "addc.cc.u32 %2, %11, %19;\n\t"
        "addc.cc.u32 %3, %12, %20;\n\t"
        "addc.cc.u32 %4, %13, %21;\n\t"
        "addc.cc.u32 %5, %14, %22;\n\t"
        "addc.cc.u32 %6, %15, %23;\n\t"
        "addc.cc.u32 %7, %16, %24;\n\t"
        "addc.u32    %8, 0, 0;"
        : "=&r"(r[0]), "=&r"(r[1]), "=&r"(r[2]), "=&r"(r[3]), "=&r"(r[4]), "=&r"(r[5]), "=&r"(r[6]), "=&r"(r[7]), "=&r"(c)
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(a[7]), "r"(b[0]), "r"(b[1]),
          "r"(b[2]), "r"(b[3]), "r"(b[4]), "r"(b[5]), "r"(b[6]), "r"(b[7]));
#else
    uint64_t k = 0;
    uint32_t t[8];
    for (int i = 0; i < 8; i++) {
        k += (uint64_t)a[i] + b[i];
        t[i] = (uint32_t)k;
        k >>= 32;
    }
    for (int i = 0; i < 8; i++) r[i] = t[i];
    c = (uint32_t)k;
#endif
    return c;
}

// r = a - b over 8 words, returns borrow (1 if a < b).
ZKB_HD uint32_t sub8(uint32_t *r, const uint32_t *a, const uint32_t *b)
{
    uint32_t brw;
#if ZKB_ASM
    asm("sub.cc.u32  %0, %9,  %17;\n\t"
        "subc.cc.u32 %1, %10, %18;\n\t"
        "subc.cc.u32 %2, %11, %19;\n\t"
        "subc.cc.u32 %3, %12, %20;\n\t"
        "subc.cc.u32 %4, %13, %21;\n\t"
        "subc.cc.u32 %5, %14, %22;\n\t"
        "subc.cc.u32 %6, %15, %23;\n\t"
        "subc.cc.u32 %7, %16, %24;\n\t"
        "subc.u32    %8, 0, 0;"
        : "=&r"(r[0]), "=&r"(r[1]), "=&r"(r[2]), "=&r"(r[3]), "=&r"(r[4]), "=&r"(r[5]), "=&r"(r[6]), "=&r"(r[7]),
          "=&r"(brw)
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(a[7]), "r"(b[0]), "r"(b[1]),
          "r"(b[2]), "r"(b[3]), "r"(b[4]), "r"(b[5]), "r"(b[6]), "r"(b[7]));
    brw &= 1u; // subc of 0-0-borrow gives 0xffffffff when the chain borrowed
#else
    int64_t k = 0;
    uint32_t t[8];
    for (int i = 0; i < 8; i++) {
        k += (int64_t)a[i] - (int64_t)b[i];
        t[i] = (uint32_t)k;
        k >>= 32; // arithmetic shift keeps the borrow as -1
    }
    for (int i = 0; i < 8; i++) r[i] = t[i];
    brw = (k != 0);
#endif
    return brw;
}

// ---------------------------------------------------------------------------------------
// Field element
// ---------------------------------------------------------------------------------------
template <class PR> struct Fp {
    uint32_t v[8];
    typedef PR Params;

    static ZKB_HD Fp zero()
    {
        Fp r;
#pragma unroll
        for (int i = 0; i < 8; i++) r.v[i] = 0;
        return r;
    }
    static ZKB_HD Fp one()
    {
        Fp r;
#pragma unroll
        for (int i = 0; i < 8; i++) r.v[i] = PR::ONE(i);
        return r;
    }
    static ZKB_HD Fp r2()
    {
        Fp r;
#pragma unroll
        for (int i = 0; i < 8; i++) r.v[i] = PR::R2(i);
        return r;
    }
    ZKB_HD bool is_zero() const
    {
        uint32_t o = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) o |= v[i];
        return o == 0;
    }
    ZKB_HD bool operator==(const Fp &b) const
    {
        uint32_t o = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) o |= v[i] ^ b.v[i];
        return o == 0;
    }
    ZKB_HD bool operator!=(const Fp &b) const { return !(*this == b); }

    // canonical reduce of a value < 2p
    ZKB_HD void reduce_once()
    {
        uint32_t p[8], t[8];
#pragma unroll
        for (int i = 0; i < 8; i++) p[i] = PR::P(i);
        const uint32_t brw = sub8(t, v, p);
#pragma unroll
        for (int i = 0; i < 8; i++) v[i] = brw ? v[i] : t[i];
    }

    friend ZKB_HD Fp operator+(const Fp &a, const Fp &b)
    {
        Fp r;
        add8(r.v, a.v, b.v); // p < 2^254: no carry out of 256 bits
        r.reduce_once();
        return r;
    }
    friend ZKB_HD Fp operator-(const Fp &a, const Fp &b)
    {
        Fp r;
        uint32_t p[8], t[8];
        const uint32_t brw = sub8(r.v, a.v, b.v);
#pragma unroll
        for (int i = 0; i < 8; i++) p[i] = PR::P(i);
        add8(t, r.v, p);
#pragma unroll
        for (int i = 0; i < 8; i++) r.v[i] = brw ? t[i] : r.v[i];
        return r;
    }
    ZKB_HD Fp neg() const
    {
        if (is_zero()) return *this;
        Fp r;
        uint32_t p[8];
#pragma unroll
        for (int i = 0; i < 8; i++) p[i] = PR::P(i);
        sub8(r.v, p, v);
        return r;
    }
    ZKB_HD Fp dbl() const { return *this + *this; }

    // Montgomery product a*b*R^-1 mod p, canonical.
    friend ZKB_HD Fp operator*(const Fp &a, const Fp &b) { return mont_mul(a, b); }

    // Translation units that are latency- rather than throughput-bound define ZKB_NOINLINE_MUL:
    // the ~300-instruction product then becomes one function per field instead of being
    // expanded at every use (ptxas time drops by an order of magnitude).
    static ZKB_MUL_ATTR Fp mont_mul(const Fp &a, const Fp &b) { return mont_mul_inline(a, b); }
    // The product itself, always expanded in place: for callers that put several INDEPENDENT products into one basic
    // block (an Fq2 product is three) so that ptxas interleaves their carry chains -- a lone warp per scheduler
    // otherwise waits out every dependent IMAD.
    static ZKB_HD Fp mont_mul_inline(const Fp &a, const Fp &b)
    {
        uint32_t x[8], y[8];
        row_first(x, y, a.v, b.v[0]);
        {
            const uint32_t m = x[0] * PR::INV;
            chain_odd(y, PR::P(1), PR::P(3), PR::P(5), PR::P(7), m);
            chain_even(x, y[7], PR::P(0), PR::P(2), PR::P(4), PR::P(6), m);
        }
        // rows 1..7; the two accumulators swap roles each row
#pragma unroll
        for (int i = 1; i < 8; i++) {
            uint32_t yn[8];
            // after the shift: new column-0 accumulator = old y, new column-1 accumulator = old x >> 64
            chain_odd_shift(yn, y[0], x, a.v[1], a.v[3], a.v[5], a.v[7], b.v[i]);
            chain_even(y, yn[7], a.v[0], a.v[2], a.v[4], a.v[6], b.v[i]);
            const uint32_t m = y[0] * PR::INV;
            chain_odd(yn, PR::P(1), PR::P(3), PR::P(5), PR::P(7), m);
            chain_even(y, yn[7], PR::P(0), PR::P(2), PR::P(4), PR::P(6), m);
            // rename: x <- y (column 0), y <- yn (column 1)
#pragma unroll
            for (int k = 0; k < 8; k++) {
                x[k] = y[k];
                y[k] = yn[k];
            }
        }
        // result = (x >> 32) + y
        Fp r;
        uint32_t xs[8];
#pragma unroll
        for (int k = 0; k < 7; k++) xs[k] = x[k + 1];
        xs[7] = 0;
        add8(r.v, xs, y);
        r.reduce_once();
        return r;
    }
    // Squaring = the general product.  A dedicated square (28 cross products formed once and doubled + 8 diagonal ones:
    // 36 + 72 = 108 wide multiply-adds instead of 136, bit-exact on every GPU test) was built and measured in round 2 and
    // REJECTED: the bucket accumulation, whose mixed addition holds 2 squares in 10 products, went from 6.17 ms to 6.29 ms
    // per launch (one column walk) and 6.37 ms (two interleaved walks) -- the ~130 extra add / shift / move instructions
    // and the longer dependent chain cost more than the 4 % of multiplier-pipe work saved; batch verification: unchanged.
    ZKB_HD Fp sqr() const { return (*this) * (*this); }

    // out of / into Montgomery form (fp.tcc:228-238 as_bigint; :190-196 constructor)
    ZKB_HD Fp from_mont() const
    {
        Fp o = zero();
        o.v[0] = 1;
        return (*this) * o;
    }
    ZKB_HD Fp to_mont() const { return (*this) * r2(); }

    // a^(p-2) (fp.tcc:642-685 computes the same value by extended gcd); 0 -> 0
    ZKB_HD Fp inverse() const
    {
        Fp res = one();
        Fp base = *this;
        uint32_t e[8];
#pragma unroll
        for (int i = 0; i < 8; i++) e[i] = PR::P(i);
        e[0] -= 2; // p is odd and P[0] >= 2 for both fields
        for (int i = 0; i < 256; i++) {
            if ((e[i >> 5] >> (i & 31)) & 1) res = res * base;
            base = base.sqr();
        }
        return res;
    }
};

typedef Fp<FqP> Fq;
typedef Fp<FrP> Fr;

// ---------------------------------------------------------------------------------------
// Fq2 = Fq[u]/(u^2+1)   (libff/algebra/fields/fp2.tcc:52-136; non-residue -1, alt_bn128_init.cpp:95)
// ---------------------------------------------------------------------------------------
struct Fq2 {
    Fq c0, c1;
    static ZKB_HD Fq2 zero() { return Fq2{Fq::zero(), Fq::zero()}; }
    static ZKB_HD Fq2 one() { return Fq2{Fq::one(), Fq::zero()}; }
    ZKB_HD bool is_zero() const { return c0.is_zero() && c1.is_zero(); }
    ZKB_HD bool operator==(const Fq2 &b) const { return c0 == b.c0 && c1 == b.c1; }
    ZKB_HD bool operator!=(const Fq2 &b) const { return !(*this == b); }
    friend ZKB_HD Fq2 operator+(const Fq2 &a, const Fq2 &b) { return Fq2{a.c0 + b.c0, a.c1 + b.c1}; }
    friend ZKB_HD Fq2 operator-(const Fq2 &a, const Fq2 &b) { return Fq2{a.c0 - b.c0, a.c1 - b.c1}; }
    ZKB_HD Fq2 neg() const { return Fq2{c0.neg(), c1.neg()}; }
    ZKB_HD Fq2 dbl() const { return Fq2{c0.dbl(), c1.dbl()}; }
    // Karatsuba, 3 base-field products (fp2.tcc:73-84)
    friend ZKB_HD Fq2 operator*(const Fq2 &a, const Fq2 &b)
    {
        const Fq aa = a.c0 * b.c0;
        const Fq bb = a.c1 * b.c1;
        const Fq s = (a.c0 + a.c1) * (b.c0 + b.c1);
        return Fq2{aa - bb, s - aa - bb};
    }
    // complex squaring, 2 base-field products (fp2.tcc:94-120)
    ZKB_HD Fq2 sqr() const
    {
        const Fq ab = c0 * c1;
        const Fq t = (c0 + c1) * (c0 - c1);
        return Fq2{t, ab.dbl()};
    }
    // (fp2.tcc:123-136)
    ZKB_HD Fq2 inverse() const
    {
        const Fq t = (c0.sqr() + c1.sqr()).inverse();
        return Fq2{c0 * t, (c1 * t).neg()};
    }
};

} // namespace zkb
