// glv.cuh -- the efficient endomorphism of alt_bn128 G1 for the two variable-base scalar multiplications of the proof
// assembly (s * g1_A and r * g1_B, r1cs_gg_ppzksnark.tcc:495; the reference runs them as plain double-and-add through
// libff's operator*, scalar_multiplication/wnaf.tcc / alt_bn128_g1.cpp:240-330).
//
// The curve y^2 = x^3 + 3 over Fq has phi(x, y) = (beta x, y) with beta^3 = 1, and on the order-r group phi acts as the
// multiplication by lambda, lambda^2 + lambda + 1 = 0 (mod r).  A scalar k < 2^256 is split as
//     k = k1 + k2 lambda (mod r),   0 <= k1 < 2^128,   -2^128 < k2 <= 0,
// by rounding (k, 0) down onto the lattice {(a, b): a + b lambda = 0 (mod r)} spanned by the two short vectors below
// (found with the extended Euclidean algorithm on (r, lambda); a1 b2 - a2 b1 = r).  k P = k1 P + |k2| (-phi(P)): one
// chain of 128 doublings instead of 254.  The result is the same group element; only its projective representation
// differs from the plain chain's (proofs are compared in affine / serialised form).
//
// Plain C++ (no CUDA types) so that tests/host/glv_host.cpp can run this very code on the CPU (tests/test_glv_host.py).
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define ZKB_GLV_FN __host__ __device__ __forceinline__
#else
#define ZKB_GLV_FN inline
#endif

namespace glv {

// v1 = (a1, b1) = (0x89d3256894d213e3, -0x6f4d8248eeb859fc8211bbeb7d4f1128), v2 = (a2, b2) = (0x6f4d8248eeb859fd0be4e1541221250b, 0x89d3256894d213e3)
// g1 = floor(2^256 b2 / r), g2 = floor(2^256 |b1| / r)
#define ZKB_GLV_G1 {0xc7e0b3d7u, 0xd91d232eu, 0x00000002u}
#define ZKB_GLV_G2 {0x391eb18du, 0x7a7bd9d4u, 0xa773d2cfu, 0x4ccef014u, 0x00000002u}
#define ZKB_GLV_A1 {0x94d213e3u, 0x89d32568u}
#define ZKB_GLV_NB1 {0x7d4f1128u, 0x8211bbebu, 0xeeb859fcu, 0x6f4d8248u}
#define ZKB_GLV_A2 {0x1221250bu, 0x0be4e154u, 0xeeb859fdu, 0x6f4d8248u}
#define ZKB_GLV_B2 {0x94d213e3u, 0x89d32568u}
// beta = 2203960485148121921418603742825762020974279258880205651966 in Montgomery form (R = 2^256)
#define ZKB_GLV_BETA_MONT {0xd782e155u, 0x71930c11u, 0xffbe3323u, 0xa6bb947cu, 0xd4741444u, 0xaa303344u, 0x26594943u, 0x2c3b3f0du}

// out[0 .. NO) = limbs [skip, skip + NO) of a * b (a: NA limbs, b: NB limbs)
template <int NA, int NB, int NO> ZKB_GLV_FN void mul_window(const uint32_t *a, const uint32_t *b, int skip, uint32_t *out)
{
    uint32_t full[NA + NB];
    for (int i = 0; i < NA + NB; i++) full[i] = 0;
    for (int i = 0; i < NA; i++) {
        uint64_t carry = 0;
        for (int j = 0; j < NB; j++) {
            const uint64_t t = (uint64_t)a[i] * b[j] + full[i + j] + carry;
            full[i + j] = (uint32_t)t;
            carry = t >> 32;
        }
        full[i + NB] = (uint32_t)carry;
    }
    for (int i = 0; i < NO; i++) out[i] = skip + i < NA + NB ? full[skip + i] : 0u;
}

// x -= y over N limbs (mod 2^(32 N))
template <int N> ZKB_GLV_FN void sub_n(uint32_t *x, const uint32_t *y)
{
    uint64_t borrow = 0;
    for (int i = 0; i < N; i++) {
        const uint64_t t = (uint64_t)x[i] - y[i] - borrow;
        x[i] = (uint32_t)t;
        borrow = (t >> 32) & 1u;
    }
}

// k (8 limbs, any value < 2^256) -> k1 and |k2| as 5 limbs each (both < 2^128: the fifth limb is zero; kept so that a
// digit loop may read one window past the top).  k = k1 - |k2| lambda (mod r).
ZKB_GLV_FN void decompose(const uint32_t k[8], uint32_t k1[5], uint32_t k2abs[5])
{
    const uint32_t g1[3] = ZKB_GLV_G1, g2[5] = ZKB_GLV_G2, a1[2] = ZKB_GLV_A1, nb1[4] = ZKB_GLV_NB1, a2[4] = ZKB_GLV_A2, b2[2] = ZKB_GLV_B2;
    uint32_t c1[3], c2[5], t[5];
    mul_window<8, 3, 3>(k, g1, 8, c1); // floor(k g1 / 2^256) < 2^66
    mul_window<8, 5, 5>(k, g2, 8, c2); // floor(k g2 / 2^256) < 2^130
    // k1 = k - c1 a1 - c2 a2: non-negative and < 2^128, so 160 bits of every term are enough
    for (int i = 0; i < 5; i++) k1[i] = k[i];
    mul_window<3, 2, 5>(c1, a1, 0, t);
    sub_n<5>(k1, t);
    mul_window<5, 4, 5>(c2, a2, 0, t);
    sub_n<5>(k1, t);
    // k2 = c1 |b1| - c2 b2 <= 0
    mul_window<5, 2, 5>(c2, b2, 0, k2abs);
    mul_window<3, 4, 5>(c1, nb1, 0, t);
    sub_n<5>(k2abs, t);
}

} // namespace glv
