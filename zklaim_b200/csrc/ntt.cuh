// ntt.cuh -- evaluation domains over Fr (see ntt.cu for the design).
#pragma once
#include "common.cuh"

namespace zkb {

// twiddles of one power-of-two size n = 2^logn: fwd[i] = w^i, inv[i] = w^-i for i < n/2,
// w = libff::get_root_of_unity<Fr>(n) (libff/algebra/fields/field_utils.tcc:44-60)
struct R2Tables {
    int logn = 0;
    Fr *fwd = nullptr, *inv = nullptr;
};

// libfqfft evaluation domain of size m as get_evaluation_domain picks it
// (libfqfft/evaluation_domain/get_evaluation_domain.tcc:33-52)
struct Domain {
    int kind = 0;
    int dev = 0;
    size_t m = 0;
    // basic_radix2: big_m = m.  step_radix2: m = big_m + small_m (step_radix2_domain.tcc:21-37)
    size_t big_m = 0, small_m = 0;
    R2Tables big, small;
    // basic: cos_fwd[i] = g^i / m and cos_inv[i] = g^-i / m (the 1/m of iFFT folded in);
    // step:  cos_fwd[i] = g^i and cos_inv[i] = g^-i;  g = Fr::multiplicative_generator, i < m
    Fr *cos_fwd = nullptr, *cos_inv = nullptr;
    Fr zinv;              // basic: 1 / (g^m - 1)                       (basic_radix2_domain.tcc:102-110)
    Fr *zinv_tab = nullptr; // step: 1 / Z(g x_i) per element of the coset (step_radix2_domain.tcc:222-246)
    Fr *om_fwd = nullptr; // step: omega^i, i < big_m   (omega = root of unity of order 2 * big_m)
    Fr *om_inv = nullptr; // step: omega^-i, i < small_m
    size_t bytes = 0;
};

int domain_select_host(size_t min_size, size_t *m_out, int *kind_out);
// host-side constants of a domain (a handful of field operations per call, portable path of fp.cuh)
Fr host_fr_from_u64(uint64_t x);
Fr host_fr_pow(Fr b, uint64_t e);
Fr host_fr_root_of_unity(int logn); // libff::get_root_of_unity<Fr>(2^logn), field_utils.tcc:44-60
int host_ceil_log2(size_t n);       // libff::log2
int domain_build(zkb_ctx *ctx, size_t m, int kind, Domain *d);
void domain_free(Domain *d);

// All operations are batched: `nvec` vectors of m elements, vector v at a + v * vstride.
// tmp: scratch of the same shape (used by the natural-order paths).
// a <- cosetFFT(iFFT(a), g): evaluations on the domain -> evaluations on the coset (r1cs_to_qap.tcc:239-268)
int domain_to_coset(zkb_ctx *ctx, const Domain *d, Fr *a, size_t vstride, int nvec, Fr *tmp);
// h <- icosetFFT(a, g) (r1cs_to_qap.tcc:306-311); a is destroyed.  divide_by_Z_on_coset (:301-303) is the caller's:
// it multiplies by Domain::zinv (basic) or Domain::zinv_tab[i] (step) in its pointwise pass.
// Writes m coefficients in natural order to h_mont + v * hstride (Montgomery) and, when
// h_plain != NULL, their plain-integer form to h_plain + v * hstride.
int domain_h_from_coset(zkb_ctx *ctx, const Domain *d, Fr *a, size_t vstride, int nvec, Fr *tmp, Fr *h_mont,
                        Fr *h_plain, size_t hstride);
// the same with the pointwise step (A B - C) / Z folded in (basic domain of >= 256 points; else ZKB_ERR_DOMAIN and nothing
// is enqueued): abc = [A | B | C], nvec vectors each, on the coset; A is destroyed
int domain_h_fused(zkb_ctx *ctx, const Domain *d, Fr *abc, int nvec, Fr *h_mont, Fr *h_plain, size_t hstride);
// single natural-order operation (ZKB_NTT_*), in place
int domain_op(zkb_ctx *ctx, const Domain *d, int op, Fr *a, size_t vstride, int nvec, Fr *tmp);

} // namespace zkb
