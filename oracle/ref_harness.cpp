/*
 * oracle/ref_harness.cpp -- C-callable harness around the UNMODIFIED reference sources.
 *
 * TEST INFRASTRUCTURE.  Nothing in the product path (zklaim_b200/, include/) links or
 * loads this.  It is compiled by oracle/Makefile together with the reference's own
 * libff / libfqfft / libsnark / zklaim sources, read in place from /root/reference,
 * into oracle/_ref/libzkref*.so.  It is the ground truth the CPU restatement
 * (oracle/zkoracle.c) and the CUDA path are pinned against, and the timed
 * "reference" CPU baseline of bench.py.
 *
 * Everything here is a thin marshalling layer: raw little-endian 64-bit limbs in and
 * out of the reference's own classes.  The only algorithmic lines are
 * `prove_with_rs` -- the reference prover's orchestration
 * (libsnark/zk_proof_systems/ppzksnark/r1cs_gg_ppzksnark/r1cs_gg_ppzksnark.tcc:390-505)
 * re-expressed with the two blinding scalars r, s (:418-419, drawn there from
 * std::random_device) as parameters, because the reference offers no way to fix them.
 * Every arithmetic call inside it is the reference's.
 *
 * Layout conventions (shared with include/zkb.h):
 *   Fr / Fq element : 4 x u64, Montgomery form, canonical (< modulus)   [fp.hpp:42]
 *   Fq2             : c0 then c1                                        [fp2.hpp]
 *   G1 Jacobian     : X, Y, Z  = 12 x u64                               [alt_bn128_g1.hpp:35]
 *   G2 Jacobian     : X, Y, Z over Fq2 = 24 x u64
 *   G1 affine       : x, y = 8 x u64; the point at infinity is encoded as x = y = 0
 *   G2 affine       : x, y over Fq2 = 16 x u64; infinity = all zero
 */
#include <chrono>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <memory>
#include <sstream>
#include <vector>

#include "zklaim/libsnark_wrapper.cpp" /* the reference bridge; pulls snark.cpp + zklaim_gadget.cpp */

#include <libff/algebra/curves/alt_bn128/alt_bn128_pp.hpp>
#include <libff/algebra/scalar_multiplication/multiexp.hpp>
#include <libff/common/rng.hpp>
#include <libfqfft/evaluation_domain/get_evaluation_domain.hpp>
#include <libfqfft/evaluation_domain/domains/basic_radix2_domain.hpp>
#include <libfqfft/evaluation_domain/domains/step_radix2_domain.hpp>
#include <libsnark/gadgetlib1/gadgets/hashes/sha256/sha256_gadget.hpp>
#include <libsnark/knowledge_commitment/kc_multiexp.hpp>
#include <libsnark/reductions/r1cs_to_qap/r1cs_to_qap.hpp>
#include <libsnark/zk_proof_systems/ppzksnark/r1cs_gg_ppzksnark/r1cs_gg_ppzksnark.hpp>

#ifdef MULTICORE
#include <omp.h>
#endif

typedef libff::alt_bn128_pp PP;
typedef libff::Fr<PP> FrT;
typedef libff::Fq<PP> FqT;
typedef libff::Fqe<PP> Fq2T;
typedef libff::G1<PP> G1T;
typedef libff::G2<PP> G2T;
typedef uint64_t u64;

namespace {

double now_s()
{
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

template <typename F> inline F ld(const u64 *p)
{
    F x;
    memcpy(x.mont_repr.data, p, 32);
    return x;
}
template <typename F> inline void st(u64 *p, const F &x) { memcpy(p, x.mont_repr.data, 32); }

inline Fq2T ld2(const u64 *p) { return Fq2T(ld<FqT>(p), ld<FqT>(p + 4)); }
inline void st2(u64 *p, const Fq2T &x)
{
    st(p, x.c0);
    st(p + 4, x.c1);
}

inline G1T ld_g1(const u64 *p) { return G1T(ld<FqT>(p), ld<FqT>(p + 4), ld<FqT>(p + 8)); }
inline void st_g1(u64 *p, const G1T &g)
{
    st(p, g.X);
    st(p + 4, g.Y);
    st(p + 8, g.Z);
}
inline G2T ld_g2(const u64 *p) { return G2T(ld2(p), ld2(p + 8), ld2(p + 16)); }
inline void st_g2(u64 *p, const G2T &g)
{
    st2(p, g.X);
    st2(p + 8, g.Y);
    st2(p + 16, g.Z);
}

inline bool all_zero(const u64 *p, size_t n)
{
    for (size_t i = 0; i < n; i++)
        if (p[i]) return false;
    return true;
}
inline G1T ld_g1_affine(const u64 *p)
{
    if (all_zero(p, 8)) return G1T::zero();
    return G1T(ld<FqT>(p), ld<FqT>(p + 4), FqT::one());
}
inline void st_g1_affine(u64 *p, const G1T &g)
{
    if (g.is_zero()) {
        memset(p, 0, 64);
        return;
    }
    G1T c(g);
    c.to_affine_coordinates();
    st(p, c.X);
    st(p + 4, c.Y);
}
inline G2T ld_g2_affine(const u64 *p)
{
    if (all_zero(p, 16)) return G2T::zero();
    return G2T(ld2(p), ld2(p + 8), Fq2T::one());
}
inline void st_g2_affine(u64 *p, const G2T &g)
{
    if (g.is_zero()) {
        memset(p, 0, 128);
        return;
    }
    G2T c(g);
    c.to_affine_coordinates();
    st2(p, c.X);
    st2(p + 8, c.Y);
}

struct CoutSilencer {
    std::streambuf *saved;
    CoutSilencer() : saved(std::cout.rdbuf(nullptr)) {}
    ~CoutSilencer()
    {
        std::cout.rdbuf(saved);
        std::cout.clear();
    }
};

bool g_inited = false;

struct Circuit {
    r1cs_constraint_system<FrT> cs;           /* as built by the gadget / example */
    std::vector<FrT> primary, auxiliary;      /* a satisfying assignment (may be empty) */
    bool has_key = false;
    r1cs_gg_ppzksnark_proving_key<PP> pk;
    r1cs_gg_ppzksnark_verification_key<PP> vk;
    const r1cs_constraint_system<FrT> &prover_cs() const { return has_key ? pk.constraint_system : cs; }
};

/* zklaim credential with `k` payloads, the values of zklaim/main.c:70-101,194-208,
 * salt derived from `seed` instead of /dev/urandom (zklaim/zklaim.c:114-120). */
struct ZkCtx {
    zklaim_ctx ctx;
    std::vector<zklaim_wrap_payload_ctx> nodes;
    ZkCtx(size_t k, u64 seed) : nodes(k)
    {
        memset(&ctx, 0, sizeof(ctx));
        for (size_t i = 0; i < k; i++) {
            zklaim_payload &pl = nodes[i].pl;
            memset(&pl, 0, sizeof(pl));
            const u64 attrs[5] = {23 + i, 1, 2, 3, 599};
            memcpy(pl.pre, attrs, 40);
            pl.salt = 0x9e3779b97f4a7c15ull * (seed + 1) + 0xbf58476d1ce4e5b9ull * (i + 1);
            memcpy(pl.pre + 40, &pl.salt, 8);
            SHA256(pl.pre, sizeof(pl.pre), pl.hash);
            pl.data_ref[0] = 20;
            pl.data_op[0] = (enum zklaim_op)(zklaim_greater | zklaim_eq);
            for (int j = 1; j < 5; j++) {
                pl.data_ref[j] = 0;
                pl.data_op[j] = zklaim_noop;
            }
            nodes[i].next = (i + 1 < k) ? &nodes[i + 1] : NULL;
        }
        ctx.num_of_payloads = k;
        ctx.pl_ctx_head = k ? &nodes[0] : NULL;
    }
};

/* The reference prover with r, s injected; see the header comment. */
struct ProveOut {
    G1T A, C, At, Bt1, Ht, Lt;
    G2T B, Bt2;
    std::vector<FrT> H;
    double t_wm, t_A, t_B, t_H, t_L, t_total;
};

ProveOut prove_with_rs(const r1cs_gg_ppzksnark_proving_key<PP> &pk, const std::vector<FrT> &primary,
                       const std::vector<FrT> &auxiliary, const FrT &r, const FrT &s)
{
    ProveOut o;
    const double t0 = now_s();
    const qap_witness<FrT> qw =
        r1cs_to_qap_witness_map(pk.constraint_system, primary, auxiliary, FrT::zero(), FrT::zero(), FrT::zero());
    const double t1 = now_s();
#ifdef MULTICORE
    const size_t chunks = omp_get_max_threads();
#else
    const size_t chunks = 1;
#endif
    const size_t nv = qw.num_variables(), ni = qw.num_inputs(), deg = qw.degree();
    std::vector<FrT> padded(1, FrT::one());
    padded.insert(padded.end(), qw.coefficients_for_ABCs.begin(), qw.coefficients_for_ABCs.end());

    o.At = libff::multi_exp_with_mixed_addition<G1T, FrT, libff::multi_exp_method_BDLO12>(
        pk.A_query.begin(), pk.A_query.begin() + nv + 1, padded.begin(), padded.begin() + nv + 1, chunks);
    const double t2 = now_s();
    const knowledge_commitment<G2T, G1T> Bt =
        kc_multi_exp_with_mixed_addition<G2T, G1T, FrT, libff::multi_exp_method_BDLO12>(
            pk.B_query, 0, nv + 1, padded.begin(), padded.begin() + nv + 1, chunks);
    const double t3 = now_s();
    o.Ht = libff::multi_exp<G1T, FrT, libff::multi_exp_method_BDLO12>(
        pk.H_query.begin(), pk.H_query.begin() + (deg - 1), qw.coefficients_for_H.begin(),
        qw.coefficients_for_H.begin() + (deg - 1), chunks);
    const double t4 = now_s();
    o.Lt = libff::multi_exp_with_mixed_addition<G1T, FrT, libff::multi_exp_method_BDLO12>(
        pk.L_query.begin(), pk.L_query.end(), padded.begin() + ni + 1, padded.begin() + nv + 1, chunks);
    const double t5 = now_s();

    o.Bt1 = Bt.h;
    o.Bt2 = Bt.g;
    o.A = pk.alpha_g1 + o.At + r * pk.delta_g1;
    const G1T g1_B = pk.beta_g1 + Bt.h + s * pk.delta_g1;
    o.B = pk.beta_g2 + Bt.g + s * pk.delta_g2;
    o.C = o.Ht + o.Lt + s * o.A + r * g1_B - (r * s) * pk.delta_g1;
    const double t6 = now_s();
    o.H = qw.coefficients_for_H;
    o.t_wm = t1 - t0;
    o.t_A = t2 - t1;
    o.t_B = t3 - t2;
    o.t_H = t4 - t3;
    o.t_L = t5 - t4;
    o.t_total = t6 - t0;
    return o;
}

template <typename F> void field_op(int op, const u64 *a, const u64 *b, u64 *out, size_t n)
{
    for (size_t i = 0; i < n; i++) {
        const F x = ld<F>(a + 4 * i);
        const F y = b ? ld<F>(b + 4 * i) : F::zero();
        F z;
        switch (op) {
        case 0: z = x + y; break;
        case 1: z = x - y; break;
        case 2: z = x * y; break;
        case 3: z = x.squared(); break;
        case 4: z = x.is_zero() ? F::zero() : x.inverse(); break;
        case 5: z = -x; break;
        default: z = F::zero();
        }
        st(out + 4 * i, z);
    }
}

} // namespace

extern "C" {

int zkref_init(void)
{
    if (!g_inited) {
        libff::inhibit_profiling_info = true;
        libff::inhibit_profiling_counters = true;
        PP::init_public_params();
        g_inited = true;
    }
    return 0;
}

int zkref_num_threads(void)
{
#ifdef MULTICORE
    return omp_get_max_threads();
#else
    return 1;
#endif
}

int zkref_build_flags(void)
{
    int f = 0;
#ifdef MULTICORE
    f |= 1;
#endif
#ifdef USE_ASM
    f |= 2;
#endif
    return f;
}

/* ---- fields ------------------------------------------------------------------ */
/* field: 0 = Fr, 1 = Fq.  op: 0 add, 1 sub, 2 mul, 3 sqr, 4 inv (0 -> 0), 5 neg. */
void zkref_field_op(int field, int op, const u64 *a, const u64 *b, u64 *out, size_t n)
{
    if (field == 0)
        field_op<FrT>(op, a, b, out, n);
    else
        field_op<FqT>(op, a, b, out, n);
}

/* Fq2 op: 0 add, 1 sub, 2 mul, 3 sqr, 4 inv (0 -> 0), 5 neg.  8 limbs per element. */
void zkref_fq2_op(int op, const u64 *a, const u64 *b, u64 *out, size_t n)
{
    for (size_t i = 0; i < n; i++) {
        const Fq2T x = ld2(a + 8 * i);
        const Fq2T y = b ? ld2(b + 8 * i) : Fq2T::zero();
        Fq2T z;
        switch (op) {
        case 0: z = x + y; break;
        case 1: z = x - y; break;
        case 2: z = x * y; break;
        case 3: z = x.squared(); break;
        case 4: z = x.is_zero() ? Fq2T::zero() : x.inverse(); break;
        case 5: z = -x; break;
        default: z = Fq2T::zero();
        }
        st2(out + 8 * i, z);
    }
}

/* Montgomery <-> plain integer limbs (fp.tcc:228-238 as_bigint; fp.tcc:190-196 ctor). */
void zkref_field_from_plain(int field, const u64 *plain, u64 *out, size_t n)
{
    for (size_t i = 0; i < n; i++) {
        libff::bigint<4> b;
        memcpy(b.data, plain + 4 * i, 32);
        if (field == 0)
            st(out + 4 * i, FrT(b));
        else
            st(out + 4 * i, FqT(b));
    }
}
void zkref_field_to_plain(int field, const u64 *a, u64 *out, size_t n)
{
    for (size_t i = 0; i < n; i++) {
        libff::bigint<4> b = (field == 0) ? ld<FrT>(a + 4 * i).as_bigint() : ld<FqT>(a + 4 * i).as_bigint();
        memcpy(out + 4 * i, b.data, 32);
    }
}

/* libff::SHA512_rng<Fr|Fq>(first + i) for i in [0, n)  (libff/common/rng.tcc:26-72). */
void zkref_sha512_rng(int field, u64 first, size_t n, u64 *out)
{
#ifdef MULTICORE
#pragma omp parallel for
#endif
    for (size_t i = 0; i < n; i++) {
        if (field == 0)
            st(out + 4 * i, libff::SHA512_rng<FrT>(first + i));
        else
            st(out + 4 * i, libff::SHA512_rng<FqT>(first + i));
    }
}

/* constants: which = 0 modulus (plain), 1 R mod p (= one, Montgomery), 2 Rsquared (plain
 * limbs of R^2 mod p), 3 multiplicative_generator (Montgomery), 4 root_of_unity (Montgomery,
 * Fr only), 5 inv (-p^-1 mod 2^64, in limb 0), 6 s (two-adicity, limb 0). */
void zkref_field_const(int field, int which, u64 *out)
{
    memset(out, 0, 32);
    if (field == 0) {
        switch (which) {
        case 0: memcpy(out, FrT::mod.data, 32); break;
        case 1: st(out, FrT::one()); break;
        case 2: memcpy(out, FrT::Rsquared.data, 32); break;
        case 3: st(out, FrT::multiplicative_generator); break;
        case 4: st(out, FrT::root_of_unity); break;
        case 5: out[0] = FrT::inv; break;
        case 6: out[0] = FrT::s; break;
        }
    } else {
        switch (which) {
        case 0: memcpy(out, FqT::mod.data, 32); break;
        case 1: st(out, FqT::one()); break;
        case 2: memcpy(out, FqT::Rsquared.data, 32); break;
        case 3: st(out, FqT::multiplicative_generator); break;
        case 5: out[0] = FqT::inv; break;
        case 6: out[0] = FqT::s; break;
        }
    }
}

/* ---- groups ------------------------------------------------------------------ */
/* op: 0 add (operator+), 1 dbl, 2 mixed_add (b must have Z = 1 or be zero), 3 neg,
 *     4 to_affine -> writes Jacobian with Z = 1 (or the reference's zero (0,1,0)),
 *     5 add via .add() (no doubling check; alt_bn128_g1.cpp:208-254). */
void zkref_g1_op(int op, const u64 *a, const u64 *b, u64 *out, size_t n)
{
    for (size_t i = 0; i < n; i++) {
        const G1T x = ld_g1(a + 12 * i);
        const G1T y = b ? ld_g1(b + 12 * i) : G1T::zero();
        G1T z;
        switch (op) {
        case 0: z = x + y; break;
        case 1: z = x.dbl(); break;
        case 2: z = x.mixed_add(y); break;
        case 3: z = -x; break;
        case 4: z = x; z.to_special(); break;
        case 5: z = x.add(y); break;
        default: z = G1T::zero();
        }
        st_g1(out + 12 * i, z);
    }
}
void zkref_g2_op(int op, const u64 *a, const u64 *b, u64 *out, size_t n)
{
    for (size_t i = 0; i < n; i++) {
        const G2T x = ld_g2(a + 24 * i);
        const G2T y = b ? ld_g2(b + 24 * i) : G2T::zero();
        G2T z;
        switch (op) {
        case 0: z = x + y; break;
        case 1: z = x.dbl(); break;
        case 2: z = x.mixed_add(y); break;
        case 3: z = -x; break;
        case 4: z = x; z.to_special(); break;
        case 5: z = x.add(y); break;
        default: z = G2T::zero();
        }
        st_g2(out + 24 * i, z);
    }
}
int zkref_g1_equal(const u64 *a, const u64 *b) { return ld_g1(a) == ld_g1(b); }
int zkref_g2_equal(const u64 *a, const u64 *b) { return ld_g2(a) == ld_g2(b); }
int zkref_g1_well_formed(const u64 *a) { return ld_g1(a).is_well_formed(); }
int zkref_g2_well_formed(const u64 *a) { return ld_g2(a).is_well_formed(); }

/* scalar (Montgomery Fr) * point, the reference's operator* (curve_utils.tcc:13-34). */
void zkref_g1_scalar_mul(const u64 *scalars, const u64 *pts, u64 *out, size_t n)
{
    for (size_t i = 0; i < n; i++) st_g1(out + 12 * i, ld<FrT>(scalars + 4 * i) * ld_g1(pts + 12 * i));
}
void zkref_g2_scalar_mul(const u64 *scalars, const u64 *pts, u64 *out, size_t n)
{
    for (size_t i = 0; i < n; i++) st_g2(out + 24 * i, ld<FrT>(scalars + 4 * i) * ld_g2(pts + 24 * i));
}
void zkref_g1_generator(u64 *out) { st_g1(out, G1T::one()); }
void zkref_g2_generator(u64 *out) { st_g2(out, G2T::one()); }

/* n distinct affine bases k_i * G, k_i = SHA512_rng<Fr>(first + i), through the
 * reference's fixed-base tables (multiexp.tcc:509-646) and batch_to_special (:683-715). */
void zkref_g1_bases(u64 first, size_t n, u64 *out_affine)
{
    std::vector<FrT> k(n);
#ifdef MULTICORE
#pragma omp parallel for
#endif
    for (size_t i = 0; i < n; i++) k[i] = libff::SHA512_rng<FrT>(first + i);
    const size_t w = libff::get_exp_window_size<G1T>(n);
    const size_t bits = FrT::size_in_bits();
    libff::window_table<G1T> tab = libff::get_window_table(bits, w, G1T::one());
    std::vector<G1T> v = libff::batch_exp(bits, w, tab, k);
    libff::batch_to_special<G1T>(v);
    for (size_t i = 0; i < n; i++) st_g1_affine(out_affine + 8 * i, v[i]);
}
void zkref_g2_bases(u64 first, size_t n, u64 *out_affine)
{
    std::vector<FrT> k(n);
    for (size_t i = 0; i < n; i++) k[i] = libff::SHA512_rng<FrT>(first + i);
    const size_t w = libff::get_exp_window_size<G2T>(n);
    const size_t bits = FrT::size_in_bits();
    libff::window_table<G2T> tab = libff::get_window_table(bits, w, G2T::one());
    std::vector<G2T> v = libff::batch_exp(bits, w, tab, k);
    libff::batch_to_special<G2T>(v);
    for (size_t i = 0; i < n; i++) st_g2_affine(out_affine + 16 * i, v[i]);
}

/* MSM through the reference's BDLO12 path.  mixed = 1: multi_exp_with_mixed_addition
 * (multiexp.tcc:443-496), 0: multi_exp (:402-441).  chunks = 0: omp_get_max_threads().
 * Returns seconds spent inside the reference call. */
double zkref_g1_msm(const u64 *bases_affine, const u64 *scalars, size_t n, int mixed, size_t chunks, u64 *out_jac)
{
    std::vector<G1T> b(n);
    std::vector<FrT> s(n);
    for (size_t i = 0; i < n; i++) {
        b[i] = ld_g1_affine(bases_affine + 8 * i);
        s[i] = ld<FrT>(scalars + 4 * i);
    }
    if (chunks == 0) chunks = zkref_num_threads();
    const double t0 = now_s();
    G1T r = mixed ? libff::multi_exp_with_mixed_addition<G1T, FrT, libff::multi_exp_method_BDLO12>(
                        b.begin(), b.end(), s.begin(), s.end(), chunks)
                  : libff::multi_exp<G1T, FrT, libff::multi_exp_method_BDLO12>(b.begin(), b.end(), s.begin(),
                                                                              s.end(), chunks);
    const double t1 = now_s();
    st_g1(out_jac, r);
    return t1 - t0;
}
double zkref_g2_msm(const u64 *bases_affine, const u64 *scalars, size_t n, int mixed, size_t chunks, u64 *out_jac)
{
    std::vector<G2T> b(n);
    std::vector<FrT> s(n);
    for (size_t i = 0; i < n; i++) {
        b[i] = ld_g2_affine(bases_affine + 16 * i);
        s[i] = ld<FrT>(scalars + 4 * i);
    }
    if (chunks == 0) chunks = zkref_num_threads();
    const double t0 = now_s();
    G2T r = mixed ? libff::multi_exp_with_mixed_addition<G2T, FrT, libff::multi_exp_method_BDLO12>(
                        b.begin(), b.end(), s.begin(), s.end(), chunks)
                  : libff::multi_exp<G2T, FrT, libff::multi_exp_method_BDLO12>(b.begin(), b.end(), s.begin(),
                                                                              s.end(), chunks);
    const double t1 = now_s();
    st_g2(out_jac, r);
    return t1 - t0;
}

/* ---- evaluation domains ------------------------------------------------------ */
/* kind: 1 basic_radix2, 2 extended_radix2, 3 step_radix2, 4 geometric, 5 arithmetic
 * (get_evaluation_domain.tcc:33-52).  Returns 0 on success. */
int zkref_domain_select(size_t min_size, size_t *m_out, int *kind_out)
{
    try {
        std::shared_ptr<libfqfft::evaluation_domain<FrT>> d = libfqfft::get_evaluation_domain<FrT>(min_size);
        *m_out = d->m;
        if (dynamic_cast<libfqfft::basic_radix2_domain<FrT> *>(d.get()))
            *kind_out = 1;
        else if (dynamic_cast<libfqfft::extended_radix2_domain<FrT> *>(d.get()))
            *kind_out = 2;
        else if (dynamic_cast<libfqfft::step_radix2_domain<FrT> *>(d.get()))
            *kind_out = 3;
        else if (dynamic_cast<libfqfft::geometric_sequence_domain<FrT> *>(d.get()))
            *kind_out = 4;
        else
            *kind_out = 5;
        return 0;
    } catch (...) {
        return -1;
    }
}

/* op: 0 FFT, 1 iFFT, 2 cosetFFT(g), 3 icosetFFT(g), 4 divide_by_Z_on_coset,
 *     5 add_poly_Z(coeff = data[m..m+4) ; data has m+1 entries followed by coeff).
 * g = Fr::multiplicative_generator.  data holds domain->m elements (in place).
 * Returns seconds, or < 0 on error. */
double zkref_domain_op(size_t min_size, int op, u64 *data)
{
    try {
        std::shared_ptr<libfqfft::evaluation_domain<FrT>> d = libfqfft::get_evaluation_domain<FrT>(min_size);
        const size_t m = d->m;
        std::vector<FrT> a(op == 5 ? m + 1 : m);
        for (size_t i = 0; i < a.size(); i++) a[i] = ld<FrT>(data + 4 * i);
        const double t0 = now_s();
        switch (op) {
        case 0: d->FFT(a); break;
        case 1: d->iFFT(a); break;
        case 2: d->cosetFFT(a, FrT::multiplicative_generator); break;
        case 3: d->icosetFFT(a, FrT::multiplicative_generator); break;
        case 4: d->divide_by_Z_on_coset(a); break;
        case 5: d->add_poly_Z(ld<FrT>(data + 4 * (m + 1)), a); break;
        default: return -2;
        }
        const double t1 = now_s();
        for (size_t i = 0; i < a.size(); i++) st(data + 4 * i, a[i]);
        return t1 - t0;
    } catch (...) {
        return -1;
    }
}

void zkref_domain_element(size_t min_size, size_t idx, u64 *out)
{
    std::shared_ptr<libfqfft::evaluation_domain<FrT>> d = libfqfft::get_evaluation_domain<FrT>(min_size);
    st(out, d->get_domain_element(idx));
}

/* ---- circuits ---------------------------------------------------------------- */
void *zkref_circuit_zklaim(size_t payloads, u64 seed)
{
    zkref_init();
    CoutSilencer quiet;
    Circuit *c = new Circuit();
    ZkCtx z(payloads, seed);
    protoboard<FrT> pb;
    zklaim_gadget<FrT> g(pb, &z.ctx);
    g.generate_r1cs_constraints();
    g.generate_r1cs_witness(&z.ctx);
    c->cs = pb.get_constraint_system();
    c->primary = pb.primary_input();
    c->auxiliary = pb.auxiliary_input();
    return c;
}

/* A fresh witness for an existing zklaim circuit (same payload count, new salts). */
int zkref_zklaim_witness(size_t payloads, u64 seed, u64 *w_out, size_t nv)
{
    CoutSilencer quiet;
    ZkCtx z(payloads, seed);
    protoboard<FrT> pb;
    zklaim_gadget<FrT> g(pb, &z.ctx);
    g.generate_r1cs_constraints();
    g.generate_r1cs_witness(&z.ctx);
    if (!pb.is_satisfied()) return 1;
    const std::vector<FrT> full = pb.full_variable_assignment();
    if (full.size() != nv) return 2;
    for (size_t i = 0; i < nv; i++) st(w_out + 4 * i, full[i]);
    return 0;
}

/* Deterministic variant of generate_r1cs_example_with_field_input
 * (r1cs_examples.tcc:25-97): the same constraint pattern, the two starting values taken
 * from SHA512_rng(seed), SHA512_rng(seed+1) instead of random_element(). */
void *zkref_circuit_example(size_t num_constraints, size_t num_inputs, u64 seed)
{
    zkref_init();
    Circuit *c = new Circuit();
    r1cs_constraint_system<FrT> &cs = c->cs;
    cs.primary_input_size = num_inputs;
    cs.auxiliary_input_size = 2 + num_constraints - num_inputs;
    std::vector<FrT> full;
    FrT a = libff::SHA512_rng<FrT>(seed), b = libff::SHA512_rng<FrT>(seed + 1);
    full.push_back(a);
    full.push_back(b);
    for (size_t i = 0; i + 1 < num_constraints; ++i) {
        linear_combination<FrT> A, B, C;
        FrT tmp;
        if (i % 2) {
            A.add_term(i + 1, 1);
            B.add_term(i + 2, 1);
            C.add_term(i + 3, 1);
            tmp = a * b;
        } else {
            B.add_term(0, 1);
            A.add_term(i + 1, 1);
            A.add_term(i + 2, 1);
            C.add_term(i + 3, 1);
            tmp = a + b;
        }
        full.push_back(tmp);
        a = b;
        b = tmp;
        cs.add_constraint(r1cs_constraint<FrT>(A, B, C));
    }
    linear_combination<FrT> A, B, C;
    FrT fin = FrT::zero();
    for (size_t i = 1; i < cs.num_variables(); ++i) {
        A.add_term(i, 1);
        B.add_term(i, 1);
        fin = fin + full[i - 1];
    }
    C.add_term(cs.num_variables(), 1);
    cs.add_constraint(r1cs_constraint<FrT>(A, B, C));
    full.push_back(fin.squared());
    c->primary.assign(full.begin(), full.begin() + num_inputs);
    c->auxiliary.assign(full.begin() + num_inputs, full.end());
    return c;
}

/* `blocks` chained sha256_two_to_one_hash_gadget (sha256_gadget.tcc:152-226): block i hashes
 * (digest_{i-1} || right_i), right_i seeded.  Public input: none (all auxiliary) except that
 * libsnark needs >= 0 inputs; we expose the final digest's 256 bits as primary inputs by
 * allocating them first. */
void *zkref_circuit_sha_chain(size_t blocks, u64 seed)
{
    zkref_init();
    Circuit *c = new Circuit();
    protoboard<FrT> pb;
    std::vector<std::shared_ptr<digest_variable<FrT>>> dig(blocks + 1), right(blocks);
    dig[blocks].reset(new digest_variable<FrT>(pb, 256, "out"));
    pb.set_input_sizes(256);
    for (size_t i = 0; i < blocks; i++) dig[i].reset(new digest_variable<FrT>(pb, 256, "d"));
    for (size_t i = 0; i < blocks; i++) right[i].reset(new digest_variable<FrT>(pb, 256, "r"));
    std::vector<std::shared_ptr<sha256_two_to_one_hash_gadget<FrT>>> h(blocks);
    for (size_t i = 0; i < blocks; i++) {
        h[i].reset(new sha256_two_to_one_hash_gadget<FrT>(pb, *dig[i], *right[i], *dig[i + 1], "h"));
        h[i]->generate_r1cs_constraints();
    }
    libff::bit_vector left(256);
    for (size_t j = 0; j < 256; j++) left[j] = (seed >> (j % 64)) & 1;
    dig[0]->generate_r1cs_witness(left);
    for (size_t i = 0; i < blocks; i++) {
        libff::bit_vector rb(256);
        u64 x = seed * 0x9e3779b97f4a7c15ull + i * 0xd1b54a32d192ed03ull + 12345;
        for (size_t j = 0; j < 256; j++) {
            x ^= x << 13; x ^= x >> 7; x ^= x << 17;
            rb[j] = x & 1;
        }
        right[i]->generate_r1cs_witness(rb);
        h[i]->generate_r1cs_witness();
    }
    c->cs = pb.get_constraint_system();
    c->primary = pb.primary_input();
    c->auxiliary = pb.auxiliary_input();
    return c;
}

void zkref_circuit_free(void *h) { delete (Circuit *)h; }

/* dims[0..8) = num_constraints, num_variables, num_inputs, nnzA, nnzB, nnzC, domain m, domain kind */
void zkref_circuit_dims(void *h, size_t *dims)
{
    Circuit *c = (Circuit *)h;
    const r1cs_constraint_system<FrT> &cs = c->prover_cs();
    size_t nnz[3] = {0, 0, 0};
    for (const auto &k : cs.constraints) {
        nnz[0] += k.a.terms.size();
        nnz[1] += k.b.terms.size();
        nnz[2] += k.c.terms.size();
    }
    dims[0] = cs.num_constraints();
    dims[1] = cs.num_variables();
    dims[2] = cs.num_inputs();
    dims[3] = nnz[0];
    dims[4] = nnz[1];
    dims[5] = nnz[2];
    int kind = 0;
    size_t m = 0;
    zkref_domain_select(cs.num_constraints() + cs.num_inputs() + 1, &m, &kind);
    dims[6] = m;
    dims[7] = kind;
}

int zkref_circuit_is_satisfied(void *h)
{
    Circuit *c = (Circuit *)h;
    return c->prover_cs().is_satisfied(c->primary, c->auxiliary);
}

/* CSR export of matrix which (0 A, 1 B, 2 C) of the PROVER's system (after the
 * generator's swap_AB_if_beneficial, r1cs_gg_ppzksnark.tcc:217-218, when a key exists).
 * col index 0 = the constant 1, j>0 = variable j-1 (variable.tcc:262-270). */
void zkref_circuit_csr(void *h, int which, u64 *row_ptr, uint32_t *col, u64 *coeff)
{
    Circuit *c = (Circuit *)h;
    const r1cs_constraint_system<FrT> &cs = c->prover_cs();
    size_t p = 0;
    for (size_t i = 0; i < cs.constraints.size(); i++) {
        row_ptr[i] = p;
        const linear_combination<FrT> &lc =
            which == 0 ? cs.constraints[i].a : (which == 1 ? cs.constraints[i].b : cs.constraints[i].c);
        for (const auto &t : lc.terms) {
            col[p] = (uint32_t)t.index;
            st(coeff + 4 * p, t.coeff);
            p++;
        }
    }
    row_ptr[cs.constraints.size()] = p;
}

void zkref_circuit_witness(void *h, u64 *w_out)
{
    Circuit *c = (Circuit *)h;
    size_t k = 0;
    for (const auto &x : c->primary) st(w_out + 4 * (k++), x);
    for (const auto &x : c->auxiliary) st(w_out + 4 * (k++), x);
}

void zkref_circuit_set_witness(void *h, const u64 *w)
{
    Circuit *c = (Circuit *)h;
    const size_t ni = c->cs.num_inputs(), nv = c->cs.num_variables();
    c->primary.resize(ni);
    c->auxiliary.resize(nv - ni);
    for (size_t i = 0; i < ni; i++) c->primary[i] = ld<FrT>(w + 4 * i);
    for (size_t i = ni; i < nv; i++) c->auxiliary[i - ni] = ld<FrT>(w + 4 * i);
}

/* Trusted setup through the reference generator (r1cs_gg_ppzksnark.tcc:211-388). */
double zkref_circuit_setup(void *h)
{
    Circuit *c = (Circuit *)h;
    const double t0 = now_s();
    r1cs_gg_ppzksnark_keypair<PP> kp = r1cs_gg_ppzksnark_generator<PP>(c->cs);
    c->pk = kp.pk;
    c->vk = kp.vk;
    c->has_key = true;
    return now_s() - t0;
}

/* The reference generator (r1cs_gg_ppzksnark.tcc:211-388) with its secret randomness injected: the reference draws
 * t, alpha, beta, gamma, delta (:221-225) and the two group generators (:284, :294) from std::random_device and
 * offers no way to fix them.  Same shape as prove_with_rs: the orchestration is re-expressed, every arithmetic call
 * (swap_AB_if_beneficial, r1cs_to_qap_instance_map_with_evaluation, get_window_table, batch_exp, kc_batch_exp,
 * batch_exp_with_coeff, batch_to_special, reduced_pairing) is the reference's own.
 * toxic = t, alpha, beta, gamma, delta (5 x 4 limbs).  Returns seconds. */
double zkref_circuit_setup_fixed(void *h, const u64 *toxic, const u64 *g1_gen, const u64 *g2_gen)
{
    Circuit *c = (Circuit *)h;
    CoutSilencer quiet;
    const double t0 = now_s();
    r1cs_constraint_system<FrT> cs(c->cs);
    cs.swap_AB_if_beneficial();
    const FrT t = ld<FrT>(toxic), alpha = ld<FrT>(toxic + 4), beta = ld<FrT>(toxic + 8), gamma = ld<FrT>(toxic + 12),
              delta = ld<FrT>(toxic + 16);
    const G1T G = ld_g1(g1_gen);
    const G2T Q = ld_g2(g2_gen);
    qap_instance_evaluation<FrT> qap = r1cs_to_qap_instance_map_with_evaluation(cs, t);
    const size_t nv = qap.num_variables(), ni = qap.num_inputs();
    size_t nzA = 0, nzB = 0;
    for (size_t i = 0; i <= nv; i++) {
        nzA += qap.At[i].is_zero() ? 0 : 1;
        nzB += qap.Bt[i].is_zero() ? 0 : 1;
    }
    const FrT ginv = gamma.inverse(), dinv = delta.inverse();
    std::vector<FrT> abc(ni + 1), Lt(nv - ni), Ht(qap.Ht.begin(), qap.Ht.end() - 2);
    for (size_t i = 0; i <= nv; i++) {
        const FrT x = beta * qap.At[i] + alpha * qap.Bt[i] + qap.Ct[i];
        if (i <= ni)
            abc[i] = x * ginv;
        else
            Lt[i - ni - 1] = x * dinv;
    }
#ifdef MULTICORE
    const size_t chunks = omp_get_max_threads();
#else
    const size_t chunks = 1;
#endif
    const size_t bits = FrT::size_in_bits();
    const size_t w1 = libff::get_exp_window_size<G1T>(nzA + nzB + nv), w2 = libff::get_exp_window_size<G2T>(nzB);
    const libff::window_table<G1T> T1 = libff::get_window_table(bits, w1, G);
    const libff::window_table<G2T> T2 = libff::get_window_table(bits, w2, Q);
    auto &pk = c->pk;
    pk.alpha_g1 = alpha * G;
    pk.beta_g1 = beta * G;
    pk.beta_g2 = beta * Q;
    pk.delta_g1 = delta * G;
    pk.delta_g2 = delta * Q;
    pk.A_query = libff::batch_exp(bits, w1, T1, qap.At);
    libff::batch_to_special<G1T>(pk.A_query);
    pk.B_query = kc_batch_exp(bits, w2, w1, T2, T1, FrT::one(), FrT::one(), qap.Bt, chunks);
    for (auto &v : pk.B_query.values) { /* special form whatever USE_MIXED_ADDITION says */
        v.g.to_special();
        v.h.to_special();
    }
    pk.H_query = libff::batch_exp_with_coeff(bits, w1, T1, qap.Zt * dinv, Ht);
    libff::batch_to_special<G1T>(pk.H_query);
    pk.L_query = libff::batch_exp(bits, w1, T1, Lt);
    libff::batch_to_special<G1T>(pk.L_query);
    pk.constraint_system = cs;
    std::vector<FrT> abc_rest(abc.begin() + 1, abc.end());
    G1T abc0 = abc[0] * G;
    std::vector<G1T> abc_vals = libff::batch_exp(bits, w1, T1, abc_rest);
    c->vk = r1cs_gg_ppzksnark_verification_key<PP>(PP::reduced_pairing(pk.alpha_g1, pk.beta_g2), gamma * Q, pk.delta_g2,
                                                   accumulation_vector<G1T>(std::move(abc0), std::move(abc_vals)));
    c->has_key = true;
    return now_s() - t0;
}

/* gamma_g2 (16), delta_g2 (16) and gamma_ABC_g1 ((num_inputs + 1) x 8) of the verification key, affine. */
void zkref_vk_export(void *h, u64 *gamma_g2, u64 *delta_g2, u64 *gamma_abc)
{
    Circuit *c = (Circuit *)h;
    st_g2_affine(gamma_g2, c->vk.gamma_g2);
    st_g2_affine(delta_g2, c->vk.delta_g2);
    st_g1_affine(gamma_abc, c->vk.gamma_ABC_g1.first);
    for (size_t i = 0; i < c->vk.gamma_ABC_g1.rest.values.size(); i++)
        st_g1_affine(gamma_abc + 8 * (i + 1), c->vk.gamma_ABC_g1.rest.values[i]);
}

/* Install a verification key made elsewhere (the GPU generator): alpha_g1 / beta_g2 for the one pairing of
 * r1cs_gg_ppzksnark.tcc:338, gamma_g2, delta_g2, gamma_ABC_g1; zkref_verify then runs the reference verifier on it. */
void zkref_vk_import(void *h, const u64 *alpha_g1, const u64 *beta_g2, const u64 *gamma_g2, const u64 *delta_g2,
                     const u64 *gamma_abc, size_t n_abc)
{
    Circuit *c = (Circuit *)h;
    std::vector<G1T> rest(n_abc - 1);
    for (size_t i = 1; i < n_abc; i++) rest[i - 1] = ld_g1_affine(gamma_abc + 8 * i);
    G1T first = ld_g1_affine(gamma_abc);
    c->vk = r1cs_gg_ppzksnark_verification_key<PP>(PP::reduced_pairing(ld_g1_affine(alpha_g1), ld_g2_affine(beta_g2)),
                                                   ld_g2_affine(gamma_g2), ld_g2_affine(delta_g2),
                                                   accumulation_vector<G1T>(std::move(first), std::move(rest)));
}

/* sizes[0..5) = |A_query|, nnz(B_query), |H_query|, |L_query|, B_query domain size */
void zkref_key_dims(void *h, size_t *sizes)
{
    Circuit *c = (Circuit *)h;
    const auto &pk = c->pk;
    sizes[0] = pk.A_query.size();
    sizes[1] = pk.B_query.indices.size();
    sizes[2] = pk.H_query.size();
    sizes[3] = pk.L_query.size();
    sizes[4] = pk.B_query.domain_size();
}

/* Key export as affine arrays (the in-memory form after operator>>, alt_bn128_g1.cpp:453-462).
 * singles: alpha_g1(8) beta_g1(8) delta_g1(8) beta_g2(16) delta_g2(16) = 56 u64. */
void zkref_key_export(void *h, u64 *A, u64 *B1, u64 *B2, uint32_t *B_idx, u64 *H, u64 *L, u64 *singles)
{
    Circuit *c = (Circuit *)h;
    const auto &pk = c->pk;
    for (size_t i = 0; i < pk.A_query.size(); i++) st_g1_affine(A + 8 * i, pk.A_query[i]);
    for (size_t i = 0; i < pk.B_query.indices.size(); i++) {
        B_idx[i] = (uint32_t)pk.B_query.indices[i];
        st_g1_affine(B1 + 8 * i, pk.B_query.values[i].h);
        st_g2_affine(B2 + 16 * i, pk.B_query.values[i].g);
    }
    for (size_t i = 0; i < pk.H_query.size(); i++) st_g1_affine(H + 8 * i, pk.H_query[i]);
    for (size_t i = 0; i < pk.L_query.size(); i++) st_g1_affine(L + 8 * i, pk.L_query[i]);
    st_g1_affine(singles, pk.alpha_g1);
    st_g1_affine(singles + 8, pk.beta_g1);
    st_g1_affine(singles + 16, pk.delta_g1);
    st_g2_affine(singles + 24, pk.beta_g2);
    st_g2_affine(singles + 40, pk.delta_g2);
}

/* Install a key from affine arrays (for fixtures / synthetic keys); the verification key is
 * left empty, so zkref_verify is not meaningful afterwards. */
void zkref_key_import(void *h, const u64 *A, size_t nA, const u64 *B1, const u64 *B2, const uint32_t *B_idx,
                      size_t nB, const u64 *H, size_t nH, const u64 *L, size_t nL, const u64 *singles)
{
    Circuit *c = (Circuit *)h;
    auto &pk = c->pk;
    pk.A_query.resize(nA);
    for (size_t i = 0; i < nA; i++) pk.A_query[i] = ld_g1_affine(A + 8 * i);
    pk.B_query.indices.resize(nB);
    pk.B_query.values.resize(nB);
    pk.B_query.domain_size_ = c->cs.num_variables() + 1;
    for (size_t i = 0; i < nB; i++) {
        pk.B_query.indices[i] = B_idx[i];
        pk.B_query.values[i] = knowledge_commitment<G2T, G1T>(ld_g2_affine(B2 + 16 * i), ld_g1_affine(B1 + 8 * i));
    }
    pk.H_query.resize(nH);
    for (size_t i = 0; i < nH; i++) pk.H_query[i] = ld_g1_affine(H + 8 * i);
    pk.L_query.resize(nL);
    for (size_t i = 0; i < nL; i++) pk.L_query[i] = ld_g1_affine(L + 8 * i);
    pk.alpha_g1 = ld_g1_affine(singles);
    pk.beta_g1 = ld_g1_affine(singles + 8);
    pk.delta_g1 = ld_g1_affine(singles + 16);
    pk.beta_g2 = ld_g2_affine(singles + 24);
    pk.delta_g2 = ld_g2_affine(singles + 40);
    pk.constraint_system = c->cs;
    c->has_key = true;
}

/* r1cs_to_qap_witness_map on the prover's system with d1 = d2 = d3 = 0; H_out has m+1 elems. */
double zkref_witness_map(void *h, const u64 *w, u64 *H_out)
{
    Circuit *c = (Circuit *)h;
    const r1cs_constraint_system<FrT> &cs = c->prover_cs();
    const size_t ni = cs.num_inputs(), nv = cs.num_variables();
    std::vector<FrT> p(ni), a(nv - ni);
    for (size_t i = 0; i < ni; i++) p[i] = ld<FrT>(w + 4 * i);
    for (size_t i = ni; i < nv; i++) a[i - ni] = ld<FrT>(w + 4 * i);
    const double t0 = now_s();
    const qap_witness<FrT> qw = r1cs_to_qap_witness_map(cs, p, a, FrT::zero(), FrT::zero(), FrT::zero());
    const double t1 = now_s();
    for (size_t i = 0; i < qw.coefficients_for_H.size(); i++) st(H_out + 4 * i, qw.coefficients_for_H[i]);
    return t1 - t0;
}

/* Reference prover with fixed r, s.
 *   proof_out : A (12) | B (24) | C (12) Jacobian limbs
 *   evals_out : At (12) | Bt.h (12) | Bt.g (24) | Ht (12) | Lt (12), may be NULL
 *   H_out     : m+1 Fr, may be NULL
 *   times_out : wm, A, B, H, L, total seconds; may be NULL */
int zkref_prove(void *h, const u64 *w, const u64 *r, const u64 *s, u64 *proof_out, u64 *evals_out, u64 *H_out,
                double *times_out)
{
    Circuit *c = (Circuit *)h;
    if (!c->has_key) return -1;
    const r1cs_constraint_system<FrT> &cs = c->pk.constraint_system;
    const size_t ni = cs.num_inputs(), nv = cs.num_variables();
    std::vector<FrT> p(ni), a(nv - ni);
    for (size_t i = 0; i < ni; i++) p[i] = ld<FrT>(w + 4 * i);
    for (size_t i = ni; i < nv; i++) a[i - ni] = ld<FrT>(w + 4 * i);
    ProveOut o = prove_with_rs(c->pk, p, a, ld<FrT>(r), ld<FrT>(s));
    st_g1(proof_out, o.A);
    st_g2(proof_out + 12, o.B);
    st_g1(proof_out + 36, o.C);
    if (evals_out) {
        st_g1(evals_out, o.At);
        st_g1(evals_out + 12, o.Bt1);
        st_g2(evals_out + 24, o.Bt2);
        st_g1(evals_out + 48, o.Ht);
        st_g1(evals_out + 60, o.Lt);
    }
    if (H_out)
        for (size_t i = 0; i < o.H.size(); i++) st(H_out + 4 * i, o.H[i]);
    if (times_out) {
        times_out[0] = o.t_wm;
        times_out[1] = o.t_A;
        times_out[2] = o.t_B;
        times_out[3] = o.t_H;
        times_out[4] = o.t_L;
        times_out[5] = o.t_total;
    }
    return 0;
}

/* The stock reference prover (random r, s), timed; for the CPU baseline. */
double zkref_prove_stock(void *h, const u64 *w, u64 *proof_out)
{
    Circuit *c = (Circuit *)h;
    const r1cs_constraint_system<FrT> &cs = c->pk.constraint_system;
    const size_t ni = cs.num_inputs(), nv = cs.num_variables();
    std::vector<FrT> p(ni), a(nv - ni);
    for (size_t i = 0; i < ni; i++) p[i] = ld<FrT>(w + 4 * i);
    for (size_t i = ni; i < nv; i++) a[i - ni] = ld<FrT>(w + 4 * i);
    const double t0 = now_s();
    r1cs_gg_ppzksnark_proof<PP> pr = r1cs_gg_ppzksnark_prover<PP>(c->pk, p, a);
    const double t1 = now_s();
    if (proof_out) {
        st_g1(proof_out, pr.g_A);
        st_g2(proof_out + 12, pr.g_B);
        st_g1(proof_out + 36, pr.g_C);
    }
    return t1 - t0;
}

/* Reference verifier (strong IC, r1cs_gg_ppzksnark.tcc:612-622) on the key made by
 * zkref_circuit_setup.  primary: num_inputs Fr.  Returns 1 = accept. */
int zkref_verify(void *h, const u64 *primary, const u64 *proof)
{
    Circuit *c = (Circuit *)h;
    const size_t ni = c->pk.constraint_system.num_inputs();
    std::vector<FrT> p(ni);
    for (size_t i = 0; i < ni; i++) p[i] = ld<FrT>(primary + 4 * i);
    r1cs_gg_ppzksnark_proof<PP> pr(ld_g1(proof), ld_g2(proof + 12), ld_g1(proof + 36));
    return r1cs_gg_ppzksnark_verifier_strong_IC<PP>(c->vk, p, pr) ? 1 : 0;
}

/* libsnark's own operator<< on the proof (r1cs_gg_ppzksnark.tcc:168-177): the bytes that
 * end up in zklaim_ctx::proof.  Returns the length (134 with BINARY_OUTPUT). */
size_t zkref_proof_serialize(const u64 *proof, unsigned char *out, size_t cap)
{
    r1cs_gg_ppzksnark_proof<PP> pr(ld_g1(proof), ld_g2(proof + 12), ld_g1(proof + 36));
    std::stringstream ss;
    ss << pr;
    const std::string s = ss.str();
    if (s.size() <= cap) memcpy(out, s.data(), s.size());
    return s.size();
}

/* ---- pairing (verifier side) ---------------------------------------------- */
static void st_fq12(u64 *p, const libff::alt_bn128_Fq12 &f)
{
    const libff::alt_bn128_Fq6 *h[2] = {&f.c0, &f.c1};
    for (int i = 0; i < 2; i++) {
        st2(p + 24 * i, h[i]->c0);
        st2(p + 24 * i + 8, h[i]->c1);
        st2(p + 24 * i + 16, h[i]->c2);
    }
}
/* alt_bn128_ate_miller_loop (reduced = 0) or alt_bn128_reduced_pairing (reduced = 1) of affine P, Q
 * (alt_bn128_pairing.cpp:371-420, 523-533); out: 12 Fq = c0.c0.c0, c0.c0.c1, c0.c1.c0, ... c1.c2.c1 */
void zkref_pairing(const u64 *g1_affine, const u64 *g2_affine, int reduced, u64 *out48)
{
    zkref_init();
    const G1T P = ld_g1_affine(g1_affine);
    const G2T Q = ld_g2_affine(g2_affine);
    libff::alt_bn128_Fq12 f = PP::miller_loop(PP::precompute_G1(P), PP::precompute_G2(Q));
    if (reduced) f = PP::final_exponentiation(f);
    st_fq12(out48, f);
}
/* vk.alpha_g1_beta_g2 (r1cs_gg_ppzksnark.tcc:338) of the circuit's key */
void zkref_vk_alpha_beta(void *h, u64 *out48) { st_fq12(out48, ((Circuit *)h)->vk.alpha_g1_beta_g2); }

/* ---- the zklaim C bridge, reference flavour (libsnark_wrapper.cpp:195-276) --------- */
/* Runs libsnark_trusted_setup / libsnark_prove / libsnark_verify of the reference on a
 * k-payload credential; returns prove seconds incl. the per-call key import, or < 0. */
double zkref_bridge_roundtrip(size_t payloads, u64 seed, int *verify_rc)
{
    ZkCtx z(payloads, seed);
    if (libsnark_trusted_setup(&z.ctx)) return -1;
    const double t0 = now_s();
    if (libsnark_prove(&z.ctx)) return -2;
    const double t1 = now_s();
    *verify_rc = libsnark_verify(&z.ctx);
    free(z.ctx.pk);
    free(z.ctx.vk);
    free(z.ctx.proof);
    return t1 - t0;
}

} /* extern "C" */
