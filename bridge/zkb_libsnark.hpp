/*
 * zkb_libsnark.hpp -- host-side adapter between libsnark's r1cs_gg_ppzksnark types and the zkb C ABI
 * (include/zkb.h).  This is the code a zklaim maintainer adds to the tree: with it,
 *
 *     zklaim/snark.cpp:126
 *         return r1cs_gg_ppzksnark_prover<ppT>(proving_key, pb.primary_input(), pb.auxiliary_input());
 * becomes
 *         return zkb_bridge::prover<ppT>(proving_key, pb.primary_input(), pb.auxiliary_input());
 *
 * with the same signature and return type (r1cs_gg_ppzksnark.hpp:373-376).  Everything else of
 * zklaim (C API, gadgets, key / proof serialisation, verifier) is untouched.
 *
 * The adapter is header-only and is compiled together with the reference's own headers; it
 * contains no arithmetic: it flattens the proving key into the arrays of zkb_key_desc
 * (affine Montgomery limbs, CSR matrices), keeps the uploaded key cached per device, draws the
 * blinding scalars exactly where the reference does (r1cs_gg_ppzksnark.tcc:418-419,
 * Fr::random_element) and wraps the three returned points into r1cs_gg_ppzksnark_proof.
 * Only alt_bn128 (the curve zklaim is built for, zklaim/snark.cpp ppT) is supported.
 */
#ifndef ZKB_LIBSNARK_HPP
#define ZKB_LIBSNARK_HPP

#include <array>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include <openssl/sha.h>

#include <libff/algebra/curves/alt_bn128/alt_bn128_pp.hpp>
#include <libsnark/zk_proof_systems/ppzksnark/r1cs_gg_ppzksnark/r1cs_gg_ppzksnark.hpp>

#include "../include/zkb.h"

namespace zkb_bridge {

typedef libff::alt_bn128_pp PP;
typedef libff::Fr<PP> FrT;
typedef libff::G1<PP> G1T;
typedef libff::G2<PP> G2T;

inline void put_fp(uint64_t *dst, const libff::alt_bn128_Fq &x) { memcpy(dst, x.mont_repr.data, 32); }
inline void put_fr(uint64_t *dst, const FrT &x) { memcpy(dst, x.mont_repr.data, 32); }
inline libff::alt_bn128_Fq get_fp(const uint64_t *src)
{
    libff::alt_bn128_Fq x;
    memcpy(x.mont_repr.data, src, 32);
    return x;
}

/* affine limbs of a key element; zero -> all-zero limbs (include/zkb.h conventions) */
inline void put_g1_affine(uint64_t *dst, const G1T &p)
{
    if (p.is_zero()) {
        memset(dst, 0, 64);
        return;
    }
    if (p.Z == libff::alt_bn128_Fq::one()) { /* the form operator>> leaves (alt_bn128_g1.cpp:453-462) */
        put_fp(dst, p.X);
        put_fp(dst + 4, p.Y);
        return;
    }
    G1T c(p);
    c.to_affine_coordinates();
    put_fp(dst, c.X);
    put_fp(dst + 4, c.Y);
}
inline void put_g2_affine(uint64_t *dst, const G2T &p)
{
    if (p.is_zero()) {
        memset(dst, 0, 128);
        return;
    }
    G2T c(p);
    if (!(c.Z == libff::alt_bn128_Fq2::one())) c.to_affine_coordinates();
    put_fp(dst, c.X.c0);
    put_fp(dst + 4, c.X.c1);
    put_fp(dst + 8, c.Y.c0);
    put_fp(dst + 12, c.Y.c1);
}
inline G1T get_g1_jacobian(const uint64_t *src) { return G1T(get_fp(src), get_fp(src + 4), get_fp(src + 8)); }
inline G2T get_g2_jacobian(const uint64_t *src)
{
    return G2T(libff::alt_bn128_Fq2(get_fp(src), get_fp(src + 4)), libff::alt_bn128_Fq2(get_fp(src + 8), get_fp(src + 12)),
               libff::alt_bn128_Fq2(get_fp(src + 16), get_fp(src + 20)));
}

inline libff::alt_bn128_Fq2 get_fp2(const uint64_t *src) { return libff::alt_bn128_Fq2(get_fp(src), get_fp(src + 4)); }
inline libff::alt_bn128_Fq12 get_fq12(const uint64_t *src) /* 48 limbs, member order c0.c0.c0 .. c1.c2.c1 */
{
    return libff::alt_bn128_Fq12(libff::alt_bn128_Fq6(get_fp2(src), get_fp2(src + 8), get_fp2(src + 16)),
                                 libff::alt_bn128_Fq6(get_fp2(src + 24), get_fp2(src + 32), get_fp2(src + 40)));
}

/* flat copy of a proving key; owns the storage the zkb_key_desc points into */
struct KeyArrays {
    std::vector<uint64_t> row_ptr[3], coeff[3];
    std::vector<uint32_t> col[3];
    std::vector<uint64_t> singles, A, B1, B2, H, L;
    std::vector<uint32_t> B_idx;
    zkb_key_desc desc;
};

inline void pack_matrix(const libsnark::r1cs_constraint_system<FrT> &cs, int which, KeyArrays &k)
{
    std::vector<uint64_t> &rp = k.row_ptr[which];
    rp.assign(1, 0);
    for (const auto &c : cs.constraints) {
        const libsnark::linear_combination<FrT> &lc = which == 0 ? c.a : (which == 1 ? c.b : c.c);
        for (const auto &t : lc.terms) {
            k.col[which].push_back((uint32_t)t.index);
            const size_t o = k.coeff[which].size();
            k.coeff[which].resize(o + 4);
            put_fr(&k.coeff[which][o], t.coeff);
        }
        rp.push_back(k.col[which].size());
    }
}

/* points the zkb_key_desc of `k` at its arrays (after they have their final sizes) */
inline void finish_desc(KeyArrays &k, size_t nc, size_t nv, size_t ni, size_t nb, size_t nh)
{
    zkb_key_desc &d = k.desc;
    memset(&d, 0, sizeof(d));
    d.num_constraints = nc;
    d.num_variables = nv;
    d.num_inputs = ni;
    zkb_csr *m[3] = {&d.A, &d.B, &d.C};
    for (int w = 0; w < 3; w++) {
        m[w]->row_ptr = k.row_ptr[w].data();
        m[w]->col = k.col[w].data();
        m[w]->coeff = k.coeff[w].data();
    }
    d.alpha_g1 = &k.singles[0];
    d.beta_g1 = &k.singles[8];
    d.delta_g1 = &k.singles[16];
    d.beta_g2 = &k.singles[24];
    d.delta_g2 = &k.singles[40];
    d.A_query = k.A.data();
    d.B_nnz = nb;
    d.B_index = k.B_idx.data();
    d.B_g1 = k.B1.data();
    d.B_g2 = k.B2.data();
    d.H_size = nh;
    d.H_query = k.H.data();
    d.L_query = k.L.data();
}

inline void pack_proving_key(const libsnark::r1cs_gg_ppzksnark_proving_key<PP> &pk, KeyArrays &k)
{
    const libsnark::r1cs_constraint_system<FrT> &cs = pk.constraint_system;
    for (int w = 0; w < 3; w++) pack_matrix(cs, w, k);
    k.singles.resize(8 * 3 + 16 * 2);
    put_g1_affine(&k.singles[0], pk.alpha_g1);
    put_g1_affine(&k.singles[8], pk.beta_g1);
    put_g1_affine(&k.singles[16], pk.delta_g1);
    put_g2_affine(&k.singles[24], pk.beta_g2);
    put_g2_affine(&k.singles[40], pk.delta_g2);
    k.A.resize(pk.A_query.size() * 8);
    for (size_t i = 0; i < pk.A_query.size(); i++) put_g1_affine(&k.A[8 * i], pk.A_query[i]);
    const size_t nb = pk.B_query.indices.size();
    k.B1.resize(nb * 8 + 8);
    k.B2.resize(nb * 16 + 16);
    k.B_idx.resize(nb + 1);
    for (size_t i = 0; i < nb; i++) {
        k.B_idx[i] = (uint32_t)pk.B_query.indices[i];
        put_g1_affine(&k.B1[8 * i], pk.B_query.values[i].h);
        put_g2_affine(&k.B2[16 * i], pk.B_query.values[i].g);
    }
    k.H.resize(pk.H_query.size() * 8 + 8);
    for (size_t i = 0; i < pk.H_query.size(); i++) put_g1_affine(&k.H[8 * i], pk.H_query[i]);
    k.L.resize(pk.L_query.size() * 8 + 8);
    for (size_t i = 0; i < pk.L_query.size(); i++) put_g1_affine(&k.L[8 * i], pk.L_query[i]);
    finish_desc(k, cs.num_constraints(), cs.num_variables(), cs.num_inputs(), nb, pk.H_query.size());
}

/* ---- key import with the square roots left to the GPU ---------------------------------
 * libsnark_import_pk (zklaim/libsnark_wrapper.cpp:160-168) runs operator>> of r1cs_gg_ppzksnark_proving_key
 * (r1cs_gg_ppzksnark.tcc:68-88), which recomputes Y = sqrt(X^3 + b) for every element of the key on one host thread
 * (alt_bn128_g1.cpp:436-450, alt_bn128_g2.cpp:449-463): 7 s for the zklaim key.  This reader walks the same byte stream
 * with the reference's framing (the build's BINARY_OUTPUT + MONTGOMERY_OUTPUT, point compression on: a G1 element is
 * is_zero, X (32 bytes), lsb(Y); sizes are decimal + '\n'; sparse_vector.tcc:290-325, knowledge_commitment.tcc:127-134),
 * collects the abscissas, and hands them to zkb_points_decompress; the constraint system at the end of the stream is
 * parsed by libsnark's own operator>>. */
struct PkCursor {
    const unsigned char *p, *end;
    bool ok;
    size_t size_line()
    {
        size_t v = 0;
        bool any = false;
        while (p < end && *p >= '0' && *p <= '9') {
            v = v * 10 + (size_t)(*p - '0');
            p++;
            any = true;
        }
        if (!any || p >= end || *p != '\n') ok = false;
        else p++;
        return v;
    }
    void point(size_t words, std::vector<uint64_t> &x, std::vector<uint8_t> &flags)
    {
        if (!ok || (size_t)(end - p) < 2 + 8 * words) {
            ok = false;
            return;
        }
        const unsigned zero = (unsigned)(p[0] - '0'), lsb = (unsigned)(p[1 + 8 * words] - '0');
        if (zero > 1 || lsb > 1) ok = false;
        const size_t o = x.size();
        x.resize(o + words);
        memcpy(&x[o], p + 1, 8 * words);
        flags.push_back((uint8_t)(lsb | (zero << 1)));
        p += 2 + 8 * words;
    }
};

/* fills `k` from the serialised proving key; *nv / *ni = variables / inputs of its constraint system */
inline void import_pk_deferred(zkb_ctx *ctx, const unsigned char *pk, size_t size, KeyArrays &k, size_t *nv_out, size_t *ni_out)
{
#if !defined(BINARY_OUTPUT) || !defined(MONTGOMERY_OUTPUT) || defined(NO_PT_COMPRESSION)
#error "the deferred key import reads the BINARY_OUTPUT + MONTGOMERY_OUTPUT + compressed-point stream zklaim is built with"
#endif
    PkCursor c{pk, pk + size, true};
    std::vector<uint64_t> x1, x2;
    std::vector<uint8_t> f1, f2;
    /* r1cs_gg_ppzksnark.tcc:68-88: alpha_g1, beta_g1, beta_g2, delta_g1, delta_g2, A, B, H, L, constraint system */
    c.point(4, x1, f1);
    c.point(4, x1, f1);
    c.point(8, x2, f2);
    c.point(4, x1, f1);
    c.point(8, x2, f2);
    const size_t nA = c.size_line();
    if (c.ok) x1.reserve(4 * (3 + nA));
    for (size_t i = 0; c.ok && i < nA; i++) c.point(4, x1, f1);
    c.size_line(); /* B_query.domain_size_ */
    const size_t nidx = c.size_line();
    k.B_idx.assign(nidx + 1, 0);
    for (size_t i = 0; c.ok && i < nidx; i++) k.B_idx[i] = (uint32_t)c.size_line();
    const size_t nB = c.size_line();
    if (nB != nidx) c.ok = false;
    for (size_t i = 0; c.ok && i < nB; i++) { /* knowledge_commitment<G2, G1>: g then h */
        c.point(8, x2, f2);
        c.point(4, x1, f1);
    }
    const size_t nH = c.size_line();
    for (size_t i = 0; c.ok && i < nH; i++) c.point(4, x1, f1);
    const size_t nL = c.size_line();
    for (size_t i = 0; c.ok && i < nL; i++) c.point(4, x1, f1);
    if (!c.ok) throw std::runtime_error("zkb key import: the proving key stream does not have libsnark's framing");
    libsnark::r1cs_constraint_system<FrT> cs;
    {
        std::istringstream in(std::string((const char *)c.p, (size_t)(c.end - c.p)));
        in >> cs;
        if (in.fail()) throw std::runtime_error("zkb key import: constraint system unreadable");
    }
    for (int w = 0; w < 3; w++) pack_matrix(cs, w, k);
    /* the roots, on the GPU */
    const size_t n1 = f1.size(), n2 = f2.size();
    std::vector<uint64_t> p1(8 * n1 + 8), p2(16 * n2 + 16);
    if (zkb_points_decompress(ctx, ZKB_G1, x1.data(), f1.data(), n1, p1.data(), NULL) != ZKB_OK ||
        zkb_points_decompress(ctx, ZKB_G2, x2.data(), f2.data(), n2, p2.data(), NULL) != ZKB_OK)
        throw std::runtime_error(std::string("zkb_points_decompress failed: ") + zkb_last_error(ctx));
    /* G1 order: alpha, beta, delta, A (nA), B.h (nB), H (nH), L (nL); G2 order: beta, delta, B.g (nB) */
    k.singles.resize(8 * 3 + 16 * 2);
    memcpy(&k.singles[0], &p1[0], 3 * 64);
    memcpy(&k.singles[24], &p2[0], 2 * 128);
    const uint64_t *q = &p1[3 * 8];
    k.A.assign(q, q + 8 * nA);
    q += 8 * nA;
    k.B1.assign(8 * nB + 8, 0);
    memcpy(k.B1.data(), q, 64 * nB);
    q += 8 * nB;
    k.H.assign(8 * nH + 8, 0);
    memcpy(k.H.data(), q, 64 * nH);
    q += 8 * nH;
    k.L.assign(8 * nL + 8, 0);
    memcpy(k.L.data(), q, 64 * nL);
    k.B2.assign(16 * nB + 16, 0);
    memcpy(k.B2.data(), &p2[2 * 16], 128 * nB);
    finish_desc(k, cs.num_constraints(), cs.num_variables(), cs.num_inputs(), nB, nH);
    *nv_out = cs.num_variables();
    *ni_out = cs.num_inputs();
}

/* identity of a serialised key: SHA-256 of the bytes and their number */
struct KeyId {
    std::array<unsigned char, 32> digest;
    size_t size;
    bool operator<(const KeyId &o) const { return size != o.size ? size < o.size : digest < o.digest; }
};
inline KeyId key_id_of_bytes(const void *p, size_t n)
{
    KeyId id;
    id.size = n;
    SHA256((const unsigned char *)p, n, id.digest.data());
    return id;
}

/* 64-bit FNV-1a */
inline uint64_t fnv1a(const void *p, size_t n, uint64_t h = 0xcbf29ce484222325ull)
{
    const unsigned char *b = (const unsigned char *)p;
    for (size_t i = 0; i < n; i++) h = (h ^ b[i]) * 0x100000001b3ull;
    return h;
}

/* identity of an in-memory proving key (the header-only drop-in zkb_bridge::prover has no byte stream to hash): SHA-256
 * over its sizes and every group element of the key */
inline KeyId key_fingerprint(const libsnark::r1cs_gg_ppzksnark_proving_key<PP> &pk)
{
    SHA256_CTX sha;
    SHA256_Init(&sha);
    const uint64_t dims[7] = {pk.constraint_system.num_constraints(), pk.constraint_system.num_variables(),
                              pk.constraint_system.num_inputs(), pk.A_query.size(), pk.B_query.indices.size(),
                              pk.H_query.size(), pk.L_query.size()};
    SHA256_Update(&sha, dims, sizeof(dims));
    auto mix_g1 = [&](const G1T &p) { /* Jacobian limbs as stored: a key is affine-normalised (Z = 1 or 0) */
        uint64_t t[12];
        memcpy(t, p.X.mont_repr.data, 32);
        memcpy(t + 4, p.Y.mont_repr.data, 32);
        memcpy(t + 8, p.Z.mont_repr.data, 32);
        SHA256_Update(&sha, t, sizeof(t));
    };
    auto mix_g2 = [&](const G2T &p) {
        uint64_t t[16];
        put_g2_affine(t, p);
        SHA256_Update(&sha, t, sizeof(t));
    };
    mix_g1(pk.alpha_g1);
    mix_g1(pk.beta_g1);
    mix_g1(pk.delta_g1);
    mix_g2(pk.beta_g2);
    mix_g2(pk.delta_g2);
    for (const G1T &p : pk.A_query) mix_g1(p);
    for (const auto &kc : pk.B_query.values) mix_g1(kc.h);
    for (const G1T &p : pk.H_query) mix_g1(p);
    for (const G1T &p : pk.L_query) mix_g1(p);
    KeyId id;
    id.size = 0; /* distinguishes object identities from byte-stream identities (size > 0) */
    SHA256_Final(id.digest.data(), &sha);
    return id;
}

/* The devices this process proves on and the keys uploaded to them, for the lifetime of the process.
 * Devices: ZKB_DEVICES="0,2,3" (a list), else ZKB_DEVICE=n (one device; what a one-process-per-GPU launcher sets), else
 * every visible device (CUDA_VISIBLE_DEVICES applies).  zkb_pool (include/zkb.h) gives one context per device.
 * Thread safety: the maps are guarded by mu(); a zkb_ctx serves one host thread at a time (zkb.h), so callers hold
 * gpu() around every zkb_* call -- host-side witness generation of concurrent callers overlaps, GPU use is serialised. */
class Session {
  public:
    static Session &get()
    {
        static Session s;
        return s;
    }
    std::mutex &gpu() { return gpu_mu_; }
    zkb_pool *pool()
    {
        std::lock_guard<std::mutex> g(mu_);
        return pool_locked();
    }
    zkb_ctx *ctx() { return zkb_pool_ctx(pool(), 0); }
    int devices() { return zkb_pool_size(pool()); }
    /* returns the cached device keys for `id`, or uploads them (every device of the pool) with `pack` */
    template <class PackFn> zkb_key_set *keys(const KeyId &id, PackFn pack)
    {
        {
            std::lock_guard<std::mutex> g(mu_);
            auto it = keys_.find(id);
            if (it != keys_.end()) return it->second;
        }
        std::unique_ptr<KeyArrays> arrays(new KeyArrays());
        pack(*arrays); /* may use the GPU (deferred import): outside mu_, under the caller's gpu() lock */
        std::lock_guard<std::mutex> g(mu_);
        auto it = keys_.find(id);
        if (it != keys_.end()) return it->second;
        zkb_key_set *k = NULL;
        if (zkb_pool_key_upload(pool_locked(), &arrays->desc, &k) != ZKB_OK)
            throw std::runtime_error(std::string("zkb_pool_key_upload failed: ") + zkb_pool_last_error(pool_));
        keys_[id] = k;
        return k;
    }
    template <class PackFn> const zkb_key *key(const KeyId &id, PackFn pack) { return zkb_pool_key(keys(id, pack), 0); }
    bool has_key(const KeyId &id)
    {
        std::lock_guard<std::mutex> g(mu_);
        return keys_.count(id) != 0;
    }
    /* first witness seen for a key: size its MSM windows for that witness's 0 / 1 statistics (zkb_key_tune) */
    void tune_once(zkb_key_set *ks, const uint64_t *w)
    {
        std::lock_guard<std::mutex> g(mu_);
        if (!tuned_.insert(ks).second) return;
        if (zkb_pool_key_tune(pool_locked(), ks, w) != ZKB_OK)
            throw std::runtime_error(std::string("zkb_key_tune failed: ") + zkb_pool_last_error(pool_));
    }
    /* SHA-256 identity of a serialised key, remembered per (address, size) of the caller's buffer: the 20 MB of a zklaim
     * key are hashed once, later calls re-check sampled probes of the buffer */
    KeyId id_of_buffer(const unsigned char *p, size_t n)
    {
        const uint64_t probe = probe_of(p, n);
        {
            std::lock_guard<std::mutex> g(mu_);
            auto it = seen_.find(std::make_pair((const void *)p, n));
            if (it != seen_.end() && it->second.first == probe) return it->second.second;
        }
        const KeyId id = key_id_of_bytes(p, n);
        std::lock_guard<std::mutex> g(mu_);
        if (seen_.size() > 64) seen_.clear();
        seen_[std::make_pair((const void *)p, n)] = std::make_pair(probe, id);
        return id;
    }
    void clear()
    {
        std::lock_guard<std::mutex> g(mu_);
        for (auto &kv : keys_) zkb_pool_key_free(kv.second);
        keys_.clear();
        tuned_.clear();
        seen_.clear();
        if (pool_) zkb_pool_shutdown(pool_);
        pool_ = NULL;
    }
    /* no CUDA call from a static destructor: at process exit the CUDA runtime may already be torn down (with several
     * devices in use that ends in a fault inside the driver).  Callers that want the memory back call zkb_bridge_reset();
     * otherwise the process exit releases it. */
    ~Session() {}

  private:
    zkb_pool *pool_locked()
    {
        if (pool_) return pool_;
        std::vector<int> ids;
        if (const char *list = getenv("ZKB_DEVICES")) {
            for (const char *q = list; *q;) {
                char *e;
                const long v = strtol(q, &e, 10);
                if (e == q) break;
                ids.push_back((int)v);
                q = *e == ',' ? e + 1 : e;
            }
        } else if (const char *one = getenv("ZKB_DEVICE")) {
            ids.push_back(atoi(one));
        }
        if (zkb_pool_init(ids.empty() ? NULL : ids.data(), (int)ids.size(), &pool_) != ZKB_OK)
            throw std::runtime_error(std::string("zkb_pool_init failed: ") + zkb_last_error(NULL));
        return pool_;
    }
    static uint64_t probe_of(const unsigned char *p, size_t n)
    {
        uint64_t h = fnv1a(&n, sizeof(n));
        const size_t edge = n < 2048 ? n : 2048;
        h = fnv1a(p, edge, h);
        h = fnv1a(p + n - edge, edge, h);
        for (size_t i = 1; i < 64; i++) h = fnv1a(p + (n / 64) * i, n - (n / 64) * i < 64 ? n - (n / 64) * i : 64, h);
        return h;
    }
    std::mutex mu_, gpu_mu_;
    zkb_pool *pool_ = NULL;
    std::map<KeyId, zkb_key_set *> keys_;
    std::set<const zkb_key_set *> tuned_;
    std::map<std::pair<const void *, size_t>, std::pair<uint64_t, KeyId>> seen_;
};

inline libsnark::r1cs_gg_ppzksnark_proof<PP> wrap_proof(const zkb_proof &p)
{
    G1T A = get_g1_jacobian(p.A);
    G2T B = get_g2_jacobian(p.B);
    G1T C = get_g1_jacobian(p.C);
    return libsnark::r1cs_gg_ppzksnark_proof<PP>(std::move(A), std::move(B), std::move(C));
}

/* ---- verifier side -------------------------------------------------------------------------------- */
/* r1cs_gg_ppzksnark_verification_key (r1cs_gg_ppzksnark.hpp:159-216) as the flat arrays of zkb_vk_desc */
struct VkArrays {
    uint64_t alpha_beta[48], gamma_g2[16], delta_g2[16];
    std::vector<uint64_t> gamma_abc;
    zkb_vk_desc desc;
};
inline void put_fq12(uint64_t *dst, const libff::alt_bn128_Fq12 &f)
{
    const libff::alt_bn128_Fq6 *h[2] = {&f.c0, &f.c1};
    for (int i = 0; i < 2; i++) {
        const libff::alt_bn128_Fq2 *c[3] = {&h[i]->c0, &h[i]->c1, &h[i]->c2};
        for (int j = 0; j < 3; j++) {
            put_fp(dst + 24 * i + 8 * j, c[j]->c0);
            put_fp(dst + 24 * i + 8 * j + 4, c[j]->c1);
        }
    }
}
inline void pack_verification_key(const libsnark::r1cs_gg_ppzksnark_verification_key<PP> &vk, VkArrays &a)
{
    put_fq12(a.alpha_beta, vk.alpha_g1_beta_g2);
    put_g2_affine(a.gamma_g2, vk.gamma_g2);
    put_g2_affine(a.delta_g2, vk.delta_g2);
    /* accumulation_vector: `first`, then the sparse `rest` over the domain 0 .. num_inputs - 1 (absent index = zero) */
    const size_t ni = vk.gamma_ABC_g1.rest.domain_size();
    a.gamma_abc.assign(8 * (ni + 1), 0);
    put_g1_affine(&a.gamma_abc[0], vk.gamma_ABC_g1.first);
    for (size_t k = 0; k < vk.gamma_ABC_g1.rest.indices.size(); k++)
        put_g1_affine(&a.gamma_abc[8 * (vk.gamma_ABC_g1.rest.indices[k] + 1)], vk.gamma_ABC_g1.rest.values[k]);
    a.desc.alpha_g1_beta_g2 = a.alpha_beta;
    a.desc.gamma_g2 = a.gamma_g2;
    a.desc.delta_g2 = a.delta_g2;
    a.desc.gamma_ABC_g1 = a.gamma_abc.data();
    a.desc.num_inputs = ni;
}
inline void pack_proof(const libsnark::r1cs_gg_ppzksnark_proof<PP> &pr, uint64_t *dst /* 48 limbs */)
{
    const libff::alt_bn128_Fq *g1[2][3] = {{&pr.g_A.X, &pr.g_A.Y, &pr.g_A.Z}, {&pr.g_C.X, &pr.g_C.Y, &pr.g_C.Z}};
    for (int k = 0; k < 3; k++) {
        put_fp(dst + 4 * k, *g1[0][k]);
        put_fp(dst + 36 + 4 * k, *g1[1][k]);
    }
    const libff::alt_bn128_Fq2 *g2[3] = {&pr.g_B.X, &pr.g_B.Y, &pr.g_B.Z};
    for (int k = 0; k < 3; k++) {
        put_fp(dst + 12 + 8 * k, g2[k]->c0);
        put_fp(dst + 12 + 8 * k + 4, g2[k]->c1);
    }
}
/* n proofs under one key: the decisions of r1cs_gg_ppzksnark_verifier_strong_IC (r1cs_gg_ppzksnark.tcc:612-622), which
 * rejects an input of the wrong length before any pairing work */
inline std::vector<bool> verifier_strong_IC_batch(const libsnark::r1cs_gg_ppzksnark_verification_key<PP> &vk,
                                                  const std::vector<std::vector<FrT>> &inputs,
                                                  const std::vector<libsnark::r1cs_gg_ppzksnark_proof<PP>> &proofs)
{
    if (inputs.size() != proofs.size()) throw std::runtime_error("verifier_strong_IC_batch: inputs / proofs differ in number");
    VkArrays a;
    pack_verification_key(vk, a);
    const size_t ni = a.desc.num_inputs, n = proofs.size();
    std::vector<bool> res(n, false);
    std::vector<size_t> live;
    for (size_t i = 0; i < n; i++)
        if (inputs[i].size() == ni) live.push_back(i);
    if (live.empty()) return res;
    std::vector<uint64_t> x(live.size() * ni * 4 + 4), pr(live.size() * 48);
    for (size_t k = 0; k < live.size(); k++) {
        for (size_t j = 0; j < ni; j++) put_fr(&x[(k * ni + j) * 4], inputs[live[k]][j]);
        pack_proof(proofs[live[k]], &pr[48 * k]);
    }
    std::vector<int32_t> acc(live.size(), 0);
    zkb_ctx *ctx = Session::get().ctx();
    std::lock_guard<std::mutex> gpu(Session::get().gpu());
    const int rc = zkb_verify_batch(ctx, &a.desc, x.data(), pr.data(), live.size(), acc.data());
    if (rc != ZKB_OK) throw std::runtime_error(std::string("zkb_verify_batch failed: ") + zkb_last_error(ctx));
    for (size_t k = 0; k < live.size(); k++) res[live[k]] = acc[k] != 0;
    return res;
}
/* Drop-in for r1cs_gg_ppzksnark_verifier_strong_IC<alt_bn128_pp> (r1cs_gg_ppzksnark.tcc:612-622; called from
 * verify_proof, zklaim/snark.cpp:53-63). */
template <typename ppT>
bool verifier_strong_IC(const libsnark::r1cs_gg_ppzksnark_verification_key<ppT> &vk,
                        const libsnark::r1cs_gg_ppzksnark_primary_input<ppT> &primary_input,
                        const libsnark::r1cs_gg_ppzksnark_proof<ppT> &proof)
{
    static_assert(std::is_same<ppT, PP>::value, "the zkb backend implements alt_bn128 only");
    return verifier_strong_IC_batch(vk, {primary_input}, {proof})[0];
}

/* full variable assignment as the contiguous limb array zkb_prove expects (Fp_model is 32 bytes, fp.hpp:42) */
inline void pack_assignment(const std::vector<FrT> &primary, const std::vector<FrT> &auxiliary, uint64_t *dst)
{
    for (size_t i = 0; i < primary.size(); i++) put_fr(dst + 4 * i, primary[i]);
    for (size_t i = 0; i < auxiliary.size(); i++) put_fr(dst + 4 * (primary.size() + i), auxiliary[i]);
}

/* the prover with the blinding scalars given (tests fix them; the drop-in below draws them) */
inline libsnark::r1cs_gg_ppzksnark_proof<PP> prover_with_rs(zkb_key_set *keys, const std::vector<FrT> &primary,
                                                           const std::vector<FrT> &auxiliary, const FrT &r, const FrT &s)
{
    size_t nv = 0;
    const zkb_key *key = zkb_pool_key(keys, 0);
    if (zkb_key_num_variables(key, &nv) != ZKB_OK || nv != primary.size() + auxiliary.size())
        throw std::runtime_error("zkb prover: the assignment does not have the key's number of variables");
    std::vector<uint64_t> w(4 * (primary.size() + auxiliary.size()));
    pack_assignment(primary, auxiliary, w.data());
    std::lock_guard<std::mutex> gpu(Session::get().gpu());
    Session::get().tune_once(keys, w.data());
    zkb_proof out;
    zkb_ctx *ctx = Session::get().ctx();
    const int rc = zkb_prove(ctx, key, w.data(), (const uint64_t *)r.mont_repr.data, (const uint64_t *)s.mont_repr.data, &out, NULL);
    if (rc != ZKB_OK) throw std::runtime_error(std::string("zkb_prove failed: ") + zkb_last_error(ctx));
    return wrap_proof(out);
}

/* Drop-in for r1cs_gg_ppzksnark_prover<alt_bn128_pp> (r1cs_gg_ppzksnark.tcc:390-505). */
template <typename ppT>
libsnark::r1cs_gg_ppzksnark_proof<ppT> prover(const libsnark::r1cs_gg_ppzksnark_proving_key<ppT> &pk,
                                              const libsnark::r1cs_gg_ppzksnark_primary_input<ppT> &primary_input,
                                              const libsnark::r1cs_gg_ppzksnark_auxiliary_input<ppT> &auxiliary_input)
{
    static_assert(std::is_same<ppT, PP>::value, "the zkb backend implements alt_bn128 only");
    zkb_key_set *keys = Session::get().keys(key_fingerprint(pk), [&](KeyArrays &k) { pack_proving_key(pk, k); });
    const FrT r = FrT::random_element(), s = FrT::random_element(); /* r1cs_gg_ppzksnark.tcc:418-419 */
    return prover_with_rs(keys, primary_input, auxiliary_input, r, s);
}

/* ---- the generator (include/zkb.h zkb_setup) ---------------------------------- */
inline G1T get_g1_affine(const uint64_t *src)
{
    bool zero = true;
    for (int i = 0; i < 8; i++) zero = zero && src[i] == 0;
    return zero ? G1T::zero() : G1T(get_fp(src), get_fp(src + 4), libff::alt_bn128_Fq::one());
}
inline G2T get_g2_affine(const uint64_t *src)
{
    bool zero = true;
    for (int i = 0; i < 16; i++) zero = zero && src[i] == 0;
    return zero ? G2T::zero()
                : G2T(libff::alt_bn128_Fq2(get_fp(src), get_fp(src + 4)), libff::alt_bn128_Fq2(get_fp(src + 8), get_fp(src + 12)),
                      libff::alt_bn128_Fq2::one());
}
inline void put_g1_jacobian(uint64_t *dst, const G1T &p)
{
    put_fp(dst, p.X);
    put_fp(dst + 4, p.Y);
    put_fp(dst + 8, p.Z);
}
inline void put_g2_jacobian(uint64_t *dst, const G2T &p)
{
    put_fp(dst, p.X.c0);
    put_fp(dst + 4, p.X.c1);
    put_fp(dst + 8, p.Y.c0);
    put_fp(dst + 12, p.Y.c1);
    put_fp(dst + 16, p.Z.c0);
    put_fp(dst + 20, p.Z.c1);
}

/* r1cs_gg_ppzksnark_generator (r1cs_gg_ppzksnark.tcc:211-388) with its randomness given (tests fix it; the drop-in
 * below draws it): every field / group computation runs in zkb_setup, the objects are libsnark's. */
inline libsnark::r1cs_gg_ppzksnark_keypair<PP> generator_with_randomness(const libsnark::r1cs_constraint_system<FrT> &r1cs, const FrT &t,
                                                                         const FrT &alpha, const FrT &beta, const FrT &gamma,
                                                                         const FrT &delta, const G1T &g1_gen, const G2T &g2_gen)
{
    KeyArrays k;
    for (int w = 0; w < 3; w++) pack_matrix(r1cs, w, k);
    const size_t nc = r1cs.num_constraints(), nv = r1cs.num_variables(), ni = r1cs.num_inputs();
    size_t m = 0;
    int kind = 0;
    if (zkb_domain_select(nc + ni + 1, &m, &kind) != ZKB_OK) throw std::runtime_error("zkb_setup: no evaluation domain for this size");
    uint64_t g1j[12], g2j[24];
    put_g1_jacobian(g1j, g1_gen);
    put_g2_jacobian(g2j, g2_gen);
    zkb_setup_desc d;
    memset(&d, 0, sizeof(d));
    d.num_constraints = nc;
    d.num_variables = nv;
    d.num_inputs = ni;
    zkb_csr *mats[3] = {&d.A, &d.B, &d.C};
    for (int w = 0; w < 3; w++) {
        mats[w]->row_ptr = k.row_ptr[w].data();
        mats[w]->col = k.col[w].data();
        mats[w]->coeff = k.coeff[w].data();
    }
    d.t = (const uint64_t *)t.mont_repr.data;
    d.alpha = (const uint64_t *)alpha.mont_repr.data;
    d.beta = (const uint64_t *)beta.mont_repr.data;
    d.gamma = (const uint64_t *)gamma.mont_repr.data;
    d.delta = (const uint64_t *)delta.mont_repr.data;
    d.g1_generator = g1j;
    d.g2_generator = g2j;
    std::vector<uint64_t> sing(3 * 8 + 3 * 16), A((nv + 1) * 8), B1((nv + 1) * 8), B2((nv + 1) * 16), H(m * 8), L((nv - ni) * 8 + 8),
        abc((ni + 1) * 8);
    zkb_setup_out o;
    o.alpha_g1 = &sing[0];
    o.beta_g1 = &sing[8];
    o.delta_g1 = &sing[16];
    o.beta_g2 = &sing[24];
    o.delta_g2 = &sing[40];
    o.gamma_g2 = &sing[56];
    o.A_query = A.data();
    o.B_g1 = B1.data();
    o.B_g2 = B2.data();
    o.H_query = H.data();
    o.L_query = L.data();
    o.gamma_ABC_g1 = abc.data();
    int swapped = 0;
    zkb_ctx *ctx = Session::get().ctx();
    std::unique_lock<std::mutex> gpu(Session::get().gpu());
    if (zkb_setup(ctx, &d, &o, &swapped) != ZKB_OK) throw std::runtime_error(std::string("zkb_setup failed: ") + zkb_last_error(ctx));

    libsnark::r1cs_constraint_system<FrT> cs(r1cs);
    if (swapped)
        for (auto &c : cs.constraints) std::swap(c.a, c.b); /* what swap_AB_if_beneficial did (r1cs.tcc:216-224) */
    libff::G1_vector<PP> Aq(nv + 1), Hq(m - 1), Lq(nv - ni), abc_rest(ni);
    for (size_t i = 0; i <= nv; i++) Aq[i] = get_g1_affine(&A[8 * i]);
    for (size_t i = 0; i + 1 < m; i++) Hq[i] = get_g1_affine(&H[8 * i]);
    for (size_t i = 0; i < nv - ni; i++) Lq[i] = get_g1_affine(&L[8 * i]);
    for (size_t i = 0; i < ni; i++) abc_rest[i] = get_g1_affine(&abc[8 * (i + 1)]);
    libsnark::knowledge_commitment_vector<G2T, G1T> Bq;
    Bq.domain_size_ = nv + 1;
    for (size_t i = 0; i <= nv; i++) {
        const G1T h = get_g1_affine(&B1[8 * i]);
        if (h.is_zero()) continue; /* kc_batch_exp keeps the non-zero scalars only (kc_multiexp.tcc:108-116) */
        Bq.indices.push_back(i);
        Bq.values.emplace_back(libsnark::knowledge_commitment<G2T, G1T>(get_g2_affine(&B2[16 * i]), h));
    }
    G1T alpha_g1 = get_g1_affine(o.alpha_g1), beta_g1 = get_g1_affine(o.beta_g1), delta_g1 = get_g1_affine(o.delta_g1);
    G2T beta_g2 = get_g2_affine(o.beta_g2), delta_g2 = get_g2_affine(o.delta_g2), gamma_g2 = get_g2_affine(o.gamma_g2);
    G1T abc0 = get_g1_affine(&abc[0]);
    libsnark::accumulation_vector<G1T> gamma_ABC_g1(std::move(abc0), std::move(abc_rest));
    /* the one pairing of the verification key (r1cs_gg_ppzksnark.tcc:338): zkb_pairing, bit-identical to reduced_pairing */
    uint64_t ab[48];
    if (zkb_pairing(ctx, o.alpha_g1, o.beta_g2, 1, 1, ab) != ZKB_OK)
        throw std::runtime_error(std::string("zkb_pairing failed: ") + zkb_last_error(ctx));
    gpu.unlock();
    libsnark::r1cs_gg_ppzksnark_verification_key<PP> vk(get_fq12(ab), gamma_g2, delta_g2, gamma_ABC_g1);
    G2T delta_g2_pk = delta_g2;
    libsnark::r1cs_gg_ppzksnark_proving_key<PP> pk(std::move(alpha_g1), std::move(beta_g1), std::move(beta_g2), std::move(delta_g1),
                                                   std::move(delta_g2_pk), std::move(Aq), std::move(Bq), std::move(Hq), std::move(Lq),
                                                   std::move(cs));
    return libsnark::r1cs_gg_ppzksnark_keypair<PP>(std::move(pk), std::move(vk));
}

/* Drop-in for r1cs_gg_ppzksnark_generator<alt_bn128_pp> (r1cs_gg_ppzksnark.tcc:211-388): the randomness is drawn with
 * the reference's own random_element() calls, in the reference's order (:221-225, :284, :294). */
template <typename ppT>
libsnark::r1cs_gg_ppzksnark_keypair<ppT> generator(const libsnark::r1cs_gg_ppzksnark_constraint_system<ppT> &r1cs)
{
    static_assert(std::is_same<ppT, PP>::value, "the zkb backend implements alt_bn128 only");
    const FrT t = FrT::random_element(), alpha = FrT::random_element(), beta = FrT::random_element(), gamma = FrT::random_element(),
              delta = FrT::random_element();
    const G1T g1 = G1T::random_element();
    const G2T g2 = G2T::random_element();
    return generator_with_randomness(r1cs, t, alpha, beta, gamma, delta, g1, g2);
}

} // namespace zkb_bridge
#endif /* ZKB_LIBSNARK_HPP */
