/*
 * zkb_zklaim_bridge.cpp -- zklaim's C++ bridge with the Groth16 prover moved to the GPU (libzkb.so).
 *
 * Built IN PLACE against the reference tree (bridge/Makefile, -I$(REF)): the unmodified reference bridge
 * zklaim/libsnark_wrapper.cpp (which pulls snark.cpp and zklaim_gadget.cpp) is included as is, so
 * gadget construction, witness generation, trusted setup, key / proof (de)serialisation and the verifier are
 * the reference's own host code.  What this file adds is the one function that changes,
 *
 *     int libsnark_prove(zklaim_ctx *ctx)                      zklaim/libsnark_wrapper.cpp:218-250
 *       -> generate_proof<ppT>(pk, ctx)                        zklaim/snark.cpp:108-127
 *            -> r1cs_gg_ppzksnark_prover<ppT>(pk, primary, auxiliary)   snark.cpp:126
 *
 * re-stated as zkb_libsnark_prove() with the prover call replaced by zkb_bridge::prover (zkb_libsnark.hpp)
 * and the 7-second key import (libsnark_import_pk, :228-230) done once per key instead of once per proof:
 * the device key is cached under a hash of ctx->pk.  Same contract: 0 = ok, 1 = witness unsatisfied;
 * ctx->proof is malloc'ed by the callee (:178-179).
 *
 * The zkbr_* functions at the end are measurement / test support (used by bench.py and tests/): they drive
 * the same code for a batch of credentials and expose the raw witnesses so that the C ABI can be timed
 * with host buffers.
 */
#include "zklaim/libsnark_wrapper.cpp" /* the reference bridge, unchanged */

#include <chrono>
#include <openssl/sha.h>

#include "zkb_libsnark.hpp"

using namespace zkb_bridge;

namespace {

struct CachedKey {
    const zkb_key *key = NULL;
    size_t num_variables = 0, num_inputs = 0;
};
std::map<uint64_t, CachedKey> g_keys_by_pk_hash;

/* device key for the serialised proving key in ctx->pk: imported and uploaded on first use only */
CachedKey &key_for_ctx(zklaim_ctx *ctx)
{
    const uint64_t h = fnv1a(ctx->pk, ctx->pk_size);
    auto it = g_keys_by_pk_hash.find(h);
    if (it != g_keys_by_pk_hash.end()) return it->second;
    r1cs_gg_ppzksnark_proving_key<ppT> pk;
    libsnark_import_pk(&pk, ctx->pk, ctx->pk_size);
    CachedKey ck;
    ck.key = Session::get().key(h, [&](KeyArrays &k) { pack_proving_key(pk, k); });
    ck.num_variables = pk.constraint_system.num_variables();
    ck.num_inputs = pk.constraint_system.num_inputs();
    return g_keys_by_pk_hash[h] = ck;
}

/* host half of generate_proof (snark.cpp:113-124): circuit + witness from the credential; false = unsatisfied */
bool build_witness(zklaim_ctx *ctx, std::vector<FrT> &primary, std::vector<FrT> &auxiliary)
{
    protoboard<FrT> pb;
    zklaim_gadget<FrT> g(pb, ctx);
    g.generate_r1cs_constraints();
    g.generate_r1cs_witness(ctx);
    if (!pb.is_satisfied()) return false;
    primary = pb.primary_input();
    auxiliary = pb.auxiliary_input();
    return true;
}

struct StdoutMute { /* what libsnark_wrapper.cpp:221-226 / :245-246 do around every call */
    int saved;
    StdoutMute()
    {
        libff::inhibit_profiling_counters = 1;
        libff::inhibit_profiling_info = 1;
        saved = dup(STDOUT_FILENO);
        fflush(stdout);
        close(STDOUT_FILENO);
    }
    ~StdoutMute()
    {
        fflush(stdout);
        dup2(saved, STDOUT_FILENO);
        close(saved);
    }
};

} // namespace

extern "C" {

/* Replacement of libsnark_prove (zklaim/zklaim.h:259, libsnark_wrapper.cpp:218-250). */
int zkb_libsnark_prove(zklaim_ctx *ctx)
{
    StdoutMute mute;
    ppT::init_public_params();
    try {
        CachedKey &ck = key_for_ctx(ctx);
        std::vector<FrT> primary, auxiliary;
        if (!build_witness(ctx, primary, auxiliary)) return 1;
        const FrT r = FrT::random_element(), s = FrT::random_element();
        const r1cs_gg_ppzksnark_proof<ppT> proof = prover_with_rs(ck.key, primary, auxiliary, r, s);
        ctx->proof_size = libsnark_export_proof(proof, &ctx->proof);
        return 0;
    } catch (const std::exception &e) {
        fprintf(stderr, "zkb_libsnark_prove: %s\n", e.what());
        return -1;
    }
}

/* `n` credentials that share one proving key: witnesses on the host, one batched GPU call. */
int zkb_libsnark_prove_batch(zklaim_ctx **ctxs, size_t n)
{
    if (!n) return 0;
    StdoutMute mute;
    ppT::init_public_params();
    try {
        CachedKey &ck = key_for_ctx(ctxs[0]);
        const size_t nv = ck.num_variables;
        std::vector<uint64_t> w(n * nv * 4), r(n * 4), s(n * 4);
        for (size_t i = 0; i < n; i++) {
            std::vector<FrT> primary, auxiliary;
            if (!build_witness(ctxs[i], primary, auxiliary)) return 1;
            if (primary.size() + auxiliary.size() != nv) return -2;
            pack_assignment(primary, auxiliary, &w[i * nv * 4]);
            put_fr(&r[4 * i], FrT::random_element());
            put_fr(&s[4 * i], FrT::random_element());
        }
        Session::get().tune_once(ck.key, w.data());
        std::vector<zkb_proof> out(n);
        zkb_ctx *z = Session::get().ctx();
        if (zkb_prove_batch(z, ck.key, n, w.data(), r.data(), s.data(), out.data()) != ZKB_OK) {
            fprintf(stderr, "zkb_prove_batch: %s\n", zkb_last_error(z));
            return -1;
        }
        for (size_t i = 0; i < n; i++) ctxs[i]->proof_size = libsnark_export_proof(wrap_proof(out[i]), &ctxs[i]->proof);
        return 0;
    } catch (const std::exception &e) {
        fprintf(stderr, "zkb_libsnark_prove_batch: %s\n", e.what());
        return -1;
    }
}

void zkb_bridge_reset(void)
{
    g_keys_by_pk_hash.clear();
    Session::get().clear();
}

} // extern "C"

/* ---------------------------------------------------------------------------------------------------
 * measurement / test support
 * --------------------------------------------------------------------------------------------------- */
namespace {

/* A credential as zklaim/main.c:70-101,194-208 builds it: `k` payloads of five uint64 attributes
 * (23, 1, 2, 3, 599) and a salt, SHA-256 commitment over the 48-byte pre-image, prover predicate
 * "attribute 0 >= 20", the others noop.  The salt comes from `seed` instead of /dev/urandom
 * (zklaim/zklaim.c:114-120) so that runs are reproducible.  No signature (zklaim_ecc.c needs libgcrypt). */
struct Credential {
    zklaim_ctx ctx;
    std::vector<zklaim_wrap_payload_ctx> chain;
    Credential(size_t k, uint64_t seed) : chain(k)
    {
        memset(&ctx, 0, sizeof(ctx));
        for (size_t i = 0; i < k; i++) {
            zklaim_payload &pl = chain[i].pl;
            memset(&pl, 0, sizeof(pl));
            const uint64_t attr[5] = {23 + i, 1, 2, 3, 599};
            uint64_t x = seed * 0x2545f4914f6cdd1dull + (i + 1) * 0x9e3779b97f4a7c15ull;
            x ^= x >> 29;
            x *= 0xbf58476d1ce4e5b9ull;
            x ^= x >> 32;
            pl.salt = x;
            memcpy(pl.pre, attr, sizeof(attr));
            memcpy(pl.pre + sizeof(attr), &pl.salt, sizeof(pl.salt));
            SHA256(pl.pre, sizeof(pl.pre), pl.hash);
            for (int j = 0; j < 5; j++) {
                pl.data_ref[j] = j == 0 ? 20 : 0;
                pl.data_op[j] = j == 0 ? (enum zklaim_op)(zklaim_greater | zklaim_eq) : zklaim_noop;
            }
            chain[i].next = i + 1 < k ? &chain[i + 1] : NULL;
        }
        ctx.num_of_payloads = k;
        ctx.pl_ctx_head = k ? &chain[0] : NULL;
    }
    ~Credential()
    {
        free(ctx.pk);
        free(ctx.vk);
        free(ctx.proof);
    }
};

struct Bench {
    size_t payloads;
    std::unique_ptr<Credential> issuer; /* holds pk / vk */
    CachedKey key;
    double setup_s = 0, import_upload_s = 0;
};

double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

} // namespace

extern "C" {

/* trusted setup for a `payloads`-payload credential through the reference bridge, then key import + upload */
void *zkbr_new(size_t payloads, uint64_t seed)
{
    try {
        Bench *b = new Bench();
        b->payloads = payloads;
        b->issuer.reset(new Credential(payloads, seed));
        double t = now_s();
        if (libsnark_trusted_setup(&b->issuer->ctx) != 0) {
            delete b;
            return NULL;
        }
        b->setup_s = now_s() - t;
        t = now_s();
        {
            StdoutMute mute;
            ppT::init_public_params();
            b->key = key_for_ctx(&b->issuer->ctx);
        }
        b->import_upload_s = now_s() - t;
        return b;
    } catch (const std::exception &e) {
        fprintf(stderr, "zkbr_new: %s\n", e.what());
        return NULL;
    }
}

void zkbr_free(void *h) { delete (Bench *)h; }

/* dims[0..6) = num_variables, num_inputs, pk bytes, vk bytes, setup seconds * 1e6, import + upload seconds * 1e6 */
void zkbr_dims(void *h, size_t *dims)
{
    Bench *b = (Bench *)h;
    dims[0] = b->key.num_variables;
    dims[1] = b->key.num_inputs;
    dims[2] = b->issuer->ctx.pk_size;
    dims[3] = b->issuer->ctx.vk_size;
    dims[4] = (size_t)(b->setup_s * 1e6);
    dims[5] = (size_t)(b->import_upload_s * 1e6);
}

const zkb_key *zkbr_key(void *h) { return ((Bench *)h)->key.key; }
zkb_ctx *zkbr_ctx(void) { return Session::get().ctx(); }

/* `count` witnesses (num_variables x 4 limbs each) for credentials with salts seed0, seed0 + 1, ...:
 * the reference's host-side gadget code (zklaim_gadget.cpp:705-783).  Returns 0, or 1 if one is unsatisfied. */
int zkbr_witnesses(void *h, size_t count, uint64_t seed0, uint64_t *w_out)
{
    Bench *b = (Bench *)h;
    StdoutMute mute;
    ppT::init_public_params();
    const size_t nv = b->key.num_variables;
    for (size_t i = 0; i < count; i++) {
        Credential c(b->payloads, seed0 + i);
        std::vector<FrT> primary, auxiliary;
        if (!build_witness(&c.ctx, primary, auxiliary)) return 1;
        if (primary.size() + auxiliary.size() != nv) return 2;
        pack_assignment(primary, auxiliary, w_out + i * nv * 4);
        if (i == 0) Session::get().tune_once(b->key.key, w_out);
    }
    return 0;
}

/* fresh blinding scalars, drawn as the reference draws them (Fr::random_element, fp.tcc:695-721) */
void zkbr_random_fr(size_t count, uint64_t *out)
{
    ppT::init_public_params();
    for (size_t i = 0; i < count; i++) put_fr(out + 4 * i, FrT::random_element());
}

/* the reference verifier (r1cs_gg_ppzksnark_verifier_strong_IC via the key in ctx->vk) on a zkb proof:
 * primary = the first num_inputs elements of the witness.  Returns 1 if accepted. */
int zkbr_verify(void *h, const uint64_t *w, const zkb_proof *proof)
{
    Bench *b = (Bench *)h;
    StdoutMute mute;
    ppT::init_public_params();
    r1cs_gg_ppzksnark_verification_key<ppT> vk;
    libsnark_import_vk(&vk, b->issuer->ctx.vk, b->issuer->ctx.vk_size);
    std::vector<FrT> primary(b->key.num_inputs);
    for (size_t i = 0; i < primary.size(); i++) memcpy(primary[i].mont_repr.data, w + 4 * i, 32);
    return r1cs_gg_ppzksnark_verifier_strong_IC<ppT>(vk, primary, wrap_proof(*proof)) ? 1 : 0;
}

/* end to end through the zklaim-facing entry points: `count` credentials -> zkb_libsnark_prove_batch ->
 * libsnark_verify (the reference's) on each.  Returns the number of accepted proofs; seconds[0] = prove wall
 * time (host witness generation + GPU), seconds[1] = verify wall time. */
int zkbr_roundtrip(void *h, size_t count, uint64_t seed0, double *seconds)
{
    Bench *b = (Bench *)h;
    std::vector<std::unique_ptr<Credential>> creds;
    std::vector<zklaim_ctx *> ptrs;
    for (size_t i = 0; i < count; i++) {
        creds.emplace_back(new Credential(b->payloads, seed0 + i));
        zklaim_ctx &c = creds.back()->ctx;
        c.pk_size = b->issuer->ctx.pk_size;
        c.pk = (unsigned char *)malloc(c.pk_size);
        memcpy(c.pk, b->issuer->ctx.pk, c.pk_size);
        c.vk_size = b->issuer->ctx.vk_size;
        c.vk = (unsigned char *)malloc(c.vk_size);
        memcpy(c.vk, b->issuer->ctx.vk, c.vk_size);
        ptrs.push_back(&c);
    }
    double t = now_s();
    const int rc = count == 1 ? zkb_libsnark_prove(ptrs[0]) : zkb_libsnark_prove_batch(ptrs.data(), count);
    seconds[0] = now_s() - t;
    if (rc != 0) return -1;
    t = now_s();
    int ok = 0;
    for (size_t i = 0; i < count; i++) ok += libsnark_verify(ptrs[i]) == 0 ? 1 : 0;
    seconds[1] = now_s() - t;
    return ok;
}

} // extern "C"
