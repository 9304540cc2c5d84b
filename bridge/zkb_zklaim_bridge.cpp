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
 * and the 7-second key import (libsnark_import_pk, :228-230) done once per key instead of once per proof -- and its
 * square roots on the GPU (import_pk_deferred): the device keys are cached under the SHA-256 of ctx->pk.
 * Same contract: 0 = ok, 1 = witness unsatisfied;
 * ctx->proof is malloc'ed by the callee (:178-179).
 *
 * The zkbr_* functions at the end are measurement / test support (used by bench.py and tests/): they drive
 * the same code for a batch of credentials and expose the raw witnesses so that the C ABI can be timed
 * with host buffers.
 */
#include "zklaim/libsnark_wrapper.cpp" /* the reference bridge, unchanged */

#include <chrono>
#include <thread>
#include <fcntl.h>
#include <openssl/sha.h>

#include <libsnark/gadgetlib1/gadgets/hashes/sha256/sha256_gadget.hpp>

#include "zkb_libsnark.hpp"

using namespace zkb_bridge;

namespace {

struct CachedKey {
    zkb_key_set *keys = NULL; /* one replica per device of the session's pool */
    size_t num_variables = 0, num_inputs = 0;
    const zkb_key *key() const { return zkb_pool_key(keys, 0); }
};
std::mutex g_cache_mu;
std::map<KeyId, CachedKey> g_keys_by_id;

bool host_import_enabled()
{
    const char *e = getenv("ZKB_HOST_IMPORT"); /* 1: the reference's own importer (one square root per point on the host) */
    return e && atoi(e) != 0;
}

/* device keys for the serialised proving key in ctx->pk: imported and uploaded on first use only.  The identity of a key
 * is the SHA-256 of its bytes (+ their number).  Import = libsnark's framing read here, the square roots of the
 * compressed points on the GPU (import_pk_deferred): 7.3 s -> a fraction of a second for the zklaim key. */
CachedKey key_for_ctx(zklaim_ctx *ctx)
{
    if (!ctx->pk || !ctx->pk_size) throw std::runtime_error("no proving key in the context");
    const KeyId id = Session::get().id_of_buffer(ctx->pk, ctx->pk_size);
    {
        std::lock_guard<std::mutex> g(g_cache_mu);
        auto it = g_keys_by_id.find(id);
        if (it != g_keys_by_id.end()) return it->second;
    }
    CachedKey ck;
    ck.keys = Session::get().keys(id, [&](KeyArrays &k) {
        if (host_import_enabled()) {
            r1cs_gg_ppzksnark_proving_key<ppT> pk;
            libsnark_import_pk(&pk, ctx->pk, ctx->pk_size);
            pack_proving_key(pk, k);
            ck.num_variables = pk.constraint_system.num_variables();
            ck.num_inputs = pk.constraint_system.num_inputs();
        } else {
            std::lock_guard<std::mutex> gpu(Session::get().gpu());
            import_pk_deferred(Session::get().ctx(), ctx->pk, ctx->pk_size, k, &ck.num_variables, &ck.num_inputs);
        }
    });
    if (!ck.num_variables) { /* another thread uploaded the key in the meantime: ask the device key */
        zkb_key_num_variables(ck.key(), &ck.num_variables);
    }
    std::lock_guard<std::mutex> g(g_cache_mu);
    auto it = g_keys_by_id.find(id);
    if (it != g_keys_by_id.end()) return it->second;
    return g_keys_by_id[id] = ck;
}

/* a key this process has just made (zkb_libsnark_trusted_setup): upload it from the objects at hand, so that the first
 * proof under it does not serialise -> parse -> decompress what is already in memory */
void register_fresh_key(zklaim_ctx *ctx, const r1cs_gg_ppzksnark_proving_key<ppT> &pk)
{
    const KeyId id = Session::get().id_of_buffer(ctx->pk, ctx->pk_size);
    CachedKey ck;
    ck.keys = Session::get().keys(id, [&](KeyArrays &k) { pack_proving_key(pk, k); });
    ck.num_variables = pk.constraint_system.num_variables();
    ck.num_inputs = pk.constraint_system.num_inputs();
    std::lock_guard<std::mutex> g(g_cache_mu);
    g_keys_by_id[id] = ck;
}

/* host half of generate_proof exactly as the reference runs it (snark.cpp:113-124): circuit + constraints + witness +
 * host-side satisfiability check per proof; false = unsatisfied.  ~110 ms per 1-payload credential on one core, of
 * which 10 ms are the witness.  Used when ZKB_HOST_CHECK=1. */
bool build_witness_reference(zklaim_ctx *ctx, std::vector<FrT> &primary, std::vector<FrT> &auxiliary)
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

/* the circuit's shape depends on the payload count only (zklaim_gadget.cpp:348-540 walks
 * the chain, predicates and values are witness data), constraint generation is not needed to compute the witness
 * (zklaim_gadget.cpp:705-783 only assigns values), and the GPU prover checks <a,z><b,z> = <c,z> for every constraint
 * itself (k_sat_check -> ZKB_ERR_UNSAT).  So a gadget is built once per payload count and thread and re-witnessed per
 * credential: ~10 ms instead of ~110 ms per proof, and credentials of a batch are processed by all host threads. */
struct WitnessEngine {
    std::vector<zklaim_wrap_payload_ctx> proto;
    zklaim_ctx pctx;
    protoboard<FrT> pb;
    std::unique_ptr<zklaim_gadget<FrT>> g;
    explicit WitnessEngine(size_t k) : proto(k)
    {
        memset(&pctx, 0, sizeof(pctx));
        for (size_t i = 0; i < k; i++) {
            memset(&proto[i].pl, 0, sizeof(proto[i].pl));
            proto[i].next = i + 1 < k ? &proto[i + 1] : NULL;
        }
        pctx.num_of_payloads = k;
        pctx.pl_ctx_head = k ? &proto[0] : NULL;
        g.reset(new zklaim_gadget<FrT>(pb, &pctx));
    }
    void run(zklaim_ctx *ctx, std::vector<FrT> &primary, std::vector<FrT> &auxiliary)
    {
        pb.clear_values(); /* a fresh protoboard starts from zeros (protoboard.tcc) */
        g->generate_r1cs_witness(ctx);
        primary = pb.primary_input();
        auxiliary = pb.auxiliary_input();
    }
};

bool host_check_enabled()
{
    const char *e = getenv("ZKB_HOST_CHECK");
    return e && atoi(e) != 0;
}

/* witness of one credential; false = the reference's host-side check (if enabled) found it unsatisfied */
bool build_witness(zklaim_ctx *ctx, std::vector<FrT> &primary, std::vector<FrT> &auxiliary)
{
    if (host_check_enabled()) return build_witness_reference(ctx, primary, auxiliary);
    static thread_local std::map<size_t, std::unique_ptr<WitnessEngine>> engines;
    std::unique_ptr<WitnessEngine> &e = engines[ctx->num_of_payloads];
    if (!e) e.reset(new WitnessEngine(ctx->num_of_payloads));
    e->run(ctx, primary, auxiliary);
    return true;
}

struct StdoutMute { /* what libsnark_wrapper.cpp:221-226 / :245-246 do around every call: the gadget prints to stdout */
    /* several host threads may be inside the bridge at once: the first one in redirects, the last one out restores */
    static std::mutex &mu()
    {
        static std::mutex m;
        return m;
    }
    static int &depth()
    {
        static int d = 0;
        return d;
    }
    static int &saved()
    {
        static int fd = -1;
        return fd;
    }
    StdoutMute()
    {
        std::lock_guard<std::mutex> g(mu());
        if (depth()++) return;
        libff::inhibit_profiling_counters = 1;
        libff::inhibit_profiling_info = 1;
        fflush(stdout);
        saved() = dup(STDOUT_FILENO);
        /* the reference close()s fd 1; pointing it at /dev/null silences the same output without leaving descriptor 1
         * free for whoever opens a file next (CUDA, a profiler, another thread), who would then lose it on restore */
        const int nul = open("/dev/null", O_WRONLY);
        if (nul >= 0) {
            dup2(nul, STDOUT_FILENO);
            close(nul);
        }
        std::cout.setstate(std::ios_base::failbit); /* the gadget streams every witness value to cout (:770, :781) */
    }
    ~StdoutMute()
    {
        std::lock_guard<std::mutex> g(mu());
        if (--depth()) return;
        fflush(stdout);
        std::cout.clear();
        if (saved() >= 0) {
            dup2(saved(), STDOUT_FILENO);
            close(saved());
            saved() = -1;
        }
    }
};

} // namespace

extern "C" {

/* Replacement of libsnark_trusted_setup (zklaim/zklaim.h:258, libsnark_wrapper.cpp:195-215): the constraint system
 * comes from the reference's gadget exactly as in generate_keypair (snark.cpp:77-92), the generator runs on the GPU
 * (zkb_bridge::generator -> zkb_setup), and both keys are serialised by the reference's own exporters. */
int zkb_libsnark_trusted_setup(zklaim_ctx *ctx)
{
    StdoutMute mute;
    ppT::init_public_params();
    try {
        protoboard<FrT> pb;
        zklaim_gadget<FrT> g(pb, ctx);
        g.generate_r1cs_constraints();
        const r1cs_constraint_system<FrT> cs = pb.get_constraint_system();
        r1cs_gg_ppzksnark_keypair<ppT> keypair = zkb_bridge::generator<ppT>(cs);
        ctx->vk_size = libsnark_export_vk(keypair.vk, &ctx->vk);
        ctx->pk_size = libsnark_export_pk(keypair.pk, &ctx->pk);
        register_fresh_key(ctx, keypair.pk);
        return 0;
    } catch (const std::exception &e) {
        fprintf(stderr, "zkb_libsnark_trusted_setup: %s\n", e.what());
        return -1;
    }
}

/* Replacement of libsnark_prove (zklaim/zklaim.h:259, libsnark_wrapper.cpp:218-250).  Callable from several host
 * threads: witness generation runs concurrently, the GPU is used by one caller at a time. */
int zkb_libsnark_prove(zklaim_ctx *ctx)
{
    StdoutMute mute;
    ppT::init_public_params();
    try {
        const CachedKey ck = key_for_ctx(ctx);
        std::vector<FrT> primary, auxiliary;
        if (!build_witness(ctx, primary, auxiliary)) return 1;
        /* a context whose payload count does not fit the circuit of its proving key (zklaim_gadget sizes the circuit from
         * num_of_payloads) would make the device read past the assignment */
        if (primary.size() + auxiliary.size() != ck.num_variables) return -2;
        const FrT r = FrT::random_element(), s = FrT::random_element();
        std::vector<uint64_t> w(4 * (primary.size() + auxiliary.size()));
        pack_assignment(primary, auxiliary, w.data());
        zkb_proof out;
        int rc;
        std::string err;
        {
            std::lock_guard<std::mutex> gpu(Session::get().gpu());
            Session::get().tune_once(ck.keys, w.data());
            zkb_ctx *z = Session::get().ctx();
            rc = zkb_prove(z, ck.key(), w.data(), (const uint64_t *)r.mont_repr.data, (const uint64_t *)s.mont_repr.data, &out, NULL);
            if (rc != ZKB_OK) err = zkb_last_error(z);
        }
        if (rc == ZKB_ERR_UNSAT) return 1; /* libsnark_wrapper.cpp:233-240: witness unsatisfied */
        if (rc != ZKB_OK) {
            fprintf(stderr, "zkb_prove: %s\n", err.c_str());
            return -1;
        }
        ctx->proof_size = libsnark_export_proof(wrap_proof(out), &ctx->proof);
        return 0;
    } catch (const std::exception &e) {
        fprintf(stderr, "zkb_libsnark_prove: %s\n", e.what());
        return -1;
    }
}

/* `n` credentials that share one proving key: witnesses on the host (all host threads), then the batch is split over
 * every device of the session (zkb_pool_prove_batch: one host thread and one proof stream per GPU).
 * results (optional, n entries): what libsnark_prove would have returned for credential i -- 0 = proof written to
 * ctxs[i]->proof, 1 = its witness does not satisfy the constraint system (no proof).  Returns 0 when every credential
 * got a proof, 1 when at least one was unsatisfied (the others still have theirs), < 0 when the call itself failed
 * (-2: a context with a different proving key or payload count than ctxs[0]). */
int zkb_libsnark_prove_batch_ex(zklaim_ctx **ctxs, size_t n, int *results)
{
    if (!n) return 0;
    StdoutMute mute;
    ppT::init_public_params();
    try {
        for (size_t i = 1; i < n; i++) /* one key for the whole batch, as verify_batch requires of the vk */
            if (ctxs[i]->pk_size != ctxs[0]->pk_size || (ctxs[i]->pk != ctxs[0]->pk && memcmp(ctxs[i]->pk, ctxs[0]->pk, ctxs[0]->pk_size)))
                return -2;
        const CachedKey ck = key_for_ctx(ctxs[0]);
        const size_t nv = ck.num_variables;
        /* 55 MB for 64 one-payload credentials: kept per caller thread (a fresh vector costs the zero-fill and the page
         * faults on every batch); entries of credentials the host check rejects are zeroed below */
        static thread_local std::vector<uint64_t> w;
        if (w.size() < n * nv * 4) w.resize(n * nv * 4);
        uint64_t *const wp = w.data(); /* the OpenMP workers below must not name `w`: each has its own (empty) thread_local */
        std::vector<uint64_t> r(n * 4), s(n * 4);
        std::vector<int> res(n, 0);
        int bad = 0;
#pragma omp parallel for schedule(dynamic, 1)
        for (size_t i = 0; i < n; i++) {
            std::vector<FrT> primary, auxiliary;
            if (!build_witness(ctxs[i], primary, auxiliary)) { /* only with ZKB_HOST_CHECK=1 */
                res[i] = 1;
                memset(wp + i * nv * 4, 0, nv * 32);
                continue;
            }
            if (primary.size() + auxiliary.size() != nv) {
#pragma omp atomic write
                bad = 2;
                continue;
            }
            pack_assignment(primary, auxiliary, wp + i * nv * 4);
        }
        if (bad) return -2;
        /* credentials the host check rejected keep a zero assignment: the device reports them unsatisfied as well */
        for (size_t i = 0; i < n; i++) {
            put_fr(&r[4 * i], FrT::random_element());
            put_fr(&s[4 * i], FrT::random_element());
        }
        std::vector<zkb_proof> out(n);
        std::vector<int32_t> status(n, 0);
        int rc;
        std::string err;
        {
            std::lock_guard<std::mutex> gpu(Session::get().gpu());
            size_t first_ok = 0;
            while (first_ok < n && res[first_ok]) first_ok++;
            if (first_ok < n) Session::get().tune_once(ck.keys, wp + first_ok * nv * 4);
            zkb_pool *pool = Session::get().pool();
            rc = zkb_pool_prove_batch(pool, ck.keys, n, wp, r.data(), s.data(), out.data(), status.data());
            if (rc != ZKB_OK) err = zkb_pool_last_error(pool);
        }
        if (rc != ZKB_OK && rc != ZKB_ERR_UNSAT) {
            fprintf(stderr, "zkb_pool_prove_batch: %s\n", err.c_str());
            return -1;
        }
        int any = 0;
        for (size_t i = 0; i < n; i++) {
            if (status[i] != ZKB_OK) res[i] = 1;
            if (res[i] == 0) ctxs[i]->proof_size = libsnark_export_proof(wrap_proof(out[i]), &ctxs[i]->proof);
            any |= res[i];
            if (results) results[i] = res[i];
        }
        return any ? 1 : 0;
    } catch (const std::exception &e) {
        fprintf(stderr, "zkb_libsnark_prove_batch: %s\n", e.what());
        return -1;
    }
}

int zkb_libsnark_prove_batch(zklaim_ctx **ctxs, size_t n) { return zkb_libsnark_prove_batch_ex(ctxs, n, NULL); }

/* devices the bridge proves on (ZKB_DEVICES / ZKB_DEVICE / all visible) */
int zkb_libsnark_devices(void)
{
    try {
        return Session::get().devices();
    } catch (const std::exception &e) {
        fprintf(stderr, "zkb_libsnark_devices: %s\n", e.what());
        return -1;
    }
}

/* Replacement of libsnark_verify (zklaim/zklaim.h:262, libsnark_wrapper.cpp:252-276) for `n` credentials that share one
 * verification key: keys and proofs are parsed by the reference's own importers, the public input is the reference's
 * zklaim_input_map (zklaim_gadget.cpp:116), the pairing work runs on the GPU.  results[i] = libsnark_verify's return
 * value for credential i (0 = accepted, 1 = rejected); returns 0, or < 0 when the call itself failed. */
int zkb_libsnark_verify_batch(zklaim_ctx **ctxs, size_t n, int *results)
{
    if (!n) return 0;
    StdoutMute mute;
    ppT::init_public_params();
    try {
        r1cs_gg_ppzksnark_verification_key<ppT> vk;
        libsnark_import_vk(&vk, ctxs[0]->vk, ctxs[0]->vk_size);
        std::vector<std::vector<FrT>> inputs(n);
        std::vector<r1cs_gg_ppzksnark_proof<ppT>> proofs(n);
        for (size_t i = 0; i < n; i++) {
            if (ctxs[i]->vk_size != ctxs[0]->vk_size || memcmp(ctxs[i]->vk, ctxs[0]->vk, ctxs[0]->vk_size)) return -2;
            libsnark_import_proof(&proofs[i], ctxs[i]->proof, ctxs[i]->proof_size);
            inputs[i] = zklaim_input_map<FrT>(ctxs[i]);
        }
        const std::vector<bool> ok = zkb_bridge::verifier_strong_IC_batch(vk, inputs, proofs);
        for (size_t i = 0; i < n; i++) results[i] = ok[i] ? 0 : 1;
        return 0;
    } catch (const std::exception &e) {
        fprintf(stderr, "zkb_libsnark_verify_batch: %s\n", e.what());
        return -1;
    }
}

/* Replacement of libsnark_verify for one credential: 0 = accepted. */
int zkb_libsnark_verify(zklaim_ctx *ctx)
{
    int res = 1;
    const int rc = zkb_libsnark_verify_batch(&ctx, 1, &res);
    return rc ? rc : res;
}

void zkb_bridge_reset(void)
{
    {
        std::lock_guard<std::mutex> g(g_cache_mu);
        g_keys_by_id.clear();
    }
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

const zkb_key *zkbr_key(void *h) { return ((Bench *)h)->key.key(); }
zkb_ctx *zkbr_ctx(void) { return Session::get().ctx(); }

/* `count` witnesses (num_variables x 4 limbs each) for credentials with salts seed0, seed0 + 1, ...:
 * the reference's host-side gadget code (zklaim_gadget.cpp:705-783).  Returns 0, or 1 if one is unsatisfied. */
int zkbr_witnesses(void *h, size_t count, uint64_t seed0, uint64_t *w_out)
{
    Bench *b = (Bench *)h;
    StdoutMute mute;
    ppT::init_public_params();
    const size_t nv = b->key.num_variables;
    int bad = 0;
#pragma omp parallel for schedule(dynamic, 1)
    for (size_t i = 0; i < count; i++) {
        Credential c(b->payloads, seed0 + i);
        std::vector<FrT> primary, auxiliary;
        if (!build_witness(&c.ctx, primary, auxiliary) || primary.size() + auxiliary.size() != nv) {
#pragma omp atomic write
            bad = 1;
            continue;
        }
        pack_assignment(primary, auxiliary, w_out + i * nv * 4);
    }
    if (bad) return bad;
    std::lock_guard<std::mutex> gpu(Session::get().gpu());
    Session::get().tune_once(b->key.keys, w_out);
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

/* a credential whose commitment does not match its pre-image (one hash byte flipped) through zkb_libsnark_prove:
 * the reference returns 1 for it (libsnark_wrapper.cpp:233-240); so must the GPU path (k_sat_check -> ZKB_ERR_UNSAT) */
int zkbr_prove_tampered(void *h, uint64_t seed)
{
    Bench *b = (Bench *)h;
    Credential c(b->payloads, seed);
    c.chain[0].pl.hash[3] ^= 0x40;
    c.ctx.pk_size = b->issuer->ctx.pk_size;
    c.ctx.pk = (unsigned char *)malloc(c.ctx.pk_size);
    memcpy(c.ctx.pk, b->issuer->ctx.pk, c.ctx.pk_size);
    return zkb_libsnark_prove(&c.ctx);
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

/* `count` credentials -> zkb_libsnark_prove_batch; credential `tamper` (if < count) gets the proof of its neighbour;
 * then BOTH verifiers: the reference's libsnark_verify per credential and zkb_libsnark_verify_batch.  Returns the
 * number of credentials on which the two agree (must be `count`), < 0 on failure; accepted[0] = how many the GPU
 * verifier accepted; seconds[0] = reference verify wall time, seconds[1] = GPU batch verify wall time. */
int zkbr_verify_roundtrip(void *h, size_t count, uint64_t seed0, size_t tamper, int *accepted, double *seconds)
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
    if ((count == 1 ? zkb_libsnark_prove(ptrs[0]) : zkb_libsnark_prove_batch(ptrs.data(), count)) != 0) return -1;
    if (tamper < count && count > 1) {
        const size_t other = (tamper + 1) % count;
        memcpy(ptrs[tamper]->proof, ptrs[other]->proof, ptrs[other]->proof_size);
    }
    std::vector<int> ref_res(count), gpu_res(count, -7);
    double t = now_s();
    for (size_t i = 0; i < count; i++) ref_res[i] = libsnark_verify(ptrs[i]);
    seconds[0] = now_s() - t;
    t = now_s();
    if (zkb_libsnark_verify_batch(ptrs.data(), count, gpu_res.data()) != 0) return -2;
    seconds[1] = now_s() - t;
    int agree = 0;
    *accepted = 0;
    for (size_t i = 0; i < count; i++) {
        agree += ref_res[i] == gpu_res[i] ? 1 : 0;
        *accepted += gpu_res[i] == 0 ? 1 : 0;
    }
    return agree;
}

/* The zklaim-facing throughput: `threads` host threads, each proving `batches` batches of
 * `batch` fresh credentials through zkb_libsnark_prove_batch.  The entry point is thread-safe: the witness generation of
 * one caller (zklaim's gadget code, all OpenMP threads) overlaps the GPU phase of another, which is how a service keeps
 * the device fed.  Returns the credentials proved (every one checked to have a proof; the last batch of thread 0 is run
 * through the reference's libsnark_verify), < 0 on failure; seconds[0] = wall time of the whole stream. */
int zkbr_stream(void *h, size_t batches, size_t batch, int threads, uint64_t seed0, double *seconds)
{
    Bench *b = (Bench *)h;
    if (threads < 1) threads = 1;
    std::vector<int> rc(threads, 0);
    std::vector<size_t> proved(threads, 0);
    std::vector<std::unique_ptr<Credential>> kept; /* the last batch of thread 0: verified after the clock stops */
    auto body = [&](int t) {
        for (size_t k = 0; k < batches && rc[t] == 0; k++) {
            std::vector<std::unique_ptr<Credential>> creds;
            std::vector<zklaim_ctx *> ptrs;
            for (size_t i = 0; i < batch; i++) {
                creds.emplace_back(new Credential(b->payloads, seed0 + ((uint64_t)t << 32) + k * batch + i));
                zklaim_ctx &c = creds.back()->ctx;
                /* the callers share the issuer's key buffer (same address: hashed once) */
                c.pk = b->issuer->ctx.pk;
                c.pk_size = b->issuer->ctx.pk_size;
                ptrs.push_back(&c);
            }
            int r = zkb_libsnark_prove_batch(ptrs.data(), batch);
            for (size_t i = 0; i < batch && r == 0; i++)
                if (!ptrs[i]->proof) r = -5;
            for (auto &c : creds) c->ctx.pk = NULL; /* not owned */
            rc[t] = r;
            if (r == 0) proved[t] += batch;
            if (r == 0 && t == 0 && k + 1 == batches) kept = std::move(creds);
        }
    };
    const double t0 = now_s();
    std::vector<std::thread> th;
    for (int t = 1; t < threads; t++) th.emplace_back(body, t);
    body(0);
    for (auto &x : th) x.join();
    seconds[0] = now_s() - t0;
    size_t total = 0;
    for (int t = 0; t < threads; t++) {
        if (rc[t]) return rc[t] < 0 ? rc[t] : -rc[t];
        total += proved[t];
    }
    for (auto &c : kept) { /* the reference's verifier on what the stream produced */
        c->ctx.vk = b->issuer->ctx.vk;
        c->ctx.vk_size = b->issuer->ctx.vk_size;
        const int v = libsnark_verify(&c->ctx);
        c->ctx.vk = NULL;
        if (v != 0) return -6;
    }
    return (int)total;
}

/* `count` credentials of which credential `tamper` has a flipped hash byte, through zkb_libsnark_prove_batch_ex: the
 * tampered one must be reported unsatisfied (1, no proof), every other one must verify under the reference's
 * libsnark_verify.  Returns the number of credentials that behaved so (must be `count`); rc_out = the call's result. */
int zkbr_batch_with_tampered(void *h, size_t count, uint64_t seed0, size_t tamper, int *rc_out)
{
    Bench *b = (Bench *)h;
    std::vector<std::unique_ptr<Credential>> creds;
    std::vector<zklaim_ctx *> ptrs;
    for (size_t i = 0; i < count; i++) {
        creds.emplace_back(new Credential(b->payloads, seed0 + i));
        zklaim_ctx &c = creds.back()->ctx;
        if (i == tamper) creds.back()->chain[0].pl.hash[5] ^= 0x10;
        c.pk_size = b->issuer->ctx.pk_size;
        c.pk = (unsigned char *)malloc(c.pk_size);
        memcpy(c.pk, b->issuer->ctx.pk, c.pk_size);
        c.vk_size = b->issuer->ctx.vk_size;
        c.vk = (unsigned char *)malloc(c.vk_size);
        memcpy(c.vk, b->issuer->ctx.vk, c.vk_size);
        ptrs.push_back(&c);
    }
    std::vector<int> res(count, -9);
    *rc_out = zkb_libsnark_prove_batch_ex(ptrs.data(), count, res.data());
    int good = 0;
    for (size_t i = 0; i < count; i++) {
        if (i == tamper)
            good += res[i] == 1 && ptrs[i]->proof == NULL ? 1 : 0;
        else
            good += res[i] == 0 && ptrs[i]->proof && libsnark_verify(ptrs[i]) == 0 ? 1 : 0;
    }
    return good;
}

/* the issuer's serialised proving key through BOTH importers -- the reference's
 * libsnark_import_pk (one Tonelli-Shanks root per point on the host) + pack_proving_key, and import_pk_deferred (roots
 * on the GPU).  Returns 0 when every array of the two flat keys is identical (points at infinity included), else the
 * index of the first array that differs (1 ..); seconds[0] = reference import, seconds[1] = deferred import. */
int zkbr_import_check(void *h, double *seconds)
{
    Bench *b = (Bench *)h;
    StdoutMute mute;
    ppT::init_public_params();
    try {
        KeyArrays ref, got;
        double t = now_s();
        {
            r1cs_gg_ppzksnark_proving_key<ppT> pk;
            libsnark_import_pk(&pk, b->issuer->ctx.pk, b->issuer->ctx.pk_size);
            pack_proving_key(pk, ref);
        }
        seconds[0] = now_s() - t;
        size_t nv = 0, ni = 0;
        t = now_s();
        {
            std::lock_guard<std::mutex> gpu(Session::get().gpu());
            import_pk_deferred(Session::get().ctx(), b->issuer->ctx.pk, b->issuer->ctx.pk_size, got, &nv, &ni);
        }
        seconds[1] = now_s() - t;
        if (nv != b->key.num_variables || ni != b->key.num_inputs) return 100;
        if (ref.singles != got.singles) return 1;
        if (ref.A != got.A) return 2;
        if (ref.B_idx != got.B_idx) return 3;
        if (ref.B1 != got.B1) return 4;
        if (ref.B2 != got.B2) return 5;
        if (ref.H != got.H) return 6;
        if (ref.L != got.L) return 7;
        for (int w = 0; w < 3; w++)
            if (ref.row_ptr[w] != got.row_ptr[w] || ref.col[w] != got.col[w] || ref.coeff[w] != got.coeff[w]) return 8 + w;
        if (ref.desc.B_nnz != got.desc.B_nnz || ref.desc.H_size != got.desc.H_size || ref.desc.num_constraints != got.desc.num_constraints)
            return 11;
        return 0;
    } catch (const std::exception &e) {
        fprintf(stderr, "zkbr_import_check: %s\n", e.what());
        return -1;
    }
}

/* the first zkb_libsnark_prove of a process under a key it has never seen (caches dropped first): key import + upload +
 * tuning + proof.  Returns zkb_libsnark_prove's result followed by libsnark_verify's (0 = proved and accepted);
 * seconds[0] = the cold call, seconds[1] = a second (warm) call. */
int zkbr_cold_prove(void *h, uint64_t seed, double *seconds)
{
    Bench *b = (Bench *)h;
    Credential c(b->payloads, seed);
    c.ctx.pk_size = b->issuer->ctx.pk_size;
    c.ctx.pk = (unsigned char *)malloc(c.ctx.pk_size);
    memcpy(c.ctx.pk, b->issuer->ctx.pk, c.ctx.pk_size);
    c.ctx.vk_size = b->issuer->ctx.vk_size;
    c.ctx.vk = (unsigned char *)malloc(c.ctx.vk_size);
    memcpy(c.ctx.vk, b->issuer->ctx.vk, c.ctx.vk_size);
    {
        std::lock_guard<std::mutex> g(g_cache_mu);
        g_keys_by_id.clear();
    }
    Session::get().clear(); /* device keys, contexts and the buffer identities: as in a fresh process */
    double t = now_s();
    int rc = zkb_libsnark_prove(&c.ctx);
    seconds[0] = now_s() - t;
    if (rc) return rc;
    rc = libsnark_verify(&c.ctx);
    if (rc) return 10 + rc;
    free(c.ctx.proof);
    c.ctx.proof = NULL;
    t = now_s();
    rc = zkb_libsnark_prove(&c.ctx);
    seconds[1] = now_s() - t;
    /* the Bench object's cached key set died with the session: re-create it for later calls on this handle */
    b->key = key_for_ctx(&b->issuer->ctx);
    return rc;
}

/* A `payloads`-payload credential end to end on a GPU-made key: zkb_libsnark_trusted_setup, then a proof by the
 * reference's own CPU prover (stock = 1: libsnark_prove, unmodified) or by zkb_libsnark_prove (stock = 0), then the
 * reference's libsnark_verify.  Returns libsnark_verify's result (0 = accepted), < 0 on an earlier failure;
 * seconds[0] = setup, [1] = prove, [2] = verify. */
int zkbr_setup_roundtrip(size_t payloads, uint64_t seed, int stock, double *seconds)
{
    Credential c(payloads, seed);
    double t = now_s();
    if (zkb_libsnark_trusted_setup(&c.ctx) != 0) return -1;
    seconds[0] = now_s() - t;
    t = now_s();
    if ((stock ? libsnark_prove(&c.ctx) : zkb_libsnark_prove(&c.ctx)) != 0) return -2;
    seconds[1] = now_s() - t;
    t = now_s();
    const int rc = libsnark_verify(&c.ctx);
    seconds[2] = now_s() - t;
    return rc;
}

/* ---- BASELINE.json config #5 for bench.py: "synthetic 2^22-constraint R1CS (chained SHA-256 gadgets)" ---------------
 * `blocks` chained libsnark sha256_two_to_one_hash_gadgets (gadgetlib1/gadgets/hashes/sha256/sha256_gadget.hpp; the
 * same gadget zklaim's commitment uses, zklaim_gadget.cpp) with a satisfying assignment from `seed`; the final digest's
 * 256 bits are the public input.  153 blocks = 4 173 840 constraints = domain 2^22.  The constraint system is handed
 * out as the CSR arrays of zkb_setup_desc / zkb_key_desc; the key comes from zkb_setup. */
namespace {
struct Synth {
    r1cs_constraint_system<FrT> cs;
    std::vector<FrT> primary, auxiliary;
};
} // namespace

void *zkbr_synth_sha_chain(size_t blocks, uint64_t seed)
{
    StdoutMute mute;
    ppT::init_public_params();
    try {
        Synth *sy = new Synth();
        protoboard<FrT> pb;
        std::vector<std::shared_ptr<digest_variable<FrT>>> dig(blocks + 1), right(blocks);
        dig[blocks].reset(new digest_variable<FrT>(pb, 256, "out")); /* allocated first: the primary input */
        pb.set_input_sizes(256);
        for (size_t i = 0; i < blocks; i++) dig[i].reset(new digest_variable<FrT>(pb, 256, "d"));
        for (size_t i = 0; i < blocks; i++) right[i].reset(new digest_variable<FrT>(pb, 256, "r"));
        std::vector<std::shared_ptr<sha256_two_to_one_hash_gadget<FrT>>> h(blocks);
        for (size_t i = 0; i < blocks; i++) {
            h[i].reset(new sha256_two_to_one_hash_gadget<FrT>(pb, *dig[i], *right[i], *dig[i + 1], "h"));
            h[i]->generate_r1cs_constraints();
        }
        uint64_t x = seed * 0x9e3779b97f4a7c15ull + 0x1234567;
        auto bits = [&]() {
            libff::bit_vector v(256);
            for (size_t j = 0; j < 256; j++) {
                x ^= x << 13;
                x ^= x >> 7;
                x ^= x << 17;
                v[j] = x & 1;
            }
            return v;
        };
        dig[0]->generate_r1cs_witness(bits());
        for (size_t i = 0; i < blocks; i++) {
            right[i]->generate_r1cs_witness(bits());
            h[i]->generate_r1cs_witness();
        }
        sy->cs = pb.get_constraint_system();
        sy->primary = pb.primary_input();
        sy->auxiliary = pb.auxiliary_input();
        return sy;
    } catch (const std::exception &e) {
        fprintf(stderr, "zkbr_synth_sha_chain: %s\n", e.what());
        return NULL;
    }
}

void zkbr_synth_free(void *h) { delete (Synth *)h; }

/* dims[0..6) = num_constraints, num_variables, num_inputs, nnz(A), nnz(B), nnz(C) */
void zkbr_synth_dims(void *h, size_t *dims)
{
    Synth *sy = (Synth *)h;
    dims[0] = sy->cs.num_constraints();
    dims[1] = sy->cs.num_variables();
    dims[2] = sy->cs.num_inputs();
    size_t nnz[3] = {0, 0, 0};
    for (const auto &c : sy->cs.constraints) {
        nnz[0] += c.a.terms.size();
        nnz[1] += c.b.terms.size();
        nnz[2] += c.c.terms.size();
    }
    dims[3] = nnz[0];
    dims[4] = nnz[1];
    dims[5] = nnz[2];
}

/* matrix `which` (0 = A, 1 = B, 2 = C) as CSR: row_ptr (num_constraints + 1), col (nnz), coeff (nnz x 4 limbs) */
void zkbr_synth_csr(void *h, int which, uint64_t *row_ptr, uint32_t *col, uint64_t *coeff)
{
    Synth *sy = (Synth *)h;
    size_t k = 0, row = 0;
    row_ptr[0] = 0;
    for (const auto &c : sy->cs.constraints) {
        const linear_combination<FrT> &lc = which == 0 ? c.a : (which == 1 ? c.b : c.c);
        for (const auto &t : lc.terms) {
            col[k] = (uint32_t)t.index;
            put_fr(coeff + 4 * k, t.coeff);
            k++;
        }
        row_ptr[++row] = k;
    }
}

void zkbr_synth_witness(void *h, uint64_t *w)
{
    Synth *sy = (Synth *)h;
    pack_assignment(sy->primary, sy->auxiliary, w);
}

/* the curve's generators as Jacobian limbs (12 + 24): what zkb_setup takes as g1_generator / g2_generator */
void zkbr_generators(uint64_t *g1_jac, uint64_t *g2_jac)
{
    ppT::init_public_params();
    put_g1_jacobian(g1_jac, G1T::one());
    put_g2_jacobian(g2_jac, G2T::one());
}

} // extern "C"
