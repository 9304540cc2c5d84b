/*
 * zkb.h -- C ABI of the B200-native Groth16 prover backend for zklaim (libzkb.so).
 *
 * Drop-in boundary: everything above this header is the unchanged
 * reference host code -- zklaim's C API (zklaim/zklaim.h:117-262), the C++ bridge
 * (zklaim/libsnark_wrapper.cpp:195-276), gadget witness generation
 * (zklaim/zklaim_gadget.cpp:705-783) and libsnark's key/proof (de)serialisation.  The single
 * call that moves to the GPU is
 *
 *     zklaim/snark.cpp:126
 *         return r1cs_gg_ppzksnark_prover<ppT>(proving_key, pb.primary_input(), pb.auxiliary_input());
 *     (signature: libsnark/zk_proof_systems/ppzksnark/r1cs_gg_ppzksnark/r1cs_gg_ppzksnark.hpp:373-376)
 *
 * which becomes zkb_key_upload() (once per proving key) + zkb_prove() (per proof); see
 * INTEGRATION.md for the bridge-side stub.  The other entry points expose the sub-steps
 * of that call for the reference's own micro-benchmarks (multi_exp, evaluation-domain
 * FFTs, r1cs_to_qap_witness_map) and for parity tests of the arithmetic layers.
 *
 * Conventions
 *   - plain C, no C++ types, no exceptions cross this boundary; nothing aborts.
 *   - return value: 0 = ok, < 0 = error (zkb_last_error() gives the text).
 *   - all host buffers are caller-owned; the library copies what it keeps.
 *   - one zkb_ctx per host thread (it owns one CUDA stream and its scratch memory);
 *     zkb_key / zkb_bases are immutable after creation and may be shared by contexts on
 *     the same device.
 *   - field elements: 4 x uint64_t little-endian limbs, Montgomery form (R = 2^256),
 *     canonical (< modulus) -- bit-identical to libff's Fp_model::mont_repr
 *     (libff/algebra/fields/fp.hpp:42).  Fq2 = c0 then c1.
 *   - G1 affine: x, y (8 limbs); G2 affine: x.c0 x.c1 y.c0 y.c1 (16 limbs); the point at
 *     infinity is all-zero limbs.  This is the in-memory form of a key after libsnark's
 *     operator>> (alt_bn128_g1.cpp:453-462: Z = 1 or zero).
 *   - G1 Jacobian: X, Y, Z (12 limbs) as alt_bn128_g1.hpp:35; G2 Jacobian 24 limbs.
 *     Infinity <=> Z == 0.  Outputs are some representative of the group element (the
 *     reference normalises to affine when it serialises, alt_bn128_g1.cpp:404-418).
 */
#ifndef ZKB_H
#define ZKB_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct zkb_ctx zkb_ctx;
typedef struct zkb_bases zkb_bases;
typedef struct zkb_key zkb_key;

enum {
    ZKB_OK = 0,
    ZKB_ERR_CUDA = -1,     /* a CUDA call failed (no device, out of memory, launch error) */
    ZKB_ERR_ARG = -2,      /* bad argument */
    ZKB_ERR_DOMAIN = -3,   /* no supported evaluation domain for this size */
    ZKB_ERR_UNSAT = -4     /* witness does not satisfy the constraint system */
};

/* field / group selectors for the element-wise entry points */
enum { ZKB_FR = 0, ZKB_FQ = 1, ZKB_FQ2 = 2 };
enum { ZKB_OP_ADD = 0, ZKB_OP_SUB = 1, ZKB_OP_MUL = 2, ZKB_OP_SQR = 3, ZKB_OP_INV = 4, ZKB_OP_NEG = 5 };
enum { ZKB_G1 = 1, ZKB_G2 = 2 };
/* group ops on Jacobian operands: a + b, 2a, a + b with b affine-normalised (Z = 1), -a,
 * to_affine_coordinates (alt_bn128_g1.cpp:60-77: Z = 1, or the zero (0, 1, 0)) */
enum { ZKB_GOP_ADD = 0, ZKB_GOP_DBL = 1, ZKB_GOP_MIXED_ADD = 2, ZKB_GOP_NEG = 3, ZKB_GOP_NORMALIZE = 4 };
/* evaluation-domain kinds, numbered as libfqfft's factory tries them
 * (libfqfft/evaluation_domain/get_evaluation_domain.tcc:33-52) */
enum { ZKB_DOMAIN_BASIC_RADIX2 = 1, ZKB_DOMAIN_EXTENDED_RADIX2 = 2, ZKB_DOMAIN_STEP_RADIX2 = 3 };
/* domain operations = evaluation_domain<Fr> members (libfqfft/evaluation_domain/evaluation_domain.hpp:36-99),
 * coset shift fixed to Fr::multiplicative_generator as in r1cs_to_qap.tcc:262-311 */
enum { ZKB_NTT_FFT = 0, ZKB_NTT_IFFT = 1, ZKB_NTT_COSET_FFT = 2, ZKB_NTT_ICOSET_FFT = 3, ZKB_NTT_DIV_Z_ON_COSET = 4 };

/* ---- lifetime ---------------------------------------------------------------------- */
/* Creates a context on CUDA device `device_id`.  Fails (ZKB_ERR_CUDA) when no GPU is
 * present: there is no CPU fallback anywhere in this library.
 * A context = its own CUDA streams and scratch memory, used by one host thread at a time.  SEVERAL contexts may live
 * on one device: a key or base set (zkb_key_upload, zkb_bases_upload) is immutable once uploaded and tuned and may be
 * used by every context of its device at once.  That is how a proving service keeps a GPU full: a few caller threads,
 * one context each, one batch in flight each -- the latency-bound tail of one batch (fix-ups, bucket reduction,
 * assembly) runs under the throughput-bound kernels of another (bench.py: 12.3 -> 11.4 ms per batch of 64 with four). */
int zkb_init(int device_id, zkb_ctx **out);
void zkb_shutdown(zkb_ctx *ctx);
const char *zkb_last_error(const zkb_ctx *ctx); /* ctx may be NULL: last error of zkb_init */
const char *zkb_version(void);

/* ---- K1-K3: element-wise arithmetic (parity tests of the field / group layers) ------- */
/* libff Fp_model / Fp2_model operators (fp.tcc, fp2.tcc); b may be NULL for unary ops;
 * n elements of 4 limbs (8 for ZKB_FQ2). */
int zkb_field_op(zkb_ctx *ctx, int field, int op, const uint64_t *a, const uint64_t *b, uint64_t *out, size_t n);
/* alt_bn128_G1 / alt_bn128_G2 operator+, dbl, mixed_add, unary minus
 * (alt_bn128_g1.cpp:139-358, alt_bn128_g2.cpp:149-365) on Jacobian limbs. */
int zkb_group_op(zkb_ctx *ctx, int group, int op, const uint64_t *a, const uint64_t *b, uint64_t *out, size_t n);
/* scalar * point, libff operator*(Fr, G) (curve_utils.tcc:13-34): scalars Montgomery Fr, points Jacobian. */
int zkb_group_scalar_mul(zkb_ctx *ctx, int group, const uint64_t *scalars, const uint64_t *points, uint64_t *out,
                         size_t n);

/* Sum of n Jacobian points -- the final loop of libff::multi_exp over its per-chunk partial results
 * (multiexp.tcc:433-438).  This is the combine step of an MSM whose point range is sharded over several GPUs:
 * every rank computes zkb_msm on its range, the 96- / 192-byte partials are gathered, one rank sums them. */
int zkb_group_sum(zkb_ctx *ctx, int group, const uint64_t *points, size_t n, uint64_t *out);

/* ---- key import: point decompression ---------------------------------------------
 * libsnark stores a key element as (is_zero, X, least significant bit of Y) and recomputes Y = sqrt(X^3 + b) on import
 * with one Tonelli-Shanks root per point (operator>>, libff/algebra/curves/alt_bn128/alt_bn128_g1.cpp:420-465,
 * alt_bn128_g2.cpp:433-478; roots: libff/algebra/fields/fp.tcc:724-776, fp2.tcc:146-198) -- 7 s for the zklaim key,
 * reached from libsnark_import_pk (zklaim/libsnark_wrapper.cpp:160-168, called by every libsnark_prove, :228-230).
 * Here: n abscissas in, n affine points out.  x: n x 4 limbs (G1) / n x 8 limbs (G2: c0, c1), Montgomery; flags[i]:
 * bit 0 = the stored lsb of Y (G2: of Y.c0) as a plain integer, bit 1 = is_zero (output: all-zero limbs); out_affine:
 * n x 8 / n x 16 limbs.  status (optional, n entries): 1 = ok, 0 = X^3 + b is not a square.  Returns ZKB_ERR_ARG if
 * any point failed.  (When Y.c0 == 0 in G2 both roots have the stored parity and the reference keeps whichever its
 * root algorithm returns; that case has probability 2^-254 for a key element and is not distinguished here.) */
int zkb_points_decompress(zkb_ctx *ctx, int group, const uint64_t *x, const uint8_t *flags, size_t n, uint64_t *out_affine,
                          int32_t *status);

/* ---- K4/K5: multi-scalar multiplication ------------------------------------------- */
/* Uploads n affine bases and (precompute_factor > 1) their 2^(j*G*c) multiples so that up
 * to `precompute_factor` windows share one bucket set (HBM for doublings; B200 has 180 GB).
 * window_bits = 0 lets the library choose c from n.  precompute_factor = 0 means "all
 * windows" (no doublings at all at MSM time), 1 means none. */
int zkb_bases_upload(zkb_ctx *ctx, int group, const uint64_t *affine, size_t n, int window_bits,
                     int precompute_factor, zkb_bases **out);
void zkb_bases_free(zkb_bases *b);
size_t zkb_bases_device_bytes(const zkb_bases *b);

/* sum_i scalars[i] * bases[offset + i], i < n -- libff::multi_exp<T, Fr, multi_exp_method_BDLO12>
 * and multi_exp_with_mixed_addition (multiexp.tcc:402-496).  scalars: host, Montgomery Fr.
 * out: Jacobian limbs (12 for G1, 24 for G2). */
int zkb_msm(zkb_ctx *ctx, const zkb_bases *bases, size_t offset, const uint64_t *scalars, size_t n, uint64_t *out);
/* Same with the scalars already in device memory (d_scalars is a CUDA device pointer);
 * the result stays on the host (it is 96 / 192 bytes). */
int zkb_msm_device(zkb_ctx *ctx, const zkb_bases *bases, size_t offset, const void *d_scalars, size_t n,
                   uint64_t *out);

/* ---- K7-K9: evaluation domains ---------------------------------------------------- */
/* libfqfft::get_evaluation_domain<Fr>(min_size) (get_evaluation_domain.tcc:33-52): the
 * domain size m and kind the reference would pick.  Host-only, needs no ctx. */
int zkb_domain_select(size_t min_size, size_t *m_out, int *kind_out);
/* In-place domain operation on m = domain size elements of host memory. */
int zkb_domain_op(zkb_ctx *ctx, size_t min_size, int op, uint64_t *data);

/* ---- an entry point that allocates nothing ---------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Natural-order radix-2 FFT (inverse = 0) or inverse FFT incl. the 1/n factor (inverse != 0) of n = 2^log2n Fr elements
 * -- the values zkb_domain_op(ZKB_NTT_FFT / ZKB_NTT_IFFT) gives on the basic radix-2 domain of size n, bit for bit.
 * d_in, d_out: DEVICE pointers on the context's device, n x 4 limbs each (Montgomery Fr); they may be equal.
 * All scratch (the working copy of the vector and the twiddle table) lives in one static device array inside
 * libzkb.so: 768 KiB per GPU that loads the library, held until exit; hence log2n <= zkb_fft_noalloc_max_log2() = 14,
 * larger sizes return ZKB_ERR_ARG (use zkb_domain_op).  The call performs no cudaMalloc / cudaFree and uses no
 * allocating library routine; the first call on a device also looks up the array's address (cudaGetSymbolAddress).
 * The array is shared by every call on that device and keeps the last call's contents (each call rewrites what it
 * reads; nothing needs zeroing): calls that overlap on the GPU -- two contexts, two host threads -- race on it.  The
 * call returns after its work on the context's stream has completed. */
int zkb_fft_noalloc_max_log2(void);
int zkb_fft_noalloc(zkb_ctx *ctx, int log2n, int inverse, const void *d_in, void *d_out);

/* ---- the proving key ---------------------------------------------------------------- */
/* CSR of one R1CS matrix: row i = constraint i; col 0 = the constant 1, col j > 0 = variable
 * j-1 of the full assignment (libsnark/relations/variable.tcc:262-270). */
typedef struct {
    const uint64_t *row_ptr; /* num_constraints + 1 */
    const uint32_t *col;     /* nnz */
    const uint64_t *coeff;   /* nnz x 4 limbs, Montgomery Fr */
} zkb_csr;

/* r1cs_gg_ppzksnark_proving_key (r1cs_gg_ppzksnark.hpp:69-150) as flat arrays. */
typedef struct {
    size_t num_constraints, num_variables, num_inputs;
    zkb_csr A, B, C;          /* pk.constraint_system (already A/B-swapped by the generator) */
    const uint64_t *alpha_g1; /* 8 limbs affine */
    const uint64_t *beta_g1;
    const uint64_t *delta_g1;
    const uint64_t *beta_g2;  /* 16 limbs affine */
    const uint64_t *delta_g2;
    const uint64_t *A_query;  /* (num_variables + 1) x 8 */
    size_t B_nnz;
    const uint32_t *B_index;  /* sparse_vector indices, ascending (sparse_vector.tcc:141-147) */
    const uint64_t *B_g1;     /* B_nnz x 8   (knowledge_commitment::h) */
    const uint64_t *B_g2;     /* B_nnz x 16  (knowledge_commitment::g) */
    size_t H_size;            /* domain m - 1 */
    const uint64_t *H_query;  /* H_size x 8 */
    const uint64_t *L_query;  /* (num_variables - num_inputs) x 8 */
} zkb_key_desc;

/* All query / key-element pointers NULL = "constraint system only": the key then serves zkb_witness_map (config #4:
 * the witness map needs the R1CS and its domain, not the queries); zkb_prove* on it returns ZKB_ERR_ARG. */
int zkb_key_upload(zkb_ctx *ctx, const zkb_key_desc *desc, zkb_key **out);
void zkb_key_free(zkb_key *key);
size_t zkb_key_device_bytes(const zkb_key *key);
/* Re-plans the MSM window widths of the key for the scalar statistics of a sample witness w (num_variables x 4
 * limbs): Groth16 witnesses are dominated by 0 / 1 values, which the reference special-cases in
 * multi_exp_with_mixed_addition (multiexp.tcc:466-481); here they cost one bucket entry whatever the window, so
 * the windows are sized for the remaining dense scalars.  Proofs are unchanged (any window gives the same group
 * element); only speed and device memory change.  Must not run concurrently with proofs that use this key. */
int zkb_key_tune(zkb_ctx *ctx, zkb_key *key, const uint64_t *w);
/* window bits currently planned for the A, B-in-G1, H + L and B-in-G2 queries */
int zkb_key_windows(const zkb_key *key, int out[4]);
/* num_variables of the key's constraint system = field elements per witness */
int zkb_key_num_variables(const zkb_key *key, size_t *nv_out);
/* domain chosen for this key (what r1cs_to_qap_witness_map would use, r1cs_to_qap.tcc:218) */
int zkb_key_domain(const zkb_key *key, size_t *m_out, int *kind_out);

/* ---- K10 + K7/K8: r1cs_to_qap_witness_map (r1cs_to_qap.tcc:205-334, d1 = d2 = d3 = 0) ---- */
/* w: num_variables x 4 limbs (primary then auxiliary).  H_out: (m + 1) x 4 limbs. */
int zkb_witness_map(zkb_ctx *ctx, const zkb_key *key, const uint64_t *w, uint64_t *H_out);

/* ---- the prover --------------------------------------------------------------------- */
typedef struct {
    uint64_t A[12]; /* g_A, G1 Jacobian */
    uint64_t B[24]; /* g_B, G2 Jacobian */
    uint64_t C[12]; /* g_C, G1 Jacobian */
} zkb_proof;

/* r1cs_gg_ppzksnark_prover (r1cs_gg_ppzksnark.tcc:390-505) with the blinding scalars r, s
 * (:418-419) supplied by the caller (Montgomery Fr).  H_out (optional, may be NULL)
 * receives coefficients_for_H, (m + 1) x 4 limbs.  Returns ZKB_ERR_UNSAT when the witness
 * does not satisfy the system: the device checks <a,z> * <b,z> == <c,z> for every constraint
 * (r1cs_constraint_system::is_satisfied, which the reference asserts, :406-408, and which
 * zklaim's generate_proof tests on the host, snark.cpp:119-124); the proof of such a witness
 * is returned as all-zero limbs.
 * Host buffers may be pageable or pinned.  A pageable assignment of 16 MB or more (the 124 MB of a 2^22-constraint
 * circuit, a whole batch) is copied into a pinned area of the context by several host threads, 2 MB pieces each followed by
 * its own asynchronous transfer (ZKB_STAGED_UPLOAD=0: plain cudaMemcpy). */
int zkb_prove(zkb_ctx *ctx, const zkb_key *key, const uint64_t *w, const uint64_t r[4], const uint64_t s[4],
              zkb_proof *out, uint64_t *H_out);

/* ---- one proof over several GPUs ------------------------------------------------------
 * The five MSMs of a proof shard the way libff::multi_exp already chunks them on the CPU (multiexp.tcc:417-438):
 * shard `rank` of `world` keeps the contiguous range [rank * n / world, ...) of the A, B, H and L queries (rank 0
 * also the fixed elements alpha, beta, delta), every shard computes the witness map itself (replicas: no NTT is
 * distributed) and its share of the four sums; the 480-byte shares are gathered (host / NCCL all_gather) and
 * zkb_prove_combine adds them up and performs the assembly of r1cs_gg_ppzksnark.tcc:487-495. */
typedef struct {
    uint64_t A[12];  /* share of g1_A            (G1 Jacobian) */
    uint64_t B1[12]; /* share of g1_B                          */
    uint64_t C[12];  /* share of H + L - rs delta              */
    uint64_t B2[24]; /* share of g2_B            (G2 Jacobian) */
} zkb_partial;
/* `desc` describes the WHOLE key (as for zkb_key_upload); only the shard's ranges go to the device. */
int zkb_key_upload_shard(zkb_ctx *ctx, const zkb_key_desc *desc, int rank, int world, zkb_key **out);
int zkb_prove_partial(zkb_ctx *ctx, const zkb_key *key, const uint64_t *w, const uint64_t r[4], const uint64_t s[4],
                      zkb_partial *out, uint64_t *H_out);
int zkb_prove_combine(zkb_ctx *ctx, const zkb_partial *parts, size_t n, const uint64_t r[4], const uint64_t s[4],
                      zkb_proof *out);

/* The same with the witness map SPLIT by polynomial instead of replicated (r1cs_to_qap.tcc:240-293 treats A, B and C
 * alike: SpMV, iFFT, coset FFT -- three independent chains).  Three shards run one chain each, one shard divides and
 * transforms back, every shard receives the H coefficients; the exchanges (two m x 32-byte vectors in, one out) are
 * the caller's -- NCCL send / recv / broadcast in zklaim_b200/sharding.py, peer copies in a single process:
 *   _begin   uploads w, r, s (w: host memory, or device memory of this context's GPU -- e.g. a witness gathered over
 *            NVLink from one range per rank), starts the shard's witness-only sums (A, B-in-G1, B-in-G2 ranges) on side streams and, for
 *            which = 0 / 1 / 2, writes the coset evaluations of polynomial A / B / C to d_part (device, m x 4 limbs);
 *            which = -1: no chain.  check != 0: this shard also evaluates all three products and tests
 *            <a,z><b,z> = <c,z> per constraint (its _end then reports ZKB_ERR_UNSAT); one shard should do so.
 *   _combine d_abc = [A' | B' | C'] on the coset (device, 3 m x 4 limbs, destroyed) -> H coefficients as plain integers
 *            at d_H_plain (device, m x 4 limbs): (A' B' - C') / Z, inverse coset transform (:270-311).
 *   _end     H (device, plain) -> the shard's share of the H + L sum; joins the side lanes; writes the partial.
 * One proof in flight per context.  zkb_last_device_ms() after _end = the device time of the three calls. */
int zkb_prove_split_begin(zkb_ctx *ctx, const zkb_key *key, const uint64_t *w, const uint64_t r[4], const uint64_t s[4], int which,
                          int check, void *d_part);
/* one more chain on a shard that has run _begin (two GPUs: shard 0 takes A and C, shard 1 takes B): the coset evaluations
 * of polynomial `which` to d_part. */
int zkb_prove_split_chain(zkb_ctx *ctx, const zkb_key *key, int which, void *d_part);
int zkb_prove_split_combine(zkb_ctx *ctx, const zkb_key *key, void *d_abc, void *d_H_plain);
int zkb_prove_split_end(zkb_ctx *ctx, const zkb_key *key, const void *d_H_plain, zkb_partial *out);

/* `count` independent proofs under one key; w, r, s, out are arrays of `count` items
 * (w: count x num_variables x 4).  The batch is pipelined on the context's device. */
int zkb_prove_batch(zkb_ctx *ctx, const zkb_key *key, size_t count, const uint64_t *w, const uint64_t *r,
                    const uint64_t *s, zkb_proof *out);

/* Per-proof outcome of the most recent zkb_prove / zkb_prove_batch* / zkb_prove_batch_collect call on this context:
 * status[i] = ZKB_OK or ZKB_ERR_UNSAT (that call returned ZKB_ERR_UNSAT if any proof's witness was unsatisfying; the
 * other proofs of the batch are valid, the offending ones are all-zero limbs). */
int zkb_prove_batch_status(const zkb_ctx *ctx, int32_t *status, size_t n);

/* Same with w (count x num_variables x 4 limbs), r and s (count x 4 limbs) already in device memory
 * on the context's device; the proofs still come back to the host (384 bytes each). */
int zkb_prove_batch_device(zkb_ctx *ctx, const zkb_key *key, size_t count, const void *d_w, const void *d_r,
                           const void *d_s, zkb_proof *out);

/* Pipelined form of zkb_prove_batch for a stream of batches (a proving service): _submit starts the host-to-device
 * copy of a batch's inputs on a copy stream and returns; _collect waits for the OLDEST submitted batch's inputs,
 * proves it and returns its proofs.  Two staging slots alternate, so at most two batches may be submitted and not
 * yet collected; with submit(k + 1) issued before collect(k) the input transfer of batch k + 1 overlaps the
 * computation of batch k.  w, r, s must stay valid (and should be pinned) until the matching _collect returns.
 * Results are those of zkb_prove_batch; ZKB_ERR_UNSAT is reported by the _collect of the offending batch. */
int zkb_prove_batch_submit(zkb_ctx *ctx, const zkb_key *key, size_t count, const uint64_t *w, const uint64_t *r,
                           const uint64_t *s);
int zkb_prove_batch_collect(zkb_ctx *ctx, zkb_proof *out);

/* ---- several GPUs from one host process -------------------------------------------------
 * A pool owns one zkb_ctx per device and drives them with one host thread per device; a C caller shaped like
 * zklaim/main_benchmark.c:174-198 (credential after credential through the prover) uses every GPU of the box
 * through these calls without Python or NCCL.  zkb_pool_init(NULL, 0, ..) takes all visible devices.
 *   zkb_pool_prove_batch     independent proofs: the batch is cut into one contiguous slice per context (key replicated
 *                            per DEVICE by zkb_pool_key_upload: a device id listed twice gives two contexts that share
 *                            one replica); results in submission order;
 *   zkb_pool_msm             one MSM, bases range-sharded the way libff::multi_exp chunks them (multiexp.tcc:417-431):
 *                            device i holds and sums chunk i, the 96- / 192-byte partials are added on device 0
 *                            (multiexp.tcc:433-438);
 *   zkb_pool_prove_sharded   ONE proof over all devices (key set from zkb_pool_key_upload_sharded): zkb_prove_partial
 *                            per device, zkb_prove_combine on device 0.
 * Errors: the code of the first failing device; zkb_pool_last_error names it. */
typedef struct zkb_pool zkb_pool;
typedef struct zkb_key_set zkb_key_set;
typedef struct zkb_bases_set zkb_bases_set;
int zkb_device_count(void);
int zkb_pool_init(const int *device_ids, int n, zkb_pool **out);
void zkb_pool_shutdown(zkb_pool *pool);
int zkb_pool_size(const zkb_pool *pool);
zkb_ctx *zkb_pool_ctx(zkb_pool *pool, int i); /* the context of device i, e.g. for zkb_last_device_ms */
const char *zkb_pool_last_error(const zkb_pool *pool);
int zkb_pool_key_upload(zkb_pool *pool, const zkb_key_desc *desc, zkb_key_set **out);
int zkb_pool_key_upload_sharded(zkb_pool *pool, const zkb_key_desc *desc, zkb_key_set **out);
int zkb_pool_key_tune(zkb_pool *pool, zkb_key_set *keys, const uint64_t *w);
const zkb_key *zkb_pool_key(const zkb_key_set *keys, int i);
void zkb_pool_key_free(zkb_key_set *keys);
/* status (optional, count entries): per-proof outcome as zkb_prove_batch_status reports it */
int zkb_pool_prove_batch(zkb_pool *pool, const zkb_key_set *keys, size_t count, const uint64_t *w, const uint64_t *r,
                         const uint64_t *s, zkb_proof *out, int32_t *status);
int zkb_pool_prove_sharded(zkb_pool *pool, const zkb_key_set *keys, const uint64_t *w, const uint64_t r[4], const uint64_t s[4],
                           zkb_proof *out);
int zkb_pool_bases_upload(zkb_pool *pool, int group, const uint64_t *affine, size_t n, int window_bits, int precompute_factor,
                          zkb_bases_set **out);
void zkb_pool_bases_free(zkb_bases_set *bases);
int zkb_pool_msm(zkb_pool *pool, const zkb_bases_set *bases, const uint64_t *scalars, size_t n, uint64_t *out);

/* ---- the generator ---------------------------------------------------------------
 * r1cs_gg_ppzksnark_generator (r1cs_gg_ppzksnark.tcc:211-388) with its secret randomness supplied by the caller,
 * the way zkb_prove takes (r, s): t, alpha, beta, gamma, delta (:221-225, Montgomery Fr) and the two random
 * group generators (:284, :294; Jacobian limbs).  On the device: the Lagrange basis of the domain at t
 * (basic_radix2_domain_aux.tcc:183-236, step_radix2_domain.tcc:161-186), A_i(t) / B_i(t) / C_i(t)
 * (r1cs_to_qap.tcc:105-170), the L and gamma_ABC scalars (:254-275), the powers of t for H (:277-282), and every
 * fixed-base exponentiation (batch_exp / kc_batch_exp / batch_exp_with_coeff, :303-333, :340-341) with 8-bit
 * window tables.  The constraint system is passed as built; swap_AB_if_beneficial (r1cs.tcc:182-231, called at
 * :217-218) is applied inside and reported in *swapped (the proving key holds the swapped system).
 * Outputs are caller-allocated affine arrays (all-zero limbs = the zero of the group, which the reference's
 * batch_to_special leaves as (0, 1, 0)); B is dense over num_variables + 1 entries, entries whose scalar is zero are
 * the zero of the group (kc_batch_exp skips them, kc_multiexp.tcc:150-160: the sparse B_query = the non-zero ones).
 * vk.alpha_g1_beta_g2 is one pairing of two outputs and stays with the caller (:338).
 * Returns ZKB_ERR_ARG if t is an element of the domain (probability ~ m / 2^254; the reference special-cases it). */
typedef struct {
    size_t num_constraints, num_variables, num_inputs;
    zkb_csr A, B, C;                                  /* as built, NOT swapped */
    const uint64_t *t, *alpha, *beta, *gamma, *delta; /* 4 limbs each */
    const uint64_t *g1_generator;                     /* 12 limbs */
    const uint64_t *g2_generator;                     /* 24 limbs */
} zkb_setup_desc;
typedef struct {
    uint64_t *alpha_g1, *beta_g1, *delta_g1; /* 8 limbs each */
    uint64_t *beta_g2, *delta_g2, *gamma_g2; /* 16 limbs each */
    uint64_t *A_query;                       /* (num_variables + 1) x 8 */
    uint64_t *B_g1;                          /* (num_variables + 1) x 8 */
    uint64_t *B_g2;                          /* (num_variables + 1) x 16 */
    uint64_t *H_query;                       /* (m - 1) x 8, m = the domain size for num_constraints + num_inputs + 1 */
    uint64_t *L_query;                       /* (num_variables - num_inputs) x 8 */
    uint64_t *gamma_ABC_g1;                  /* (num_inputs + 1) x 8 */
} zkb_setup_out;
int zkb_setup(zkb_ctx *ctx, const zkb_setup_desc *desc, zkb_setup_out *out, int *swapped);

/* ---- verifier side: pairings and batch verification ---------------------- */
/* alt_bn128_ate_miller_loop(precompute_G1(P), precompute_G2(Q)) (reduced = 0) or alt_bn128_reduced_pairing(P, Q)
 * (reduced = 1) for n affine pairs (libff/algebra/curves/alt_bn128/alt_bn128_pairing.cpp:371-420, 523-533).
 * out: n x 12 Fq (48 limbs each) in libff's member order c0.c0.c0, c0.c0.c1, c0.c1.c0, ... c1.c2.c1 --
 * bit-identical to the reference's alt_bn128_Fq12 / alt_bn128_GT. */
int zkb_pairing(zkb_ctx *ctx, const uint64_t *g1_affine, const uint64_t *g2_affine, size_t n, int reduced, uint64_t *out);

/* r1cs_gg_ppzksnark_verification_key (r1cs_gg_ppzksnark.hpp:159-216) as flat host arrays */
typedef struct zkb_vk_desc {
    const uint64_t *alpha_g1_beta_g2; /* 48 limbs: the GT element e(alpha, beta), member order as zkb_pairing */
    const uint64_t *gamma_g2;         /* 16 limbs, affine */
    const uint64_t *delta_g2;         /* 16 limbs, affine */
    const uint64_t *gamma_ABC_g1;     /* (num_inputs + 1) x 8 limbs, affine: accumulation_vector first, then rest */
    uint64_t num_inputs;
} zkb_vk_desc;
/* r1cs_gg_ppzksnark_verifier_strong_IC (r1cs_gg_ppzksnark.tcc:523-622, reached from libsnark_verify,
 * zklaim/libsnark_wrapper.cpp:252-276) for n proofs under one key.  primary: n x num_inputs Fr (Montgomery);
 * proofs: n x 48 limbs in the layout zkb_prove writes (A 12, B 24, C 12: Jacobian).  accept[i] = 1 when the
 * reference verifier returns true for proof i, else 0 (a point off its curve, or the pairing equation failing).  A proof
 * coordinate >= q or an input >= r is not a canonical residue: such a proof is rejected (accept[i] = 0).  n x max(num_inputs, 1)
 * must stay below 2^29 per call (the input-sum kernels index their lanes with 32 bits): larger batches are refused. */
int zkb_verify_batch(zkb_ctx *ctx, const zkb_vk_desc *vk, const uint64_t *primary, const uint64_t *proofs, size_t n,
                     int32_t *accept);

/* ---- measurement helpers (used by bench.py only) ------------------------------------ */
/* n distinct G1 bases k_i * G generated ON the device (k_i = 252 bits of a counter-based hash of (seed, i)) and held as
 * zkb_bases_upload would hold them: the MSM sweep of BASELINE config #3 needs 2^24 random points, which are 1 GB to
 * make on the host and push.  Parity of the MSM itself is tested on host-made bases (tests/test_gpu_large.py). */
int zkb_bases_synthetic(zkb_ctx *ctx, size_t n, uint64_t seed, int window_bits, int precompute_factor, zkb_bases **out);
/* Sustained integer multiply-add rate of the device, measured with chains of dependent Fq Montgomery
 * products (136 multiply-adds of 32 x 32 + 64 -> 64 bits each, the instruction mix of every kernel here) on
 * all SMs; `iters` = products per chain / 2 (0 = default).  Returns multiply-adds per second, < 0 on error. */
double zkb_measure_imad_peak(zkb_ctx *ctx, int iters);
/* Device milliseconds (CUDA events on the context's stream) of the kernels launched by the most
 * recent zkb_msm* / zkb_witness_map / zkb_prove* call, and how many kernels that was. */
/* Live timing of the dominant kernel (bucket accumulation) with CUDA events on the context's stream.
 * zkb_profile(ctx, 1) resets and enables it; out = {G1 ms, G1 launches, G1 pairs, G2 ms, G2 launches, G2 pairs}
 * summed over the calls since; a "pair" is one mixed addition of a base into a bucket. */
int zkb_profile(zkb_ctx *ctx, int enable);
int zkb_profile_read(const zkb_ctx *ctx, double out[6]);
double zkb_last_device_ms(const zkb_ctx *ctx);
long zkb_last_kernel_launches(const zkb_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* ZKB_H */
