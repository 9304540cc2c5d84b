"""Golden values for the BASELINE.json configurations that are too large to recompute with the CPU reference on every
test run ("2^22 / 2^24 once, cached digest").  Made ONCE, in the build container, from the
UNMODIFIED reference (oracle/_ref/libzkref_mt.so = libff / libfqfft / libsnark compiled in place):

  msm_g1_{22,24}   libff::multi_exp<G1, Fr, multi_exp_method_BDLO12> (multiexp.tcc:402-441) over n = 2^22 / 2^24 distinct
                   bases k_i G, k_i = SHA512_rng<Fr>(MSM_BASE_FIRST + i), scalars SHA512_rng<Fr>(MSM_SCALAR_FIRST + i);
                   stored: the affine-normalised result.  Cross-checked here in closed form: (sum s_i k_i mod r) G.
  s22_*            config #5: 153 chained sha256_two_to_one_hash_gadgets (4 173 840 constraints, domain 2^22): key by the
                   reference generator with injected randomness (test_oracle_generator.randomness(ref, S22_SEED)),
                   proof by the reference prover with injected (r, s); stored: serialised proof bytes and SHA-256 of H.

    python tests/golden/make_golden_large.py [msm22] [msm24] [s22]      (about an hour on 8 cores for all three)

tests/test_gpu_large.py regenerates the same inputs on the GPU box and compares the CUDA path with these values.
"""
import hashlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import zkref  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "large.npz")
G1 = 1
R = 21888242871839275222246405745257275088548364400416034343698204186575808495617
MSM_BASE_FIRST, MSM_SCALAR_FIRST = 1 << 44, 3 << 44
S22_BLOCKS, S22_CIRCUIT_SEED, S22_SEED, S22_R, S22_S = 153, 7, 2222, 1001, 1002


def _ints(limbs):
    b = np.ascontiguousarray(limbs, dtype="<u8").tobytes()
    return [int.from_bytes(b[32 * i:32 * i + 32], "little") for i in range(len(limbs))]


def msm(lg):
    n = 1 << lg
    t = time.time()
    bases = zkref.bases(G1, MSM_BASE_FIRST, n, mt=True)
    s = zkref.sha512_rng(MSM_SCALAR_FIRST, n, mt=True)
    print(f"2^{lg}: inputs in {time.time() - t:.0f}s", flush=True)
    res, sec = zkref.msm(G1, bases, s, chunks=0, mt=True)
    aff = zkref.to_affine_jac(G1, res[None])[0]
    print(f"2^{lg}: libff::multi_exp {sec:.1f}s", flush=True)
    # closed form: bases are k_i G
    k = zkref.sha512_rng(MSM_BASE_FIRST, n, mt=True)
    acc = 0
    step = 1 << 18
    for lo in range(0, n, step):
        ki = _ints(zkref.to_plain(0, k[lo:lo + step]))
        si = _ints(zkref.to_plain(0, s[lo:lo + step]))
        acc = (acc + sum(a * b for a, b in zip(ki, si))) % R
    e = np.array([[(acc >> (64 * j)) & ((1 << 64) - 1) for j in range(4)]], dtype=np.uint64)
    chk = zkref.to_affine_jac(G1, zkref.scalar_mul(G1, zkref.from_plain(0, e), zkref.generator(G1)[None]))[0]
    assert np.array_equal(chk, aff), "closed form disagrees with libff::multi_exp"
    return {f"msm_g1_{lg}": aff, f"msm_g1_{lg}_ref_seconds": np.array([sec])}


def s22():
    from test_oracle_generator import randomness
    saved = os.dup(1)
    os.dup2(2, 1)  # libsnark prints to fd 1
    try:
        circ = zkref.Circuit.sha_chain(S22_BLOCKS, S22_CIRCUIT_SEED, mt=True)
        d = circ.dims()
        toxic, g1, g2 = randomness(zkref, S22_SEED)
        t_setup = circ.setup_fixed(toxic, g1, g2)
        w = circ.witness()
        r = zkref.sha512_rng(S22_R, 1, mt=True)[0]
        s = zkref.sha512_rng(S22_S, 1, mt=True)[0]
        t = time.time()
        exp = circ.prove(w, r, s)
        t_prove = time.time() - t
        # the verifier of the checker build (no asm: SURVEY.md section 0.4) on the timing build's key and proof; the
        # verifier needs the verification key and the number of inputs only, so a one-block circuit object carries it
        kd, vk = circ.export_key(), circ.export_vk()
        v = zkref.Circuit.sha_chain(1, S22_CIRCUIT_SEED)
        v.setup()
        v.import_vk(kd["alpha_g1"], kd["beta_g2"], vk["gamma_g2"], vk["delta_g2"], vk["gamma_ABC_g1"])
        ok = v.verify(w[:d["num_inputs"]], exp["proof"])
        bad = exp["proof"].copy()
        bad[36:48] = bad[0:12]
        ok_bad = v.verify(w[:d["num_inputs"]], bad)
    finally:
        os.dup2(saved, 1)
    print(f"s22: {d}, reference generator {t_setup:.0f}s, prover {t_prove:.0f}s, verifier {ok} / tampered {ok_bad}", flush=True)
    assert ok and not ok_bad
    return {"s22_dims": np.array([d["num_constraints"], d["num_variables"], d["num_inputs"], d["m"], d["kind"]], dtype=np.uint64),
            "s22_proof_bytes": np.frombuffer(zkref.proof_bytes(exp["proof"]), dtype=np.uint8),
            "s22_H_sha256": np.frombuffer(hashlib.sha256(np.ascontiguousarray(exp["H"]).tobytes()).digest(), dtype=np.uint8),
            "s22_witness_sha256": np.frombuffer(hashlib.sha256(np.ascontiguousarray(w).tobytes()).digest(), dtype=np.uint8),
            "s22_ref_seconds": np.array([t_setup, t_prove])}


if __name__ == "__main__":
    what = sys.argv[1:] or ["msm22", "msm24", "s22"]
    have = dict(np.load(OUT)) if os.path.exists(OUT) else {}
    for w in what:
        have.update(msm(int(w[3:])) if w.startswith("msm") else s22())
        np.savez_compressed(OUT, **have)
        print("wrote", w, flush=True)
