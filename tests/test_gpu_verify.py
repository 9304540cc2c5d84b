"""Verifier-side parity: zkb_pairing / zkb_verify_batch through the C ABI against
  * golden vectors made from the unmodified reference (tests/golden/pairing.npz, verify.npz),
  * libff's alt_bn128_ate_miller_loop / alt_bn128_reduced_pairing (alt_bn128_pairing.cpp:371-420, 523-533) and
    libsnark's r1cs_gg_ppzksnark_verifier_strong_IC (r1cs_gg_ppzksnark.tcc:523-622) run on the same inputs
    (oracle/_ref): Fq12 limbs bit-identical, accept / reject decisions identical,
  * the zklaim credential circuit: a batch of 64 GPU proofs verified on the GPU, valid and tampered.
"""
import os
import time

import numpy as np
import pytest

import zklaim_b200 as zk

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("reduced", [False, True])
def test_pairing_matches_golden(zctx, reduced):
    g = np.load(os.path.join(GOLD, "pairing.npz"))
    got = zctx.pairing(g["P"], g["Q"], reduced)
    assert np.array_equal(got, g["reduced" if reduced else "miller"])


def test_pairing_matches_reference_on_fresh_points(zctx, ref):
    n = 40  # more than one warp of lanes
    P = ref.bases(zk.G1, (1 << 43) + 7, n, mt=False)
    Q = ref.bases(zk.G2, (1 << 43) + 9, n, mt=False)
    P[3] = 0
    for reduced in (False, True):
        got = zctx.pairing(P, Q, reduced)
        for i in (0, 3, 17, 31, 32, 39):
            assert np.array_equal(got[i], ref.pairing(P[i], Q[i], reduced)), (reduced, i)
    # bilinearity at batch scale: e(P_i, Q_i) values are pairwise distinct, and e(P, Q) of a repeated pair is stable
    red = zctx.pairing(P, Q, True)
    assert len({r.tobytes() for r in red}) == n


def test_verify_batch_matches_golden_decisions(zctx):
    g = np.load(os.path.join(GOLD, "verify.npz"))
    vk = {k: g[k] for k in ("alpha_g1_beta_g2", "gamma_g2", "delta_g2", "gamma_ABC_g1")}
    acc = zctx.verify_batch(vk, g["primary"], g["proofs"])
    assert acc.astype(np.int32).tolist() == g["accept"].tolist() == [1, 1, 1, 1, 0, 0, 0, 0]
    # a single proof, and the empty batch
    assert zctx.verify_batch(vk, g["primary"][:1], g["proofs"][:1]).tolist() == [True]
    assert zctx.verify_batch(vk, g["primary"][:0], g["proofs"][:0]).tolist() == []


def test_non_canonical_encodings_are_rejected(zctx):
    """A coordinate >= q or a public input >= r is not something the reference could have serialised (Fp_model keeps
    residues canonical, fp.tcc): adding the modulus to a limb group of a VALID proof must turn accept into reject, for
    the proof concerned only."""
    g = np.load(os.path.join(GOLD, "verify.npz"))
    vk = {k: g[k] for k in ("alpha_g1_beta_g2", "gamma_g2", "delta_g2", "gamma_ABC_g1")}
    Q = 21888242871839275222246405745257275088696311157297823662689037894645226208583
    R = 21888242871839275222246405745257275088548364400416034343698204186575808495617

    def plus(limbs, m):
        v = int.from_bytes(np.ascontiguousarray(limbs, dtype="<u8").tobytes(), "little") + m
        assert v < 1 << 256
        return np.frombuffer(v.to_bytes(32, "little"), dtype="<u8")

    proofs, primary = g["proofs"][:4].copy(), g["primary"][:4].copy()
    assert zctx.verify_batch(vk, primary, proofs).tolist() == [True] * 4
    proofs[1, 0:4] = plus(proofs[1, 0:4], Q)            # A.X + q
    proofs[2, 44:48] = plus(proofs[2, 44:48], Q)        # C.Z + q
    primary[3, 0] = plus(primary[3, 0], R)              # first public input + r
    assert zctx.verify_batch(vk, primary, proofs).tolist() == [True, False, False, False]


def test_zklaim_batch_prove_then_verify_on_gpu(zctx, ref):
    """64 credential proofs made by zkb_prove_batch, verified by zkb_verify_batch; every decision equals the
    reference verifier's (valid proofs, a swapped C, a wrong public input)."""
    circ = ref.Circuit.zklaim(1, 1)
    circ.setup()
    key = zctx.upload_key(circ.export_key())
    d = circ.dims()
    ni = d["num_inputs"]
    n = 64
    ws = np.stack([ref.zklaim_witness(1, 300 + i, d["num_variables"]) for i in range(n)])
    proofs = zctx.prove_batch(key, ws, ref.sha512_rng(7000, n), ref.sha512_rng(8000, n))
    vk = circ.export_vk()
    vk["alpha_g1_beta_g2"] = circ.vk_alpha_beta()
    primary = ws[:, :ni].copy()
    proofs[5, 36:] = proofs[6, 36:]          # C of another proof
    primary[9, 0] = ref.sha512_rng(1, 1)[0]  # wrong public input
    proofs[12, 12] ^= np.uint64(2)           # B off the twist
    t0 = time.time()
    acc = zctx.verify_batch(vk, primary, proofs)
    wall = time.time() - t0
    exp = np.ones(n, dtype=bool)
    exp[[5, 9, 12]] = False
    assert acc.tolist() == exp.tolist()
    for i in (0, 5, 9, 12, 63):
        assert circ.verify(primary[i], proofs[i]) == bool(exp[i]), i
    print(f"\nverify_batch: {n} proofs, {ni} public inputs: {zctx.last_device_ms():.2f} ms device, {wall * 1e3:.2f} ms wall")
    key.free()
    circ.free()


def test_key_table_cache_follows_the_verification_key(zctx, ref):
    """zkb_verify_batch keeps the window table of gamma_ABC_g1 while the key stays the same: alternate two keys with
    the same number of inputs and check every call's decisions (a stale table would reject the valid proofs)."""
    g = np.load(os.path.join(GOLD, "verify.npz"))
    vk1 = {k: g[k] for k in ("alpha_g1_beta_g2", "gamma_g2", "delta_g2", "gamma_ABC_g1")}
    ni = len(vk1["gamma_ABC_g1"]) - 1
    circ = ref.Circuit.example(41, ni, seed=41)
    circ.setup()
    w = circ.witness()
    pr2 = np.stack([circ.prove(w, ref.sha512_rng(70 + k, 1)[0], ref.sha512_rng(80 + k, 1)[0], want_H=False)["proof"]
                    for k in range(3)]).reshape(3, 48)
    vk2 = circ.export_vk()
    vk2["alpha_g1_beta_g2"] = circ.vk_alpha_beta()
    x2 = np.tile(w[:ni], (3, 1, 1))
    exp1 = g["accept"].astype(bool).tolist()
    for _ in range(2):
        assert zctx.verify_batch(vk1, g["primary"], g["proofs"]).tolist() == exp1
        assert zctx.verify_batch(vk1, g["primary"], g["proofs"]).tolist() == exp1  # table reused
        assert zctx.verify_batch(vk2, x2, pr2).tolist() == [True, True, True]
        # proofs of key 2 under key 1: rejected
        assert zctx.verify_batch(vk1, x2, pr2).tolist() == [False, False, False]
    circ.free()


@pytest.mark.parametrize("n", [200, 1100])
def test_all_kernel_forms_decide_like_the_reference(zctx, ref, n):
    """zkb_verify_batch / zkb_pairing pick one of three kernel forms by batch size (18 lanes per pairing up to 128
    proofs, 6 lanes up to 1 024, one lane above): the golden batch tiled to n = 200 (6-lane form) and n = 1 100 (one-lane
    form) with tampered A / B / C (a point off its curve, the point at infinity, elements swapped between proofs) and a
    wrong public input.  Every decision must equal r1cs_gg_ppzksnark_verifier_strong_IC's (r1cs_gg_ppzksnark.tcc:523-622)
    on the same proof; pairing values at the same batch sizes must equal the golden Miller / reduced values."""
    g = np.load(os.path.join(GOLD, "verify.npz"))
    vk = {k: g[k] for k in ("alpha_g1_beta_g2", "gamma_g2", "delta_g2", "gamma_ABC_g1")}
    ni = len(vk["gamma_ABC_g1"]) - 1
    reps = (n + 7) // 8
    primary = np.tile(g["primary"], (reps, 1, 1))[:n].copy()
    proofs = np.tile(g["proofs"], (reps, 1))[:n].copy()
    exp = np.tile(g["accept"].astype(bool), reps)[:n].copy()
    # tamper with proofs that are valid in the golden batch (indices = 0 .. 3 mod 8), spread over the batch
    valid = [i for i in range(n) if exp[i]]
    pick = [valid[k * len(valid) // 9] for k in range(9)]
    one_q = ref.field_const(zk.FQ, 1)
    proofs[pick[0], 0] ^= np.uint64(4)                           # A off the curve
    proofs[pick[1], 0:12] = 0                                    # A = infinity (0, 1, 0)
    proofs[pick[1], 4:8] = one_q
    proofs[pick[2], 12] ^= np.uint64(2)                          # B off the twist
    proofs[pick[3], 12:36] = 0                                   # B = infinity
    proofs[pick[3], 20:24] = one_q
    proofs[pick[4], 36] ^= np.uint64(8)                          # C off the curve
    proofs[pick[5], 36:48] = 0                                   # C = infinity
    proofs[pick[5], 40:44] = one_q
    a6, c6 = proofs[pick[6], 0:12].copy(), proofs[pick[6], 36:48].copy()
    proofs[pick[6], 0:12], proofs[pick[6], 36:48] = c6, a6       # A and C swapped inside a proof
    proofs[pick[7], 36:48] = proofs[pick[8], 36:48]              # C of another (valid) proof with other randomness
    primary[pick[8], ni - 1] = ref.sha512_rng(99, 1)[0]          # a wrong public input
    acc = zctx.verify_batch(vk, primary, proofs)
    # untouched proofs keep the golden decision; every tampering above invalidates a proof that was valid
    want = exp.copy()
    want[pick] = False
    same_c = np.array_equal(g["proofs"][pick[7] % 8, 36:48], g["proofs"][pick[8] % 8, 36:48])
    if same_c:
        want[pick[7]] = True  # tiling made both proofs identical: replacing C by itself changes nothing
    assert acc.tolist() == want.tolist()
    # pairing values at this batch size (the same three kernel forms)
    gp = np.load(os.path.join(GOLD, "pairing.npz"))
    reps = (n + 5) // 6
    P, Q = np.tile(gp["P"], (reps, 1))[:n], np.tile(gp["Q"], (reps, 1))[:n]
    for reduced in (False, True):
        got = zctx.pairing(P, Q, reduced)
        assert np.array_equal(got, np.tile(gp["reduced" if reduced else "miller"], (reps, 1))[:n])


@pytest.mark.parametrize("n", [200, 1100])
def test_kernel_forms_against_the_live_reference_verifier(zctx, ref, n):
    """the same three forms on proofs of a circuit the reference holds in full (its verifier runs on each tampered proof)"""
    ni = 5
    circ = ref.Circuit.example(41, ni, seed=41)
    circ.setup()
    w = circ.witness()
    base = np.stack([circ.prove(w, ref.sha512_rng(170 + k, 1)[0], ref.sha512_rng(180 + k, 1)[0], want_H=False)["proof"]
                     for k in range(4)]).reshape(4, 48)
    vk = circ.export_vk()
    vk["alpha_g1_beta_g2"] = circ.vk_alpha_beta()
    proofs = np.tile(base, ((n + 3) // 4, 1))[:n].copy()
    primary = np.tile(w[:ni], (n, 1, 1)).copy()
    one_q = ref.field_const(zk.FQ, 1)
    t = [3, n // 7, n // 5, n // 3, n // 2, n - 70, n - 33, n - 2, n - 1]
    proofs[t[0], 1] ^= np.uint64(1)
    proofs[t[1], 0:12] = 0
    proofs[t[1], 4:8] = one_q
    proofs[t[2], 13] ^= np.uint64(1)
    # (B = infinity is left to the golden-based test above: libff's alt_bn128_ate_precompute_G2 ASSERTS on it,
    #  alt_bn128_pairing.cpp: assert(Q1.Z == alt_bn128_Fq2::one()), so the reference aborts instead of rejecting)
    proofs[t[3], 13] ^= np.uint64(4)
    proofs[t[4], 37] ^= np.uint64(1)
    proofs[t[5], 36:48] = 0
    proofs[t[5], 40:44] = one_q
    proofs[t[6], 36:48] = proofs[(t[6] + 1), 36:48]   # C of the neighbouring proof (other r, s)
    primary[t[7], 0] = ref.sha512_rng(5, 1)[0]
    a, c = proofs[t[8], 0:12].copy(), proofs[t[8], 36:48].copy()
    proofs[t[8], 0:12], proofs[t[8], 36:48] = c, a
    acc = zctx.verify_batch(vk, primary, proofs)
    want = np.ones(n, dtype=bool)
    for i in t:
        want[i] = circ.verify(primary[i], proofs[i])
    assert not want[t].any()
    for i in (0, 1, n // 4, n - 3):
        assert circ.verify(primary[i], proofs[i])
    assert acc.tolist() == want.tolist()
    circ.free()
