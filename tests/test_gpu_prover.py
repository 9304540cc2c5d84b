"""Full-prover parity: zkb_prove / zkb_prove_batch / zkb_witness_map against the reference prover
(r1cs_gg_ppzksnark.tcc:390-505, r1cs_to_qap.tcc:205-334) run with the same injected (r, s):
  * coefficients_for_H bit-identical,
  * proof elements (A, B, C) bit-identical after the reference's own affine normalisation, and the
    serialised proof (r1cs_gg_ppzksnark.tcc:168-189) byte-identical,
  * the reference verifier (r1cs_gg_ppzksnark.tcc:523-576) accepts the GPU proof.
Circuits: the zklaim credential circuit of zklaim/main.c (1 payload; 5 payloads = step domain),
libsnark's generate_r1cs_example_with_field_input (dense witnesses; basic and step domains).
"""
import numpy as np
import pytest

import zklaim_b200 as zk

pytestmark = pytest.mark.gpu


def _norm_proof(ref, pr):
    a = ref.to_affine_jac(zk.G1, pr[None, 0:12])[0]
    b = ref.to_affine_jac(zk.G2, pr[None, 12:36])[0]
    c = ref.to_affine_jac(zk.G1, pr[None, 36:48])[0]
    return np.concatenate([a, b, c])


def _check_against_reference(zctx, ref, circ, key, w, seed):
    d = circ.dims()
    r = ref.sha512_rng(seed, 1)[0]
    s = ref.sha512_rng(seed + 1, 1)[0]
    exp = circ.prove(w, r, s)
    got, H = zctx.prove(key, w, r, s, want_H=True)
    assert np.array_equal(H, exp["H"]), "coefficients_for_H differ"
    assert np.array_equal(_norm_proof(ref, got), _norm_proof(ref, exp["proof"])), "proof elements differ"
    assert ref.proof_bytes(got) == ref.proof_bytes(exp["proof"])
    assert circ.verify(w[: d["num_inputs"]], got), "reference verifier rejects the GPU proof"
    return got


@pytest.fixture(scope="module")
def zklaim1(zctx, ref):
    circ = ref.Circuit.zklaim(1, 1)
    assert circ.is_satisfied()
    circ.setup()
    key = zctx.upload_key(circ.export_key())
    yield circ, key
    key.free()
    circ.free()


def test_zklaim_circuit_proof_bit_exact(zctx, ref, zklaim1):
    circ, key = zklaim1
    d = circ.dims()
    assert key.domain() == (d["m"], d["kind"]) == (32768, zk.DOMAIN_BASIC_RADIX2)
    w = circ.witness()
    _check_against_reference(zctx, ref, circ, key, w, 1000)
    # witness map alone
    Hexp, _ = circ.witness_map(w)
    assert np.array_equal(zctx.witness_map(key, w), Hexp)


def test_zklaim_batch_of_distinct_witnesses(zctx, ref, zklaim1):
    circ, key = zklaim1
    d = circ.dims()
    n = 5
    ws = np.stack([ref.zklaim_witness(1, 100 + i, d["num_variables"]) for i in range(n)])
    r = ref.sha512_rng(5000, n)
    s = ref.sha512_rng(6000, n)
    got = zctx.prove_batch(key, ws, r, s)
    for i in range(n):
        exp = circ.prove(ws[i], r[i], s[i], want_H=False)
        assert np.array_equal(_norm_proof(ref, got[i]), _norm_proof(ref, exp["proof"])), i
        assert circ.verify(ws[i][: d["num_inputs"]], got[i])


def test_key_tune_keeps_proofs_bit_exact(zctx, ref):
    """zkb_key_tune re-plans the MSM windows from a sample witness: narrower windows for the 0 / 1-dominated
    A / B queries, proofs unchanged (single proof, batch, and a dense witness under the tuned key)."""
    circ = ref.Circuit.zklaim(1, 1)
    circ.setup()
    key = zctx.upload_key(circ.export_key())
    d = circ.dims()
    w = circ.witness()
    before = key.windows()
    key.tune(w)
    after = key.windows()
    assert after[0] < before[0] and after[1] < before[1] and after[3] < before[3], (before, after)
    assert after[2] >= before[2] - 1  # the H query stays dense
    _check_against_reference(zctx, ref, circ, key, w, 4242)
    n = 3
    ws = np.stack([ref.zklaim_witness(1, 900 + i, d["num_variables"]) for i in range(n)])
    r = ref.sha512_rng(5100, n)
    s = ref.sha512_rng(6100, n)
    got = zctx.prove_batch(key, ws, r, s)
    for i in range(n):
        exp = circ.prove(ws[i], r[i], s[i], want_H=False)
        assert ref.proof_bytes(got[i]) == ref.proof_bytes(exp["proof"]), i
    key.tune(w)  # idempotent
    assert key.windows() == after
    key.free()
    circ.free()


def test_key_tune_dense_witness(zctx, ref):
    circ = ref.Circuit.example(1013, 10, seed=5)
    circ.setup()
    key = zctx.upload_key(circ.export_key())
    key.tune(circ.witness())
    _check_against_reference(zctx, ref, circ, key, circ.witness(), 99)
    key.free()
    circ.free()


def test_unsatisfied_witness_is_reported(zctx, ref, zklaim1):
    circ, key = zklaim1
    w = circ.witness().copy()
    w[100] = ref.sha512_rng(1, 1)[0]
    r = ref.sha512_rng(1, 1)[0]
    with pytest.raises(zk.ZkbError) as e:
        zctx.prove(key, w, r, r)
    assert e.value.code == zk.ERR_UNSAT


@pytest.mark.parametrize("nc,ni", [(20, 3), (1013, 10), (2048 - 11, 10), (2048 + 256 - 11, 10), (5000, 100)])
def test_example_circuits_dense_witness(zctx, ref, nc, ni):
    circ = ref.Circuit.example(nc, ni, seed=nc)
    assert circ.is_satisfied()
    circ.setup()
    d = circ.dims()
    key = zctx.upload_key(circ.export_key())
    assert key.domain() == (d["m"], d["kind"])
    _check_against_reference(zctx, ref, circ, key, circ.witness(), 77 + nc)
    key.free()
    circ.free()


def test_zklaim_five_payloads_step_domain(zctx, ref):
    circ = ref.Circuit.zklaim(5, 3)
    circ.setup()
    d = circ.dims()
    assert d["kind"] == zk.DOMAIN_STEP_RADIX2 and d["m"] == 147456
    key = zctx.upload_key(circ.export_key())
    _check_against_reference(zctx, ref, circ, key, circ.witness(), 31)
    key.free()
    circ.free()
