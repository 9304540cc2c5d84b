"""K11 parity: zkb_setup == r1cs_gg_ppzksnark_generator (r1cs_gg_ppzksnark.tcc:211-388) for the same secret randomness,
bit for bit over every element of the proving key and of the verification key's group elements; proofs made by the
GPU prover under a GPU-made key are accepted by the reference verifier.  The oracle is the reference generator with
injected randomness (oracle/ref_harness.cpp zkref_circuit_setup_fixed, pinned by tests/test_oracle_generator.py)."""
import numpy as np
import pytest

import zklaim_b200 as zk
from test_oracle_generator import randomness

pytestmark = pytest.mark.gpu

KEY_FIELDS = ("alpha_g1", "beta_g1", "delta_g1", "beta_g2", "delta_g2", "A_query", "B_index", "B_g1", "B_g2", "H_query", "L_query")


def _cs(circ):
    d = circ.dims()
    return dict(num_constraints=d["num_constraints"], num_variables=d["num_variables"], num_inputs=d["num_inputs"],
                csr_A=circ.csr(0), csr_B=circ.csr(1), csr_C=circ.csr(2)), d


def _check(zctx, ref, circ, seed):
    cs, d = _cs(circ)          # before the key exists: the system as built
    toxic, g1, g2 = randomness(ref, seed)
    circ.setup_fixed(toxic, g1, g2)
    want, vk = circ.export_key(), circ.export_vk()
    got = zctx.setup(cs, toxic, g1, g2)
    for f in KEY_FIELDS:
        assert got[f].shape == want[f].shape and np.array_equal(got[f], want[f]), f
    for k in "ABC":  # swap_AB_if_beneficial (r1cs.tcc:182-231) decided the same way
        for a, b in zip(got["csr_" + k], want["csr_" + k]):
            assert np.array_equal(a, b)
    assert np.array_equal(got["gamma_g2"], vk["gamma_g2"])
    assert np.array_equal(got["gamma_ABC_g1"], vk["gamma_ABC_g1"])
    # GPU key -> GPU prover -> reference verifier, on a verification key assembled from the GPU's elements
    circ.import_vk(got["alpha_g1"], got["beta_g2"], got["gamma_g2"], got["delta_g2"], got["gamma_ABC_g1"])
    key = zctx.upload_key(got)
    w = circ.witness()
    rs = ref.sha512_rng(seed + 9, 2)
    proof = zctx.prove(key, w, rs[0], rs[1])
    assert circ.verify(w[:d["num_inputs"]], proof)
    assert np.array_equal(ref.to_affine_jac(zk.G1, proof[None, :12]), ref.to_affine_jac(zk.G1, circ.prove(w, rs[0], rs[1], False)["proof"][None, :12]))
    key.free()
    return got["swapped"]


@pytest.mark.parametrize("nc,ni", [(20, 3), (1013, 10), (2048 + 256 - 11, 10), (5000, 100)])
def test_generator_matches_reference_on_examples(zctx, ref, nc, ni):
    circ = ref.Circuit.example(nc, ni, seed=nc)
    _check(zctx, ref, circ, nc)
    circ.free()


def test_generator_matches_reference_on_the_zklaim_circuit(zctx, ref):
    circ = ref.Circuit.zklaim(1, 3)
    _check(zctx, ref, circ, 4242)
    circ.free()


def test_generator_swaps_a_and_b_like_the_reference(zctx, ref):
    """a system whose B side touches more variables than its A side: the generator must exchange them"""
    circ = ref.Circuit.sha_chain(1, seed=5)
    swapped = _check(zctx, ref, circ, 77)
    circ.free()
    circ = ref.Circuit.example(50, 4, seed=3)
    swapped2 = _check(zctx, ref, circ, 78)
    circ.free()
    assert isinstance(swapped, bool) and isinstance(swapped2, bool)


def test_generator_argument_errors(zctx, ref):
    circ = ref.Circuit.example(20, 3, seed=1)
    cs, d = _cs(circ)
    toxic, g1, g2 = randomness(ref, 5)
    bad = toxic.copy()
    bad[4] = 0  # delta = 0 has no inverse
    with pytest.raises(zk.ZkbError) as e:
        zctx.setup(cs, bad, g1, g2)
    assert e.value.code == zk.ERR_ARG
    bad = toxic.copy()
    bad[0] = ref.field_const(zk.FR, 1)  # t = 1 = omega^0 is an element of every radix-2 domain
    with pytest.raises(zk.ZkbError) as e:
        zctx.setup(cs, bad, g1, g2)
    assert e.value.code == zk.ERR_ARG
    circ.free()
