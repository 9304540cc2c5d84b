"""CPU tier: the generator with injected randomness (oracle/ref_harness.cpp zkref_circuit_setup_fixed) is pinned against
the reference itself: a key it makes lets the reference's STOCK prover produce proofs its own verifier accepts
(r1cs_gg_ppzksnark.tcc:390-505, :612-622), its fixed elements are the injected scalars times the injected generators,
and it is deterministic.  This is the oracle of tests/test_gpu_setup.py."""
import numpy as np
import pytest

G1, G2, FR = 1, 2, 0


def randomness(ref, seed):
    toxic = ref.sha512_rng(seed << 20, 5)
    k = ref.sha512_rng((seed << 20) + 100, 2)
    g1 = ref.scalar_mul(G1, k[:1], ref.generator(G1)[None])[0]
    g2 = ref.scalar_mul(G2, k[1:], ref.generator(G2)[None])[0]
    return toxic, g1, g2


@pytest.mark.parametrize("nc,ni", [(20, 3), (300, 10)])
def test_fixed_generator_is_a_valid_reference_key(ref, nc, ni):
    circ = ref.Circuit.example(nc, ni, seed=nc)
    toxic, g1, g2 = randomness(ref, nc)
    circ.setup_fixed(toxic, g1, g2)
    key = circ.export_key()
    w = circ.witness()
    proof = np.zeros(48, dtype=np.uint64)
    circ.L.zkref_prove_stock(circ.h, w.ctypes.data, proof.ctypes.data)
    assert circ.verify(w[:ni], proof)
    bad = w[:ni].copy()
    bad[0, 0] ^= np.uint64(1)
    assert not circ.verify(bad, proof)
    # fixed elements = injected scalar * injected generator
    for name, idx, grp, gen in (("alpha_g1", 1, G1, g1), ("beta_g1", 2, G1, g1), ("delta_g1", 4, G1, g1),
                                ("beta_g2", 2, G2, g2), ("delta_g2", 4, G2, g2)):
        exp = ref.to_affine_jac(grp, ref.scalar_mul(grp, toxic[idx:idx + 1], gen[None]))[0]
        assert np.array_equal(exp[:len(key[name])], key[name]), name
    vk = circ.export_vk()
    exp = ref.to_affine_jac(G2, ref.scalar_mul(G2, toxic[3:4], g2[None]))[0]
    assert np.array_equal(exp[:16], vk["gamma_g2"])
    # deterministic
    circ2 = ref.Circuit.example(nc, ni, seed=nc)
    circ2.setup_fixed(toxic, g1, g2)
    key2 = circ2.export_key()
    for f in ("A_query", "B_index", "B_g1", "B_g2", "H_query", "L_query"):
        assert np.array_equal(key[f], key2[f]), f
    circ.free()
    circ2.free()
