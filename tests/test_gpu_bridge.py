"""The zklaim-facing side of the drop-in (bridge/zkb_zklaim_bridge.cpp): credentials -> zkb_libsnark_prove(_batch) ->
the reference's own libsnark_verify; the cached, constraint-free witness engine against the
reference's per-proof pipeline; the reference's return-code contract for an unsatisfied credential."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cred():
    from zklaim_b200 import bridge
    if not bridge.available():
        pytest.skip("bridge/_build/libzkb_zklaim.so not built (make -C bridge, needs the reference tree)")
    c = bridge.Credentials(1, seed=3)
    yield c
    c.close()


def test_roundtrip_through_the_zklaim_entry_points(cred):
    ok, t_prove, t_verify = cred.roundtrip(1, 4000)
    assert ok == 1
    ok, t_prove, t_verify = cred.roundtrip(6, 4100)
    assert ok == 6


def test_cached_witness_engine_equals_reference_pipeline(cred):
    """the engine skips generate_r1cs_constraints / is_satisfied and re-uses one gadget per thread: the assignment must
    be the one the reference's snark.cpp:113-124 computes (ZKB_HOST_CHECK=1 selects that pipeline in the bridge)"""
    fast = cred.witnesses(5, 8000)
    os.environ["ZKB_HOST_CHECK"] = "1"
    try:
        slow = cred.witnesses(5, 8000)
    finally:
        del os.environ["ZKB_HOST_CHECK"]
    assert np.array_equal(fast, slow)
    assert not np.array_equal(fast[0], fast[1])


def test_unsatisfied_credential_returns_one(cred):
    assert cred.prove_tampered(91) == 1
    os.environ["ZKB_HOST_CHECK"] = "1"
    try:
        assert cred.prove_tampered(92) == 1
    finally:
        del os.environ["ZKB_HOST_CHECK"]
    ok, _, _ = cred.roundtrip(2, 4200)  # the session is still healthy afterwards
    assert ok == 2


@pytest.mark.parametrize("stock_prover", [False, True])
def test_trusted_setup_on_the_gpu_gives_a_key_the_reference_accepts(cred, stock_prover):
    """SURVEY 8f row 2 through the zklaim entry points: zkb_libsnark_trusted_setup (zkb_setup on the GPU, keys
    serialised by the reference's exporters) -> a proof by the reference's UNMODIFIED CPU libsnark_prove (or by
    zkb_libsnark_prove) under that key -> the reference's libsnark_verify accepts (0)."""
    from zklaim_b200 import bridge
    rc, t_setup, t_prove, t_verify = bridge.setup_roundtrip(1, seed=11, stock_prover=stock_prover)
    assert rc == 0


def test_batch_verification_through_the_zklaim_entry_point(cred):
    """zkb_libsnark_verify_batch (keys / proofs parsed by the reference's importers, zklaim_input_map for the public
    input, pairings on the GPU) returns what the reference's libsnark_verify returns, credential by credential --
    including one that carries another credential's proof."""
    agree, accepted, t_ref, t_gpu = cred.verify_roundtrip(16, seed0=9000, tamper=3)
    assert agree == 16
    assert accepted == 15
    agree, accepted, _, _ = cred.verify_roundtrip(1, seed0=9100)
    assert (agree, accepted) == (1, 1)
    print(f"\nverify 16 credentials: reference libsnark_verify {t_ref * 1e3:.1f} ms, zkb_libsnark_verify_batch {t_gpu * 1e3:.1f} ms")


def test_one_unsatisfied_credential_does_not_fail_the_batch(cred):
    """zkb_libsnark_prove_batch_ex: per-credential results as libsnark_prove would give them (libsnark_wrapper.cpp:233-240):
    the tampered credential gets 1 and no proof, the others get proofs the reference verifier accepts."""
    good, rc = cred.batch_with_tampered(7, 4400, 2)
    assert (good, rc) == (7, 1)
    good, rc = cred.batch_with_tampered(3, 4500, 99)  # nobody tampered
    assert (good, rc) == (3, 0)


def test_deferred_key_import_equals_the_reference_importer(cred):
    """libsnark_import_pk (square root per point on the host, libsnark_wrapper.cpp:160-168) and the
    deferred importer (libsnark's framing, roots by zkb_points_decompress) give the same flat key, element for element"""
    rc, t_ref, t_gpu = cred.import_check()
    assert rc == 0
    print(f"\nkey import: reference {t_ref:.2f} s, deferred (GPU roots) {t_gpu:.3f} s")
    assert t_gpu < t_ref


def test_cold_prove_on_a_new_key(cred):
    """the first zkb_libsnark_prove of a process under a key it has never seen: import + upload + tuning + proof"""
    rc, cold, warm = cred.cold_prove(31)
    assert rc == 0
    print(f"\ncold zkb_libsnark_prove {cold:.3f} s, warm {warm * 1e3:.1f} ms")
    ok, _, _ = cred.roundtrip(2, 4600)  # the rebuilt session serves later calls
    assert ok == 2
