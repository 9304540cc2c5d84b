"""Several devices behind the C ABI (zkb_pool, include/zkb.h): one context and one host thread per device.
On a box with one GPU the pool is built over the SAME device three times (three contexts, three host threads), which
exercises the splitting / gathering logic; on a multi-GPU box it spans every visible device as well.
  zkb_pool_prove_batch    == zkb_prove_batch proof for proof (and per-proof status for an unsatisfying witness),
  zkb_pool_msm            == libff::multi_exp over the whole range (multiexp.tcc:417-438 chunking + final sum),
  zkb_pool_prove_sharded  == the reference proof, bytes identical."""
import numpy as np
import pytest

import zklaim_b200 as zk

pytestmark = pytest.mark.gpu


def _pools():
    n = zk.device_count()
    cfgs = [[0, 0, 0]]
    if n >= 2:
        cfgs.append(list(range(n)))
    return cfgs


@pytest.mark.parametrize("devices", _pools() if zk.device_count() else [[0, 0, 0]])
def test_pool_batch_msm_and_sharded_proof(ref, devices):
    pool = zk.Pool(devices)
    assert pool.size == len(devices)
    circ = ref.Circuit.example(1013, 10, seed=21)
    circ.setup()
    kd = circ.export_key()
    w = circ.witness()
    n = 7  # not a multiple of the pool size
    ws = np.stack([w] * n)
    bad = w.copy()
    bad[50] = ref.sha512_rng(3, 1)[0]
    ws[4] = bad
    r, s = ref.sha512_rng(100, n), ref.sha512_rng(200, n)
    keys = pool.upload_key(kd)
    pool.tune_key(keys, w)
    proofs, status = pool.prove_batch(keys, ws, r, s)
    assert status.tolist() == [0, 0, 0, 0, zk.ERR_UNSAT, 0, 0]
    assert not proofs[4].any()
    for i in (0, 3, 5, 6):
        exp = circ.prove(w, r[i], s[i], want_H=False)
        assert ref.proof_bytes(proofs[i]) == ref.proof_bytes(exp["proof"]), i
    pool.free_key(keys)
    # one proof over all devices of the pool
    keys = pool.upload_key(kd, sharded=True)
    got = pool.prove_sharded(keys, w, r[0], s[0])
    assert ref.proof_bytes(got) == ref.proof_bytes(circ.prove(w, r[0], s[0], want_H=False)["proof"])
    with pytest.raises(zk.ZkbError):
        pool.prove_batch(keys, ws, r, s)  # a sharded key set does not serve whole proofs
    pool.free_key(keys)
    # range-sharded MSM
    m = 5001
    bases = ref.bases(zk.G1, 1 << 36, m)
    sc = ref.sha512_rng(777, m)
    B = pool.upload_bases(zk.G1, bases)
    got = pool.msm(B, sc)
    exp, _ = ref.msm(zk.G1, bases, sc)
    assert np.array_equal(ref.to_affine_jac(zk.G1, got[None]), ref.to_affine_jac(zk.G1, exp[None]))
    pool.free_bases(B)
    circ.free()
    pool.close()
