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
    try:  # collection must not need the library (the CPU tier collects this file too)
        n = zk.device_count()
    except Exception:
        n = 0
    cfgs = [[0, 0, 0]]
    if n >= 2:
        cfgs.append(list(range(n)))
    return cfgs


@pytest.mark.parametrize("devices", _pools())
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


def test_c_only_caller_proves_a_batch_on_every_device():
    """bridge/_build/zkb_batch_demo (plain C, no Python / torch in the process): 64 credentials through
    zkb_libsnark_prove_batch, split over the bridge's devices, every proof accepted by the reference's libsnark_verify
    (zklaim/libsnark_wrapper.cpp:252-276).  One GPU: two contexts on device 0 (ZKB_DEVICES=0,0); more: all of them."""
    import json
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "bridge", "_build", "zkb_batch_demo")
    if not os.path.exists(exe):
        pytest.skip("bridge/_build/zkb_batch_demo not built (make -C bridge all, needs the reference tree)")
    n = zk.device_count()
    env = dict(os.environ)
    env.pop("ZKB_DEVICE", None)
    env["ZKB_DEVICES"] = "0,0" if n < 2 else ",".join(str(i) for i in range(n))
    out = subprocess.run([exe, "64", "1"], env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    line = json.loads([l for l in out.stdout.splitlines() if l.startswith("{")][-1])
    assert line["devices"] == max(n, 2) and line["accepted_by_libsnark_verify"] == 64
    print("\n" + json.dumps(line))
