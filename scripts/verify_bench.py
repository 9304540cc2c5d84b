"""Batch verification timing: zkb_verify_batch (GPU) next to the reference verifier r1cs_gg_ppzksnark_verifier_strong_IC
(oracle/_ref; single-threaded, as libsnark_verify runs it) on proofs of the zklaim credential circuit made by
zkb_prove_batch; every GPU decision compared with the reference's on a sample.  Development / measurement aid
(bench.py is the contract).  Usage: python scripts/verify_bench.py [batches=64,512,4096]"""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import zklaim_b200 as zk
from oracle import zkref as ref

batches = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [64, 512, 4096]
ctx = zk.Context(0)
mt = ref.available(mt=True)
saved = os.dup(1); os.dup2(2, 1)
circ = ref.Circuit.zklaim(1, 1, mt=mt)
circ.setup()
key = ctx.upload_key(circ.export_key())
d = circ.dims()
ni = d["num_inputs"]
base = 64
ws = np.stack([ref.zklaim_witness(1, 300 + i, d["num_variables"]) for i in range(base)])
proofs = ctx.prove_batch(key, ws, ref.sha512_rng(7000, base), ref.sha512_rng(8000, base))
primary = ws[:, :ni].copy()
vk = circ.export_vk()
vk["alpha_g1_beta_g2"] = circ.vk_alpha_beta()
# reference: per-proof seconds (median of 8), decisions on valid and tampered proofs
bad = proofs[0].copy(); bad[36:] = proofs[1][36:]
t_ref = []
for i in range(8):
    t0 = time.time(); ok = circ.verify(primary[i], proofs[i]); t_ref.append(time.time() - t0)
    assert ok
assert not circ.verify(primary[0], bad)
ref_ms = float(np.median(t_ref) * 1e3)
rows = []
for n in batches:
    reps = (n + base - 1) // base
    P = np.tile(proofs, (reps, 1))[:n].copy()
    X = np.tile(primary, (reps, 1, 1))[:n].copy()
    P[n // 2] = bad; X[n // 2] = primary[0]
    best_dev, best_wall = 1e9, 1e9
    for it in range(3):
        t0 = time.time(); acc = ctx.verify_batch(vk, X, P); best_wall = min(best_wall, time.time() - t0)
        best_dev = min(best_dev, ctx.last_device_ms())
    exp = np.ones(n, dtype=bool); exp[n // 2] = False
    rows.append({"batch": n, "public_inputs": ni, "gpu_device_ms": best_dev, "gpu_wall_ms": best_wall * 1e3,
                 "gpu_proofs_per_s": n / best_wall, "ref_ms_per_proof_1_thread": ref_ms, "ref_batch_ms_1_thread": ref_ms * n,
                 "speedup_vs_ref_1_thread": ref_ms * n / (best_wall * 1e3), "decisions_equal": bool((acc == exp).all())})
    print(rows[-1], file=sys.stderr, flush=True)
os.dup2(saved, 1)
print(json.dumps({"workload": "Groth16 batch verification, zklaim 1-payload circuit (6 public inputs)", "rows": rows}))
sys.exit(0 if all(r["decisions_equal"] for r in rows) else 1)
