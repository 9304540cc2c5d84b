"""Does a batch gain from two half-batches in flight on ONE GPU (two contexts = two stream sets and scratch sets, the
latency-bound tails of one half overlapping the throughput-bound kernels of the other)?  Development aid.
Usage: python scripts/pool_overlap.py [batch=64] [reps=8]"""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import zklaim_b200 as zk
from zklaim_b200 import bridge

batch = int(sys.argv[1]) if len(sys.argv) > 1 else 64
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 8
saved = os.dup(1); os.dup2(2, 1)
cred = bridge.Credentials(1, seed=1)
w = torch.from_numpy(cred.witnesses(batch, 1000).view(np.int64)).pin_memory().numpy().view(np.uint64)
r = torch.from_numpy(cred.random_fr(batch).view(np.int64)).pin_memory().numpy().view(np.uint64)
s = torch.from_numpy(cred.random_fr(batch).view(np.int64)).pin_memory().numpy().view(np.uint64)
res = {}
# single context, blocking call (what e2e.serial_value measures)
for i in range(3): cred.ctx.prove_batch(cred.key, w, r, s)
t = time.perf_counter()
for i in range(reps): ref = cred.ctx.prove_batch(cred.key, w, r, s)
res["one_context_ms"] = 1e3 * (time.perf_counter() - t) / reps
kd = None
from oracle import zkref
circ = zkref.Circuit.zklaim(1, 1)
for nctx in (1, 2, 3, 4):
    pool = zk.Pool([0] * nctx)
    # the bridge's key arrays are not exposed: take an equivalent key from the reference generator for the timing
    if kd is None:
        circ.setup(); kd = circ.export_key(); wz = circ.witness()
        wb = torch.from_numpy(np.stack([wz] * batch).view(np.int64)).pin_memory().numpy().view(np.uint64)
    keys = pool.upload_key(kd); pool.tune_key(keys, wz)
    for i in range(3): pool.prove_batch(keys, wb, r, s)
    t = time.perf_counter()
    for i in range(reps): proofs, st = pool.prove_batch(keys, wb, r, s)
    res[f"pool_{nctx}_contexts_ms"] = 1e3 * (time.perf_counter() - t) / reps
    assert not st.any()
    pool.free_key(keys); pool.close()
os.dup2(saved, 1)
print(json.dumps(res))
