"""Host side of the zklaim-facing path: how fast does zklaim's own gadget code produce witnesses on this box?
(bridge WitnessEngine, all OpenMP threads).  Development aid.  Usage: python scripts/host_witness_rate.py"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from zklaim_b200 import bridge
saved = os.dup(1); os.dup2(2, 1)
cred = bridge.Credentials(1, seed=1)
cred.witnesses(64, 1)
res = {"host_threads": len(os.sched_getaffinity(0))}
for n in (64, 256):
    t = time.perf_counter(); cred.witnesses(n, 100 + n); dt = time.perf_counter() - t
    res[f"witnesses_{n}_per_s"] = n / dt
for thr in (1, 2, 3):
    cred.stream(1, 64, thr)
    n, sec = cred.stream(4, 64, thr)
    res[f"stream_{thr}_callers_proofs_per_s"] = n / sec
t = time.perf_counter(); ok, tp, tv = cred.roundtrip(64, 777); res["roundtrip_64_prove_ms"] = tp * 1e3
os.dup2(saved, 1)
print(json.dumps(res))
