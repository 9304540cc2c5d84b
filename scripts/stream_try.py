"""How many caller threads / batches per caller does the zklaim-facing stream (zkb_libsnark_prove_batch, bridge) need?
Development aid behind bench.py's e2e_zklaim (3 callers x 10 batches).  Usage: python scripts/stream_try.py"""
import os, sys, json
sys.path.insert(0, os.getcwd())
from zklaim_b200 import bridge
saved = os.dup(1); os.dup2(2, 1)
cred = bridge.Credentials(1, seed=1)
cred.stream(1, 64, 3)
res = {}
for thr in (2, 3, 4):
    for nb in (4, 10):
        n, sec = cred.stream(nb, 64, thr)
        res[f"{thr}x{nb}"] = round(n / sec, 1)
os.dup2(saved, 1)
print(json.dumps(res))
