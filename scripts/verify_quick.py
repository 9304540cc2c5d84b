"""One zkb_verify_batch call on the golden verifier fixtures replicated to a batch (no reference needed): for ncu launch
lists.  Usage: python scripts/verify_quick.py [batch=64] [reps=3]"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import zklaim_b200 as zk

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
g = np.load(os.path.join(ROOT, "tests", "golden", "verify.npz"))
vk = {k: g[k] for k in ("alpha_g1_beta_g2", "gamma_g2", "delta_g2", "gamma_ABC_g1")}
k = (n + 7) // 8
P = np.tile(g["proofs"], (k, 1))[:n].copy()
X = np.tile(g["primary"], (k, 1, 1))[:n].copy()
exp = np.tile(g["accept"], k)[:n].astype(bool)
ctx = zk.Context(0)
for _ in range(reps):
    acc = ctx.verify_batch(vk, X, P)
    print(f"batch {n}: {ctx.last_device_ms():.3f} ms device, decisions ok: {bool((acc == exp).all())}")
