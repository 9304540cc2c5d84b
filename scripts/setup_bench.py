"""Generator timing: zkb_setup (GPU) next to the reference generator (oracle/_ref, all host threads) on the zklaim circuit
and on libsnark example systems up to 2^20 constraints; every key element compared bit for bit.  Development /
measurement aid (bench.py is the contract).  Usage: python scripts/setup_bench.py [logs=16,18,20]"""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import zklaim_b200 as zk
from oracle import zkref as ref
from test_oracle_generator import randomness

logs = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [16, 18, 20]
ctx = zk.Context(0)
mt = ref.available(mt=True)
saved = os.dup(1); os.dup2(2, 1)
rows = []
cases = [("zklaim 1 payload", lambda: ref.Circuit.zklaim(1, 3, mt=mt))] + \
        [(f"example 2^{lg}", (lambda lg=lg: ref.Circuit.example((1 << lg) - 11, 10, seed=lg, mt=mt))) for lg in logs]
for name, make in cases:
    circ = make()
    d = circ.dims()
    cs = dict(num_constraints=d["num_constraints"], num_variables=d["num_variables"], num_inputs=d["num_inputs"],
              csr_A=circ.csr(0), csr_B=circ.csr(1), csr_C=circ.csr(2))
    toxic, g1, g2 = randomness(ref, 1234)
    best_wall, best_dev = 1e9, 1e9
    for it in range(2):
        t0 = time.time()
        got = ctx.setup(cs, toxic, g1, g2)
        best_wall = min(best_wall, time.time() - t0)
        best_dev = min(best_dev, ctx.last_device_ms())
    t_ref = circ.setup_fixed(toxic, g1, g2)
    want, vk = circ.export_key(), circ.export_vk()
    same = all(np.array_equal(got[f], want[f]) for f in ("alpha_g1", "beta_g1", "delta_g1", "beta_g2", "delta_g2", "A_query", "B_index",
                                                          "B_g1", "B_g2", "H_query", "L_query")) and \
        np.array_equal(got["gamma_g2"], vk["gamma_g2"]) and np.array_equal(got["gamma_ABC_g1"], vk["gamma_ABC_g1"])
    n_g1 = 3 * (d["num_variables"] + 1) + d["m"] - 1 + 3
    n_g2 = d["num_variables"] + 4
    rows.append({"circuit": name, "constraints": d["num_constraints"], "variables": d["num_variables"], "m": d["m"],
                 "g1_exponentiations": n_g1, "g2_exponentiations": n_g2, "gpu_device_ms": best_dev, "gpu_wall_s": best_wall,
                 "ref_s": t_ref, "ref_threads": int(ref.lib(mt).zkref_num_threads()), "key_bit_exact": bool(same)})
    print(rows[-1], file=sys.stderr, flush=True)
    circ.free()
os.dup2(saved, 1)
print(json.dumps({"workload": "Groth16 generator, injected randomness", "rows": rows}))
sys.exit(0 if all(r["key_bit_exact"] for r in rows) else 1)
