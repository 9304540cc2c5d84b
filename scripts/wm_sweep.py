"""BASELINE.json configs[3]: r1cs_to_qap_witness_map at domain sizes 2^18 .. 2^22 (basic radix-2) and one step-domain size,
on libsnark's generate_r1cs_example_with_field_input(m - 11, 10) instances; H coefficients compared bit for bit with the
reference (oracle/_ref) and timed next to it.  Usage: python scripts/wm_sweep.py [logs=18,19,20] [step=1]
Development / measurement aid (bench.py is the contract).  Work model: WM(m) = 3.5 m log2 m + 9 m Fr products."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import zklaim_b200 as zk
from oracle import zkref

logs = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [18, 19, 20]
step = int(sys.argv[2]) if len(sys.argv) > 2 else 1
sizes = [(1 << lg) - 11 for lg in logs] + ([(1 << 17) + (1 << 14) - 11] if step else [])
ctx = zk.Context(0)
peak = ctx.imad_peak(512)
saved = os.dup(1); os.dup2(2, 1)
rows = []
for nc in sizes:
    mt = zkref.available(mt=True)
    circ = zkref.Circuit.example(nc, 10, seed=nc, mt=mt)
    d = circ.dims()
    w = circ.witness()
    # the witness map needs the constraint system only: a key whose queries are dummies (one generator point each)
    kd = {"num_constraints": d["num_constraints"], "num_variables": d["num_variables"], "num_inputs": d["num_inputs"],
          "csr_A": circ.csr(0), "csr_B": circ.csr(1), "csr_C": circ.csr(2)}
    g1 = zkref.bases(zk.G1, 1, 1)[0]; g2 = zkref.bases(zk.G2, 1, 1)[0]
    nv, ni, m = d["num_variables"], d["num_inputs"], d["m"]
    kd.update(alpha_g1=g1, beta_g1=g1, delta_g1=g1, beta_g2=g2, delta_g2=g2, A_query=np.tile(g1, (nv + 1, 1)),
              B_index=np.zeros(1, dtype=np.uint32), B_g1=g1[None], B_g2=g2[None], H_query=np.tile(g1, (m - 1, 1)),
              L_query=np.tile(g1, (nv - ni, 1)))
    key = ctx.upload_key(kd)
    best = 1e9
    for it in range(4):
        H = ctx.witness_map(key, w)
        best = min(best, ctx.last_device_ms())
    Href, t_ref = circ.witness_map(w)
    same = bool(np.array_equal(H, Href))
    lg = float(np.log2(m))
    work = (3.5 * m * lg + 9 * m) * 136
    rows.append({"m": m, "kind": d["kind"], "constraints": nc, "gpu_ms": best, "ref_s": t_ref, "ref_threads": int(zkref.lib(mt).zkref_num_threads()),
                 "H_bit_exact": same, "imad_frac": work / (best * 1e-3) / peak, "hbm_GBps": 7 * 64 * m / (best * 1e-3) / 1e9})
    print(rows[-1], file=sys.stderr, flush=True)
    key.free(); circ.free()
os.dup2(saved, 1)
print(json.dumps({"workload": "witness map sweep", "imad_peak": peak, "rows": rows}))
sys.exit(0 if all(r["H_bit_exact"] for r in rows) else 1)
