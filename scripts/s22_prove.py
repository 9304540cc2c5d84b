"""BASELINE.json configs[4]: synthetic chained-SHA-256 R1CS (153 x sha256_two_to_one_hash_gadget = 4.17 M constraints,
domain 2^22), full Groth16 prove on one B200 next to the reference's multicore libsnark prover with the same (r, s).
Usage: python scripts/s22_prove.py [blocks=153] [reps=3]      (development / measurement aid; bench.py is the contract)
Prints one JSON line; bit-exactness = serialised proof bytes identical to the reference prover's."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import zklaim_b200 as zk
from oracle import zkref

blocks = int(sys.argv[1]) if len(sys.argv) > 1 else 153
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
saved = os.dup(1); os.dup2(2, 1)          # libsnark prints to fd 1
t = time.time(); circ = zkref.Circuit.sha_chain(blocks, 7, mt=True); t_build = time.time() - t
d = circ.dims()
print(f"circuit: {d}, built in {t_build:.1f}s", file=sys.stderr, flush=True)
t_setup = circ.setup()
print(f"reference setup {t_setup:.1f}s", file=sys.stderr, flush=True)
t = time.time(); kd = circ.export_key(); t_export = time.time() - t
ctx = zk.Context(0)
t = time.time(); key = ctx.upload_key(kd); t_upload = time.time() - t
w = circ.witness()
before = key.windows()
t = time.time(); key.tune(w); t_tune = time.time() - t
print(f"key upload {t_upload:.1f}s, tune {t_tune:.1f}s: windows {before} -> {key.windows()}, {key.device_bytes()/1e9:.2f} GB on device", file=sys.stderr, flush=True)
r = zkref.sha512_rng(1001, 1, mt=True)[0]; s = zkref.sha512_rng(1002, 1, mt=True)[0]
lat, dev = [], []
for i in range(reps + 1):
    t = time.time(); got, H = ctx.prove(key, w, r, s, want_H=True); lat.append(time.time() - t); dev.append(ctx.last_device_ms())
    print(f"gpu prove {lat[-1]*1e3:.1f} ms wall, {dev[-1]:.1f} ms device", file=sys.stderr, flush=True)
t = time.time(); exp = circ.prove(w, r, s); t_ref = time.time() - t
print(f"reference prove {t_ref:.1f}s ({zkref.lib(True).zkref_num_threads()} threads)", file=sys.stderr, flush=True)
same_H = bool(np.array_equal(H, exp["H"]))
same_proof = zkref.proof_bytes(got) == zkref.proof_bytes(exp["proof"])
os.dup2(saved, 1)
print(json.dumps({"workload": f"sha256 chain x{blocks}", **{k: d[k] for k in ("num_constraints", "num_variables", "num_inputs", "m", "kind")},
                  "H_bit_exact": same_H, "proof_bit_exact": same_proof, "gpu_prove_ms_wall": min(lat[1:]) * 1e3, "gpu_prove_ms_device": min(dev[1:]),
                  "reference_prove_s": t_ref, "reference_threads": int(zkref.lib(True).zkref_num_threads()), "reference_setup_s": t_setup,
                  "key_device_gb": key.device_bytes() / 1e9, "windows": key.windows(), "key_upload_s": t_upload}), flush=True)
sys.exit(0 if same_H and same_proof else 1)
