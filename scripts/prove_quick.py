"""Quick prover timing on the GPU (development aid): zklaim 1-payload circuit, single proof and batch."""
import sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import zklaim_b200 as zk
from oracle import zkref

batch = int(sys.argv[1]) if len(sys.argv) > 1 else 64
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 5
ctx = zk.Context(0)
t = time.time()
circ = zkref.Circuit.zklaim(1, 1, mt=True)
circ.setup()
kd = circ.export_key()
print(f"setup {time.time()-t:.1f}s", flush=True)
t = time.time()
key = ctx.upload_key(kd)
print(f"key upload {time.time()-t:.2f}s, {key.device_bytes()/1e6:.0f} MB on device", flush=True)
d = circ.dims()
w1 = circ.witness()
if os.environ.get("ZKB_TUNE", "1") == "1":
    print("windows before tune", key.windows(), flush=True)
    key.tune(w1)
    print(f"windows after tune {key.windows()}, {key.device_bytes()/1e6:.0f} MB on device", flush=True)
ws = np.stack([w1] * batch)
r = zkref.sha512_rng(5000, batch)
s = zkref.sha512_rng(6000, batch)
for it in range(iters):
    t = time.time()
    ctx.prove(key, w1, r[0], s[0])
    wall = time.time() - t
    print(f"single: device {ctx.last_device_ms():.3f} ms wall {wall*1e3:.3f} ms", flush=True)
for it in range(iters):
    t = time.time()
    ctx.prove_batch(key, ws, r, s)
    wall = time.time() - t
    print(f"batch {batch}: device {ctx.last_device_ms():.3f} ms wall {wall*1e3:.3f} ms -> {batch/wall:.0f} proofs/s e2e, {batch/ctx.last_device_ms()*1e3:.0f} proofs/s device", flush=True)
