"""End to end through the zklaim-facing entry points (zkb_libsnark_prove_batch: host witness generation by zklaim's own
gadget code + GPU prove) and the reference's libsnark_verify.  Development aid."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from zklaim_b200 import bridge
n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
saved = os.dup(1); os.dup2(2, 1)
c = bridge.Credentials(1, seed=1)
res = []
for it in range(3):
    ok, tp, tv = c.roundtrip(n, 100 + 1000 * it)
    res.append((ok, tp, tv))
os.dup2(saved, 1)
for ok, tp, tv in res:
    print(f"{n} credentials: accepted {ok}, prove (witness + GPU) {tp*1e3:.1f} ms = {n/tp:.0f} proofs/s, reference verify {tv*1e3:.1f} ms", flush=True)
