"""Two FULL batches in flight on ONE GPU: two contexts (two stream sets, two scratch sets) sharing one device key, each
proving batch after batch from witnesses resident in HBM -- do the latency-bound tails of one step (fix-ups, bucket
reduction, assembly) hide under the throughput-bound kernels of the other?  Development aid.
Usage: python scripts/two_streams.py [batch=64] [steps=10]"""
import json, os, sys, threading, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import zklaim_b200 as zk
from zklaim_b200 import bridge

batch = int(sys.argv[1]) if len(sys.argv) > 1 else 64
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
saved = os.dup(1); os.dup2(2, 1)
cred = bridge.Credentials(1, seed=1)
ctxs = [cred.ctx, zk.Context(0), zk.Context(0)]
inp = []
for k in range(3):
    w = torch.from_numpy(cred.witnesses(batch, 1000 + 100 * k).view(np.int64)).cuda()
    r = torch.from_numpy(cred.random_fr(batch).view(np.int64)).cuda()
    s = torch.from_numpy(cred.random_fr(batch).view(np.int64)).cuda()
    inp.append((w, r, s))
torch.cuda.synchronize()

def run(k, n):
    w, r, s = inp[k]
    for i in range(n):
        ctxs[k].prove_batch_device(cred.key, w.data_ptr(), r.data_ptr(), s.data_ptr(), batch)

res = {}
for nctx in (1, 2, 3):
    for k in range(nctx): run(k, 3)
    torch.cuda.synchronize()
    t = time.perf_counter()
    th = [threading.Thread(target=run, args=(k, steps)) for k in range(nctx)]
    for x in th: x.start()
    for x in th: x.join()
    dt = time.perf_counter() - t
    res[f"{nctx}_contexts"] = {"proofs_per_s": round(nctx * steps * batch / dt, 1), "ms_per_batch": round(1e3 * dt / (nctx * steps), 3)}
os.dup2(saved, 1)
print(json.dumps(res))
