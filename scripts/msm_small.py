"""Small dense G1 MSMs (2^14 .. 2^18), synthetic bases on the GPU: device / wall ms and the plan the library chose.
Development aid for the latency regime.  Usage: python scripts/msm_small.py [logs] [window_bits]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import zklaim_b200 as zk
logs = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [14, 15, 16, 17, 18]
ctx = zk.Context(0)
for lg in logs:
    n = 1 << lg
    B = ctx.synthetic_bases(n, seed=5000)
    sc = torch.randint(-(1 << 63), (1 << 63) - 1, (n, 4), dtype=torch.int64, device="cuda")
    sc[:, 3] &= (1 << 60) - 1
    best_d, best_w = 1e9, 1e9
    for it in range(6):
        torch.cuda.synchronize()
        t = time.perf_counter()
        ctx.msm_device(B, sc.data_ptr(), n)
        best_w = min(best_w, 1e3 * (time.perf_counter() - t))
        best_d = min(best_d, ctx.last_device_ms())
    print(json.dumps({"log2n": lg, "ms_device": round(best_d, 4), "ms_wall": round(best_w, 4), "launches": ctx.kernel_launches()}))
    B.free()
