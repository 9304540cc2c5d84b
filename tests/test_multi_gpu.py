"""Two real GPUs, NCCL (runs only where torch sees >= 2 CUDA devices: `gpurun --gpus 2`): the MSM range-sharding
path of zklaim_b200/sharding.py end to end -- each rank uploads its range to ITS GPU, zkb_msm there, all_gather of
the 96-byte partials over NCCL, zkb_group_sum -- against the reference's multi_exp on the whole input."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r"""
import os, sys
import numpy as np
sys.path.insert(0, {root!r})
import torch, torch.distributed as dist
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
import zklaim_b200 as zk
from zklaim_b200 import sharding
from oracle import zkref
ctx = zk.Context(rank)
ok = True
for group, n in ((zk.G1, 20011), (zk.G2, 1500)):
    bases = zkref.bases(group, 1 << 40, n)
    s = zkref.sha512_rng(9 << 32, n)
    lo, hi = sharding.shard_range(n, rank, world)
    B = ctx.upload_bases(group, bases[lo:hi])
    total = sharding.sharded_msm(ctx, group, B, s[lo:hi], device=f"cuda:{{rank}}")
    exp, _ = zkref.msm(group, bases, s)
    ok = ok and np.array_equal(zkref.to_affine_jac(group, total[None]), zkref.to_affine_jac(group, exp[None]))
    B.free()
dist.barrier()
dist.destroy_process_group()
print("RANK", rank, "OK" if ok else "MISMATCH", flush=True)
sys.exit(0 if ok else 1)
"""


def test_sharded_msm_two_gpus_nccl(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    script = tmp_path / "worker.py"
    script.write_text(WORKER.format(root=ROOT))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29531", str(script)]
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert p.returncode == 0 and "RANK 0 OK" in p.stdout and "RANK 1 OK" in p.stdout, p.stdout[-4000:]
