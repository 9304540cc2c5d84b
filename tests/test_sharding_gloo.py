"""Multi-GPU host logic on the CPU (world_size 2, gloo): the range partition of an MSM (libff's own chunking,
multiexp.tcc:402-441), the all_gather of one Jacobian partial per rank and the combine step.  The GPU is replaced
by the oracle here (tests may use it): each rank evaluates ITS range with oracle/zkoracle.c, so what is checked is
zklaim_b200/sharding.py -- the partition covers every item exactly once, the gather returns every rank's partial in
rank order on every rank, and the summed result equals the single-range MSM."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from zklaim_b200 import G1, G2, sharding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("total,world", [(0, 1), (1, 1), (5, 2), (10, 3), (3, 4), (1 << 20, 8), (29363, 8), (7, 8)])
def test_shard_range_partitions_like_libff(total, world):
    rs = [sharding.shard_range(total, r, world) for r in range(world)]
    covered = sorted((lo, hi) for lo, hi in rs if hi > lo)
    assert sum(hi - lo for lo, hi in covered) == total
    pos = 0
    for lo, hi in covered:
        assert lo == pos
        pos = hi
    if total >= world and world > 1:  # multiexp.tcc:417-431: total // chunks each, the last takes the remainder
        one = total // world
        assert rs[0] == (0, one) and rs[-1] == ((world - 1) * one, total)
    else:                             # multiexp.tcc:411-416: not split at all
        assert rs[0] == (0, total)


def test_gather_without_process_group_is_identity():
    p = np.arange(12, dtype=np.uint64)
    assert np.array_equal(sharding.gather_partials(p), p[None])


WORKER = r"""
import os, sys
import numpy as np
sys.path.insert(0, {root!r})
import torch.distributed as dist
from zklaim_b200 import G1, G2, sharding
from oracle import zkport

class OracleDevice:                      # stands in for zklaim_b200.Context on the CPU box
    def __init__(self, bases): self.bases = bases
    def msm(self, bases_local, scalars): return zkport.msm(self.g, bases_local, scalars)
    def group_sum(self, g, parts):
        acc = np.zeros(12 if g == G1 else 24, dtype=np.uint64)
        acc[4 if g == G1 else 8:8 if g == G1 else 12] = ONE_Q      # the zero (0, 1, 0); in G2 Y = (1, 0)
        for p in parts:
            acc = zkport.group_op(g, 0, acc[None], p[None])[0]
        return acc

ONE_Q = np.array([0xd35d438dc58f0d9d, 0x0a78eb28f5c70b3d, 0x666ea36f7879462c, 0x0e0a77c19a07df2f], dtype=np.uint64)
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:{port}", rank=int(sys.argv[1]), world_size={world})
rank, world = dist.get_rank(), dist.get_world_size()
d = np.load({fixture!r})
ok = True
for g, name, n in ((G1, "g1", {n}), (G2, "g2", {n2})):
    bases, scalars = d[name + "_msm_bases"][:n], d[name + "_msm_scalars"][:n]
    lo, hi = sharding.shard_range(n, rank, world)
    dev = OracleDevice(bases)
    dev.g = g
    total = sharding.sharded_msm(dev, g, bases[lo:hi], scalars[lo:hi])
    parts = sharding.gather_partials(zkport.msm(g, bases[lo:hi], scalars[lo:hi]) if hi > lo else dev.group_sum(g, []))
    assert parts.shape[0] == world
    full = zkport.msm(g, bases, scalars)
    norm = zkport.group_op(g, 4, total[None])[0]
    same = norm.tolist() == zkport.group_op(g, 4, full[None])[0].tolist()
    if n == len(d[name + "_msm_bases"]):    # the whole fixture: also the reference's own answer
        same = same and norm.tolist() == d[name + "_msm"].tolist()
    ok = ok and same
dist.barrier()
dist.destroy_process_group()
print("RANK", rank, "OK" if ok else "MISMATCH", flush=True)
sys.exit(0 if ok else 1)
"""


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world,n,n2", [(2, 96, 24), (2, 1, 1), (3, 37, 5)])
def test_sharded_msm_world_size_n_gloo(tmp_path, world, n, n2):
    """ranks evaluate their ranges with the oracle; gather over gloo; every rank must obtain the full MSM"""
    from oracle import zkport  # noqa: F401  (the C restatement must be built: make -C oracle port)
    fixture = os.path.join(ROOT, "tests", "golden", "groups.npz")
    assert os.path.exists(fixture), "run tests/golden/make_golden.py"
    script = tmp_path / "worker.py"
    script.write_text(WORKER.format(root=ROOT, port=_free_port(), world=world, fixture=fixture, n=n, n2=n2))
    procs = [subprocess.Popen([sys.executable, str(script), str(r)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(world)]
    outs = []
    for p in procs:
        try:
            out, _ = p.communicate(timeout=120)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        outs.append(out)
    for r, (p, out) in enumerate(zip(procs, outs)):
        assert p.returncode == 0 and f"RANK {r} OK" in out, out


SPLIT_WORKER = r"""
import ctypes, sys
import numpy as np
sys.path.insert(0, {root!r})
import torch.distributed as dist
from zklaim_b200 import sharding

M = 64
def view(ptr, n):                        # the "device" buffers are host tensors under gloo
    return np.ctypeslib.as_array(ctypes.cast(ptr, ctypes.POINTER(ctypes.c_uint64)), shape=(n,))

class FakeKey:
    def domain(self): return (M, 1)

class FakeDevice:                        # records the data flow of zkb_prove_split_* instead of doing arithmetic
    def __init__(self, rank): self.rank = rank; self.checked = False
    def prove_split_begin(self, key, w, r, s, which, check, d_part=0):
        self.checked = check
        assert (which >= 0) == (d_part != 0)
        if which >= 0: view(d_part, M * 4)[:] = 1000 * (which + 1) + np.arange(M * 4, dtype=np.uint64)
    def prove_split_chain(self, key, which, d_part):
        assert self.rank == 0 and which == 2  # two ranks: rank 0 also takes C
        view(d_part, M * 4)[:] = 1000 * (which + 1) + np.arange(M * 4, dtype=np.uint64)
    def prove_split_combine(self, key, d_abc, d_H):
        abc = view(d_abc, 3 * M * 4).reshape(3, M * 4)
        for k in range(3): assert abc[k][0] == 1000 * (k + 1) and abc[k][5] == 1000 * (k + 1) + 5, abc[k][:8]
        view(d_H, M * 4)[:] = abc[0] + 2 * abc[1] + 3 * abc[2]
    def prove_split_end(self, key, d_H):
        out = np.zeros(60, dtype=np.uint64)
        out[0], out[1] = self.rank, view(d_H, M * 4).sum()
        return out
    def prove_combine(self, parts, r, s): return parts
    def prove_partial(self, key, w, r, s): raise AssertionError("the split form was expected")

dist.init_process_group("gloo", init_method="tcp://127.0.0.1:{port}", rank=int(sys.argv[1]), world_size={world})
rank, world = dist.get_rank(), dist.get_world_size()
dev = FakeDevice(rank)
z = np.zeros((4, 4), dtype=np.uint64)
parts = sharding.sharded_prove_split(dev, FakeKey(), z, z[0], z[0])
want_H = sum((k + 1) * (1000 * (k + 1) * M * 4 + (M * 4) * (M * 4 - 1) // 2) for k in range(3))
ok = parts.shape == (world, 60) and [int(x) for x in parts[:, 0]] == list(range(world)) and all(int(x) == want_H for x in parts[:, 1])
ok = ok and dev.checked == (rank == world - 1)
dist.barrier()
dist.destroy_process_group()
print("RANK", rank, "OK" if ok else "MISMATCH", flush=True)
sys.exit(0 if ok else 1)
"""


@pytest.mark.parametrize("world", [2, 3, 4])
def test_split_witness_map_orchestration_gloo(tmp_path, world):
    """sharding.sharded_prove_split over gloo with a device that records the data flow: ranks 0 / 1 / 2 produce the three
    coset vectors (two ranks: rank 0 produces A and C), rank 0 receives the others in order and combines, every rank receives the same H and contributes its
    share in rank order; only the last rank is asked to check satisfiability."""
    script = tmp_path / "split_worker.py"
    script.write_text(SPLIT_WORKER.format(root=ROOT, port=_free_port(), world=world))
    procs = [subprocess.Popen([sys.executable, str(script), str(r)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(world)]
    outs = []
    for p in procs:
        try:
            out, _ = p.communicate(timeout=120)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        outs.append(out)
    for r, (p, out) in enumerate(zip(procs, outs)):
        assert p.returncode == 0 and f"RANK {r} OK" in out, out
