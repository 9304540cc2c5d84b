"""Range sharding of a multi-scalar multiplication over the GPUs of one node.

libff::multi_exp already has this shape on the CPU (multiexp.tcc:402-441): the point / scalar vectors are cut into
`chunks` contiguous ranges of `total / chunks` items (the last one takes the remainder), every range is evaluated
independently (`#pragma omp parallel for`), and the per-range results are added up (:433-438).  Here a range is
a GPU: one process per GPU (torch.distributed), every rank uploads the bases of ITS range only and evaluates it
with libzkb.so; the only exchange is an all_gather of one Jacobian point per rank (96 bytes in G1, 192 in G2),
after which every rank holds all partials and the sum is n - 1 point additions (zkb_group_sum).  NCCL's reduce
operators do not apply -- the group law is not a limb-wise sum -- so the collective is a gather, and the data
path itself needs none: this is the "tiny reduction of partial sums" of the north star.

No arithmetic lives in this file; without libzkb.so / a GPU the combine step raises like everything else.
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import numpy as np

from . import G1, G2


def shard_range(total: int, rank: int, world: int) -> Tuple[int, int]:
    """[lo, hi) of rank `rank`: libff's chunking (multiexp.tcc:417-431) -- `total // world` items each, the last
    rank also takes the remainder; with fewer items than ranks rank 0 takes everything (:411-416)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank / world size")
    if total < world or world == 1:
        return (0, total) if rank == 0 else (0, 0)
    one = total // world
    lo = rank * one
    hi = total if rank == world - 1 else lo + one
    return lo, hi


def gather_partials(partial: np.ndarray, group=None, device: Optional[str] = None) -> np.ndarray:
    """all_gather of one Jacobian point per rank -> (world, 12 | 24) uint64, identical on every rank.
    `group` is a torch.distributed process group (None = the default one; no process group = world of one).
    With the nccl backend the 96 / 192 bytes travel through a device tensor on `device`."""
    import torch
    import torch.distributed as dist

    p = np.ascontiguousarray(partial, dtype=np.uint64)
    if not (dist.is_available() and dist.is_initialized()):
        return p[None].copy()
    world = dist.get_world_size(group)
    backend = dist.get_backend(group)
    dev = torch.device(device if device is not None else ("cuda" if backend == "nccl" else "cpu"))
    mine = torch.from_numpy(p.view(np.int64).copy()).to(dev)
    flat = mine.reshape(-1)
    out = torch.empty(world * flat.numel(), dtype=mine.dtype, device=dev)
    dist.all_gather_into_tensor(out, flat, group=group)  # one tensor, one read-back (world small copies cost 20 us each)
    return out.cpu().numpy().view(np.uint64).reshape((world,) + tuple(mine.shape))


def sharded_msm(ctx, group_id: int, bases_local, scalars_local: np.ndarray, group=None, device: Optional[str] = None,
                combine: Optional[Callable[[int, np.ndarray], np.ndarray]] = None) -> np.ndarray:
    """sum_i s_i P_i over ALL ranks' ranges.  `bases_local` = this rank's zkb Bases (uploaded from its
    shard_range), `scalars_local` the matching scalars.  Returns the Jacobian limbs of the total on every rank.
    `combine(group_id, partials)` defaults to the GPU (ctx.group_sum)."""
    if group_id not in (G1, G2):
        raise ValueError("group_id must be G1 or G2")
    if hasattr(scalars_local, "data_ptr"):  # a CUDA tensor of n x 4 int64 / uint64 limbs already on this rank's GPU
        n = int(scalars_local.shape[0])
        part = ctx.msm_device(bases_local, scalars_local.data_ptr(), n) if n else None
    else:
        part = ctx.msm(bases_local, scalars_local) if len(scalars_local) else None
    if part is None:  # an empty range contributes the zero of the group, (0, 1, 0) in Montgomery form (alt_bn128_init.cpp:145-147)
        part = ctx.group_sum(group_id, np.zeros((0, 12 if group_id == G1 else 24), dtype=np.uint64))
    parts = gather_partials(part, group, device)
    return (combine or ctx.group_sum)(group_id, parts)


def sharded_prove(ctx, key_shard, w: np.ndarray, r: np.ndarray, s: np.ndarray, group=None, device: Optional[str] = None) -> np.ndarray:
    """ONE Groth16 proof over all ranks: `key_shard` = ctx.upload_key(key, rank, world) (this rank's range of the A, B, H
    and L queries; rank 0 also alpha, beta, delta).  Every rank runs the witness map itself (replicas: a distributed NTT
    would need an all-to-all of the whole vector per transform) and the MSMs over its ranges; the exchange is an
    all_gather of one 480-byte share per rank; zkb_prove_combine adds the shares and does the assembly.  Returns the
    proof (48 uint64: A, B, C Jacobian) on every rank."""
    part = ctx.prove_partial(key_shard, w, r, s)
    parts = gather_partials(part, group, device)
    return ctx.prove_combine(parts, r, s)


def sharded_prove_split(ctx, key_shard, w: np.ndarray, r: np.ndarray, s: np.ndarray, group=None, device: Optional[str] = None,
                        trace: Optional[dict] = None) -> np.ndarray:
    """ONE Groth16 proof over all ranks with the witness map SPLIT by polynomial instead of replicated (needs >= 2 ranks;
    one rank falls back to sharded_prove).  r1cs_to_qap_witness_map treats A, B and C alike (r1cs_to_qap.tcc:240-293): ranks
    0 / 1 / 2 run SpMV + iFFT + coset FFT of one polynomial each (zkb_prove_split_begin; with two ranks, rank 0 takes A and
    C -- zkb_prove_split_chain -- and rank 1 takes B), ranks 1 and 2 send their m x 32
    bytes to rank 0 (NCCL send / recv), rank 0 divides by Z and transforms back (zkb_prove_split_combine), broadcasts the H
    coefficients, and every rank computes its share of the H + L sum (zkb_prove_split_end); the witness-only sums run on
    side streams from the start.  The last rank also checks satisfiability.  Exchanges: 2 x 32 m bytes in, 32 m bytes out.
    On GPUs the witness itself travels the same way: every rank copies ONE range of it over PCIe and an all_gather over
    NVLink completes it on every GPU (the full copy per rank costs 124 MB / PCIe rate for the 2^22 circuit).
    `trace` (optional dict) receives this rank's wall time per phase in ms (the syncs between phases are there anyway)."""
    import time

    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) < 2:
        return sharded_prove(ctx, key_shard, w, r, s, group, device)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dev = torch.device(device if device is not None else ("cuda" if dist.get_backend(group) == "nccl" else "cpu"))

    def sync():  # the library works on its own streams and returns with its results complete; torch's stream carries NCCL
        if dev.type == "cuda":
            torch.cuda.synchronize(dev)

    m = key_shard.domain()[0]
    which = rank if rank < 3 else -1
    senders = [1, 2] if world >= 3 else [1]  # ranks whose polynomial travels to rank 0
    abc = torch.empty((3, m, 4), dtype=torch.int64, device=dev) if rank == 0 else None
    part = abc[0] if rank == 0 else (torch.empty((m, 4), dtype=torch.int64, device=dev) if which >= 0 else None)
    H = torch.empty((m, 4), dtype=torch.int64, device=dev)
    t_last = [time.perf_counter()]

    def mark(name):
        if trace is not None:
            now = time.perf_counter()
            trace[name] = trace.get(name, 0.0) + 1e3 * (now - t_last[0])
            t_last[0] = now

    w_arg = w
    if dev.type == "cuda":
        w2 = np.ascontiguousarray(w, dtype=np.uint64).reshape(-1, 4)
        nv = w2.shape[0]
        per = -(-nv // world)
        W = torch.empty((per * world, 4), dtype=torch.int64, device=dev)
        lo, hi = min(rank * per, nv), min((rank + 1) * per, nv)
        mine = W[rank * per:(rank + 1) * per]
        if hi > lo:
            mine[:hi - lo].copy_(torch.from_numpy(w2[lo:hi].view(np.int64)), non_blocking=True)
        dist.all_gather_into_tensor(W, mine, group=group)
        w_arg = W.data_ptr()
    sync()
    mark("witness: one range per rank over PCIe, all_gather over NVLink")
    ctx.prove_split_begin(key_shard, w_arg, r, s, which, check=(rank == world - 1), d_part=part.data_ptr() if part is not None else 0)
    if rank == 0 and world == 2:
        ctx.prove_split_chain(key_shard, 2, abc[2].data_ptr())
    mark("begin: own polynomial to the coset (ranks 0-2)")
    if rank == 0:
        for q in dist.batch_isend_irecv([dist.P2POp(dist.irecv, abc[k], k, group) for k in senders]):
            q.wait()
        sync()
        mark("recv 2 x 32 m bytes")
        ctx.prove_split_combine(key_shard, abc.data_ptr(), H.data_ptr())
        mark("combine: (A B - C) / Z, coset back to coefficients")
    elif rank in senders:
        for q in dist.batch_isend_irecv([dist.P2POp(dist.isend, part, 0, group)]):
            q.wait()
    dist.broadcast(H, src=0, group=group)
    sync()
    mark("broadcast of H (32 m bytes)")
    share = ctx.prove_split_end(key_shard, H.data_ptr())
    mark("end: H + L sums over this rank's ranges, witness-only sums joined")
    parts = gather_partials(share, group, device)
    mark("all_gather of the 480-byte shares")
    out = ctx.prove_combine(parts, r, s)
    mark("assembly")
    return out
