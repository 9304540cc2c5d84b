#!/usr/bin/env python
"""bench.py -- the driver's measurement contract for the zklaim Groth16 prover backend.

Workload (BASELINE.json configs[1], "zklaim main_benchmark"): a batch of 64 single-payload zklaim credential
proofs (SHA-256 commitment + range predicate circuit: 29 356 constraints, 26 891 variables, domain 2^15) under
one proving key, one proof stream per GPU.  A step = one batch of 64 proofs on every rank (weak scaling).

  value : proofs/s, witnesses already resident in HBM (zkb_prove_batch_device), whole job
  e2e   : proofs/s through the C ABI with HOST buffers (zkb_prove_batch: pinned witnesses in, proofs out)
  roofline     : the dominant kernel (G1 bucket accumulation of the H + L query, k_accumulate<Fq>) against the
                 measured integer multiply-add (IMAD.WIDE) rate of the device -- this path is integer work, neither
                 HBM- nor tensor-bound; its HBM fraction is reported beside it
  cpu_baseline : the reference's own libsnark prover (oracle/_ref, -DMULTICORE -DUSE_ASM) on the host cores

The same JSON line also carries the other BASELINE.json configurations (device times, CUDA events on the library's
stream; parity of each is in tests/test_gpu_large.py):
  msm_g1       : config #3, G1 MSM at 2^20 / 2^22 / 2^24 random points and dense scalars (points/s, fraction of the IMAD roofline)
  witness_map  : config #4, r1cs_to_qap_witness_map at 2^18 / 2^20 / 2^22 on a synthetic R1CS
  s22          : config #5, one full Groth16 proof of 153 chained SHA-256 gadgets (4.17 M constraints, domain 2^22)
  sharded      : (N > 1) the strong-scaling curve: one 2^24 MSM and one S22 proof over all N GPUs (zklaim_b200/sharding.py)

`--impl reference` times the reference CPU prover instead (same workload and config, bounded sample per step).
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

BATCH = 64
# proof streams per GPU: independent contexts (own CUDA streams and scratch) that share the device key, one batch in flight
# each -- the latency-bound tail of one step (fix-ups, bucket reduction, assembly) runs under the throughput-bound kernels
# of another (measured, ms per batch of 64 with 1 / 2 / 3 / 4 / 5 / 6 streams: 12.28 / 11.53 / 11.43 / 11.37 / 11.33 / 11.31)
STREAMS = max(1, int(os.environ.get("ZKB_BENCH_STREAMS", "4")))
REF_SAMPLE = 8  # proofs per step of the CPU arm: a bounded sample of the 64-proof batch (the metric is a per-proof rate)
WORKLOAD = "zklaim main_benchmark: batch of 64 single-payload credential proofs (Groth16, alt_bn128, 29356 constraints, domain 2^15)"
# one Fq multiplication = 136 IMAD; a mixed addition in the reference's formulas = 11 Fq-mul
IMAD_PER_MUL = 136
MUL_PER_MADD_REF = 11   # madd-2007-bl as libff uses it (alt_bn128_g1.cpp:256-323)
MUL_PER_MADD_XYZZ = 10  # madd-2008-s as the kernel executes it (ec.cuh): squares are general products (fp.cuh Fp::sqr)
IMAD_PER_MADD_EXECUTED = MUL_PER_MADD_XYZZ * IMAD_PER_MUL  # = 1360
BYTES_PER_PAIR = 64 + 8  # one affine base (two 32-byte sectors) + one (entry, bucket) pair of the sorted list


class ClockSampler(threading.Thread):
    """samples SM clock and throttle reasons of one GPU while the timed region runs (pynvml)"""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self.stop_flag = index, [], set(), None, False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if not self.nv:
            return
        nv = self.nv
        names = {
            nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
            nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
        }
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.02)

    def result(self):
        self.stop_flag = True
        if self.is_alive():
            self.join(1.0)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback"


class stdout_to_stderr:
    """libsnark prints progress to fd 1; the JSON line must stay alone on stdout"""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        os.dup2(2, 1)

    def __exit__(self, *a):
        sys.stdout.flush()
        os.dup2(self.saved, 1)
        os.close(self.saved)


def common_config(world):
    """the `config` object: identical in both arms (the driver compares them)"""
    return {"workload": WORKLOAD, "batch_per_gpu": BATCH,
            "parallelism": f"{STREAMS} proof streams per GPU (contexts sharing one device key, a batch of {BATCH} each in flight) x{world} GPUs",
            "l2": "2 input sets per stream alternate; key + scratch working set (> 1 GB) exceeds the 126 MB L2, no flush",
            "inputs": "seeded synthetic credentials (salts), witnesses generated by zklaim's gadget code on the host",
            "reference_arm": f"a step of --impl reference proves {REF_SAMPLE} credentials of this workload on the host CPU (bounded sample; "
                             "the metric is a per-proof rate)"}


def reference_prover(sample):
    """the reference's CPU prover (oracle/_ref/libzkref_mt.so: libsnark built -DMULTICORE -DUSE_ASM, all host
    threads) on `sample` proofs of the workload.  Returns (callable running one sample -> seconds, cores, kind)"""
    from oracle import zkref
    import numpy as np
    mt = zkref.available(mt=True)
    if not (mt or zkref.available()):
        raise RuntimeError("oracle/_ref is not built")
    with stdout_to_stderr():
        circ = zkref.Circuit.zklaim(1, 1, mt=mt)
        circ.setup()
        nv = circ.dims()["num_variables"]
        ws = [circ.witness()]
        if zkref.available():
            ws += [zkref.zklaim_witness(1, 2000 + i, nv) for i in range(min(sample, 4) - 1)]
        cores = int(zkref.lib(mt).zkref_num_threads())

    def run():
        tot = 0.0
        with stdout_to_stderr():
            for i in range(sample):
                tot += circ.prove_stock(ws[i % len(ws)])
        return tot

    return run, cores, "reference"


def _side_by_side_worker(n_proofs, start_evt, q):
    """one single-thread libsnark prover (its own key): part of the throughput-style CPU baseline"""
    os.environ["OMP_NUM_THREADS"] = "1"
    os.dup2(2, 1)
    from oracle import zkref
    mt = zkref.available(mt=True)
    circ = zkref.Circuit.zklaim(1, 1, mt=mt)
    circ.setup()
    w = circ.witness()
    circ.prove_stock(w)  # warm
    q.put("ready")
    start_evt.wait()
    t0 = time.perf_counter()
    for _ in range(n_proofs):
        circ.prove_stock(w)
    q.put(time.perf_counter() - t0)


def reference_throughput(per_worker=2):
    """nproc single-thread provers side by side (the shape of zklaim/main_benchmark.c: independent credentials): proofs/s
    of the whole socket.  Returns (proofs/s, workers) or None."""
    import multiprocessing as mp
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    ctx = mp.get_context("spawn")
    evt, q = ctx.Event(), ctx.Queue()
    procs = [ctx.Process(target=_side_by_side_worker, args=(per_worker, evt, q), daemon=True) for _ in range(n)]
    try:
        for p in procs:
            p.start()
        for _ in procs:
            if q.get(timeout=240) != "ready":
                return None
        t0 = time.perf_counter()
        evt.set()
        for _ in procs:
            q.get(timeout=240)
        dt = time.perf_counter() - t0
        return n * per_worker / dt, n
    except Exception:
        return None
    finally:
        for p in procs:
            p.join(timeout=5)
            if p.is_alive():
                p.terminate()


def run_reference(args):
    rank, _, world = dist_env()
    if rank != 0:
        return
    sample = REF_SAMPLE
    run, cores, kind = reference_prover(sample)
    for _ in range(args.warmup):
        run()
    t = [run() for _ in range(args.steps)]
    total = sum(t)
    value = sample * args.steps / total
    thr = None if args.no_cpu_throughput else reference_throughput()
    line = {
        "impl": "reference", "metric": "groth16_proofs_per_s", "value": value, "unit": "proofs/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u32x8 Montgomery (254-bit integer)", "data": "synthetic",
        "config": common_config(world),
        "cpu_baseline": {"value": value, "unit": "proofs/s", "cores": cores, "kind": kind,
                         "sample": f"{sample * args.steps} proofs of the workload ({sample} per step), libsnark r1cs_gg_ppzksnark_prover "
                                   f"(-DMULTICORE -DUSE_ASM), one proof at a time on OpenMP {cores} threads",
                         "throughput_value": thr[0] if thr else None,
                         "throughput_note": (f"{thr[1]} single-thread provers side by side, 2 proofs each" if thr else "not measured")},
        "e2e": {"value": value, "unit": "proofs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "prove_ms": 1e3 * total / (sample * args.steps),
    }
    print(json.dumps(line), flush=True)


IMAD_WORK_MSM = {16: 2.318e9, 18: 7.879e9, 20: 2.738e10, 22: 9.840e10, 24: 3.559e11}  # G1 MSM IMAD totals, SURVEY.md section 8(d)


def _rand_fr_host(np, rng, n):
    a = rng.integers(0, 1 << 64, size=(n, 4), dtype=np.uint64)
    a[:, 3] &= np.uint64((1 << 60) - 1)  # < 2^252 < r: a canonical residue, taken as Montgomery form
    return a


def extra_msm(ctx, torch, imad_peak, logs=(20, 22, 24)):
    """config #3: G1 MSM over random points (made on the device) and dense random scalars resident in HBM"""
    out = {}
    for lg in logs:
        n = 1 << lg
        B = ctx.synthetic_bases(n, seed=1000 + lg)
        sc = torch.randint(-(1 << 63), (1 << 63) - 1, (n, 4), dtype=torch.int64, device="cuda")
        sc[:, 3] &= (1 << 60) - 1
        best_dev, best_wall = 1e9, 1e9
        for it in range(4):
            torch.cuda.synchronize()
            t = time.perf_counter()
            ctx.msm_device(B, sc.data_ptr(), n)
            best_wall = min(best_wall, 1e3 * (time.perf_counter() - t))
            best_dev = min(best_dev, ctx.last_device_ms())
        out[str(lg)] = {"ms": best_dev, "ms_wall": best_wall, "mpts_per_s": n / best_dev / 1e3,
                        "imad_frac": IMAD_WORK_MSM[lg] / (best_dev * 1e-3) / imad_peak}
        B.free()
        del sc
    out["note"] = "device ms = CUDA events around the whole MSM (recode, sort, accumulate, reduce); imad_frac = SURVEY 8(d) work table / time / measured IMAD peak"
    return out


def synthetic_r1cs(np, zk, ctx, m, seed):
    """m - 2 constraints (x_i + x_{i+1}) * y_i = z_i over 3 (m - 2) + 1 variables, one public input: domain size m"""
    nc = m - 2
    rng = np.random.default_rng(seed)
    x, y = _rand_fr_host(np, rng, nc + 1), _rand_fr_host(np, rng, nc)
    z = ctx.field_op(zk.FR, zk.OP_MUL, ctx.field_op(zk.FR, zk.OP_ADD, x[:-1], x[1:]), y)
    w = np.concatenate([x, y, z])
    one = np.array([0xac96341c4ffffffb, 0x36fc76959f60cd29, 0x666ea36f7879462e, 0x0e0a77c19a07df2f], dtype=np.uint64)
    i = np.arange(nc, dtype=np.uint32)
    colA = np.stack([1 + i, 2 + i], axis=1).reshape(-1)
    cs = {"num_constraints": nc, "num_variables": 3 * nc + 1, "num_inputs": 1,
          "csr_A": (2 * np.arange(nc + 1, dtype=np.uint64), colA, np.tile(one, (2 * nc, 1))),
          "csr_B": (np.arange(nc + 1, dtype=np.uint64), (nc + 2 + i).astype(np.uint32), np.tile(one, (nc, 1))),
          "csr_C": (np.arange(nc + 1, dtype=np.uint64), (2 * nc + 2 + i).astype(np.uint32), np.tile(one, (nc, 1)))}
    return cs, w


def extra_witness_map(ctx, np, zk, imad_peak, logs=(18, 20, 22)):
    """config #4: r1cs_to_qap_witness_map (3 iFFT, 3 coset FFT, pointwise (A B - C) / Z, 1 inverse coset FFT + SpMV)"""
    out = {}
    for lg in logs:
        m = 1 << lg
        cs, w = synthetic_r1cs(np, zk, ctx, m, lg)
        key = ctx.upload_key(cs)  # constraint system only
        assert key.domain() == (m, zk.DOMAIN_BASIC_RADIX2)
        best = 1e9
        for it in range(3):
            ctx.witness_map(key, w)  # raises if the witness were unsatisfying
            best = min(best, ctx.last_device_ms())
        out[str(lg)] = {"ms": best, "imad_frac": (3.5 * m * lg + 9 * m) * 136 / (best * 1e-3) / imad_peak}
        key.free()
    out["note"] = "device ms; work model SURVEY 8(d): (3.5 m log2 m + 9 m) Fr products of 136 multiply-adds"
    return out


def s22_inputs(np, bridge, toxic=None):
    cs, w = bridge.sha_chain(153, seed=7)
    g1, g2 = bridge.generators()
    if toxic is None:
        toxic = bridge.random_fr(5)
    rs = bridge.random_fr(2)
    return cs, w, toxic, g1, g2, rs


def extra_s22(ctx, np, bridge):
    """config #5: 153 chained SHA-256 gadgets (4 173 840 constraints, domain 2^22): key by zkb_setup, one full proof"""
    t0 = time.perf_counter()
    cs, w, toxic, g1, g2, rs = s22_inputs(np, bridge)
    t_build = time.perf_counter() - t0
    t0 = time.perf_counter()
    kd = ctx.setup(cs, toxic, g1, g2)
    t_setup = time.perf_counter() - t0
    key = ctx.upload_key(kd)
    key.tune(w)
    dev, wall = [], []
    for it in range(4):
        t = time.perf_counter()
        ctx.prove(key, w, rs[0], rs[1])
        wall.append(1e3 * (time.perf_counter() - t))
        dev.append(ctx.last_device_ms())
    res = {"constraints": cs["num_constraints"], "domain": 1 << 22, "prove_ms_device": min(dev[1:]), "prove_ms_wall": min(wall[1:]),
           "setup_s_gpu": t_setup, "circuit_build_s_host": t_build, "key_device_gb": key.device_bytes() / 1e9, "windows": key.windows(),
           "note": "wall = zkb_prove with the 124 MB witness in pageable host memory"}
    key.free()
    return res


def extra_sharded(ctx, np, torch, dist, zk, bridge, rank, world, local, imad_peak):
    """N > 1: the strong-scaling curve of ONE job over all ranks (zklaim_b200/sharding.py: ranges per rank, an all_gather
    of one partial per rank, no other collective)"""
    from zklaim_b200 import sharding
    out = {}

    def tmax(x):
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # one G1 MSM of 2^20 / 2^24 points, point ranges split per rank
    for lg in (20, 24):
        lo, hi = sharding.shard_range(1 << lg, rank, world)
        B = ctx.synthetic_bases(hi - lo, seed=5000 + rank)
        sc = torch.randint(-(1 << 63), (1 << 63) - 1, (hi - lo, 4), dtype=torch.int64, device="cuda")
        sc[:, 3] &= (1 << 60) - 1
        best, best_dev = 1e9, 1e9
        for it in range(5):
            dist.barrier()
            torch.cuda.synchronize()
            t = time.perf_counter()
            sharding.sharded_msm(ctx, zk.G1, B, sc, device=f"cuda:{local}")
            best = min(best, tmax(1e3 * (time.perf_counter() - t)))
            best_dev = min(best_dev, tmax(ctx.last_device_ms()))
        out[f"msm_2^{lg}"] = {"ms_wall": best, "ms_device": best_dev, "mpts_per_s": (1 << lg) / best / 1e3,
                              "imad_frac_per_gpu": IMAD_WORK_MSM[lg] / (best_dev * 1e-3) / imad_peak / world}
        B.free()
        del sc
    # one S22 proof: every rank holds one range of the four queries (rank 0 also alpha, beta, delta)
    toxic = torch.from_numpy(bridge.random_fr(7).view(np.int64)).cuda()
    dist.broadcast(toxic, 0)
    rnd = toxic.cpu().numpy().view(np.uint64)
    cs, w, _, g1, g2, _ = s22_inputs(np, bridge, toxic=rnd[:5])
    kd = ctx.setup(cs, rnd[:5], g1, g2)
    key = ctx.upload_key(kd, rank, world)
    key.tune(w)
    del kd
    wp = torch.from_numpy(w.view(np.int64)).pin_memory().numpy().view(np.uint64)  # pinned: the 124 MB travel at PCIe speed
    def norm(p):  # (A, B, C) as affine points: two runs may differ in the Jacobian representation (order of additions)
        p = np.ascontiguousarray(p).reshape(48)
        return np.concatenate([ctx.group_op(zk.G1, zk.GOP_NORMALIZE, p[0:12][None])[0], ctx.group_op(zk.G2, zk.GOP_NORMALIZE, p[12:36][None])[0],
                               ctx.group_op(zk.G1, zk.GOP_NORMALIZE, p[36:48][None])[0]])

    def one_proof(fn):
        best, best_dev, proof = 1e9, 1e9, None
        for it in range(4):
            dist.barrier()
            torch.cuda.synchronize()
            t = time.perf_counter()
            proof = fn()
            best = min(best, tmax(1e3 * (time.perf_counter() - t)))
            best_dev = min(best_dev, tmax(ctx.last_device_ms()))
        mine = torch.from_numpy(proof.view(np.int64).copy()).cuda()
        ref0 = mine.clone()
        dist.broadcast(ref0, 0)
        return best, best_dev, proof, bool(torch.equal(mine, ref0))

    wall_r, dev_r, proof_r, same_r = one_proof(lambda: sharding.sharded_prove(ctx, key, wp, rnd[5], rnd[6], device=f"cuda:{local}"))
    out["s22_one_proof_replicated_witness_map"] = {"ms_device": dev_r, "ms_wall": wall_r, "same_proof_on_every_rank": same_r}
    wall_s, dev_s, proof_s, same_s = one_proof(lambda: sharding.sharded_prove_split(ctx, key, wp, rnd[5], rnd[6], device=f"cuda:{local}"))
    phases = {}
    dist.barrier()
    torch.cuda.synchronize()
    sharding.sharded_prove_split(ctx, key, wp, rnd[5], rnd[6], device=f"cuda:{local}", trace=phases)
    out["s22_one_proof"] = {"ms_device": dev_s, "ms_wall": wall_s, "same_proof_on_every_rank": same_s,
                            "equals_replicated_form": bool(np.array_equal(norm(proof_r), norm(proof_s))),
                            "key_shard_device_gb": key.device_bytes() / 1e9,
                            "phases_ms_wall_rank0": {k: round(v, 3) for k, v in phases.items()},
                            "form": "witness map split by polynomial over ranks 0-2 (two ranks: rank 0 takes A and C), H broadcast; "
                                    "ms_device = max over ranks of the library's CUDA-event time (NCCL exchanges are in ms_wall only)"}
    key.free()
    return out


def run_ours(args):
    import numpy as np
    import torch
    rank, local, world = dist_env()
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    os.environ["ZKB_DEVICE"] = str(local)
    import zklaim_b200 as zk
    from zklaim_b200 import bridge

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- inputs: issuer setup + host-side witness generation through zklaim's own code (the bridge) ----
    with stdout_to_stderr():
        cred = bridge.Credentials(1, seed=1 + rank)
        ctx, key, nv = cred.ctx, cred.key, cred.num_variables
        nsets = 2 * STREAMS  # stream j alternates between the sets 2 j and 2 j + 1
        w_host = [torch.from_numpy(cred.witnesses(BATCH, 1000 + 100 * k + 10000 * rank).view(np.int64)).pin_memory() for k in range(nsets)]
        r_host = [torch.from_numpy(cred.random_fr(BATCH).view(np.int64)).pin_memory() for _ in range(nsets)]
        s_host = [torch.from_numpy(cred.random_fr(BATCH).view(np.int64)).pin_memory() for _ in range(nsets)]
    w_dev = [t.cuda() for t in w_host]
    r_dev = [t.cuda() for t in r_host]
    s_dev = [t.cuda() for t in s_host]
    as_u64 = lambda t: t.numpy().view(np.uint64)

    ctxs = [ctx] + [zk.Context(local) for _ in range(STREAMS - 1)]  # the key is read-only: any context of its device may use it

    def step_device(k, c=None):
        k %= nsets
        return (c or ctx).prove_batch_device(key, w_dev[k].data_ptr(), r_dev[k].data_ptr(), s_dev[k].data_ptr(), BATCH)

    def step_host(k, c=None):
        k %= nsets
        return (c or ctx).prove_batch(key, as_u64(w_host[k]), as_u64(r_host[k]), as_u64(s_host[k]))

    def run_streams(step, n):
        """n steps in total over the STREAMS contexts: stream j takes the steps j, j + STREAMS, ...; returns its device ms"""
        import threading
        dev = [0.0] * STREAMS
        err = []

        def body(j):
            try:
                for i in range(j, n, STREAMS):
                    step(2 * j + (i // STREAMS) % 2, ctxs[j])
                    dev[j] += ctxs[j].last_device_ms()
            except Exception as e:  # surfaced after the join
                err.append(e)

        th = [threading.Thread(target=body, args=(j,)) for j in range(1, STREAMS)]
        for t in th:
            t.start()
        body(0)
        for t in th:
            t.join()
        if err:
            raise err[0]
        return sum(dev)

    # sanity (outside the timed region): the reference verifier accepts what this path produces
    proofs = step_device(0)
    with stdout_to_stderr():
        assert cred.verify(as_u64(w_host[0])[0], proofs[0]) and cred.verify(as_u64(w_host[0])[BATCH - 1], proofs[BATCH - 1]), \
            "reference verifier rejected a GPU proof"

    for i in range(args.warmup):
        step_device(i)
    run_streams(step_device, STREAMS * args.warmup)  # every stream: scratch grown, kernels loaded
    # one stream alone (explains the roofline: its kernels are not time-sliced with another batch's)
    barrier()
    t0 = time.perf_counter()
    one_dev_ms = 0.0
    for i in range(min(args.steps, 10)):
        step_device(i)
        one_dev_ms += ctx.last_device_ms()
    torch.cuda.synchronize()
    one_stream_ms = 1e3 * (time.perf_counter() - t0) / min(args.steps, 10)
    one_dev_ms /= min(args.steps, 10)
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = sum(c.kernel_launches() for c in ctxs)
    barrier()
    t0 = time.perf_counter()
    dev_ms = run_streams(step_device, args.steps)
    barrier()
    dt = max_over_ranks(time.perf_counter() - t0)
    launches = sum(c.kernel_launches() for c in ctxs) - launches0
    clocks = sampler.result()
    dev_ms = max_over_ranks(dev_ms)

    # ---- e2e: the same steps through the C ABI with host buffers ----
    # A proving service keeps two batches in flight (zkb_prove_batch_submit / _collect): the pinned inputs of step
    # k + 1 travel to the device while step k computes.  Every step's host-to-device copy and proof read-back
    # happen inside the timed region; the first copy is not overlapped with anything.
    def e2e_run(n, first):
        k0 = first % nsets
        ctx.prove_batch_submit(key, as_u64(w_host[k0]), as_u64(r_host[k0]), as_u64(s_host[k0]))
        for i in range(n):
            if i + 1 < n:
                k = (first + i + 1) % nsets
                ctx.prove_batch_submit(key, as_u64(w_host[k]), as_u64(r_host[k]), as_u64(s_host[k]))
            ctx.prove_batch_collect()

    for i in range(min(args.warmup, 3)):
        step_host(i)
    e2e_run(2, 0)
    run_streams(step_host, STREAMS * 2)
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        step_host(i)
    barrier()
    dt_e2e_serial = max_over_ranks(time.perf_counter() - t0)
    barrier()
    t0 = time.perf_counter()
    e2e_run(args.steps, 0)
    barrier()
    dt_e2e_pipe = max_over_ranks(time.perf_counter() - t0)
    barrier()
    t0 = time.perf_counter()
    run_streams(step_host, args.steps)  # STREAMS callers, each a blocking zkb_prove_batch per step: H2D, kernels, D2H inside
    barrier()
    dt_e2e = max_over_ranks(time.perf_counter() - t0)

    # ---- single-proof latency (prove ms) ----
    lat = []
    w1, r1, s1 = as_u64(w_host[0])[0], as_u64(r_host[0])[0], as_u64(s_host[0])[0]
    for i in range(7):
        torch.cuda.synchronize()
        t = time.perf_counter()
        ctx.prove(key, w1, r1, s1)
        lat.append(1e3 * (time.perf_counter() - t))
    prove_ms = sorted(lat)[len(lat) // 2]

    # ---- roofline of the dominant kernel: CUDA events around k_accumulate<Fq> on the library's stream ----
    ctx.profile(True)
    for i in range(3):
        step_device(i)
    prof = ctx.profile_read()
    ctx.profile(False)
    prof_c = None
    if STREAMS > 1:  # the same launch while the other streams' kernels share the SMs
        ctx.profile(True)
        run_streams(step_device, 3 * STREAMS)
        prof_c = ctx.profile_read()
        ctx.profile(False)

    imad_peak = ctx.imad_peak(1024)
    extras = {}
    if world == 1 and not args.no_extras:
        # the zklaim-facing entry point: credentials in, proofs out (host witness generation by zklaim's gadget code on all
        # host threads + GPU); three caller threads, so that the witnesses of two batches are generated while a third is on the GPU
        try:
            with stdout_to_stderr():
                cred.stream(1, BATCH, 3)  # warm
                n_z, sec_z = cred.stream(10, BATCH, 3)
            extras["e2e_zklaim"] = {"value": n_z / sec_z, "unit": "proofs/s", "api": "zkb_libsnark_prove_batch (bridge), 3 caller threads x 10 batches of 64 credentials (a stream: 4 batches per caller leave 12 % of the time to filling and draining the pipeline)",
                                    "host_threads": int(os.environ.get("OMP_NUM_THREADS", 0)) or len(os.sched_getaffinity(0)),
                                    "note": "bounded by the host: zklaim's gadget code produces ~2.6-2.9 k witnesses/s on 16 threads (scripts/host_witness_rate.py)"}
        except Exception as e:
            extras["e2e_zklaim"] = {"error": repr(e)}
    if not args.no_extras:
        import zklaim_b200 as zk_mod
        if world > 1:
            try:
                extras["sharded"] = extra_sharded(ctx, np, torch, dist, zk_mod, bridge, rank, world, local, imad_peak)
            except Exception as e:
                extras["sharded"] = {"error": repr(e)}
        if rank == 0:
            for name, fn in (("msm_g1", lambda: extra_msm(ctx, torch, imad_peak)),
                             ("witness_map", lambda: extra_witness_map(ctx, np, zk_mod, imad_peak)),
                             ("s22", lambda: extra_s22(ctx, np, bridge))):
                try:
                    with stdout_to_stderr():
                        extras[name] = fn()
                except Exception as e:
                    extras[name] = {"error": repr(e)}
    if world > 1:
        dist.barrier()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peaks, peak_kind = measured_peaks()
    try:  # DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture (profiles/)
        with open(os.path.join(ROOT, "profiles", "k_accumulate_traffic.json")) as f:
            tj = json.load(f)
            traffic, traffic_src = tj["traffic_bytes_per_launch"], tj.get("source", "profiles/k_accumulate_traffic.json (ncu --set full, replayed)")
    except Exception:
        traffic, traffic_src = None, None
    acc_s = prof["g1_ms"] * 1e-3 / max(prof["g1_launches"], 1)
    pairs = prof["g1_pairs"] / max(prof["g1_launches"], 1)
    executed = pairs * IMAD_PER_MADD_EXECUTED / acc_s             # what the kernel issues: 10 products per mixed addition
    contract = pairs * MUL_PER_MADD_REF * IMAD_PER_MUL / acc_s    # the reference's formula, 11 products
    sm_mhz = clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 1965
    nominal = 148 * 32 * sm_mhz * 1e6                               # 32 IMAD.WIDE per clock per SM at the sampled SM clock
    roofline = {
        "bound": "imad", "kernel": "k_accumulate<Fq> (G1 bucket accumulation)",
        "achieved": executed / 1e9, "peak": imad_peak / 1e9, "unit": "GIMAD/s", "frac": executed / imad_peak,
        "frac_basis": "executed work: 10 Fq products per mixed addition (XYZZ madd-2008-s) x 136 multiply-adds",
        "achieved_contract": contract / 1e9, "frac_contract": contract / imad_peak,
        "contract_basis": "SURVEY 8(d) algorithmic work: 11 Fq products per mixed addition (the reference's madd-2007-bl)",
        "peak_source": "chains of dependent Montgomery products (136 wide multiply-adds each) on all SMs, measured live (zkb_measure_imad_peak)",
        "peak_nominal": nominal / 1e9, "frac_nominal": executed / nominal,
        "peak_nominal_basis": f"148 SM x 32 IMAD.WIDE/clk x {sm_mhz} MHz (SM clock sampled during the timed region); independent of this library's multiplier",
        "units_per_launch": pairs, "ms_per_launch": acc_s * 1e3, "traffic": traffic, "traffic_source": traffic_src,
        "hbm": {"achieved": pairs * BYTES_PER_PAIR / acc_s / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                "frac": pairs * BYTES_PER_PAIR / acc_s / 1e9 / peaks["hbm_gbs"], "peak_source": peak_kind},
        "step_share": prof["g1_ms"] / max(prof["g1_launches"], 1) / one_dev_ms,
        "timed_alone": "the launch is timed in single-stream steps of the same workload (3 steps, CUDA events on the library's stream); "
                       "step_share = launch time / device time of a single-stream step",
    }
    if prof_c and prof_c.get("g1_launches"):
        acc_c = prof_c["g1_ms"] * 1e-3 / prof_c["g1_launches"]
        roofline["in_timed_region"] = {"ms_per_launch": acc_c * 1e3, "frac": (prof_c["g1_pairs"] / prof_c["g1_launches"]) * IMAD_PER_MADD_EXECUTED / acc_c / imad_peak,
                                       "note": f"the same launch with {STREAMS} streams in flight: other batches' kernels share the SMs, the launch takes longer "
                                               "while the device as a whole does more work per second"}
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        try:
            run, cores, kind = reference_prover(REF_SAMPLE)
            sec = run()
            thr = None if args.no_cpu_throughput else reference_throughput()
            cpu = {"value": REF_SAMPLE / sec, "unit": "proofs/s", "cores": cores, "kind": kind,
                   "sample": f"{REF_SAMPLE} proofs of the workload, libsnark r1cs_gg_ppzksnark_prover (-DMULTICORE -DUSE_ASM), one proof at a time on OpenMP {cores} threads",
                   "prove_ms": 1e3 * sec / REF_SAMPLE,
                   "throughput_value": thr[0] if thr else None,
                   "throughput_note": (f"{thr[1]} single-thread provers side by side, 2 proofs each" if thr else "not measured")}
        except Exception as e:  # the GPU numbers stand on their own
            cpu = {"unavailable": str(e)}
    total = BATCH * args.steps * world
    line = {
        "metric": "groth16_proofs_per_s", "value": total / dt, "unit": "proofs/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32x8 Montgomery (254-bit integer)", "data": "synthetic",
        "config": common_config(world),
        "device_ms_per_step": one_dev_ms,  # CUDA events around one batch with ONE stream in flight (comparable with round 1)
        "batch_latency_ms": dev_ms / args.steps,  # the same events with STREAMS batches in flight: what one batch waits
        "streams_per_gpu": STREAMS,
        "one_stream": {"value": 1e3 * BATCH / one_stream_ms * world, "ms_per_step": one_stream_ms, "device_ms_per_step": one_dev_ms,
                       "note": "one batch in flight per GPU (round 1's configuration)"},
        "prove_ms": prove_ms,
        "e2e": {"value": total / dt_e2e, "unit": "proofs/s", "h2d_bytes_per_step": BATCH * (nv * 32 + 64),
                "d2h_bytes_per_step": BATCH * (384 + 4), "ms_per_step": 1e3 * dt_e2e / args.steps,
                "api": f"{STREAMS} caller threads per GPU, each one blocking zkb_prove_batch per step on its own context (pinned host buffers; "
                       "the copies of one stream overlap the kernels of the others)",
                "pipelined_value": total / dt_e2e_pipe, "pipelined_api": "one context, zkb_prove_batch_submit / zkb_prove_batch_collect, two batches in flight",
                "serial_value": total / dt_e2e_serial, "serial_api": "one context, zkb_prove_batch, one blocking call per step"},
        "gpu_launches": launches,
        "clocks": clocks,
        "roofline": roofline,
        "cpu_baseline": cpu,
        "key_device_mb": key.device_bytes() / 1e6,
    }
    line.update(extras)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def host_threads(impl):
    """torch.distributed.run exports OMP_NUM_THREADS=1 to every rank when the variable is unset.  The reference arm
    must use all host threads it can (only rank 0 works there); our arm's host-side witness generation (outside
    the timed region) shares the cores between the ranks.  Must run before anything loads an OpenMP runtime.
    ZKB_HOST_THREADS overrides."""
    if "TORCHELASTIC_RUN_ID" not in os.environ and "ZKB_HOST_THREADS" not in os.environ:
        return
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    world = int(os.environ.get("WORLD_SIZE", 1))
    threads = n if impl == "reference" else max(1, n // world)
    os.environ["OMP_NUM_THREADS"] = os.environ.get("ZKB_HOST_THREADS", str(threads))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-cpu-throughput", action="store_true", help="skip the side-by-side single-thread CPU provers")
    ap.add_argument("--no-extras", action="store_true", help="headline workload only (no MSM / witness-map / S22 / sharded legs)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3
    host_threads(args.impl)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
