"""Quick standalone G1 MSM timing on the GPU (development aid; bench.py is the contract)."""
import sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import zklaim_b200 as zk

R = 21888242871839275222246405745257275088548364400416034343698204186575808495617
G1_GEN = None


def rand_fr(n, seed):
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 1 << 64, size=(n, 4), dtype=np.uint64)
    a[:, 3] &= np.uint64((1 << 60) - 1)   # < 2^252 < r: a valid canonical residue, taken as Montgomery form
    return a


def gen_bases(ctx, n, seed=7):
    one_q = np.array([0xd35d438dc58f0d9d, 0x0a78eb28f5c70b3d, 0x666ea36f7879462c, 0x0e0a77c19a07df2f], dtype=np.uint64)
    two_q = ctx.field_op(zk.FQ, zk.OP_ADD, one_q[None], one_q[None])[0]
    g = np.concatenate([one_q, two_q, one_q])          # (1, 2, 1) alt_bn128_init.cpp:148-150
    out = np.zeros((n, 8), dtype=np.uint64)
    step = 1 << 20
    for lo in range(0, n, step):
        m = min(step, n - lo)
        k = rand_fr(m, seed + lo)
        P = ctx.scalar_mul(zk.G1, k, np.tile(g, (m, 1)))
        P = ctx.group_op(zk.G1, zk.GOP_NORMALIZE, P)
        out[lo:lo + m] = P[:, :8]
    return out


if __name__ == "__main__":
    ctx = zk.Context(0)
    logs = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [16, 20]
    pf = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    cs = [int(x) for x in sys.argv[3].split(",")] if len(sys.argv) > 3 else [0]
    nmax = 1 << max(logs)
    t = time.time()
    bases = gen_bases(ctx, nmax)
    print(f"bases 2^{max(logs)} generated in {time.time()-t:.1f}s", flush=True)
    for lg in logs:
        n = 1 << lg
        s = rand_fr(n, 1000 + lg)
        for c in cs:
            t = time.time()
            B = ctx.upload_bases(zk.G1, bases[:n], window_bits=c, precompute_factor=pf)
            tu = time.time() - t
            best = 1e9
            for it in range(4):
                ctx.msm(B, s)
                best = min(best, ctx.last_device_ms())
            print(f"n=2^{lg} c={c} pf={pf} upload {tu:.2f}s bytes {B.device_bytes()/1e6:.0f}MB  msm {best:.3f} ms  {n/best/1e3:.2f} Mpts/s", flush=True)
            B.free()
