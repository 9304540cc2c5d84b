"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/libzkref.so = libff / libfqfft /
libsnark / zklaim sources compiled in place by oracle/Makefile).  Run in the build container
(needs /root/reference):   python tests/golden/make_golden.py
The reference holds no known-answer vectors for alt_bn128; these outputs of the
reference itself are what pins the oracle restatement (oracle/zkoracle.c) and, through it, the CUDA path.
Inputs are seeded (libff::SHA512_rng, libff/common/rng.tcc:26-72).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import zkref  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
FR, FQ, G1, G2 = 0, 1, 1, 2


def edge(field):
    p = zkref.field_const(field, 0)
    one = zkref.field_const(field, 1)
    pm1 = zkref.from_plain(field, (p - np.array([1, 0, 0, 0], dtype=np.uint64))[None])[0]
    return np.stack([np.zeros(4, dtype=np.uint64), one, pm1])


def fields():
    d = {}
    for field, name in ((FR, "fr"), (FQ, "fq")):
        a = zkref.sha512_rng(100, 48, field)
        b = zkref.sha512_rng(900, 48, field)
        e = edge(field)
        a[:9] = np.repeat(e, 3, axis=0)
        b[:9] = np.tile(e, (3, 1))
        d[name + "_a"], d[name + "_b"] = a, b
        for op, on in enumerate(("add", "sub", "mul", "sqr", "inv", "neg")):
            d[f"{name}_{on}"] = zkref.field_op(field, op, a, b if op < 3 else None)
        d[name + "_plain"] = zkref.to_plain(field, a)
        d[name + "_modulus"] = zkref.field_const(field, 0)
    a = zkref.sha512_rng(5000, 64, FQ).reshape(32, 8)
    b = zkref.sha512_rng(7000, 64, FQ).reshape(32, 8)
    a[0] = 0
    b[1] = 0
    a[2, 4:] = 0
    d["fq2_a"], d["fq2_b"] = a, b
    for op, on in enumerate(("add", "sub", "mul", "sqr", "inv", "neg")):
        d[f"fq2_{on}"] = zkref.fq2_op(op, a, b if op < 3 else None)
    np.savez_compressed(os.path.join(OUT, "fields.npz"), **d)


def groups():
    d = {}
    for group, name in ((G1, "g1"), (G2, "g2")):
        n = 12
        g = np.tile(zkref.generator(group), (n, 1))
        P = zkref.scalar_mul(group, zkref.sha512_rng(11, n), g)
        Q = zkref.scalar_mul(group, zkref.sha512_rng(7777, n), g)
        zero = zkref.group_op(group, 0, P[:1], zkref.group_op(group, 3, P[:1]))[0]
        Q[0] = P[0]
        Q[1] = zkref.group_op(group, 3, P[1:2])[0]
        P[2] = zero
        Q[3] = zero
        P[4] = zero
        Q[4] = zero
        Q[5] = zkref.to_affine_jac(group, P[5:6])[0]
        Qa = zkref.to_affine_jac(group, Q)
        s = zkref.sha512_rng(4242, n)
        s[0] = 0
        s[1] = zkref.field_const(FR, 1)
        d.update({name + "_P": P, name + "_Q": Q, name + "_Qa": Qa, name + "_s": s,
                  name + "_add": zkref.group_op(group, 0, P, Q), name + "_dbl": zkref.group_op(group, 1, P),
                  name + "_madd": zkref.group_op(group, 2, P, Qa), name + "_neg": zkref.group_op(group, 3, P),
                  name + "_affine": zkref.to_affine_jac(group, P), name + "_smul": zkref.scalar_mul(group, s, P)})
        nb = 96 if group == G1 else 24
        bases = zkref.bases(group, (1 << 40) + group, nb, mt=False)
        sc = zkref.sha512_rng(31337, nb)
        sc[::5] = 0
        sc[1::7] = zkref.field_const(FR, 1)
        bases[3] = 0
        bases[10] = bases[9]
        res, _ = zkref.msm(group, bases, sc, mixed=True)
        d.update({name + "_msm_bases": bases, name + "_msm_scalars": sc,
                  name + "_msm": zkref.to_affine_jac(group, res[None])[0]})
    np.savez_compressed(os.path.join(OUT, "groups.npz"), **d)


def domains():
    d = {}
    sizes = list(range(2, 300)) + [29363, 58722, 146799, (1 << 20) + 1, (1 << 22) - 11]
    d["select_sizes"] = np.array(sizes, dtype=np.uint64)
    d["select"] = np.array([zkref.domain_select(s) for s in sizes], dtype=np.uint64)
    for ms in (2, 16, 64, 3, 12, 24, 40):
        m, kind = zkref.domain_select(ms)
        a = zkref.sha512_rng(ms * 13, m)
        d[f"dom{ms}_in"] = a
        for op in range(5):
            d[f"dom{ms}_op{op}"] = zkref.domain_op(ms, op, a)[0]
    np.savez_compressed(os.path.join(OUT, "domains.npz"), **d)


def provers():
    for nc, ni in ((20, 3), (53, 10)):
        c = zkref.Circuit.example(nc, ni, seed=nc)
        c.setup()
        key = c.export_key()
        w = c.witness()
        r = zkref.sha512_rng(77, 1)[0]
        s = zkref.sha512_rng(78, 1)[0]
        out = c.prove(w, r, s)
        assert c.verify(w[:ni], out["proof"])
        pr = out["proof"]
        norm = np.concatenate([zkref.to_affine_jac(G1, pr[None, :12])[0], zkref.to_affine_jac(G2, pr[None, 12:36])[0],
                               zkref.to_affine_jac(G1, pr[None, 36:])[0]])
        d = {"num_constraints": nc + 0, "num_variables": key["num_variables"], "num_inputs": ni, "w": w, "r": r, "s": s,
             "proof": norm, "H": out["H"], "proof_bytes": np.frombuffer(zkref.proof_bytes(pr), dtype=np.uint8),
             "dims": np.array([c.dims()["m"], c.dims()["kind"]], dtype=np.uint64)}
        d["num_constraints"] = key["num_constraints"]
        for name in "ABC":
            rp, col, co = key["csr_" + name]
            d[f"csr_{name}_rowptr"], d[f"csr_{name}_col"], d[f"csr_{name}_coeff"] = rp, col, co
        for f in ("alpha_g1", "beta_g1", "delta_g1", "beta_g2", "delta_g2", "A_query", "B_index", "B_g1", "B_g2", "H_query", "L_query"):
            d[f] = key[f]
        np.savez_compressed(os.path.join(OUT, f"prove_nc{nc}.npz"), **d)
        c.free()


if __name__ == "__main__":
    fields()
    groups()
    domains()
    provers()
    print("golden vectors written to", OUT)
    os.system(f"ls -la {OUT}")
