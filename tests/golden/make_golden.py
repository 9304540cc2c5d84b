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


def pairings():
    """alt_bn128_ate_miller_loop / alt_bn128_reduced_pairing (alt_bn128_pairing.cpp:371-420, 523-533) on seeded points,
    incl. the degenerate input the reference accepts (P at infinity enters the loop as (0, 1))."""
    n = 6
    P = zkref.bases(G1, (1 << 42) + 1, n, mt=False)
    Q = zkref.bases(G2, (1 << 42) + 2, n, mt=False)
    P[4] = 0  # (Q at infinity trips an assert of the reference, alt_bn128_pairing.cpp:348: not a defined input)
    d = {"P": P, "Q": Q,
         "miller": np.stack([zkref.pairing(P[i], Q[i], False) for i in range(n)]),
         "reduced": np.stack([zkref.pairing(P[i], Q[i], True) for i in range(n)])}
    np.savez_compressed(os.path.join(OUT, "pairing.npz"), **d)


def verifier():
    """r1cs_gg_ppzksnark_verifier_strong_IC (r1cs_gg_ppzksnark.tcc:523-622) decisions on valid and tampered proofs."""
    nc, ni = 37, 5
    c = zkref.Circuit.example(nc, ni, seed=nc)
    c.setup()
    vk = c.export_vk()
    w = c.witness()
    proofs, primary = [], []
    for k in range(4):
        out = c.prove(w, zkref.sha512_rng(500 + k, 1)[0], zkref.sha512_rng(600 + k, 1)[0], want_H=False)
        proofs.append(np.asarray(out["proof"], dtype=np.uint64).reshape(48))
        primary.append(w[:ni].copy())
    # Jacobian representatives with Z != 1 must be accepted as well: proof 1 keeps the prover's raw output
    norm = np.concatenate([zkref.to_affine_jac(G1, proofs[0][None, :12])[0], zkref.to_affine_jac(G2, proofs[0][None, 12:36])[0],
                           zkref.to_affine_jac(G1, proofs[0][None, 36:])[0]])
    proofs[0] = norm
    t = proofs[2].copy()                      # C of another proof: pairing equation fails
    t[36:] = proofs[3][36:]
    proofs.append(t); primary.append(w[:ni].copy())
    t = norm.copy()                           # A off the curve
    t[0] ^= np.uint64(1)
    proofs.append(t); primary.append(w[:ni].copy())
    x = w[:ni].copy()                         # wrong public input
    x[2] = zkref.sha512_rng(9, 1)[0]
    proofs.append(proofs[1].copy()); primary.append(x)
    t = norm.copy()                           # A = zero (well formed, equation fails)
    t[:12] = zkref.group_op(G1, 0, norm[None, :12], zkref.group_op(G1, 3, norm[None, :12]))[0]
    proofs.append(t); primary.append(w[:ni].copy())
    proofs, primary = np.stack(proofs), np.stack(primary)
    accept = np.array([c.verify(primary[i], proofs[i]) for i in range(len(proofs))], dtype=np.int32)
    assert accept.tolist() == [1, 1, 1, 1, 0, 0, 0, 0], accept
    d = {"alpha_g1_beta_g2": c.vk_alpha_beta(), "gamma_g2": vk["gamma_g2"], "delta_g2": vk["delta_g2"],
         "gamma_ABC_g1": vk["gamma_ABC_g1"], "primary": primary, "proofs": proofs, "accept": accept}
    np.savez_compressed(os.path.join(OUT, "verify.npz"), **d)
    c.free()


if __name__ == "__main__":
    which = sys.argv[1:] or ["fields", "groups", "domains", "provers", "pairings", "verifier"]
    for name in which:
        globals()[name]()
    print("golden vectors written to", OUT)
    os.system(f"ls -la {OUT}")
