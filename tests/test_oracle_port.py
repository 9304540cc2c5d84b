"""Pins the oracle restatement (oracle/zkoracle.c via oracle/zkport.py) against the golden vectors in
tests/golden/*.npz -- outputs of the UNMODIFIED reference (libff / libfqfft / libsnark compiled in place,
tests/golden/make_golden.py).  CPU only; this is what makes the oracle trustworthy on the GPU box, where
/root/reference does not exist.

Reference behaviour pinned here:
  Fp_model / Fp2_model operators      libff/algebra/fields/fp.tcc:23-685, fp2.tcc:52-136
  alt_bn128_G1 / G2 group law         libff/algebra/curves/alt_bn128/alt_bn128_g1.cpp:139-358, alt_bn128_g2.cpp:149-365
  scalar_mul                          libff/algebra/curves/curve_utils.tcc:13-34
  multi_exp_with_mixed_addition       libff/algebra/scalar_multiplication/multiexp.tcc:402-496
  get_evaluation_domain + domains     libfqfft/evaluation_domain/get_evaluation_domain.tcc:33-52,
                                      domains/basic_radix2_domain.tcc:41-112, domains/step_radix2_domain.tcc:21-248
  r1cs_to_qap_witness_map + prover    libsnark/reductions/r1cs_to_qap/r1cs_to_qap.tcc:205-334,
                                      zk_proof_systems/ppzksnark/r1cs_gg_ppzksnark/r1cs_gg_ppzksnark.tcc:390-505
"""
import os

import numpy as np
import pytest

from oracle import zkport as P

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
OPS = ("add", "sub", "mul", "sqr", "inv", "neg")


def gold(name):
    return np.load(os.path.join(GOLD, name + ".npz"))


@pytest.mark.parametrize("name,field", [("fr", 0), ("fq", 1)])
def test_field_ops(name, field):
    g = gold("fields")
    a, b = g[name + "_a"], g[name + "_b"]
    for op, on in enumerate(OPS):
        got = P.field_op(field, op, a, b if op < 3 else None)
        assert np.array_equal(got, g[f"{name}_{on}"]), on
    # Montgomery <-> plain (Fp_model::as_bigint, fp.tcc:228-238)
    assert np.array_equal(P.field_op(field, 6, a), g[name + "_plain"])
    assert np.array_equal(P.field_op(field, 7, g[name + "_plain"]), a)


def test_fq2_ops():
    g = gold("fields")
    a, b = g["fq2_a"], g["fq2_b"]
    for op, on in enumerate(OPS):
        got = P.fq2_op(op, a, b if op < 3 else None)
        assert np.array_equal(got, g[f"fq2_{on}"]), on


@pytest.mark.parametrize("name,group", [("g1", 1), ("g2", 2)])
def test_group_ops(name, group):
    """results compared as group elements (normalised), which is what the reference serialises
    (alt_bn128_g1.cpp:404-418); includes P + P, P + (-P), zero operands (make_golden.groups)"""
    g = gold("groups")
    Pp, Q, Qa = g[name + "_P"], g[name + "_Q"], g[name + "_Qa"]
    norm = lambda x: P.group_op(group, 4, x)
    ref_norm = lambda x: P.group_op(group, 4, x)  # golden values are raw Jacobian from libff: normalise both
    assert np.array_equal(norm(P.group_op(group, 0, Pp, Q)), ref_norm(g[name + "_add"]))
    assert np.array_equal(norm(P.group_op(group, 1, Pp)), ref_norm(g[name + "_dbl"]))
    assert np.array_equal(norm(P.group_op(group, 2, Pp, Qa)), ref_norm(g[name + "_madd"]))
    assert np.array_equal(norm(P.group_op(group, 3, Pp)), ref_norm(g[name + "_neg"]))
    # to_affine_coordinates itself is pinned bit-for-bit (alt_bn128_g1.cpp:60-77)
    assert np.array_equal(norm(Pp), g[name + "_affine"])
    assert np.array_equal(norm(P.scalar_mul(group, g[name + "_s"], Pp)), ref_norm(g[name + "_smul"]))


@pytest.mark.parametrize("name,group", [("g1", 1), ("g2", 2)])
def test_msm(name, group):
    g = gold("groups")
    got = P.msm(group, g[name + "_msm_bases"], g[name + "_msm_scalars"])
    assert np.array_equal(P.group_op(group, 4, got[None])[0], g[name + "_msm"])


def test_domain_select():
    g = gold("domains")
    for s, (m, kind) in zip(g["select_sizes"], g["select"]):
        assert P.domain_select(int(s)) == (int(m), int(kind)), int(s)


@pytest.mark.parametrize("ms", [2, 16, 64, 3, 12, 24, 40])
def test_domain_ops(ms):
    """basic_radix2 (2, 16, 64) and step_radix2 (3, 12, 24, 40) domains: FFT, iFFT, cosetFFT, icosetFFT,
    divide_by_Z_on_coset"""
    g = gold("domains")
    a = g[f"dom{ms}_in"]
    for op in range(5):
        assert np.array_equal(P.domain_op(ms, op, a), g[f"dom{ms}_op{op}"]), (ms, op)


def load_key(g):
    key = {k: g[k] for k in ("alpha_g1", "beta_g1", "delta_g1", "beta_g2", "delta_g2", "A_query", "B_index", "B_g1",
                             "B_g2", "H_query", "L_query")}
    for n in "ABC":
        key["csr_" + n] = (g[f"csr_{n}_rowptr"], g[f"csr_{n}_col"], g[f"csr_{n}_coeff"])
    for k in ("num_constraints", "num_variables", "num_inputs"):
        key[k] = int(g[k])
    return key


@pytest.mark.parametrize("nc", [20, 53])
def test_prover(nc):
    """whole r1cs_gg_ppzksnark_prover with fixed (r, s): H coefficients and (A, B, C) bit-exact"""
    g = gold(f"prove_nc{nc}")
    key = load_key(g)
    proof, H, sat = P.prove(key, g["w"], g["r"], g["s"])
    assert sat
    assert np.array_equal(H, g["H"])
    assert np.array_equal(proof, g["proof"])
    assert np.array_equal(P.witness_map(key, g["w"]), g["H"])
    # a corrupted witness is reported, not proved (r1cs_gg_ppzksnark.tcc:406-408 asserts)
    w = g["w"].copy()
    w[-1, 0] ^= 1
    assert not P.prove(key, w, g["r"], g["s"])[2]
