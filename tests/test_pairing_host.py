"""CPU tier for the verifier side: zklaim_b200/csrc/pairing.cuh compiled by the HOST compiler
(tests/host/pairing_host.cpp) against golden vectors made from the unmodified reference
(tests/golden/make_golden.py pairings / verifier).  The same source is the device code of verify.cu, so what
is pinned here is the algorithm the GPU runs: tower arithmetic, Miller loop, final exponentiation and the
verifier's decision.

Reference behaviour pinned here:
  alt_bn128_ate_miller_loop / precompute_G2      libff/algebra/curves/alt_bn128/alt_bn128_pairing.cpp:234-420
  alt_bn128_final_exponentiation                 alt_bn128_pairing.cpp:109-240
  r1cs_gg_ppzksnark_verifier_strong_IC           libsnark/.../r1cs_gg_ppzksnark/r1cs_gg_ppzksnark.tcc:523-622
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")


@pytest.fixture(scope="module")
def host(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("pairhost") / "libpairhost.so")
    subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", so, os.path.join(HERE, "host", "pairing_host.cpp")],
                   check=True)
    return C.CDLL(so)


def _p(a):
    return C.c_void_p(a.ctypes.data)


@pytest.mark.parametrize("reduced", [0, 1])
def test_pairing_matches_reference(host, reduced):
    g = np.load(os.path.join(GOLD, "pairing.npz"))
    exp = g["reduced" if reduced else "miller"]
    for i in range(len(exp)):
        p, q = np.ascontiguousarray(g["P"][i]), np.ascontiguousarray(g["Q"][i])
        if not p.any():  # infinity enters the reference's loop as (0, 1) (to_affine_coordinates of zero)
            p = p.copy()
            p[4:] = np.array([0xd35d438dc58f0d9d, 0x0a78eb28f5c70b3d, 0x666ea36f7879462c, 0x0e0a77c19a07df2f], dtype=np.uint64)
        out = np.zeros(48, dtype=np.uint64)
        host.hostpair_pairing(_p(p), _p(q), reduced, _p(out))
        assert np.array_equal(out, exp[i]), i


@pytest.mark.parametrize("lanes", [6, 18])
@pytest.mark.parametrize("reduced", [0, 1])
def test_cooperative_lane_program_matches_reference(host, reduced, lanes):
    """pairing_coop.cuh (6 or 18 lanes per pairing; on the host the lanes of a phase run one after the other) gives the
    reference's Miller-loop and reduced-pairing values"""
    g = np.load(os.path.join(GOLD, "pairing.npz"))
    exp = g["reduced" if reduced else "miller"]
    for i in range(len(exp)):
        p, q = np.ascontiguousarray(g["P"][i]), np.ascontiguousarray(g["Q"][i])
        if not p.any():
            p = p.copy()
            p[4:] = np.array([0xd35d438dc58f0d9d, 0x0a78eb28f5c70b3d, 0x666ea36f7879462c, 0x0e0a77c19a07df2f], dtype=np.uint64)
        out = np.zeros(48, dtype=np.uint64)
        host.hostpair_pairing_coop(_p(p), _p(q), reduced, lanes, _p(out))
        assert np.array_equal(out, exp[i]), i


def test_pairing_of_negated_point_is_the_inverse(host):
    """e(P, Q) * e(-P, Q) == 1 (bilinearity, the property of libff's own test_bilinearity.cpp), with the product taken
    by the Fq12 multiplication of the same header"""
    from oracle import zkport

    g = np.load(os.path.join(GOLD, "pairing.npz"))
    p, q = np.ascontiguousarray(g["P"][0]), np.ascontiguousarray(g["Q"][0])
    pn = p.copy()
    pn[4:] = zkport.field_op(1, 5, p[None, 4:])[0]  # Fq negation of y
    e1, e2, prod = (np.zeros(48, dtype=np.uint64) for _ in range(3))
    host.hostpair_pairing(_p(p), _p(q), 1, _p(e1))
    host.hostpair_pairing(_p(pn), _p(q), 1, _p(e2))
    host.hostpair_fq12_op(0, _p(e1), _p(e2), _p(prod))
    one = np.zeros(48, dtype=np.uint64)
    one[:4] = np.array([0xd35d438dc58f0d9d, 0x0a78eb28f5c70b3d, 0x666ea36f7879462c, 0x0e0a77c19a07df2f], dtype=np.uint64)
    assert np.array_equal(prod, one)
    assert not np.array_equal(e1, one)


def test_verifier_decisions_match_reference(host):
    g = np.load(os.path.join(GOLD, "verify.npz"))
    ab, gam, dlt = (np.ascontiguousarray(g[k]) for k in ("alpha_g1_beta_g2", "gamma_g2", "delta_g2"))
    abc = np.ascontiguousarray(g["gamma_ABC_g1"])
    ni = len(abc) - 1
    got = []
    for x, pr in zip(g["primary"], g["proofs"]):
        x, pr = np.ascontiguousarray(x), np.ascontiguousarray(pr)
        got.append(host.hostpair_verify(_p(ab), _p(gam), _p(dlt), _p(abc), ni, _p(x), _p(pr)))
    assert got == g["accept"].tolist()
    assert got == [1, 1, 1, 1, 0, 0, 0, 0]


def test_portable_field_path_matches_reference(tmp_path):
    """fp.cuh's plain-C++ path (tests/host/fp_host.cpp) -- the arithmetic under the host builds above -- against the
    reference's Fp_model / Fp2_model outputs (fp.tcc:23-685, fp2.tcc:52-136) in tests/golden/fields.npz"""
    so = str(tmp_path / "libfphost.so")
    subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", so, os.path.join(HERE, "host", "fp_host.cpp")], check=True)
    L = C.CDLL(so)
    g = np.load(os.path.join(GOLD, "fields.npz"))
    ops = ("add", "sub", "mul", "sqr", "inv", "neg")
    for field, name in ((0, "fr"), (1, "fq")):
        a, b = np.ascontiguousarray(g[name + "_a"]), np.ascontiguousarray(g[name + "_b"])
        for op, on in enumerate(ops):
            out = np.zeros_like(a)
            L.hostfp_op(field, op, _p(a), _p(b) if op < 3 else None, _p(out), C.c_size_t(len(a)))
            assert np.array_equal(out, g[f"{name}_{on}"]), (name, on)
    a, b = np.ascontiguousarray(g["fq2_a"]), np.ascontiguousarray(g["fq2_b"])
    for op, on in enumerate(ops):
        out = np.zeros_like(a)
        L.hostfq2_op(op, _p(a), _p(b) if op < 3 else None, _p(out), C.c_size_t(len(a)))
        assert np.array_equal(out, g[f"fq2_{on}"]), on
