"""CPU tier for the scalar split used by k_smul (zklaim_b200/csrc/glv.cuh, compiled here by the host compiler): for
k < 2^256, k = k1 - |k2| lambda (mod r) with both halves below 2^128; lambda acts on alt_bn128 G1 as
(x, y) -> (beta x, y) (checked with plain integer curve arithmetic: y^2 = x^3 + 3, alt_bn128_g1.cpp:21-25 / alt_bn128_init.cpp:129-147)."""
import ctypes as C
import os
import random
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
R = 21888242871839275222246405745257275088548364400416034343698204186575808495617
Q = 21888242871839275222246405745257275088696311157297823662689037894645226208583
LAMBDA = 4407920970296243842393367215006156084916469457145843978461
BETA = 2203960485148121921418603742825762020974279258880205651966


@pytest.fixture(scope="module")
def host(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("glvhost") / "libglvhost.so")
    subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", so, os.path.join(HERE, "host", "glv_host.cpp")], check=True)
    return C.CDLL(so)


def _add(P, S):
    if P is None:
        return S
    if S is None:
        return P
    if P[0] == S[0]:
        if (P[1] + S[1]) % Q == 0:
            return None
        l = 3 * P[0] * P[0] * pow(2 * P[1], -1, Q) % Q
    else:
        l = (S[1] - P[1]) * pow(S[0] - P[0], -1, Q) % Q
    x = (l * l - P[0] - S[0]) % Q
    return x, (l * (P[0] - x) - P[1]) % Q


def _mul(k, P):
    acc = None
    while k:
        if k & 1:
            acc = _add(acc, P)
        P = _add(P, P)
        k >>= 1
    return acc


def test_constants():
    assert (LAMBDA * LAMBDA + LAMBDA + 1) % R == 0 and pow(BETA, 3, Q) == 1 and BETA != 1
    P = _mul(987654321987654321, (1, 2))
    assert _mul(LAMBDA, P) == (BETA * P[0] % Q, P[1])


def test_beta_is_in_montgomery_form(host):
    out = np.zeros(8, dtype=np.uint32)
    host.hostglv_beta_mont(out.ctypes.data_as(C.c_void_p))
    assert int.from_bytes(out.tobytes(), "little") == BETA * (1 << 256) % Q


def test_decompose(host):
    rnd = random.Random(7)
    ks = [0, 1, 2, 15, 16, R - 1, R - 2, R, R + 1, LAMBDA, R - LAMBDA, LAMBDA + 1, (R - 1) // 2, 1 << 128, (1 << 128) - 1, 1 << 253,
          (1 << 256) - 1]
    ks += [rnd.randrange(R) for _ in range(20000)] + [rnd.randrange(1 << 256) for _ in range(5000)] + [rnd.randrange(1 << 130) for _ in range(2000)]
    k = np.frombuffer(b"".join(x.to_bytes(32, "little") for x in ks), dtype=np.uint32).copy()
    k1 = np.zeros(5 * len(ks), dtype=np.uint32)
    k2 = np.zeros(5 * len(ks), dtype=np.uint32)
    host.hostglv_decompose(k.ctypes.data_as(C.c_void_p), C.c_ulong(len(ks)), k1.ctypes.data_as(C.c_void_p), k2.ctypes.data_as(C.c_void_p))
    b1, b2 = k1.tobytes(), k2.tobytes()
    for i, x in enumerate(ks):
        a = int.from_bytes(b1[20 * i:20 * i + 20], "little")
        b = int.from_bytes(b2[20 * i:20 * i + 20], "little")
        assert a < 1 << 128 and b < 1 << 128, hex(x)
        assert (a - b * LAMBDA - x) % R == 0, hex(x)
