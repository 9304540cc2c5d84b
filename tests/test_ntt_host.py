"""CPU tier for the register-radix NTT pass: zklaim_b200/csrc/ntt_tile.cuh compiled by the HOST compiler
(tests/host/ntt_host.cpp; the threads of every phase run one after the other, which is exactly the barrier semantics
of the kernel k_ntt_r8) against the oracle's restatement of libfqfft's basic_radix2_domain FFT
(libfqfft/evaluation_domain/domains/basic_radix2_domain_aux.tcc:45-79): decimation in frequency (natural in,
bit-reversed out) and decimation in time (bit-reversed in, natural out), both tile sizes, the fused coset
post-multiplication, and the bank mapping of the round plans (no shared-memory conflicts from 2^10 points on)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from oracle import zkport

HERE = os.path.dirname(os.path.abspath(__file__))
R2_FR = np.array([0x1bb8e645ae216da7, 0x53fe3ab1e35c59e3, 0x8c49833d53bb8085, 0x0216d0b17f4e44a5], dtype=np.uint64)


@pytest.fixture(scope="module")
def host(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("ntthost") / "libntthost.so")
    subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", so, os.path.join(HERE, "host", "ntt_host.cpp")], check=True)
    return C.CDLL(so)


def _brev(L):
    return np.array([int(format(i, f"0{L}b")[::-1], 2) if L else 0 for i in range(1 << L)])


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


@pytest.mark.parametrize("tile", [10, 11])
@pytest.mark.parametrize("L", [8, 9, 10, 12, 13, 15, 16])
def test_register_radix_pass_matches_reference_fft(host, L, tile):
    m = 1 << L
    rng = np.random.default_rng(L)
    a = rng.integers(0, 1 << 62, size=(m, 4), dtype=np.uint64)
    a[:, 3] &= np.uint64((1 << 60) - 1)
    one = zkport.field_op(0, 2, np.array([[1, 0, 0, 0]], dtype=np.uint64), R2_FR[None])[0]  # Montgomery one
    delta = np.zeros((m, 4), dtype=np.uint64)
    delta[1] = one
    w = zkport.domain_op(m, 0, delta)[1].copy()  # FFT of the unit vector e_1: the domain's root of unity
    exp = zkport.domain_op(m, 0, a)
    perm = _brev(L)
    d = a.copy()
    assert host.hostntt_levels(L, 0, _p(d), _p(w), None, tile) == 0
    assert np.array_equal(d[perm], exp)
    d = np.ascontiguousarray(a[perm])
    assert host.hostntt_levels(L, 1, _p(d), _p(w), None, tile) == 0
    assert np.array_equal(d, exp)
    # the fused post-multiplication: out[i] *= post[bitrev(i)]
    post = rng.integers(0, 1 << 62, size=(m, 4), dtype=np.uint64)
    post[:, 3] &= np.uint64((1 << 60) - 1)
    d = a.copy()
    assert host.hostntt_levels(L, 0, _p(d), _p(w), _p(post), tile) == 0
    assert np.array_equal(d[perm], zkport.field_op(0, 2, exp, post))
    if L >= 10:
        assert host.hostntt_worst_bank_conflict(L, 0, tile) == 1 and host.hostntt_worst_bank_conflict(L, 1, tile) == 1
