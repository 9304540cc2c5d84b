"""K7-K9 parity: evaluation-domain operations == libfqfft's basic_radix2_domain / step_radix2_domain
(libfqfft/evaluation_domain/domains/basic_radix2_domain.tcc:41-112, step_radix2_domain.tcc:39-246),
bit for bit, on SHA512_rng vectors, through zkb_domain_op.  Sizes cover one-pass, two-pass and
three-pass transforms, the smallest domains and the step domain incl. small_m = 1.
Mirrors libfqfft/tests/evaluation_domain_test.cpp:40-236 (which runs on other fields).
"""
import numpy as np
import pytest

import zklaim_b200 as zk

pytestmark = pytest.mark.gpu

OPS = [zk.NTT_FFT, zk.NTT_IFFT, zk.NTT_COSET_FFT, zk.NTT_ICOSET_FFT, zk.NTT_DIV_Z_ON_COSET]


@pytest.mark.parametrize("min_size", [2, 4, 8, 32, 500, 1 << 10, 29363, 1 << 17])
def test_basic_radix2_ops_bit_exact(zctx, ref, min_size):
    m, kind = ref.domain_select(min_size)
    assert kind == zk.DOMAIN_BASIC_RADIX2
    a = ref.sha512_rng(min_size * 7 + 1, m)
    a[1] = 0
    for op in OPS:
        exp, _ = ref.domain_op(min_size, op, a)
        got = zctx.domain_op(min_size, op, a)
        assert np.array_equal(got, exp), f"m={m} op={op}"


@pytest.mark.parametrize("min_size", [3, 5, 6, 12, 36, 1 << 10 | 1 << 7, (1 << 14) + (1 << 11), 146799])
def test_step_radix2_ops_bit_exact(zctx, ref, min_size):
    m, kind = ref.domain_select(min_size)
    assert kind == zk.DOMAIN_STEP_RADIX2
    a = ref.sha512_rng(min_size * 3 + 5, m)
    for op in OPS:
        exp, _ = ref.domain_op(min_size, op, a)
        got = zctx.domain_op(min_size, op, a)
        assert np.array_equal(got, exp), f"m={m} op={op}"


def test_fft_roundtrip_at_scale(zctx, ref):
    """size-independent property at 2^22 (BASELINE config #4's largest size): iFFT(FFT(a)) == a,
    icosetFFT(cosetFFT(a)) == a."""
    m = 1 << 22
    rng = np.random.default_rng(5)
    a = rng.integers(0, 1 << 64, size=(m, 4), dtype=np.uint64)
    a[:, 3] &= np.uint64((1 << 60) - 1)
    b = zctx.domain_op(m, zk.NTT_FFT, a)
    assert not np.array_equal(a, b)
    assert np.array_equal(zctx.domain_op(m, zk.NTT_IFFT, b), a)
    b = zctx.domain_op(m, zk.NTT_COSET_FFT, a)
    assert np.array_equal(zctx.domain_op(m, zk.NTT_ICOSET_FFT, b), a)
