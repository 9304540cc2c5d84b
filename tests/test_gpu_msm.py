"""K4-K6 parity: the Pippenger MSM kernels == libff::multi_exp<..., multi_exp_method_BDLO12>
(libff/algebra/scalar_multiplication/multiexp.tcc:165-282, :402-496) as a group element,
bit-identical after the reference's affine normalisation.  Inputs as in multiexp_profile.cpp:45
(SHA512_rng scalars) but with DISTINCT bases k_i * G; edge cases: empty, single, zero / one /
r-1 scalars, infinity bases, repeated bases (bucket collisions incl. P + P and P + (-P)).
"""
import numpy as np
import pytest

import zklaim_b200 as zk

pytestmark = pytest.mark.gpu


def _check(ref, group, got, exp):
    assert np.array_equal(ref.to_affine_jac(group, got[None]), ref.to_affine_jac(group, exp[None]))


@pytest.mark.parametrize("n,c,pf", [(1, 0, 0), (7, 4, 1), (300, 5, 1), (300, 6, 0), (5000, 0, 1), (5000, 0, 0),
                                    (5000, 9, 3), (40000, 0, 0), (40000, 18, 0), (3000, 20, 2)])
def test_g1_msm_matches_reference(zctx, ref, n, c, pf):
    bases = ref.bases(zk.G1, 1 << 40, n)
    s = ref.sha512_rng(77 << 32, n)
    B = zctx.upload_bases(zk.G1, bases, window_bits=c, precompute_factor=pf)
    got = zctx.msm(B, s)
    exp, _ = ref.msm(zk.G1, bases, s)
    _check(ref, zk.G1, got, exp)
    B.free()


def test_g1_msm_edge_scalars_and_bases(zctx, ref):
    n = 2000
    bases = ref.bases(zk.G1, 1 << 40, n)
    s = ref.sha512_rng(123, n)
    one = ref.field_const(zk.FR, 1)
    r = ref.field_const(zk.FR, 0)
    rm1 = ref.from_plain(zk.FR, (r - np.array([1, 0, 0, 0], dtype=np.uint64))[None])[0]
    s[::3] = 0            # zero scalars are skipped (multiexp.tcc:466-470)
    s[1::3] = one         # ones go through the same buckets here (multiexp.tcc:471-481 special-cases them)
    s[5] = rm1
    s[8] = rm1
    bases[10] = 0         # infinity bases exist in real keys (alt_bn128_g1.cpp:409)
    bases[11] = 0
    bases[20:40] = bases[20]   # repeated base: bucket collisions, P + P
    negy = ref.field_op(zk.FQ, 5, bases[50:51, 4:8])[0]
    bases[51, :4] = bases[50, :4]
    bases[51, 4:] = negy       # -P next to P
    s[51] = s[50]
    for pf in (1, 0):
        B = zctx.upload_bases(zk.G1, bases, precompute_factor=pf)
        got = zctx.msm(B, s)
        exp, _ = ref.msm(zk.G1, bases, s, mixed=True)
        _check(ref, zk.G1, got, exp)
        # sub-range (offset) MSM, as used for the L query (r1cs_gg_ppzksnark.tcc:477-484)
        got = zctx.msm(B, s[100:900], offset=100)
        exp, _ = ref.msm(zk.G1, bases[100:900], s[100:900])
        _check(ref, zk.G1, got, exp)
        B.free()


def test_g1_msm_all_zero_and_empty(zctx, ref):
    n = 64
    bases = ref.bases(zk.G1, 1 << 40, n)
    B = zctx.upload_bases(zk.G1, bases)
    z = zctx.msm(B, np.zeros((n, 4), dtype=np.uint64))
    assert not z[8:12].any()          # Z == 0: infinity
    e = zctx.msm(B, np.zeros((0, 4), dtype=np.uint64))
    assert not e[8:12].any()
    B.free()


@pytest.mark.parametrize("n,pf", [(1, 0), (500, 1), (3000, 0)])
def test_g2_msm_matches_reference(zctx, ref, n, pf):
    bases = ref.bases(zk.G2, 1 << 41, n)
    s = ref.sha512_rng(99 << 32, n)
    s[::5] = 0
    B = zctx.upload_bases(zk.G2, bases, precompute_factor=pf)
    got = zctx.msm(B, s)
    exp, _ = ref.msm(zk.G2, bases, s)
    _check(ref, zk.G2, got, exp)
    B.free()


def test_msm_linearity_at_scale(zctx, ref):
    """size-independent property at 2^18: MSM(a) + MSM(b) == MSM(a + b)."""
    n = 1 << 18
    bases = ref.bases(zk.G1, 1 << 40, n)
    a = ref.sha512_rng(1 << 33, n)
    b = ref.sha512_rng(2 << 33, n)
    ab = ref.field_op(zk.FR, 0, a, b)
    B = zctx.upload_bases(zk.G1, bases)
    ra, rb, rab = zctx.msm(B, a), zctx.msm(B, b), zctx.msm(B, ab)
    s = ref.group_op(zk.G1, 0, ra[None], rb[None])[0]
    _check(ref, zk.G1, s, rab)
    B.free()


@pytest.mark.parametrize("group,n", [(zk.G1, 5003), (zk.G2, 700)])
def test_range_sharded_msm_equals_whole(zctx, ref, group, n):
    """contiguous ranges (libff's chunking, multiexp.tcc:417-431) evaluated separately, partials
    summed with zkb_group_sum (multiexp.tcc:433-438) == the whole MSM == the reference.  One GPU plays every rank."""
    from zklaim_b200 import sharding
    bases = ref.bases(group, 1 << 40, n)
    s = ref.sha512_rng(5 << 32, n)
    exp, _ = ref.msm(group, bases, s)
    for world in (1, 2, 3, 8):
        parts = []
        for rank in range(world):
            lo, hi = sharding.shard_range(n, rank, world)
            B = zctx.upload_bases(group, bases[lo:hi])
            parts.append(sharding.sharded_msm(zctx, group, B, s[lo:hi]))  # no process group: world of one per call
            B.free()
        _check(ref, group, zctx.group_sum(group, np.stack(parts)), exp)
    w = 12 if group == zk.G1 else 24
    zero = zctx.group_sum(group, np.zeros((0, w), dtype=np.uint64))
    _check(ref, group, zero, ref.group_op(group, 0, zero[None], zero[None])[0])  # 0 + 0 == 0: it is the identity
    _check(ref, group, zctx.group_sum(group, np.stack([exp, zero])), exp)


def test_msm_properties_at_full_size(zctx, ref):
    """BASELINE.json configs[2] size (2^22 points; multiexp_profile.cpp:45 runs 2^10 .. 2^20) through properties that
    need no CPU run of that size: linearity MSM(a) + MSM(b) == MSM(a + b), the sum over contiguous ranges == the whole
    (multiexp.tcc:417-438), and a prefix small enough for the reference == the reference.  Bases k_i * G come from
    the GPU's own scalar multiplication (parity-tested against curve_utils.tcc:13-34 in test_gpu_ec.py)."""
    n = 1 << 22
    rng = np.random.default_rng(22)
    one_q = ref.field_const(zk.FQ, 1)
    g = np.concatenate([one_q, zctx.field_op(zk.FQ, zk.OP_ADD, one_q[None], one_q[None])[0], one_q])  # (1, 2, 1)

    def rand_fr(m):
        a = rng.integers(0, 1 << 64, size=(m, 4), dtype=np.uint64)
        a[:, 3] &= np.uint64((1 << 60) - 1)
        return a

    bases = np.empty((n, 8), dtype=np.uint64)
    step = 1 << 20
    for lo in range(0, n, step):
        P = zctx.scalar_mul(zk.G1, rand_fr(step), np.tile(g, (step, 1)))
        bases[lo:lo + step] = zctx.group_op(zk.G1, zk.GOP_NORMALIZE, P)[:, :8]
    a, b = rand_fr(n), rand_fr(n)
    ab = zctx.field_op(zk.FR, zk.OP_ADD, a, b)
    B = zctx.upload_bases(zk.G1, bases)
    ra, rb, rab = zctx.msm(B, a), zctx.msm(B, b), zctx.msm(B, ab)
    _check(ref, zk.G1, zctx.group_sum(zk.G1, np.stack([ra, rb])), rab)
    cut = [0, n // 3, n // 2 + 12345, n]
    parts = [zctx.msm(B, a[lo:hi], offset=lo) for lo, hi in zip(cut[:-1], cut[1:])]
    _check(ref, zk.G1, zctx.group_sum(zk.G1, np.stack(parts)), ra)
    k = 3000
    exp, _ = ref.msm(zk.G1, bases[n - k:], a[n - k:])
    _check(ref, zk.G1, zctx.msm(B, a[n - k:], offset=n - k), exp)
    B.free()
