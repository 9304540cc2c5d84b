"""K1-K3 parity: device Montgomery field and group arithmetic == the reference's libff, bit for bit.

Mirrors libff's own property tests (libff/algebra/fields/tests/test_fields.cpp:20-63,
libff/algebra/curves/tests/test_groups.cpp:20-113) but against reference OUTPUTS on seeded inputs
(SHA512_rng, libff/common/rng.tcc:26-72), through the C ABI.
"""
import numpy as np
import pytest

import zklaim_b200 as zk

pytestmark = pytest.mark.gpu


def _edge_field(ref, field):
    p = ref.field_const(field, 0)
    one = ref.field_const(field, 1)
    zero = np.zeros(4, dtype=np.uint64)
    pm1 = ref.from_plain(field, (p - np.array([1, 0, 0, 0], dtype=np.uint64))[None])[0]
    return np.stack([zero, one, pm1, ref.field_op(field, 0, one[None], one[None])[0]])


@pytest.mark.parametrize("field", [zk.FR, zk.FQ])
def test_field_ops_bit_exact(zctx, ref, field):
    n = 4096
    a = ref.sha512_rng(1000, n, field)
    b = ref.sha512_rng(900000, n, field)
    e = _edge_field(ref, field)
    # all pairs of edge values at the front
    k = len(e)
    a[: k * k] = np.repeat(e, k, axis=0)
    b[: k * k] = np.tile(e, (k, 1))
    for op in (zk.OP_ADD, zk.OP_SUB, zk.OP_MUL):
        got = zctx.field_op(field, op, a, b)
        assert np.array_equal(got, ref.field_op(field, op, a, b)), f"field {field} op {op}"
    for op in (zk.OP_SQR, zk.OP_NEG):
        assert np.array_equal(zctx.field_op(field, op, a), ref.field_op(field, op, a)), f"field {field} op {op}"
    m = 256
    assert np.array_equal(zctx.field_op(field, zk.OP_INV, a[:m]), ref.field_op(field, zk.OP_INV, a[:m]))


def test_fq2_ops_bit_exact(zctx, ref):
    n = 2048
    a = ref.sha512_rng(5000, 2 * n, zk.FQ).reshape(n, 8)
    b = ref.sha512_rng(70000, 2 * n, zk.FQ).reshape(n, 8)
    a[0] = 0
    b[1] = 0
    a[2, 4:] = 0
    b[3, :4] = 0
    for op in (zk.OP_ADD, zk.OP_SUB, zk.OP_MUL):
        assert np.array_equal(zctx.field_op(zk.FQ2, op, a, b), ref.fq2_op(op, a, b)), f"fq2 op {op}"
    for op in (zk.OP_SQR, zk.OP_NEG):
        assert np.array_equal(zctx.field_op(zk.FQ2, op, a), ref.fq2_op(op, a)), f"fq2 op {op}"
    assert np.array_equal(zctx.field_op(zk.FQ2, zk.OP_INV, a[:128]), ref.fq2_op(zk.OP_INV, a[:128]))


def _points(ref, group, n, first):
    """n Jacobian points with non-trivial Z: k_i * G computed by the reference"""
    g = np.tile(ref.generator(group), (n, 1))
    return ref.scalar_mul(group, ref.sha512_rng(first, n), g)


@pytest.mark.parametrize("group", [zk.G1, zk.G2])
def test_group_ops_match_reference(zctx, ref, group):
    n = 96
    w = 12 if group == zk.G1 else 24
    P = _points(ref, group, n, 11)
    Q = _points(ref, group, n, 7777)
    zero = ref.group_op(group, 0, P[:1], ref.group_op(group, 3, P[:1]))[0]  # P + (-P)
    # corner cases of test_groups.cpp:20-54: zero operands, P + P, P + (-P)
    Q[0] = P[0]
    Q[1] = ref.group_op(group, 3, P[1:2])[0]
    P[2] = zero
    Q[3] = zero
    P[4] = zero
    Q[4] = zero
    Q[5] = ref.to_affine_jac(group, P[5:6])[0]  # same point, different representative
    for op in (zk.GOP_ADD, zk.GOP_DBL, zk.GOP_NEG):
        got = zctx.group_op(group, op, P, Q if op == zk.GOP_ADD else None)
        exp = ref.group_op(group, {zk.GOP_ADD: 0, zk.GOP_DBL: 1, zk.GOP_NEG: 3}[op], P, Q if op == zk.GOP_ADD else None)
        # same group element, and bit-identical after the reference's own affine normalisation
        assert np.array_equal(ref.to_affine_jac(group, got), ref.to_affine_jac(group, exp)), f"group {group} op {op}"
    Qa = ref.to_affine_jac(group, Q)
    got = zctx.group_op(group, zk.GOP_MIXED_ADD, P, Qa)
    exp = ref.group_op(group, 2, P, Qa)
    assert np.array_equal(ref.to_affine_jac(group, got), ref.to_affine_jac(group, exp))
    assert got.shape == (n, w)


@pytest.mark.parametrize("group", [zk.G1, zk.G2])
def test_scalar_mul_matches_reference(zctx, ref, group):
    n = 24
    P = _points(ref, group, n, 31337)
    s = ref.sha512_rng(4242, n)
    s[0] = 0
    s[1] = ref.field_const(zk.FR, 1)
    got = zctx.scalar_mul(group, s, P)
    exp = ref.scalar_mul(group, s, P)
    assert np.array_equal(ref.to_affine_jac(group, got), ref.to_affine_jac(group, exp))
