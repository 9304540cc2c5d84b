"""f1 parity: zkb_points_decompress == the point import of the reference (operator>> of alt_bn128_G1 / alt_bn128_G2,
libff/algebra/curves/alt_bn128/alt_bn128_g1.cpp:420-465, alt_bn128_g2.cpp:433-478: Y = +-sqrt(X^3 + b), the root whose
plain-integer lsb -- of Y.c0 in G2 -- equals the stored bit).  Ground truth: the elements of a proving key made by the
reference generator (they went through the reference's own square roots when libsnark imported / created them) and
seeded multiples of the generators; infinity entries (is_zero) and abscissas that are on no curve point."""
import numpy as np
import pytest

import zklaim_b200 as zk

pytestmark = pytest.mark.gpu


def _flags(ref, group, pts):
    """(x limbs, flags) of affine points as the reference serialises them (alt_bn128_g1.cpp:404-418)"""
    w = 4 if group == zk.G1 else 8
    x = np.ascontiguousarray(pts[:, :w])
    y0 = np.ascontiguousarray(pts[:, w:w + 4])          # Y (G1) or Y.c0 (G2)
    lsb = (ref.to_plain(zk.FQ, y0)[:, 0] & np.uint64(1)).astype(np.uint8)
    inf = (~pts.any(axis=1)).astype(np.uint8)
    return x, (lsb | (inf << 1)).astype(np.uint8)


@pytest.mark.parametrize("group", [zk.G1, zk.G2])
def test_decompress_seeded_points(zctx, ref, group):
    n = 300
    pts = ref.bases(group, 1 << 33, n)
    pts[7] = 0
    pts[n - 1] = 0
    x, fl = _flags(ref, group, pts)
    got, st = zctx.points_decompress(group, x, fl)
    assert st.all() and np.array_equal(got, pts)
    # the other parity selects the negated point
    fl2 = fl ^ np.uint8(1)
    got2, _ = zctx.points_decompress(group, x, fl2)
    w = 4 if group == zk.G1 else 8
    for i in (0, 1, 100):
        neg = ref.field_op(zk.FQ, 5, pts[i, w:].reshape(-1, 4)).reshape(-1)
        assert np.array_equal(got2[i, :w], pts[i, :w]) and np.array_equal(got2[i, w:], neg)


def test_decompress_rejects_abscissa_off_the_curve(zctx, ref):
    pts = ref.bases(zk.G1, 1 << 34, 40)
    x, fl = _flags(ref, zk.G1, pts)
    # find an x with x^3 + 3 a non-residue: about half of all field elements
    cand = ref.sha512_rng(4242, 16, zk.FQ)
    bad = 0
    for c in cand:
        xx = x.copy()
        xx[5] = c
        try:
            zctx.points_decompress(zk.G1, xx, fl)
        except zk.ZkbError as e:
            assert e.code == zk.ERR_ARG
            bad += 1
    assert 0 < bad < 16


def test_decompress_the_zklaim_proving_key(zctx, ref):
    circ = ref.Circuit.zklaim(1, 1)
    circ.setup()
    kd = circ.export_key()
    for group, names in ((zk.G1, ("A_query", "B_g1", "H_query", "L_query")), (zk.G2, ("B_g2",))):
        for name in names:
            pts = np.ascontiguousarray(kd[name])
            x, fl = _flags(ref, group, pts)
            got, st = zctx.points_decompress(group, x, fl)
            assert st.all() and np.array_equal(got, pts), name
    assert (~kd["A_query"].any(axis=1)).sum() > 0   # the zklaim key does hold points at infinity
    circ.free()
