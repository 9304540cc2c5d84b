"""ctypes binding of oracle/libzkoracle.so -- the plain-C restatement of the reference's prover path
(oracle/zkoracle.c; built by `make -C oracle port`, gcc only).

TEST INFRASTRUCTURE: imported only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg.
Pinned against the golden vectors in tests/golden/ (outputs of the reference itself, see
tests/golden/make_golden.py) by tests/test_oracle_port.py.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libzkoracle.so")
_vp, _sz, _i = C.c_void_p, C.c_size_t, C.c_int
_lib = None


class _CSR(C.Structure):
    _fields_ = [("row_ptr", _vp), ("col", _vp), ("coeff", _vp)]


class _KeyDesc(C.Structure):
    _fields_ = [
        ("num_constraints", _sz), ("num_variables", _sz), ("num_inputs", _sz),
        ("A", _CSR), ("B", _CSR), ("C", _CSR),
        ("alpha_g1", _vp), ("beta_g1", _vp), ("delta_g1", _vp), ("beta_g2", _vp), ("delta_g2", _vp), ("A_query", _vp),
        ("B_nnz", _sz), ("B_index", _vp), ("B_g1", _vp), ("B_g2", _vp),
        ("H_size", _sz), ("H_query", _vp), ("L_query", _vp),
    ]


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            subprocess.run(["make", "-C", _HERE, "port"], check=True, capture_output=True)
        L = C.CDLL(LIB_PATH)
        L.zko_domain_select.restype = _i
        L.zko_domain_op.restype = _i
        L.zko_witness_map.restype = _i
        L.zko_prove.restype = _i
        _lib = L
    return _lib


def _u(a, last=None):
    a = np.ascontiguousarray(a, dtype=np.uint64)
    if last is not None:
        assert a.shape[-1] == last, (a.shape, last)
    return a


def field_op(field: int, op: int, a, b=None) -> np.ndarray:
    """field 0 = Fr, 1 = Fq; op 0 add 1 sub 2 mul 3 sqr 4 inv 5 neg 6 to-plain 7 from-plain"""
    a = _u(a, 4)
    out = np.empty_like(a)
    bp = _u(b, 4).ctypes.data if b is not None else None
    lib().zko_field_op(_i(field), _i(op), _vp(a.ctypes.data), _vp(bp), _vp(out.ctypes.data), _sz(a.size // 4))
    return out


def fq2_op(op: int, a, b=None) -> np.ndarray:
    a = _u(a, 8)
    out = np.empty_like(a)
    bp = _u(b, 8).ctypes.data if b is not None else None
    lib().zko_fq2_op(_i(op), _vp(a.ctypes.data), _vp(bp), _vp(out.ctypes.data), _sz(a.size // 8))
    return out


def group_op(group: int, op: int, a, b=None) -> np.ndarray:
    """op 0 add 1 dbl 2 mixed_add 3 neg 4 to_affine_coordinates; Jacobian limbs"""
    w = 12 if group == 1 else 24
    a = _u(a, w)
    out = np.empty_like(a)
    bp = _u(b, w).ctypes.data if b is not None else None
    f = lib().zko_g1_op if group == 1 else lib().zko_g2_op
    f(_i(op), _vp(a.ctypes.data), _vp(bp), _vp(out.ctypes.data), _sz(a.size // w))
    return out


def scalar_mul(group: int, scalars, points) -> np.ndarray:
    w = 12 if group == 1 else 24
    s, p = _u(scalars, 4), _u(points, w)
    out = np.empty_like(p)
    f = lib().zko_g1_scalar_mul if group == 1 else lib().zko_g2_scalar_mul
    f(_vp(s.ctypes.data), _vp(p.ctypes.data), _vp(out.ctypes.data), _sz(p.size // w))
    return out


def msm(group: int, bases_affine, scalars) -> np.ndarray:
    """sum_i s_i P_i, normalised (Z = 1 or the zero (0, 1, 0))"""
    w = 8 if group == 1 else 16
    b, s = _u(bases_affine, w), _u(scalars, 4)
    out = np.zeros(12 if group == 1 else 24, dtype=np.uint64)
    f = lib().zko_g1_msm if group == 1 else lib().zko_g2_msm
    f(_vp(b.ctypes.data), _vp(s.ctypes.data), _sz(s.size // 4), _vp(out.ctypes.data))
    return out


def domain_select(min_size: int):
    m, k = C.c_size_t(), C.c_int()
    if lib().zko_domain_select(_sz(min_size), C.byref(m), C.byref(k)):
        raise RuntimeError("no supported domain")
    return int(m.value), int(k.value)


def domain_op(min_size: int, op: int, data) -> np.ndarray:
    d = np.array(_u(data, 4), copy=True)
    if lib().zko_domain_op(_sz(min_size), _i(op), _vp(d.ctypes.data)):
        raise RuntimeError("domain op failed")
    return d


def _desc(key: dict):
    keep = []

    def ptr(a, dtype=np.uint64):
        a = np.ascontiguousarray(a, dtype=dtype)
        keep.append(a)
        return a.ctypes.data

    d = _KeyDesc()
    d.num_constraints, d.num_variables, d.num_inputs = key["num_constraints"], key["num_variables"], key["num_inputs"]
    for name in "ABC":
        rp, col, co = key["csr_" + name]
        setattr(d, name, _CSR(ptr(rp), ptr(col, np.uint32), ptr(co)))
    for f in ("alpha_g1", "beta_g1", "delta_g1", "beta_g2", "delta_g2", "A_query", "B_g1", "B_g2", "H_query", "L_query"):
        setattr(d, f, ptr(key[f]))
    d.B_index = ptr(key["B_index"], np.uint32)
    d.B_nnz = len(key["B_index"])
    d.H_size = len(key["H_query"])
    return d, keep


def witness_map(key: dict, w) -> np.ndarray:
    d, keep = _desc(key)
    w = _u(w, 4)
    m, _ = domain_select(key["num_constraints"] + key["num_inputs"] + 1)
    H = np.zeros((m + 1, 4), dtype=np.uint64)
    mm = C.c_size_t()
    rc = lib().zko_witness_map(C.byref(d), _vp(w.ctypes.data), _vp(H.ctypes.data), C.byref(mm))
    if rc < 0:
        raise RuntimeError("witness map failed")
    return H


def prove(key: dict, w, r, s):
    """-> (proof (48 limbs, every point normalised), H (m + 1, 4), satisfied)"""
    d, keep = _desc(key)
    w, r, s = _u(w, 4), _u(r, 4), _u(s, 4)
    m, _ = domain_select(key["num_constraints"] + key["num_inputs"] + 1)
    H = np.zeros((m + 1, 4), dtype=np.uint64)
    proof = np.zeros(48, dtype=np.uint64)
    rc = lib().zko_prove(C.byref(d), _vp(w.ctypes.data), _vp(r.ctypes.data), _vp(s.ctypes.data), _vp(proof.ctypes.data),
                         _vp(H.ctypes.data))
    if rc < 0:
        raise RuntimeError("prove failed")
    return proof, H, rc == 0
