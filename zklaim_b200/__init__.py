"""zklaim_b200 -- host-side (Python) mirror of the zkb C ABI (include/zkb.h).

The product is ``libzkb.so`` (hand-written sm_100a CUDA behind a plain C ABI, see
``include/zkb.h``); this module is the thin ctypes layer used by the tests and by
``bench.py``.  There is no CPU fallback: importing works without a GPU (so that the
symbol table can be checked), but creating a :class:`Context` raises when no CUDA device
is present or when the library has not been built.

Array conventions (numpy, dtype uint64, C-contiguous), identical to libff's memory layout:
  Fr / Fq      (..., 4)   Montgomery limbs     (libff/algebra/fields/fp.hpp:42)
  Fq2          (..., 8)
  G1 affine    (..., 8)   x, y; all-zero = infinity
  G2 affine    (..., 16)
  G1 Jacobian  (..., 12)  X, Y, Z              (alt_bn128_g1.hpp:35)
  G2 Jacobian  (..., 24)
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libzkb.so")

FR, FQ, FQ2 = 0, 1, 2
G1, G2 = 1, 2
OP_ADD, OP_SUB, OP_MUL, OP_SQR, OP_INV, OP_NEG = range(6)
GOP_ADD, GOP_DBL, GOP_MIXED_ADD, GOP_NEG, GOP_NORMALIZE = range(5)
NTT_FFT, NTT_IFFT, NTT_COSET_FFT, NTT_ICOSET_FFT, NTT_DIV_Z_ON_COSET = range(5)
DOMAIN_BASIC_RADIX2, DOMAIN_EXTENDED_RADIX2, DOMAIN_STEP_RADIX2 = 1, 2, 3
OK, ERR_CUDA, ERR_ARG, ERR_DOMAIN, ERR_UNSAT = 0, -1, -2, -3, -4

# every symbol include/zkb.h declares: (name, restype, argtypes)
_vp, _sz, _i, _u64p = C.c_void_p, C.c_size_t, C.c_int, C.c_void_p
SYMBOLS = {
    "zkb_init": (_i, [_i, C.POINTER(_vp)]),
    "zkb_shutdown": (None, [_vp]),
    "zkb_last_error": (C.c_char_p, [_vp]),
    "zkb_version": (C.c_char_p, []),
    "zkb_field_op": (_i, [_vp, _i, _i, _u64p, _u64p, _u64p, _sz]),
    "zkb_group_op": (_i, [_vp, _i, _i, _u64p, _u64p, _u64p, _sz]),
    "zkb_group_scalar_mul": (_i, [_vp, _i, _u64p, _u64p, _u64p, _sz]),
    "zkb_group_sum": (_i, [_vp, _i, _u64p, _sz, _u64p]),
    "zkb_points_decompress": (_i, [_vp, _i, _u64p, _vp, _sz, _u64p, _vp]),
    "zkb_bases_upload": (_i, [_vp, _i, _u64p, _sz, _i, _i, C.POINTER(_vp)]),
    "zkb_bases_synthetic": (_i, [_vp, _sz, C.c_uint64, _i, _i, C.POINTER(_vp)]),
    "zkb_bases_free": (None, [_vp]),
    "zkb_bases_device_bytes": (_sz, [_vp]),
    "zkb_msm": (_i, [_vp, _vp, _sz, _u64p, _sz, _u64p]),
    "zkb_msm_device": (_i, [_vp, _vp, _sz, _vp, _sz, _u64p]),
    "zkb_domain_select": (_i, [_sz, C.POINTER(_sz), C.POINTER(_i)]),
    "zkb_domain_op": (_i, [_vp, _sz, _i, _u64p]),
    "zkb_fft_noalloc_max_log2": (_i, []),
    "zkb_fft_noalloc": (_i, [_vp, _i, _i, _vp, _vp]),
    "zkb_key_upload": (_i, [_vp, _vp, C.POINTER(_vp)]),
    "zkb_key_free": (None, [_vp]),
    "zkb_key_device_bytes": (_sz, [_vp]),
    "zkb_key_domain": (_i, [_vp, C.POINTER(_sz), C.POINTER(_i)]),
    "zkb_key_num_variables": (_i, [_vp, C.POINTER(_sz)]),
    "zkb_device_count": (_i, []),
    "zkb_pool_init": (_i, [_vp, _i, C.POINTER(_vp)]),
    "zkb_pool_shutdown": (None, [_vp]),
    "zkb_pool_size": (_i, [_vp]),
    "zkb_pool_ctx": (_vp, [_vp, _i]),
    "zkb_pool_last_error": (C.c_char_p, [_vp]),
    "zkb_pool_key_upload": (_i, [_vp, _vp, C.POINTER(_vp)]),
    "zkb_pool_key_upload_sharded": (_i, [_vp, _vp, C.POINTER(_vp)]),
    "zkb_pool_key_tune": (_i, [_vp, _vp, _u64p]),
    "zkb_pool_key": (_vp, [_vp, _i]),
    "zkb_pool_key_free": (None, [_vp]),
    "zkb_pool_prove_batch": (_i, [_vp, _vp, _sz, _u64p, _u64p, _u64p, _vp, _vp]),
    "zkb_prove_batch_status": (_i, [_vp, _vp, _sz]),
    "zkb_pool_prove_sharded": (_i, [_vp, _vp, _u64p, _u64p, _u64p, _vp]),
    "zkb_pool_bases_upload": (_i, [_vp, _i, _u64p, _sz, _i, _i, C.POINTER(_vp)]),
    "zkb_pool_bases_free": (None, [_vp]),
    "zkb_pool_msm": (_i, [_vp, _vp, _u64p, _sz, _u64p]),
    "zkb_key_upload_shard": (_i, [_vp, _vp, _i, _i, C.POINTER(_vp)]),
    "zkb_prove_partial": (_i, [_vp, _vp, _u64p, _u64p, _u64p, _u64p, _u64p]),
    "zkb_prove_combine": (_i, [_vp, _u64p, _sz, _u64p, _u64p, _vp]),
    "zkb_prove_split_begin": (_i, [_vp, _vp, _u64p, _u64p, _u64p, _i, _i, _vp]),
    "zkb_prove_split_chain": (_i, [_vp, _vp, _i, _vp]),
    "zkb_prove_split_combine": (_i, [_vp, _vp, _vp, _vp]),
    "zkb_prove_split_end": (_i, [_vp, _vp, _vp, _u64p]),
    "zkb_key_tune": (_i, [_vp, _vp, _u64p]),
    "zkb_key_windows": (_i, [_vp, C.POINTER(_i)]),
    "zkb_setup": (_i, [_vp, _vp, _vp, C.POINTER(_i)]),
    "zkb_witness_map": (_i, [_vp, _vp, _u64p, _u64p]),
    "zkb_prove": (_i, [_vp, _vp, _u64p, _u64p, _u64p, _vp, _u64p]),
    "zkb_prove_batch": (_i, [_vp, _vp, _sz, _u64p, _u64p, _u64p, _vp]),
    "zkb_prove_batch_device": (_i, [_vp, _vp, _sz, _vp, _vp, _vp, _vp]),
    "zkb_prove_batch_submit": (_i, [_vp, _vp, _sz, _u64p, _u64p, _u64p]),
    "zkb_prove_batch_collect": (_i, [_vp, _vp]),
    "zkb_pairing": (_i, [_vp, _u64p, _u64p, _sz, _i, _u64p]),
    "zkb_verify_batch": (_i, [_vp, _vp, _u64p, _u64p, _sz, _vp]),
    "zkb_measure_imad_peak": (C.c_double, [_vp, _i]),
    "zkb_profile": (_i, [_vp, _i]),
    "zkb_profile_read": (_i, [_vp, C.POINTER(C.c_double)]),
    "zkb_last_device_ms": (C.c_double, [_vp]),
    "zkb_last_kernel_launches": (C.c_long, [_vp]),
}


class ZkbError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"zkb error {code}: {msg}")
        self.code = code


_lib = None


def load_library(path: Optional[str] = None):
    """dlopen libzkb.so and bind every symbol of include/zkb.h.  Raises if it is missing."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise ZkbError(-1, f"{p} not built (run zklaim_b200/build.sh or __graft_entry__.build()); "
                           "there is no CPU fallback")
    lib = C.CDLL(p)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    if path is None:
        _lib = lib
    return lib


def _arr(a, last: int) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.uint64)
    if a.shape[-1] != last:
        raise ValueError(f"expected last dimension {last}, got {a.shape}")
    return a


class _CSR(C.Structure):
    _fields_ = [("row_ptr", C.c_void_p), ("col", C.c_void_p), ("coeff", C.c_void_p)]


class _KeyDesc(C.Structure):
    _fields_ = [
        ("num_constraints", C.c_size_t), ("num_variables", C.c_size_t), ("num_inputs", C.c_size_t),
        ("A", _CSR), ("B", _CSR), ("C", _CSR),
        ("alpha_g1", C.c_void_p), ("beta_g1", C.c_void_p), ("delta_g1", C.c_void_p),
        ("beta_g2", C.c_void_p), ("delta_g2", C.c_void_p),
        ("A_query", C.c_void_p), ("B_nnz", C.c_size_t), ("B_index", C.c_void_p),
        ("B_g1", C.c_void_p), ("B_g2", C.c_void_p),
        ("H_size", C.c_size_t), ("H_query", C.c_void_p), ("L_query", C.c_void_p),
    ]


class _VkDesc(C.Structure):
    _fields_ = [
        ("alpha_g1_beta_g2", C.c_void_p), ("gamma_g2", C.c_void_p), ("delta_g2", C.c_void_p),
        ("gamma_ABC_g1", C.c_void_p), ("num_inputs", C.c_uint64),
    ]


class _SetupDesc(C.Structure):
    _fields_ = [
        ("num_constraints", C.c_size_t), ("num_variables", C.c_size_t), ("num_inputs", C.c_size_t),
        ("A", _CSR), ("B", _CSR), ("C", _CSR),
        ("t", C.c_void_p), ("alpha", C.c_void_p), ("beta", C.c_void_p), ("gamma", C.c_void_p), ("delta", C.c_void_p),
        ("g1_generator", C.c_void_p), ("g2_generator", C.c_void_p),
    ]


_SETUP_OUT = ("alpha_g1", "beta_g1", "delta_g1", "beta_g2", "delta_g2", "gamma_g2", "A_query", "B_g1", "B_g2", "H_query",
              "L_query", "gamma_ABC_g1")


class _SetupOut(C.Structure):
    _fields_ = [(f, C.c_void_p) for f in _SETUP_OUT]


def _key_desc(key: dict):
    """zkb_key_desc over the numpy arrays of `key`; returns (desc, the arrays it points into)"""
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
    if "A_query" in key:
        for f in ("alpha_g1", "beta_g1", "delta_g1", "beta_g2", "delta_g2", "A_query", "B_g1", "B_g2", "H_query",
                  "L_query"):
            setattr(d, f, ptr(key[f]))
        d.B_index = ptr(key["B_index"], np.uint32)
        d.B_nnz = len(key["B_index"])
        d.H_size = len(key["H_query"])
    # else: constraint system only (all query pointers NULL): the key serves witness_map alone
    return d, keep


class Proof(C.Structure):
    _fields_ = [("A", C.c_uint64 * 12), ("B", C.c_uint64 * 24), ("C", C.c_uint64 * 12)]

    def as_array(self) -> np.ndarray:
        return np.concatenate([np.array(self.A, dtype=np.uint64), np.array(self.B, dtype=np.uint64),
                               np.array(self.C, dtype=np.uint64)])


class Context:
    """One zkb_ctx: a CUDA stream plus scratch memory on one device."""

    def __init__(self, device: int = 0):
        self.lib = load_library()
        h = C.c_void_p()
        rc = self.lib.zkb_init(device, C.byref(h))
        if rc:
            raise ZkbError(rc, self.lib.zkb_last_error(None).decode())
        self.h = h
        self.device = device

    @classmethod
    def from_handle(cls, handle) -> "Context":
        """View of a zkb_ctx that somebody else owns (the bridge's); never shut down from here."""
        self = cls.__new__(cls)
        self.lib = load_library()
        self.h = C.c_void_p(handle)
        self.device = -1
        self.borrowed = True
        return self

    def close(self):
        if getattr(self, "h", None) and not getattr(self, "borrowed", False):
            self.lib.zkb_shutdown(self.h)
        self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _chk(self, rc: int):
        if rc:
            raise ZkbError(rc, self.lib.zkb_last_error(self.h).decode())

    # ---- element-wise -----------------------------------------------------------------
    def field_op(self, field: int, op: int, a, b=None) -> np.ndarray:
        w = 8 if field == FQ2 else 4
        a = _arr(a, w)
        out = np.empty_like(a)
        bp = None
        if b is not None:
            b = _arr(b, w)
            assert b.shape == a.shape
            bp = b.ctypes.data
        self._chk(self.lib.zkb_field_op(self.h, field, op, a.ctypes.data, bp, out.ctypes.data, a.size // w))
        return out

    def group_op(self, group: int, op: int, a, b=None) -> np.ndarray:
        w = 12 if group == G1 else 24
        a = _arr(a, w)
        out = np.empty_like(a)
        bp = None
        if b is not None:
            b = _arr(b, w)
            assert b.shape == a.shape
            bp = b.ctypes.data
        self._chk(self.lib.zkb_group_op(self.h, group, op, a.ctypes.data, bp, out.ctypes.data, a.size // w))
        return out

    def scalar_mul(self, group: int, scalars, points) -> np.ndarray:
        w = 12 if group == G1 else 24
        s = _arr(scalars, 4)
        p = _arr(points, w)
        out = np.empty_like(p)
        self._chk(self.lib.zkb_group_scalar_mul(self.h, group, s.ctypes.data, p.ctypes.data, out.ctypes.data,
                                                p.size // w))
        return out

    # ---- MSM --------------------------------------------------------------------------
    def group_sum(self, group: int, points) -> np.ndarray:
        """sum of Jacobian points (multiexp.tcc:433-438): the combine step of a range-sharded MSM"""
        w = 12 if group == G1 else 24
        p = _arr(points, w)
        out = np.zeros(w, dtype=np.uint64)
        self._chk(self.lib.zkb_group_sum(self.h, group, p.ctypes.data, p.size // w, out.ctypes.data))
        return out

    def points_decompress(self, group: int, x, flags):
        """y = +-sqrt(x^3 + b) with the reference's parity rule (alt_bn128_g1.cpp:436-450, alt_bn128_g2.cpp:449-463).
        x: n x 4 (G1) / n x 8 (G2) limbs; flags: n bytes, bit 0 = lsb of Y, bit 1 = is_zero.
        Returns (affine points n x 8 / n x 16, status n: 1 = ok, 0 = x is not on the curve); raises on a bad x."""
        w = 4 if group == G1 else 8
        xa = _arr(x, w)
        n = xa.size // w
        fl = np.ascontiguousarray(flags, dtype=np.uint8)
        assert fl.size == n
        out = np.zeros((n, 2 * w), dtype=np.uint64)
        st = np.zeros(n, dtype=np.int32)
        self._chk(self.lib.zkb_points_decompress(self.h, group, xa.ctypes.data, fl.ctypes.data, n, out.ctypes.data, st.ctypes.data))
        return out, st

    def upload_bases(self, group: int, affine, window_bits: int = 0, precompute_factor: int = 0) -> "Bases":
        w = 8 if group == G1 else 16
        a = _arr(affine, w)
        h = C.c_void_p()
        self._chk(self.lib.zkb_bases_upload(self.h, group, a.ctypes.data, a.size // w, window_bits,
                                            precompute_factor, C.byref(h)))
        return Bases(self, h, group, a.size // w)

    def synthetic_bases(self, n: int, seed: int = 1, window_bits: int = 0, precompute_factor: int = 0) -> "Bases":
        """n distinct G1 bases made on the device (bench support, zkb_bases_synthetic)"""
        h = C.c_void_p()
        self._chk(self.lib.zkb_bases_synthetic(self.h, n, seed, window_bits, precompute_factor, C.byref(h)))
        return Bases(self, h, G1, n)

    def msm(self, bases: "Bases", scalars, offset: int = 0) -> np.ndarray:
        s = _arr(scalars, 4)
        out = np.zeros(12 if bases.group == G1 else 24, dtype=np.uint64)
        self._chk(self.lib.zkb_msm(self.h, bases.h, offset, s.ctypes.data, s.size // 4, out.ctypes.data))
        return out

    def msm_device(self, bases: "Bases", d_scalars_ptr: int, n: int, offset: int = 0) -> np.ndarray:
        out = np.zeros(12 if bases.group == G1 else 24, dtype=np.uint64)
        self._chk(self.lib.zkb_msm_device(self.h, bases.h, offset, d_scalars_ptr, n, out.ctypes.data))
        return out

    # ---- domains ----------------------------------------------------------------------
    def domain_op(self, min_size: int, op: int, data) -> np.ndarray:
        d = np.array(_arr(data, 4), copy=True)
        self._chk(self.lib.zkb_domain_op(self.h, min_size, op, d.ctypes.data))
        return d

    def fft_noalloc(self, log2n: int, inverse: bool, d_in: int, d_out: int) -> None:
        """zkb_fft_noalloc: FFT / iFFT of 2^log2n Fr elements between DEVICE buffers (CUDA pointers, e.g. torch data_ptr()),
        with all scratch in a static device array of libzkb.so: no allocation happens in the call"""
        self._chk(self.lib.zkb_fft_noalloc(self.h, log2n, 1 if inverse else 0, C.c_void_p(d_in), C.c_void_p(d_out)))

    # ---- key / prover -----------------------------------------------------------------
    def upload_key(self, key: dict, rank: int = 0, world: int = 1) -> "Key":
        """key: dict with the fields of zkb_key_desc as numpy arrays (bridge: zkb_libsnark.hpp flattens a libsnark proving key
        into them).  world > 1: only shard `rank`'s range of every query goes to the device (zkb_key_upload_shard).
        A dict without the query arrays uploads the constraint system only (witness_map)."""
        d, keep = _key_desc(key)
        h = C.c_void_p()
        self._chk(self.lib.zkb_key_upload_shard(self.h, C.byref(d), rank, world, C.byref(h)))
        return Key(self, h, key["num_variables"])

    def setup(self, cs: dict, toxic, g1_gen, g2_gen) -> dict:
        """zkb_setup: the Groth16 generator (r1cs_gg_ppzksnark.tcc:211-388) for the given secret randomness.
        cs: num_constraints / num_variables / num_inputs + csr_A / csr_B / csr_C as built (not swapped);
        toxic: (5, 4) limbs = t, alpha, beta, gamma, delta; generators as Jacobian limbs.
        Returns the fields of zkb_key_desc (constraint system swapped when `swapped`, B_query sparse like the
        reference's) plus gamma_g2, gamma_ABC_g1 of the verification key and `swapped`."""
        keep = []

        def ptr(a, dtype=np.uint64):
            a = np.ascontiguousarray(a, dtype=dtype)
            keep.append(a)
            return a.ctypes.data

        d = _SetupDesc()
        nc, nv, ni = cs["num_constraints"], cs["num_variables"], cs["num_inputs"]
        d.num_constraints, d.num_variables, d.num_inputs = nc, nv, ni
        for name in "ABC":
            rp, col, co = cs["csr_" + name]
            setattr(d, name, _CSR(ptr(rp), ptr(col, np.uint32), ptr(co)))
        toxic = _arr(toxic, 4)
        if toxic.shape != (5, 4):
            raise ValueError("toxic: expected 5 x 4 limbs")
        for i, f in enumerate(("t", "alpha", "beta", "gamma", "delta")):
            setattr(d, f, ptr(toxic[i]))
        d.g1_generator, d.g2_generator = ptr(_arr(g1_gen, 12)), ptr(_arr(g2_gen, 24))
        m = C.c_size_t()
        kind = C.c_int()
        self._chk(self.lib.zkb_domain_select(nc + ni + 1, C.byref(m), C.byref(kind)))
        shapes = {"alpha_g1": (8,), "beta_g1": (8,), "delta_g1": (8,), "beta_g2": (16,), "delta_g2": (16,), "gamma_g2": (16,),
                  "A_query": (nv + 1, 8), "B_g1": (nv + 1, 8), "B_g2": (nv + 1, 16), "H_query": (m.value - 1, 8),
                  "L_query": (nv - ni, 8), "gamma_ABC_g1": (ni + 1, 8)}
        res = {f: np.zeros(shapes[f], dtype=np.uint64) for f in _SETUP_OUT}
        o = _SetupOut()
        for f in _SETUP_OUT:
            a = res[f] if res[f].size else np.zeros(8, dtype=np.uint64)
            keep.append(a)
            setattr(o, f, a.ctypes.data)
        swapped = C.c_int()
        self._chk(self.lib.zkb_setup(self.h, C.byref(d), C.byref(o), C.byref(swapped)))
        nz = np.flatnonzero(res["B_g1"].any(axis=1))  # kc_batch_exp keeps the non-zero entries (kc_multiexp.tcc:108-116)
        res["B_index"] = nz.astype(np.uint32)
        res["B_g1"], res["B_g2"] = res["B_g1"][nz], res["B_g2"][nz]
        a, b = ("B", "A") if swapped.value else ("A", "B")
        res.update(num_constraints=nc, num_variables=nv, num_inputs=ni, csr_A=cs["csr_" + a], csr_B=cs["csr_" + b],
                   csr_C=cs["csr_C"], swapped=bool(swapped.value))
        return res

    def prove_partial(self, key: "Key", w, r, s) -> np.ndarray:
        """this shard's share of the four sums of a proof: 60 uint64 = A (12), B1 (12), C (12), B2 (24) Jacobian limbs"""
        w, r, s = _arr(w, 4), _arr(r, 4), _arr(s, 4)
        out = np.zeros(60, dtype=np.uint64)
        self._chk(self.lib.zkb_prove_partial(self.h, key.h, w.ctypes.data, r.ctypes.data, s.ctypes.data, out.ctypes.data, None))
        return out

    def prove_split_begin(self, key: "Key", w, r, s, which: int, check: bool, d_part: int = 0) -> None:
        """zkb_prove_split_begin: d_part = CUDA device pointer (m x 4 limbs) when which in (0, 1, 2); w = host array, or an
        int = CUDA device pointer to the num_variables x 4 limbs on this context's GPU"""
        r, s = _arr(r, 4), _arr(s, 4)
        if isinstance(w, int):
            wp = C.c_void_p(w)
        else:
            w = _arr(w, 4)
            wp = C.c_void_p(w.ctypes.data)
        self._chk(self.lib.zkb_prove_split_begin(self.h, key.h, wp, r.ctypes.data, s.ctypes.data, which, 1 if check else 0,
                                                 C.c_void_p(d_part)))

    def prove_split_chain(self, key: "Key", which: int, d_part: int) -> None:
        """zkb_prove_split_chain: one more polynomial (0 / 1 / 2) to the coset on a context that has run prove_split_begin"""
        self._chk(self.lib.zkb_prove_split_chain(self.h, key.h, which, C.c_void_p(d_part)))

    def prove_split_combine(self, key: "Key", d_abc: int, d_H: int) -> None:
        self._chk(self.lib.zkb_prove_split_combine(self.h, key.h, C.c_void_p(d_abc), C.c_void_p(d_H)))

    def prove_split_end(self, key: "Key", d_H: int) -> np.ndarray:
        out = np.zeros(60, dtype=np.uint64)
        self._chk(self.lib.zkb_prove_split_end(self.h, key.h, C.c_void_p(d_H), out.ctypes.data))
        return out

    def prove_combine(self, parts, r, s) -> np.ndarray:
        """adds up the shares of all shards and assembles the proof (r1cs_gg_ppzksnark.tcc:487-495)"""
        parts, r, s = _arr(parts, 60), _arr(r, 4), _arr(s, 4)
        pr = Proof()
        self._chk(self.lib.zkb_prove_combine(self.h, parts.ctypes.data, parts.size // 60, r.ctypes.data, s.ctypes.data, C.byref(pr)))
        return pr.as_array()

    def witness_map(self, key: "Key", w) -> np.ndarray:
        w = _arr(w, 4)
        m, _ = key.domain()
        H = np.zeros((m + 1, 4), dtype=np.uint64)
        self._chk(self.lib.zkb_witness_map(self.h, key.h, w.ctypes.data, H.ctypes.data))
        return H

    def prove(self, key: "Key", w, r, s, want_H: bool = False):
        w = _arr(w, 4)
        r = _arr(r, 4)
        s = _arr(s, 4)
        pr = Proof()
        H = None
        hp = None
        if want_H:
            m, _ = key.domain()
            H = np.zeros((m + 1, 4), dtype=np.uint64)
            hp = H.ctypes.data
        self._chk(self.lib.zkb_prove(self.h, key.h, w.ctypes.data, r.ctypes.data, s.ctypes.data, C.byref(pr), hp))
        return (pr.as_array(), H) if want_H else pr.as_array()

    def prove_batch(self, key: "Key", w, r, s) -> np.ndarray:
        w = _arr(w, 4)
        r = _arr(r, 4)
        s = _arr(s, 4)
        count = r.shape[0]
        assert w.shape[0] == count and s.shape[0] == count
        out = (Proof * count)()
        self._chk(self.lib.zkb_prove_batch(self.h, key.h, count, w.ctypes.data, r.ctypes.data, s.ctypes.data, out))
        return np.frombuffer(out, dtype=np.uint64).reshape(count, 48).copy()

    def prove_batch_device(self, key: "Key", d_w: int, d_r: int, d_s: int, count: int) -> np.ndarray:
        """w, r, s are CUDA device pointers (e.g. torch tensors' data_ptr()) on this context's device."""
        out = (Proof * count)()
        self._chk(self.lib.zkb_prove_batch_device(self.h, key.h, count, d_w, d_r, d_s, out))
        return np.frombuffer(out, dtype=np.uint64).reshape(count, 48).copy()

    def prove_batch_submit(self, key: "Key", w, r, s) -> None:
        """starts the host-to-device copy of one batch (zkb_prove_batch_submit); at most two batches in flight.
        The arrays are kept alive here until the matching prove_batch_collect."""
        w, r, s = _arr(w, 4), _arr(r, 4), _arr(s, 4)
        count = r.shape[0]
        assert w.shape[0] == count and s.shape[0] == count
        self._chk(self.lib.zkb_prove_batch_submit(self.h, key.h, count, w.ctypes.data, r.ctypes.data, s.ctypes.data))
        self._inflight = getattr(self, "_inflight", [])
        self._inflight.append((w, r, s, count))

    def prove_batch_collect(self) -> np.ndarray:
        """proofs of the oldest submitted batch (zkb_prove_batch_collect)"""
        if not getattr(self, "_inflight", None):
            raise ZkbError(ERR_ARG, "nothing submitted")
        count = self._inflight[0][3]
        out = (Proof * count)()
        try:
            self._chk(self.lib.zkb_prove_batch_collect(self.h, out))
        finally:
            self._inflight.pop(0)
        return np.frombuffer(out, dtype=np.uint64).reshape(count, 48).copy()

    # ---- measurement ------------------------------------------------------------------
    def imad_peak(self, iters: int = 512) -> float:
        return float(self.lib.zkb_measure_imad_peak(self.h, iters))

    def profile(self, enable: bool = True):
        self._chk(self.lib.zkb_profile(self.h, int(enable)))

    def profile_read(self) -> dict:
        out = (C.c_double * 6)()
        self._chk(self.lib.zkb_profile_read(self.h, out))
        return {"g1_ms": out[0], "g1_launches": int(out[1]), "g1_pairs": out[2],
                "g2_ms": out[3], "g2_launches": int(out[4]), "g2_pairs": out[5]}

    # ---- verifier side -------------------------------------------------------------------
    def pairing(self, g1_affine, g2_affine, reduced: bool = True) -> np.ndarray:
        """alt_bn128_ate_miller_loop / alt_bn128_reduced_pairing (alt_bn128_pairing.cpp:371-420, 523-533) of n affine
        pairs -> (n, 48) limbs, bit-identical to libff's Fq12"""
        p = _arr(g1_affine, 8).reshape(-1, 8)
        q = _arr(g2_affine, 16).reshape(-1, 16)
        assert len(p) == len(q)
        out = np.zeros((len(p), 48), dtype=np.uint64)
        self._chk(self.lib.zkb_pairing(self.h, p.ctypes.data, q.ctypes.data, len(p), int(reduced), out.ctypes.data))
        return out

    def verify_batch(self, vk: dict, primary, proofs) -> np.ndarray:
        """r1cs_gg_ppzksnark_verifier_strong_IC (r1cs_gg_ppzksnark.tcc:523-622) for n proofs under one key.
        vk: alpha_g1_beta_g2 (48), gamma_g2 (16), delta_g2 (16), gamma_ABC_g1 (ni + 1, 8); primary (n, ni, 4);
        proofs (n, 48) as zkb_prove writes them.  Returns n booleans."""
        ab = _arr(vk["alpha_g1_beta_g2"], 48)
        g, d = _arr(vk["gamma_g2"], 16), _arr(vk["delta_g2"], 16)
        abc = _arr(vk["gamma_ABC_g1"], 8).reshape(-1, 8)
        ni = len(abc) - 1
        pr = _arr(proofs, 48).reshape(-1, 48)
        n = len(pr)
        x = _arr(primary, 4).reshape(n, ni, 4) if ni else np.zeros((n, 0, 4), dtype=np.uint64)
        desc = _VkDesc(ab.ctypes.data, g.ctypes.data, d.ctypes.data, abc.ctypes.data, ni)
        acc = np.zeros(n, dtype=np.int32)
        self._chk(self.lib.zkb_verify_batch(self.h, C.byref(desc), x.ctypes.data if ni else None, pr.ctypes.data, n,
                                            acc.ctypes.data))
        return acc.astype(bool)

    def last_device_ms(self) -> float:
        return float(self.lib.zkb_last_device_ms(self.h))

    def kernel_launches(self) -> int:
        return int(self.lib.zkb_last_kernel_launches(self.h))


class Bases:
    def __init__(self, ctx: Context, h, group: int, n: int):
        self.ctx, self.h, self.group, self.n = ctx, h, group, n

    def device_bytes(self) -> int:
        return int(self.ctx.lib.zkb_bases_device_bytes(self.h))

    def free(self):
        if self.h:
            self.ctx.lib.zkb_bases_free(self.h)
            self.h = None


class Key:
    def __init__(self, ctx: Context, h, nv: int):
        self.ctx, self.h, self.num_variables = ctx, h, nv

    def domain(self):
        m, k = C.c_size_t(), C.c_int()
        self.ctx._chk(self.ctx.lib.zkb_key_domain(self.h, C.byref(m), C.byref(k)))
        return int(m.value), int(k.value)

    def device_bytes(self) -> int:
        return int(self.ctx.lib.zkb_key_device_bytes(self.h))

    def tune(self, w) -> "Key":
        """zkb_key_tune: re-plan the MSM windows for the 0 / 1 statistics of the sample witness w (proofs unchanged)."""
        w = _arr(w, 4)
        if w.shape[0] != self.num_variables:
            raise ValueError("witness length != num_variables")
        self.ctx._chk(self.ctx.lib.zkb_key_tune(self.ctx.h, self.h, w.ctypes.data))
        return self

    def windows(self):
        """window bits planned for the A, B-in-G1, H + L and B-in-G2 queries"""
        out = (C.c_int * 4)()
        self.ctx._chk(self.ctx.lib.zkb_key_windows(self.h, out))
        return list(out)

    def free(self):
        if self.h:
            self.ctx.lib.zkb_key_free(self.h)
            self.h = None


class Pool:
    """zkb_pool: one context per device, one host thread per device (include/zkb.h).  devices=None: every visible GPU;
    a device may be listed more than once (several contexts on one GPU: how the 1-GPU test tier exercises this path)."""

    def __init__(self, devices=None):
        self.lib = load_library()
        ids = (C.c_int * len(devices))(*devices) if devices else None
        h = C.c_void_p()
        rc = self.lib.zkb_pool_init(ids, len(devices) if devices else 0, C.byref(h))
        if rc != OK:
            raise ZkbError(rc, (self.lib.zkb_last_error(None) or b"").decode())
        self.h = h
        self.size = int(self.lib.zkb_pool_size(self.h))

    def _chk(self, rc: int):
        if rc != OK:
            raise ZkbError(rc, (self.lib.zkb_pool_last_error(self.h) or b"").decode())

    def ctx(self, i: int) -> Context:
        return Context.from_handle(self.lib.zkb_pool_ctx(self.h, i))

    def upload_key(self, key: dict, sharded: bool = False):
        d, keep = _key_desc(key)
        h = C.c_void_p()
        fn = self.lib.zkb_pool_key_upload_sharded if sharded else self.lib.zkb_pool_key_upload
        self._chk(fn(self.h, C.byref(d), C.byref(h)))
        return h

    def tune_key(self, keys, w):
        w = _arr(w, 4)
        self._chk(self.lib.zkb_pool_key_tune(self.h, keys, w.ctypes.data))

    def free_key(self, keys):
        self.lib.zkb_pool_key_free(keys)

    def prove_batch(self, keys, w, r, s):
        """returns (proofs count x 48, status count: 0 or ERR_UNSAT); raises on any error but ERR_UNSAT"""
        w, r, s = _arr(w, 4), _arr(r, 4), _arr(s, 4)
        count = r.shape[0]
        assert w.shape[0] == count and s.shape[0] == count
        out = (Proof * count)()
        st = np.zeros(count, dtype=np.int32)
        rc = self.lib.zkb_pool_prove_batch(self.h, keys, count, w.ctypes.data, r.ctypes.data, s.ctypes.data, out, st.ctypes.data)
        if rc not in (OK, ERR_UNSAT):
            self._chk(rc)
        return np.frombuffer(out, dtype=np.uint64).reshape(count, 48).copy(), st

    def prove_sharded(self, keys, w, r, s) -> np.ndarray:
        w, r, s = _arr(w, 4), _arr(r, 4), _arr(s, 4)
        pr = Proof()
        self._chk(self.lib.zkb_pool_prove_sharded(self.h, keys, w.ctypes.data, r.ctypes.data, s.ctypes.data, C.byref(pr)))
        return pr.as_array()

    def upload_bases(self, group: int, affine, window_bits: int = 0, precompute_factor: int = 0):
        w = 8 if group == G1 else 16
        a = _arr(affine, w)
        h = C.c_void_p()
        self._chk(self.lib.zkb_pool_bases_upload(self.h, group, a.ctypes.data, a.size // w, window_bits, precompute_factor, C.byref(h)))
        return h, group

    def msm(self, bases, scalars) -> np.ndarray:
        h, group = bases
        s = _arr(scalars, 4)
        out = np.zeros(12 if group == G1 else 24, dtype=np.uint64)
        self._chk(self.lib.zkb_pool_msm(self.h, h, s.ctypes.data, s.size // 4, out.ctypes.data))
        return out

    def free_bases(self, bases):
        self.lib.zkb_pool_bases_free(bases[0])

    def close(self):
        if self.h:
            self.lib.zkb_pool_shutdown(self.h)
            self.h = None


def device_count() -> int:
    return int(load_library().zkb_device_count())


def domain_select(min_size: int):
    """libfqfft::get_evaluation_domain's choice (m, kind) for min_size; host only."""
    lib = load_library()
    m, k = C.c_size_t(), C.c_int()
    rc = lib.zkb_domain_select(min_size, C.byref(m), C.byref(k))
    if rc:
        raise ZkbError(rc, "no supported evaluation domain")
    return int(m.value), int(k.value)
