"""ctypes binding of oracle/_ref/libzkref*.so -- the UNMODIFIED reference (libff / libfqfft /
libsnark / zklaim sources compiled in place by oracle/Makefile) behind oracle/ref_harness.cpp.

TEST INFRASTRUCTURE: imported only by tests/, __graft_entry__.smoke() and the cpu_baseline /
--impl reference legs of bench.py, as the checker and the timed CPU baseline.  Never by the
product path (zklaim_b200/).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_libs = {}
_vp, _sz, _i, _u64, _d = C.c_void_p, C.c_size_t, C.c_int, C.c_uint64, C.c_double

_SIG = {
    "zkref_init": (_i, []),
    "zkref_num_threads": (_i, []),
    "zkref_build_flags": (_i, []),
    "zkref_field_op": (None, [_i, _i, _vp, _vp, _vp, _sz]),
    "zkref_fq2_op": (None, [_i, _vp, _vp, _vp, _sz]),
    "zkref_field_from_plain": (None, [_i, _vp, _vp, _sz]),
    "zkref_field_to_plain": (None, [_i, _vp, _vp, _sz]),
    "zkref_sha512_rng": (None, [_i, _u64, _sz, _vp]),
    "zkref_field_const": (None, [_i, _i, _vp]),
    "zkref_g1_op": (None, [_i, _vp, _vp, _vp, _sz]),
    "zkref_g2_op": (None, [_i, _vp, _vp, _vp, _sz]),
    "zkref_g1_equal": (_i, [_vp, _vp]),
    "zkref_g2_equal": (_i, [_vp, _vp]),
    "zkref_g1_well_formed": (_i, [_vp]),
    "zkref_g2_well_formed": (_i, [_vp]),
    "zkref_g1_scalar_mul": (None, [_vp, _vp, _vp, _sz]),
    "zkref_g2_scalar_mul": (None, [_vp, _vp, _vp, _sz]),
    "zkref_g1_generator": (None, [_vp]),
    "zkref_g2_generator": (None, [_vp]),
    "zkref_g1_bases": (None, [_u64, _sz, _vp]),
    "zkref_g2_bases": (None, [_u64, _sz, _vp]),
    "zkref_g1_msm": (_d, [_vp, _vp, _sz, _i, _sz, _vp]),
    "zkref_g2_msm": (_d, [_vp, _vp, _sz, _i, _sz, _vp]),
    "zkref_domain_select": (_i, [_sz, C.POINTER(_sz), C.POINTER(_i)]),
    "zkref_domain_op": (_d, [_sz, _i, _vp]),
    "zkref_domain_element": (None, [_sz, _sz, _vp]),
    "zkref_circuit_zklaim": (_vp, [_sz, _u64]),
    "zkref_zklaim_witness": (_i, [_sz, _u64, _vp, _sz]),
    "zkref_circuit_example": (_vp, [_sz, _sz, _u64]),
    "zkref_circuit_sha_chain": (_vp, [_sz, _u64]),
    "zkref_circuit_free": (None, [_vp]),
    "zkref_circuit_dims": (None, [_vp, _vp]),
    "zkref_circuit_is_satisfied": (_i, [_vp]),
    "zkref_circuit_csr": (None, [_vp, _i, _vp, _vp, _vp]),
    "zkref_circuit_witness": (None, [_vp, _vp]),
    "zkref_circuit_set_witness": (None, [_vp, _vp]),
    "zkref_circuit_setup": (_d, [_vp]),
    "zkref_key_dims": (None, [_vp, _vp]),
    "zkref_key_export": (None, [_vp] * 8),
    "zkref_key_import": (None, [_vp, _vp, _sz, _vp, _vp, _vp, _sz, _vp, _sz, _vp, _sz, _vp]),
    "zkref_circuit_setup_fixed": (_d, [_vp, _vp, _vp, _vp]),
    "zkref_vk_export": (None, [_vp, _vp, _vp, _vp]),
    "zkref_vk_import": (None, [_vp, _vp, _vp, _vp, _vp, _vp, _sz]),
    "zkref_witness_map": (_d, [_vp, _vp, _vp]),
    "zkref_prove": (_i, [_vp] * 8),
    "zkref_prove_stock": (_d, [_vp, _vp, _vp]),
    "zkref_verify": (_i, [_vp, _vp, _vp]),
    "zkref_proof_serialize": (_sz, [_vp, _vp, _sz]),
    "zkref_bridge_roundtrip": (_d, [_sz, _u64, C.POINTER(_i)]),
    "zkref_pairing": (None, [_vp, _vp, _i, _vp]),
    "zkref_vk_alpha_beta": (None, [_vp, _vp]),
}


def available(mt: bool = False) -> bool:
    return os.path.exists(os.path.join(_HERE, "_ref", "libzkref_mt.so" if mt else "libzkref.so"))


def lib(mt: bool = False):
    """mt=False: deterministic single-thread build without asm (checker);
    mt=True: the reference's fast build (-DUSE_ASM -DMULTICORE), for timing."""
    if mt not in _libs:
        path = os.path.join(_HERE, "_ref", "libzkref_mt.so" if mt else "libzkref.so")
        L = C.CDLL(path)
        for name, (res, args) in _SIG.items():
            f = getattr(L, name)
            f.restype, f.argtypes = res, args
        L.zkref_init()
        _libs[mt] = L
    return _libs[mt]


def _u(a, last=None):
    a = np.ascontiguousarray(a, dtype=np.uint64)
    if last is not None:
        assert a.shape[-1] == last, (a.shape, last)
    return a


def sha512_rng(first: int, n: int, field: int = 0, mt=False) -> np.ndarray:
    out = np.zeros((n, 4), dtype=np.uint64)
    lib(mt).zkref_sha512_rng(field, first, n, out.ctypes.data)
    return out


def field_op(field: int, op: int, a, b=None) -> np.ndarray:
    a = _u(a, 4)
    out = np.empty_like(a)
    bp = _u(b, 4).ctypes.data if b is not None else None
    lib().zkref_field_op(field, op, a.ctypes.data, bp, out.ctypes.data, a.size // 4)
    return out


def fq2_op(op: int, a, b=None) -> np.ndarray:
    a = _u(a, 8)
    out = np.empty_like(a)
    bp = _u(b, 8).ctypes.data if b is not None else None
    lib().zkref_fq2_op(op, a.ctypes.data, bp, out.ctypes.data, a.size // 8)
    return out


def field_const(field: int, which: int) -> np.ndarray:
    out = np.zeros(4, dtype=np.uint64)
    lib().zkref_field_const(field, which, out.ctypes.data)
    return out


def from_plain(field: int, plain) -> np.ndarray:
    p = _u(plain, 4)
    out = np.empty_like(p)
    lib().zkref_field_from_plain(field, p.ctypes.data, out.ctypes.data, p.size // 4)
    return out


def to_plain(field: int, a) -> np.ndarray:
    a = _u(a, 4)
    out = np.empty_like(a)
    lib().zkref_field_to_plain(field, a.ctypes.data, out.ctypes.data, a.size // 4)
    return out


def group_op(group: int, op: int, a, b=None) -> np.ndarray:
    w = 12 if group == 1 else 24
    a = _u(a, w)
    out = np.empty_like(a)
    bp = _u(b, w).ctypes.data if b is not None else None
    f = lib().zkref_g1_op if group == 1 else lib().zkref_g2_op
    f(op, a.ctypes.data, bp, out.ctypes.data, a.size // w)
    return out


def group_equal(group: int, a, b) -> bool:
    w = 12 if group == 1 else 24
    a, b = _u(a, w), _u(b, w)
    f = lib().zkref_g1_equal if group == 1 else lib().zkref_g2_equal
    return bool(f(a.ctypes.data, b.ctypes.data))


def to_affine_jac(group: int, a) -> np.ndarray:
    """canonical representative (Z = 1, or the reference's zero) -- op 4 of zkref_g?_op"""
    return group_op(group, 4, a)


def scalar_mul(group: int, scalars, points) -> np.ndarray:
    w = 12 if group == 1 else 24
    s, p = _u(scalars, 4), _u(points, w)
    out = np.empty_like(p)
    f = lib().zkref_g1_scalar_mul if group == 1 else lib().zkref_g2_scalar_mul
    f(s.ctypes.data, p.ctypes.data, out.ctypes.data, p.size // w)
    return out


def generator(group: int) -> np.ndarray:
    out = np.zeros(12 if group == 1 else 24, dtype=np.uint64)
    (lib().zkref_g1_generator if group == 1 else lib().zkref_g2_generator)(out.ctypes.data)
    return out


def bases(group: int, first: int, n: int, mt=True) -> np.ndarray:
    """n distinct affine bases k_i * G, k_i = SHA512_rng<Fr>(first + i)"""
    L = lib(mt and available(True))
    out = np.zeros((n, 8 if group == 1 else 16), dtype=np.uint64)
    (L.zkref_g1_bases if group == 1 else L.zkref_g2_bases)(first, n, out.ctypes.data)
    return out


def msm(group: int, bases_affine, scalars, mixed: bool = False, chunks: int = 1, mt=False):
    """returns (Jacobian result, seconds)"""
    w = 8 if group == 1 else 16
    b, s = _u(bases_affine, w), _u(scalars, 4)
    n = s.size // 4
    out = np.zeros(12 if group == 1 else 24, dtype=np.uint64)
    L = lib(mt)
    f = L.zkref_g1_msm if group == 1 else L.zkref_g2_msm
    t = f(b.ctypes.data, s.ctypes.data, n, int(mixed), chunks, out.ctypes.data)
    return out, t


def domain_select(min_size: int):
    m, k = C.c_size_t(), C.c_int()
    rc = lib().zkref_domain_select(min_size, C.byref(m), C.byref(k))
    if rc:
        raise RuntimeError("reference found no domain")
    return int(m.value), int(k.value)


def domain_op(min_size: int, op: int, data, mt=False):
    d = np.array(_u(data, 4), copy=True)
    t = lib(mt).zkref_domain_op(min_size, op, d.ctypes.data)
    if t < 0:
        raise RuntimeError("reference domain op failed")
    return d, t


class Circuit:
    """A constraint system + satisfying assignment (+ optionally a Groth16 key) held by the reference."""

    def __init__(self, handle, mt=False):
        self.h = handle
        self.L = lib(mt)

    @staticmethod
    def zklaim(payloads: int = 1, seed: int = 1, mt=False) -> "Circuit":
        return Circuit(lib(mt).zkref_circuit_zklaim(payloads, seed), mt)

    @staticmethod
    def example(num_constraints: int, num_inputs: int, seed: int = 1, mt=False) -> "Circuit":
        return Circuit(lib(mt).zkref_circuit_example(num_constraints, num_inputs, seed), mt)

    @staticmethod
    def sha_chain(blocks: int, seed: int = 1, mt=False) -> "Circuit":
        return Circuit(lib(mt).zkref_circuit_sha_chain(blocks, seed), mt)

    def free(self):
        if self.h:
            self.L.zkref_circuit_free(self.h)
            self.h = None

    def dims(self) -> dict:
        d = (C.c_size_t * 8)()
        self.L.zkref_circuit_dims(self.h, d)
        names = ["num_constraints", "num_variables", "num_inputs", "nnzA", "nnzB", "nnzC", "m", "kind"]
        return dict(zip(names, [int(x) for x in d]))

    def is_satisfied(self) -> bool:
        return bool(self.L.zkref_circuit_is_satisfied(self.h))

    def csr(self, which: int):
        d = self.dims()
        nnz = d[["nnzA", "nnzB", "nnzC"][which]]
        rp = np.zeros(d["num_constraints"] + 1, dtype=np.uint64)
        col = np.zeros(max(nnz, 1), dtype=np.uint32)
        co = np.zeros((max(nnz, 1), 4), dtype=np.uint64)
        self.L.zkref_circuit_csr(self.h, which, rp.ctypes.data, col.ctypes.data, co.ctypes.data)
        return rp, col[:nnz], co[:nnz]

    def witness(self) -> np.ndarray:
        w = np.zeros((self.dims()["num_variables"], 4), dtype=np.uint64)
        self.L.zkref_circuit_witness(self.h, w.ctypes.data)
        return w

    def setup(self) -> float:
        return float(self.L.zkref_circuit_setup(self.h))

    def setup_fixed(self, toxic, g1_gen, g2_gen) -> float:
        """the reference generator with injected randomness: toxic = (t, alpha, beta, gamma, delta) as 5 x 4 limbs,
        generators as Jacobian limbs (12 / 24)"""
        t, g1, g2 = _u(toxic, 4), _u(g1_gen), _u(g2_gen)
        assert t.shape == (5, 4) and g1.size == 12 and g2.size == 24
        return float(self.L.zkref_circuit_setup_fixed(self.h, t.ctypes.data, g1.ctypes.data, g2.ctypes.data))

    def export_vk(self) -> dict:
        ni = self.dims()["num_inputs"]
        g, d, abc = np.zeros(16, dtype=np.uint64), np.zeros(16, dtype=np.uint64), np.zeros((ni + 1, 8), dtype=np.uint64)
        self.L.zkref_vk_export(self.h, g.ctypes.data, d.ctypes.data, abc.ctypes.data)
        return {"gamma_g2": g, "delta_g2": d, "gamma_ABC_g1": abc}

    def vk_alpha_beta(self) -> np.ndarray:
        """vk.alpha_g1_beta_g2 as 12 Fq (48 limbs, libff member order)"""
        out = np.zeros(48, dtype=np.uint64)
        self.L.zkref_vk_alpha_beta(self.h, out.ctypes.data)
        return out

    def import_vk(self, alpha_g1, beta_g2, gamma_g2, delta_g2, gamma_abc):
        a, b, g, d, abc = _u(alpha_g1), _u(beta_g2), _u(gamma_g2), _u(delta_g2), _u(gamma_abc, 8)
        self.L.zkref_vk_import(self.h, a.ctypes.data, b.ctypes.data, g.ctypes.data, d.ctypes.data, abc.ctypes.data, len(abc))

    def export_key(self) -> dict:
        """the proving key as the flat arrays of zkb_key_desc"""
        d = self.dims()
        s = (C.c_size_t * 5)()
        self.L.zkref_key_dims(self.h, s)
        nA, nB, nH, nL, _ = [int(x) for x in s]
        A = np.zeros((nA, 8), dtype=np.uint64)
        B1 = np.zeros((max(nB, 1), 8), dtype=np.uint64)
        B2 = np.zeros((max(nB, 1), 16), dtype=np.uint64)
        Bi = np.zeros(max(nB, 1), dtype=np.uint32)
        H = np.zeros((nH, 8), dtype=np.uint64)
        Lq = np.zeros((nL, 8), dtype=np.uint64)
        sg = np.zeros(56, dtype=np.uint64)
        self.L.zkref_key_export(self.h, A.ctypes.data, B1.ctypes.data, B2.ctypes.data, Bi.ctypes.data, H.ctypes.data,
                                Lq.ctypes.data, sg.ctypes.data)
        return {
            "num_constraints": d["num_constraints"], "num_variables": d["num_variables"], "num_inputs": d["num_inputs"],
            "csr_A": self.csr(0), "csr_B": self.csr(1), "csr_C": self.csr(2),
            "alpha_g1": sg[0:8].copy(), "beta_g1": sg[8:16].copy(), "delta_g1": sg[16:24].copy(),
            "beta_g2": sg[24:40].copy(), "delta_g2": sg[40:56].copy(),
            "A_query": A, "B_index": Bi[:nB], "B_g1": B1[:nB], "B_g2": B2[:nB], "H_query": H, "L_query": Lq,
        }

    def import_key(self, key: dict):
        A, B1, B2, Bi, H, Lq = (_u(key["A_query"]), _u(key["B_g1"]), _u(key["B_g2"]),
                                np.ascontiguousarray(key["B_index"], dtype=np.uint32), _u(key["H_query"]),
                                _u(key["L_query"]))
        sg = np.concatenate([_u(key[k]).ravel() for k in ("alpha_g1", "beta_g1", "delta_g1", "beta_g2", "delta_g2")])
        self.L.zkref_key_import(self.h, A.ctypes.data, len(A), B1.ctypes.data, B2.ctypes.data, Bi.ctypes.data, len(Bi),
                                H.ctypes.data, len(H), Lq.ctypes.data, len(Lq), sg.ctypes.data)

    def witness_map(self, w):
        w = _u(w, 4)
        H = np.zeros((self.dims()["m"] + 1, 4), dtype=np.uint64)
        t = self.L.zkref_witness_map(self.h, w.ctypes.data, H.ctypes.data)
        return H, t

    def prove(self, w, r, s, want_H=True):
        """reference prover with fixed r, s -> dict(proof(48), evals(72), H, times)"""
        w, r, s = _u(w, 4), _u(r, 4), _u(s, 4)
        proof = np.zeros(48, dtype=np.uint64)
        evals = np.zeros(72, dtype=np.uint64)
        times = np.zeros(6, dtype=np.float64)
        H = np.zeros((self.dims()["m"] + 1, 4), dtype=np.uint64) if want_H else None
        rc = self.L.zkref_prove(self.h, w.ctypes.data, r.ctypes.data, s.ctypes.data, proof.ctypes.data,
                                evals.ctypes.data, H.ctypes.data if want_H else None, times.ctypes.data)
        if rc:
            raise RuntimeError("reference prove failed (no key?)")
        return {"proof": proof, "evals": evals, "H": H, "times": times}

    def prove_stock(self, w) -> float:
        w = _u(w, 4)
        return float(self.L.zkref_prove_stock(self.h, w.ctypes.data, None))

    def verify(self, primary, proof) -> bool:
        p, pr = _u(primary, 4), _u(proof)
        return bool(self.L.zkref_verify(self.h, p.ctypes.data, pr.ctypes.data))


def pairing(g1_affine, g2_affine, reduced: bool = True) -> np.ndarray:
    """libff's alt_bn128_ate_miller_loop / alt_bn128_reduced_pairing of one affine pair -> 48 limbs"""
    p, q = _u(g1_affine), _u(g2_affine)
    assert p.size == 8 and q.size == 16
    out = np.zeros(48, dtype=np.uint64)
    lib().zkref_pairing(p.ctypes.data, q.ctypes.data, int(reduced), out.ctypes.data)
    return out


def proof_bytes(proof) -> bytes:
    pr = _u(proof)
    buf = (C.c_ubyte * 512)()
    n = lib().zkref_proof_serialize(pr.ctypes.data, buf, 512)
    return bytes(buf[:n])


def zklaim_witness(payloads: int, seed: int, nv: int) -> np.ndarray:
    w = np.zeros((nv, 4), dtype=np.uint64)
    rc = lib().zkref_zklaim_witness(payloads, seed, w.ctypes.data, nv)
    if rc:
        raise RuntimeError(f"zklaim witness generation failed ({rc})")
    return w
