"""ctypes binding of bridge/_build/libzkb_zklaim.so -- zklaim's own host code (credential gadgets, trusted
setup, (de)serialisation, verifier: compiled in place from the reference tree by bridge/Makefile) with the
Groth16 prover replaced by libzkb.so.  See bridge/zkb_zklaim_bridge.cpp and INTEGRATION.md.

This is the host side of the drop-in: the GPU never sees anything but what zkb.h describes.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import Context, Key, Proof, ZkbError, load_library

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "bridge", "_build", "libzkb_zklaim.so")
_vp, _sz, _u64, _i = C.c_void_p, C.c_size_t, C.c_uint64, C.c_int
_SIG = {
    "zkb_libsnark_prove": (_i, [_vp]),
    "zkb_libsnark_prove_batch": (_i, [_vp, _sz]),
    "zkb_bridge_reset": (None, []),
    "zkbr_new": (_vp, [_sz, _u64]),
    "zkbr_free": (None, [_vp]),
    "zkbr_dims": (None, [_vp, _vp]),
    "zkbr_key": (_vp, [_vp]),
    "zkbr_ctx": (_vp, []),
    "zkbr_witnesses": (_i, [_vp, _sz, _u64, _vp]),
    "zkbr_random_fr": (None, [_sz, _vp]),
    "zkbr_verify": (_i, [_vp, _vp, _vp]),
    "zkbr_verify_roundtrip": (_i, [_vp, _sz, _u64, _sz, C.POINTER(_i), C.POINTER(C.c_double)]),
    "zkbr_roundtrip": (_i, [_vp, _sz, _u64, _vp]),
    "zkbr_prove_tampered": (_i, [_vp, _u64]),
    "zkbr_setup_roundtrip": (_i, [_sz, _u64, _i, _vp]),
    "zkbr_import_check": (_i, [_vp, _vp]),
    "zkbr_stream": (_i, [_vp, _sz, _sz, _i, _u64, _vp]),
    "zkbr_batch_with_tampered": (_i, [_vp, _sz, _u64, _sz, C.POINTER(_i)]),
    "zkbr_cold_prove": (_i, [_vp, _u64, _vp]),
    "zkb_libsnark_devices": (_i, []),
    "zkbr_synth_sha_chain": (_vp, [_sz, _u64]),
    "zkbr_synth_free": (None, [_vp]),
    "zkbr_synth_dims": (None, [_vp, _vp]),
    "zkbr_synth_csr": (None, [_vp, _i, _vp, _vp, _vp]),
    "zkbr_synth_witness": (None, [_vp, _vp]),
    "zkbr_generators": (None, [_vp, _vp]),
}
_lib = None


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise ZkbError(-1, f"{LIB_PATH} not built (make -C bridge; needs the reference tree)")
        load_library()  # libzkb.so first, so that the bridge binds to the same copy
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIG.items():
            f = getattr(L, name)
            f.restype, f.argtypes = res, args
        _lib = L
    return _lib


class Credentials:
    """A zklaim deployment: one issuer setup for `payloads`-payload credentials, its key resident on the GPU."""

    def __init__(self, payloads: int = 1, seed: int = 1):
        self.L = lib()
        self.h = self.L.zkbr_new(payloads, seed)
        if not self.h:
            raise ZkbError(-1, "bridge setup failed (no GPU, or trusted setup error)")
        d = (C.c_size_t * 6)()
        self.L.zkbr_dims(self.h, d)
        self.num_variables, self.num_inputs, self.pk_bytes, self.vk_bytes = [int(x) for x in d[:4]]
        self.setup_s, self.import_upload_s = d[4] / 1e6, d[5] / 1e6
        self.payloads = payloads
        # the bridge's own zkb context / device key, viewed through the Python mirror
        self.ctx = Context.from_handle(self.L.zkbr_ctx())
        self.key = Key(self.ctx, C.c_void_p(self.L.zkbr_key(self.h)), self.num_variables)

    def witnesses(self, count: int, seed0: int = 1000) -> np.ndarray:
        w = np.zeros((count, self.num_variables, 4), dtype=np.uint64)
        rc = self.L.zkbr_witnesses(self.h, count, seed0, w.ctypes.data)
        if rc:
            raise ZkbError(rc, "witness generation failed")
        return w

    def random_fr(self, count: int) -> np.ndarray:
        out = np.zeros((count, 4), dtype=np.uint64)
        self.L.zkbr_random_fr(count, out.ctypes.data)
        return out

    def verify(self, w: np.ndarray, proof: np.ndarray) -> bool:
        w = np.ascontiguousarray(w, dtype=np.uint64)
        p = np.ascontiguousarray(proof, dtype=np.uint64)
        assert p.size == 48
        return bool(self.L.zkbr_verify(self.h, w.ctypes.data, p.ctypes.data))

    def roundtrip(self, count: int, seed0: int = 5000):
        """count credentials through zkb_libsnark_prove(_batch) and the reference's libsnark_verify.
        Returns (accepted, prove_seconds, verify_seconds)."""
        t = (C.c_double * 2)()
        ok = self.L.zkbr_roundtrip(self.h, count, seed0, t)
        return ok, t[0], t[1]

    def verify_roundtrip(self, count: int, seed0: int = 9000, tamper: int = -1):
        """count credentials proved by zkb_libsnark_prove_batch, then verified by the reference's libsnark_verify AND by
        zkb_libsnark_verify_batch (GPU); credential `tamper` carries its neighbour's proof.
        Returns (credentials on which both verifiers agree, accepted by the GPU verifier, reference s, GPU s)."""
        t = (C.c_double * 2)()
        acc = C.c_int(0)
        agree = self.L.zkbr_verify_roundtrip(self.h, count, seed0, tamper if tamper >= 0 else count, C.byref(acc), t)
        return agree, acc.value, t[0], t[1]

    def prove_tampered(self, seed: int = 77) -> int:
        """zkb_libsnark_prove on a credential whose hash was tampered with: the reference's contract says 1"""
        return int(self.L.zkbr_prove_tampered(self.h, seed))

    def batch_with_tampered(self, count: int, seed0: int, tamper: int):
        """count credentials, credential `tamper` with a flipped hash byte, through zkb_libsnark_prove_batch_ex.
        Returns (credentials that behaved as the reference's libsnark_prove / libsnark_verify would, the call's result)"""
        rc = C.c_int(0)
        good = int(self.L.zkbr_batch_with_tampered(self.h, count, seed0, tamper, C.byref(rc)))
        return good, rc.value

    def stream(self, batches: int, batch: int = 64, threads: int = 2, seed0: int = 70000):
        """`threads` host threads x `batches` batches of `batch` credentials through zkb_libsnark_prove_batch (host witness
        generation by zklaim's gadget code + GPU).  Returns (credentials proved, wall seconds)."""
        t = (C.c_double * 1)()
        n = int(self.L.zkbr_stream(self.h, batches, batch, threads, seed0, t))
        if n < 0:
            raise ZkbError(n, "zkbr_stream failed")
        return n, t[0]

    def import_check(self):
        """the issuer's serialised proving key through the reference importer and through the deferred (GPU) importer.
        Returns (0 if the flat keys are identical, reference import s, deferred import s)"""
        t = (C.c_double * 2)()
        rc = int(self.L.zkbr_import_check(self.h, t))
        return rc, t[0], t[1]

    def cold_prove(self, seed: int = 31):
        """first zkb_libsnark_prove under a never-seen key (caches dropped).  NOTE: invalidates self.ctx / self.key
        (the session is rebuilt); returns (rc, cold s, warm s)"""
        t = (C.c_double * 2)()
        rc = int(self.L.zkbr_cold_prove(self.h, seed, t))
        self.ctx = Context.from_handle(self.L.zkbr_ctx())
        self.key = Key(self.ctx, C.c_void_p(self.L.zkbr_key(self.h)), self.num_variables)
        return rc, t[0], t[1]

    def close(self):
        if self.h:
            self.L.zkbr_free(self.h)
            self.h = None


def setup_roundtrip(payloads: int = 1, seed: int = 1, stock_prover: bool = False):
    """zkb_libsnark_trusted_setup (generator on the GPU) -> a proof under that key (the reference's unmodified CPU
    libsnark_prove when stock_prover, else zkb_libsnark_prove) -> the reference's libsnark_verify.
    Returns (libsnark_verify's result: 0 = accepted, setup s, prove s, verify s)."""
    import ctypes as C
    t = (C.c_double * 3)()
    rc = int(lib().zkbr_setup_roundtrip(payloads, seed, 1 if stock_prover else 0, t))
    return rc, t[0], t[1], t[2]


def sha_chain(blocks: int, seed: int = 1):
    """BASELINE config #5: `blocks` chained libsnark SHA-256 compression gadgets with a satisfying assignment.
    Returns (constraint system as the dict zkb setup / key upload take, witness num_variables x 4)."""
    L = lib()
    h = L.zkbr_synth_sha_chain(blocks, seed)
    if not h:
        raise ZkbError(-1, "sha chain generation failed")
    try:
        d = (C.c_size_t * 6)()
        L.zkbr_synth_dims(h, d)
        nc, nv, ni = int(d[0]), int(d[1]), int(d[2])
        cs = {"num_constraints": nc, "num_variables": nv, "num_inputs": ni}
        for which, name in enumerate("ABC"):
            nnz = int(d[3 + which])
            rp = np.zeros(nc + 1, dtype=np.uint64)
            col = np.zeros(max(nnz, 1), dtype=np.uint32)
            co = np.zeros((max(nnz, 1), 4), dtype=np.uint64)
            L.zkbr_synth_csr(h, which, rp.ctypes.data, col.ctypes.data, co.ctypes.data)
            cs["csr_" + name] = (rp, col[:nnz], co[:nnz])
        w = np.zeros((nv, 4), dtype=np.uint64)
        L.zkbr_synth_witness(h, w.ctypes.data)
        return cs, w
    finally:
        L.zkbr_synth_free(h)


def generators():
    """(G1 generator 12 limbs, G2 generator 24 limbs), Jacobian, Montgomery"""
    g1, g2 = np.zeros(12, dtype=np.uint64), np.zeros(24, dtype=np.uint64)
    lib().zkbr_generators(g1.ctypes.data, g2.ctypes.data)
    return g1, g2


def random_fr(count: int) -> np.ndarray:
    out = np.zeros((count, 4), dtype=np.uint64)
    lib().zkbr_random_fr(count, out.ctypes.data)
    return out
