"""ctypes binding of include/b200prover.h (numpy arrays of u64 limbs in, numpy arrays out).

Layouts (identical to gnark-crypto's in-memory ones, see the header): Fr/Fp element = 4 u64
Montgomery limbs; G1 affine (n, 8); G1 Jacobian (12,); G2 affine (n, 16); G2 Jacobian (24,).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

BN254 = 0
DIF = 0
DIT = 1

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class B200Error(RuntimeError):
    pass


def library_path() -> str:
    return os.environ.get("B200PROVER_LIB", os.path.join(_HERE, "libb200prover.so"))


class _Timings(C.Structure):
    _fields_ = [(k, C.c_float) for k in
                ("h2d_ms", "compute_h_ms", "msm_a_ms", "msm_b1_ms", "msm_k_ms", "msm_z_ms", "msm_b2_ms", "total_ms")]


class Groth16MsmOut(C.Structure):
    """b200_groth16_out: the five MultiExp results before the fix-ups (prove.go:201-290)."""
    _fields_ = [("ar", C.c_uint64 * 12), ("bs1", C.c_uint64 * 12), ("krs", C.c_uint64 * 12),
                ("krs2", C.c_uint64 * 12), ("bs2", C.c_uint64 * 24)]

    def as_dict(self):
        return {k: np.array(getattr(self, k), dtype=np.uint64) for k, _ in self._fields_}


class _PkDesc(C.Structure):
    _fields_ = [
        ("log_n", C.c_uint), ("generator", C.c_void_p), ("coset_gen", C.c_void_p),
        ("g1_a", C.c_void_p), ("n_a", C.c_size_t), ("g1_b", C.c_void_p), ("n_b", C.c_size_t),
        ("g1_k", C.c_void_p), ("n_k", C.c_size_t), ("g1_z", C.c_void_p), ("n_z", C.c_size_t),
        ("g2_b", C.c_void_p), ("n_g2b", C.c_size_t),
        ("infinity_a", C.c_void_p), ("infinity_b", C.c_void_p),
        ("n_wires", C.c_size_t), ("nb_public", C.c_size_t),
        ("shard_index", C.c_int), ("shard_count", C.c_int),
    ]


def load_library():
    """Load libb200prover.so; fails loudly (no fallback of any kind)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = library_path()
    if not os.path.exists(path):
        raise B200Error(f"{path} not found: build it with `make -C gnark-fork_b200/csrc` "
                        "(or __graft_entry__.build()); there is no CPU fallback")
    lib = C.CDLL(path)
    vp, sz, u, i = C.c_void_p, C.c_size_t, C.c_uint, C.c_int
    sig = {
        "b200_device_count": (i, []),
        "b200_init": (i, [i, C.POINTER(vp)]),
        "b200_destroy": (None, [vp]),
        "b200_last_error": (C.c_char_p, [vp]),
        "b200_version": (C.c_char_p, []),
        "b200_host_alloc": (vp, [sz]),
        "b200_host_free": (None, [vp]),
        "b200_dev_alloc": (i, [vp, sz, C.POINTER(vp)]),
        "b200_dev_free": (i, [vp, vp]),
        "b200_h2d": (i, [vp, vp, vp, sz]),
        "b200_d2h": (i, [vp, vp, vp, sz]),
        "b200_sync": (i, [vp]),
        "b200_fr_vec_mul": (i, [vp, i, vp, vp, vp, sz]),
        "b200_fp_vec_mul": (i, [vp, i, vp, vp, vp, sz]),
        "b200_imad_peak": (i, [vp, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double)]),
        "b200_ntt": (i, [vp, i, vp, u, i, i, i, vp, vp]),
        "b200_ntt_dev": (i, [vp, i, vp, u, i, i, i, vp, vp, C.POINTER(C.c_float)]),
        "b200_groth16_compute_h": (i, [vp, i, vp, vp, vp, sz, u, vp, vp, vp]),
        "b200_msm_g1": (i, [vp, i, vp, vp, sz, vp]),
        "b200_msm_g2": (i, [vp, i, vp, vp, sz, vp]),
        "b200_bases_upload": (i, [vp, i, i, vp, sz, C.POINTER(vp)]),
        "b200_bases_free": (i, [vp, vp]),
        "b200_msm_resident": (i, [vp, vp, vp, sz, vp]),
        "b200_msm_dev": (i, [vp, vp, vp, sz, vp, C.POINTER(C.c_float)]),
        "b200_g1_jac_add": (i, [vp, vp, vp]),
        "b200_g2_jac_add": (i, [vp, vp, vp]),
        "b200_g1_jac_to_affine": (i, [vp, vp]),
        "b200_g2_jac_to_affine": (i, [vp, vp]),
        "b200_batch_scalar_mul": (i, [vp, i, i, vp, vp, sz, vp]),
        "b200_groth16_pk_upload": (i, [vp, i, C.POINTER(_PkDesc), C.POINTER(vp)]),
        "b200_groth16_pk_free": (i, [vp, vp]),
        "b200_groth16_prove_core": (i, [vp, vp, vp, vp, vp, sz, vp, vp, sz, C.POINTER(Groth16MsmOut), C.POINTER(_Timings)]),
        "b200_groth16_finalize": (i, [C.POINTER(Groth16MsmOut), vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
        "b200_groth16_proof_bytes": (i, [vp, vp, vp, vp, sz]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)  # AttributeError if the library lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    _LIB = lib
    return lib


EXPORTED_SYMBOLS = None  # filled lazily by tests from the header


def _u64(a, shape=None) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.uint64)
    if shape is not None:
        a = a.reshape(shape)
    return a


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def pinned_empty(shape, dtype=np.uint64) -> np.ndarray:
    """numpy array over page-locked memory from b200_host_alloc (prove.go:119-125 buffers)."""
    lib = load_library()
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    p = lib.b200_host_alloc(max(n, 1))
    if not p:
        raise B200Error("b200_host_alloc failed")
    buf = (C.c_uint8 * max(n, 1)).from_address(p)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    _PINNED[arr.ctypes.data] = p
    return arr


_PINNED: dict = {}
