"""ctypes loader for oracle/liboracle.so (TEST INFRASTRUCTURE ONLY; see cpu_ref.cpp header).

numpy layout conventions shared with the product's C ABI: an Fr/Fp element is 4 little-endian
u64 limbs in Montgomery form; G1 affine = (n, 8) u64; G1 Jacobian = 12 u64; G2 affine = (n, 16);
G2 Jacobian = 24 u64.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

from . import bn254 as C

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build() -> str:
    subprocess.check_call(["make", "-C", _HERE, "-s"])
    return os.path.join(_HERE, "liboracle.so")


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        _LIB = ctypes.CDLL(path)
        _LIB.oracle_max_threads.restype = ctypes.c_int
    return _LIB


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


# ---- int <-> limb arrays ----------------------------------------------------------------
def ints_to_mont(xs, mod) -> np.ndarray:
    """list of canonical ints -> (n,4) u64 Montgomery limbs."""
    out = np.empty((len(xs), 4), dtype=np.uint64)
    buf = b"".join(((x * C.MONT_R) % mod).to_bytes(32, "little") for x in xs)
    out[:] = np.frombuffer(buf, dtype=np.uint64).reshape(-1, 4)
    return out


def mont_to_ints(a: np.ndarray, mod) -> list[int]:
    rinv = pow(C.MONT_R, -1, mod)
    b = np.ascontiguousarray(a, dtype=np.uint64).reshape(-1, 4).tobytes()
    return [(int.from_bytes(b[i:i + 32], "little") * rinv) % mod for i in range(0, len(b), 32)]


def fr(x: int) -> np.ndarray:
    return ints_to_mont([x % C.R], C.R)[0]


def g1_points_to_array(pts) -> np.ndarray:
    flat = []
    for p in pts:
        flat += [0, 0] if p is None else [p[0], p[1]]
    return ints_to_mont(flat, C.P).reshape(-1, 8)


def g2_points_to_array(pts) -> np.ndarray:
    flat = []
    for p in pts:
        flat += [0, 0, 0, 0] if p is None else [p[0][0], p[0][1], p[1][0], p[1][1]]
    return ints_to_mont(flat, C.P).reshape(-1, 16)


def g1_jac_array_to_affine(j: np.ndarray):
    X, Y, Z = mont_to_ints(j.reshape(3, 4), C.P)
    return C.jac_to_affine_g1(X, Y, Z)


def g2_jac_array_to_affine(j: np.ndarray):
    v = mont_to_ints(j.reshape(6, 4), C.P)
    return C.jac_to_affine_g2((v[0], v[1]), (v[2], v[3]), (v[4], v[5]))


def g1_array_to_points(a: np.ndarray):
    v = mont_to_ints(a.reshape(-1, 4), C.P)
    return [None if (v[i] == 0 and v[i + 1] == 0) else (v[i], v[i + 1]) for i in range(0, len(v), 2)]


def g2_array_to_points(a: np.ndarray):
    v = mont_to_ints(a.reshape(-1, 4), C.P)
    out = []
    for i in range(0, len(v), 4):
        out.append(None if not any(v[i:i + 4]) else ((v[i], v[i + 1]), (v[i + 2], v[i + 3])))
    return out


# ---- wrapped entry points ---------------------------------------------------------------
def fr_mul(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    out = np.empty_like(a)
    lib().oracle_fr_mul(_p(a), _p(b), _p(out), ctypes.c_size_t(a.shape[0]))
    return out


def fp_mul(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    out = np.empty_like(a)
    lib().oracle_fp_mul(_p(a), _p(b), _p(out), ctypes.c_size_t(a.shape[0]))
    return out


def ntt(data: np.ndarray, log_n: int, inverse: bool, decimation: int, on_coset: bool,
        gen: np.ndarray, coset_gen: np.ndarray) -> np.ndarray:
    out = np.ascontiguousarray(data.copy())
    lib().oracle_ntt(_p(out), ctypes.c_uint(log_n), int(inverse), int(decimation), int(on_coset),
                     _p(gen), _p(coset_gen))
    return out


def compute_h(a, b, c, log_n: int, gen, coset_gen) -> np.ndarray:
    n = 1 << log_n
    def pad(x):
        o = np.zeros((n, 4), dtype=np.uint64)
        o[: x.shape[0]] = x
        return o
    A, B, Cc = pad(a), pad(b), pad(c)
    lib().oracle_compute_h(_p(A), _p(B), _p(Cc), ctypes.c_uint(log_n), _p(gen), _p(coset_gen))
    return A


def msm_g1(scalars: np.ndarray, bases: np.ndarray) -> np.ndarray:
    out = np.zeros(12, dtype=np.uint64)
    lib().oracle_msm_g1(_p(scalars), _p(bases), ctypes.c_size_t(scalars.shape[0]), _p(out))
    return out


def msm_g2(scalars: np.ndarray, bases: np.ndarray) -> np.ndarray:
    out = np.zeros(24, dtype=np.uint64)
    lib().oracle_msm_g2(_p(scalars), _p(bases), ctypes.c_size_t(scalars.shape[0]), _p(out))
    return out


def g1_batch_mul(base: np.ndarray, scalars: np.ndarray) -> np.ndarray:
    out = np.zeros((scalars.shape[0], 8), dtype=np.uint64)
    lib().oracle_g1_batch_mul(_p(base), _p(scalars), ctypes.c_size_t(scalars.shape[0]), _p(out))
    return out


def g2_batch_mul(base: np.ndarray, scalars: np.ndarray) -> np.ndarray:
    out = np.zeros((scalars.shape[0], 16), dtype=np.uint64)
    lib().oracle_g2_batch_mul(_p(base), _p(scalars), ctypes.c_size_t(scalars.shape[0]), _p(out))
    return out


def g1_jac_to_aff(j: np.ndarray) -> np.ndarray:
    out = np.zeros(8, dtype=np.uint64)
    lib().oracle_g1_jac_to_aff(_p(j), _p(out))
    return out


def g2_jac_to_aff(j: np.ndarray) -> np.ndarray:
    out = np.zeros(16, dtype=np.uint64)
    lib().oracle_g2_jac_to_aff(_p(j), _p(out))
    return out
