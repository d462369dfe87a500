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


_LIB377 = None


def build() -> str:
    subprocess.check_call(["make", "-C", _HERE, "-s"])
    return os.path.join(_HERE, "liboracle.so")


def _load(name):
    path = os.path.join(_HERE, name)
    if not os.path.exists(path):
        build()
    l = ctypes.CDLL(path)
    l.oracle_max_threads.restype = ctypes.c_int
    return l


def lib():
    global _LIB
    if _LIB is None:
        _LIB = _load("liboracle.so")
    return _LIB


def lib377():
    """The same restatement compiled with the BLS12-377 constants (oracle/Makefile): Fr 4 u64, Fp 6 u64."""
    global _LIB377
    if _LIB377 is None:
        _LIB377 = _load("liboracle377.so")
    return _LIB377


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
class Oracle:
    """The C++ restatement of one curve.  fp_words = u64 limbs of a base-field element (4 BN254, 6 BLS12-377);
    Fr is 4 limbs on both curves."""

    def __init__(self, get_lib, fp_words):
        self._get, self.fp_words = get_lib, fp_words

    @property
    def lib(self):
        return self._get()

    def set_threads(self, n):
        self.lib.oracle_set_threads(int(n))

    def max_threads(self):
        return int(self.lib.oracle_max_threads())

    def fr_mul(self, a, b):
        out = np.empty_like(a)
        self.lib.oracle_fr_mul(_p(a), _p(b), _p(out), ctypes.c_size_t(a.shape[0]))
        return out

    def fp_mul(self, a, b):
        out = np.empty_like(a)
        self.lib.oracle_fp_mul(_p(a), _p(b), _p(out), ctypes.c_size_t(a.shape[0]))
        return out

    def ntt(self, data, log_n, inverse, decimation, on_coset, gen, coset_gen, inplace=False):
        out = data if inplace else np.ascontiguousarray(data.copy())
        self.lib.oracle_ntt(_p(out), ctypes.c_uint(log_n), int(inverse), int(decimation), int(on_coset),
                            _p(gen), _p(coset_gen))
        return out

    def compute_h(self, a, b, c, log_n, gen, coset_gen):
        n = 1 << log_n

        def pad(x):
            o = np.zeros((n, 4), dtype=np.uint64)
            o[: x.shape[0]] = x
            return o
        A, B, Cc = pad(a), pad(b), pad(c)
        self.lib.oracle_compute_h(_p(A), _p(B), _p(Cc), ctypes.c_uint(log_n), _p(gen), _p(coset_gen))
        return A

    def msm_g1(self, scalars, bases):
        out = np.zeros(3 * self.fp_words, dtype=np.uint64)
        self.lib.oracle_msm_g1(_p(scalars), _p(bases), ctypes.c_size_t(scalars.shape[0]), _p(out))
        return out

    def msm_g2(self, scalars, bases):
        out = np.zeros(6 * self.fp_words, dtype=np.uint64)
        self.lib.oracle_msm_g2(_p(scalars), _p(bases), ctypes.c_size_t(scalars.shape[0]), _p(out))
        return out

    def g1_batch_mul(self, base, scalars):
        out = np.zeros((scalars.shape[0], 2 * self.fp_words), dtype=np.uint64)
        self.lib.oracle_g1_batch_mul(_p(base), _p(scalars), ctypes.c_size_t(scalars.shape[0]), _p(out))
        return out

    def g2_batch_mul(self, base, scalars):
        out = np.zeros((scalars.shape[0], 4 * self.fp_words), dtype=np.uint64)
        self.lib.oracle_g2_batch_mul(_p(base), _p(scalars), ctypes.c_size_t(scalars.shape[0]), _p(out))
        return out

    def g1_jac_to_aff(self, j):
        out = np.zeros(2 * self.fp_words, dtype=np.uint64)
        self.lib.oracle_g1_jac_to_aff(_p(j), _p(out))
        return out

    def g2_jac_to_aff(self, j):
        out = np.zeros(4 * self.fp_words, dtype=np.uint64)
        self.lib.oracle_g2_jac_to_aff(_p(j), _p(out))
        return out


BN254 = Oracle(lib, 4)
BLS12_377 = Oracle(lib377, 6)

# module-level names = the BN254 instance (what every BN254 test imports)
fr_mul, fp_mul, ntt, compute_h = BN254.fr_mul, BN254.fp_mul, BN254.ntt, BN254.compute_h
msm_g1, msm_g2, g1_batch_mul, g2_batch_mul = BN254.msm_g1, BN254.msm_g2, BN254.g1_batch_mul, BN254.g2_batch_mul
g1_jac_to_aff, g2_jac_to_aff = BN254.g1_jac_to_aff, BN254.g2_jac_to_aff
