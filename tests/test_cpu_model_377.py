"""BLS12-377: the device headers compiled for the host with -DB2_CURVE_BLS12_377 (12-limb Fp, Fp2 with
u^2 = -5, the same NTT plans and MSM pipeline) against the BLS12-377 oracle.  Runs without a GPU."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from oracle import bls12_377 as B
from oracle.suite377 import ntt as N
from tests import util377 as U
from tests.conftest import ROOT


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


@pytest.fixture(scope="module")
def model377():
    d = os.path.join(ROOT, "tests", "cpu_model")
    so = os.path.join(d, "libmodel377.so")
    src = os.path.join(d, "model.cpp")
    hdrs = [os.path.join(ROOT, "gnark-fork_b200", "csrc", h) for h in ("ntt.cuh", "msm.cuh", "field.cuh", "ec.cuh")]
    if not os.path.exists(so) or any(os.path.getmtime(p) > os.path.getmtime(so) for p in [src] + hdrs):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-DB2_CURVE_BLS12_377",
                               "-I/usr/local/cuda/include", "-o", so, src])
    return ctypes.CDLL(so)


def test_oracle_constants():
    assert B.g1_on_curve(B.G1_GEN) and B.g2_on_curve(B.G2_GEN)
    assert B._aff_mul(B._Fp, B.G1_GEN, B.R) is None and B._aff_mul(B._Fp2, B.G2_GEN, B.R) is None
    assert B.fp2_mul((0, 1), (0, 1)) == (B.P - 5, 0)                  # u^2 = -5 (fields_bls12377/e12.go:79)
    assert pow(B.FR_MULT_GEN, (B.R - 1) // 2, B.R) == B.R - 1           # 22 is a non-residue: a valid coset shift
    d = N.Domain(1 << 10)
    assert pow(d.generator, 1 << 10, B.R) == 1 and pow(d.generator, 1 << 9, B.R) == B.R - 1


def test_field_mul_host_path(model377):
    rng = np.random.default_rng(1)
    n = 500
    a, b = U.rand_fr(rng, n), U.rand_fr(rng, n)
    edge = U.fr_arr([0, 1, B.R - 1, 2, B.R - 2])
    a[:5], b[:5] = edge, edge[::-1]
    out = np.empty_like(a)
    model377.model_fr_mul(_p(a), _p(b), _p(out), ctypes.c_size_t(n))
    want = [x * y % B.R for x, y in zip(U.fr_ints(a), U.fr_ints(b))]
    assert U.fr_ints(out) == want
    a, b = U.rand_fp(rng, n), U.rand_fp(rng, n)
    edge = U.fp_arr([0, 1, B.P - 1, 2, B.P - 2])
    a[:5], b[:5] = edge, edge[::-1]
    out = np.empty_like(a)
    model377.model_fp_mul(_p(a), _p(b), _p(out), ctypes.c_size_t(n))
    want = [x * y % B.P for x, y in zip(U.fp_ints(a), U.fp_ints(b))]
    assert U.fp_ints(out) == want


def test_field_inverse_host_path(model377):
    """Fermat inverse x^(p-2): both BLS12-377 moduli end in ...00000001, so p - 2 borrows across limbs."""
    rng = np.random.default_rng(2)
    a = U.rand_fr(rng, 20)
    a[0] = U.fr(0)
    out = np.empty_like(a)
    model377.model_fr_inv(_p(a), _p(out), ctypes.c_size_t(20))
    assert U.fr_ints(out) == [pow(x, -1, B.R) if x else 0 for x in U.fr_ints(a)]
    a = U.rand_fp(rng, 20)
    out = np.empty_like(a)
    model377.model_fp_inv(_p(a), _p(out), ctypes.c_size_t(20))
    assert U.fp_ints(out) == [pow(x, -1, B.P) for x in U.fp_ints(a)]


@pytest.mark.parametrize("log_n,tl,smax", [(3, 12, 12), (6, 3, 2), (9, 4, 2), (10, 6, 4)])
@pytest.mark.parametrize("inverse", [0, 1])
@pytest.mark.parametrize("dit", [0, 1])
def test_ntt_pass_plan_matches_oracle(model377, log_n, tl, smax, inverse, dit):
    rng = np.random.default_rng(100 + log_n)
    n = 1 << log_n
    x = U.rand_fr(rng, n)
    d = N.Domain(n)
    got = x.copy()
    model377.model_ntt(_p(got), log_n, inverse, dit, tl, smax, _p(U.fr(d.generator)))
    want = U.fr_ints(x)
    (N.fft_inverse if inverse else N.fft)(want, d, dit)
    got = U.fr_ints(got)
    if inverse:  # the model leaves out the 1/n scaling (the kernel applies it as a post-scale)
        got = [v * d.cardinality_inv % B.R for v in got]
    assert got == want


@pytest.mark.parametrize("n,c", [(1, 0), (7, 0), (300, 5), (600, 0)])
def test_msm_model_g1(model377, n, c):
    rng = np.random.default_rng(7 + n)
    pts, ks = U.known_dlog_bases(1, n, seed=n)
    if n > 3:
        pts[n // 2] = None   # a point at infinity
        pts[3] = pts[2]      # repeated base
    scalars = U.rand_fr(rng, n)
    scalars[0] = 0
    if n > 3:
        scalars[1] = U.fr(B.R - 1)
    out = np.zeros(18, dtype=np.uint64)
    model377.model_msm_g1(_p(scalars), _p(U.g1_arr(pts)), ctypes.c_size_t(n), c, _p(out))
    assert U.g1_jac(out) == B.msm_g1(pts, U.fr_ints(scalars))


def test_msm_model_all_equal_bases(model377):
    rng = np.random.default_rng(3)
    n = 130
    pts = [B.G1_GEN] * n
    scalars = U.rand_fr(rng, n)
    scalars[:8] = U.fr(5)
    out = np.zeros(18, dtype=np.uint64)
    model377.model_msm_g1(_p(scalars), _p(U.g1_arr(pts)), ctypes.c_size_t(n), 0, _p(out))
    assert U.g1_jac(out) == B.g1_mul(B.G1_GEN, sum(U.fr_ints(scalars)) % B.R)


@pytest.mark.parametrize("n", [1, 65, 200])
def test_msm_model_g2(model377, n):
    rng = np.random.default_rng(11 + n)
    pts, ks = U.known_dlog_bases(2, n, seed=n)
    scalars = U.rand_fr(rng, n)
    out = np.zeros(36, dtype=np.uint64)
    model377.model_msm_g2(_p(scalars), _p(U.g2_arr(pts)), ctypes.c_size_t(n), 0, _p(out))
    want = B.g2_mul(B.G2_GEN, sum(k * s for k, s in zip(ks, U.fr_ints(scalars))) % B.R)
    assert U.g2_jac(out) == want
