"""The device algorithms replayed on the CPU (same headers, host path) against the oracle:
tile geometry / twiddle exponents of every NTT pass plan, Montgomery multiplication, signed-digit
recoding + XYZZ buckets + reduction tree of the MSM.  Runs without a GPU."""
import ctypes

import numpy as np
import pytest

from oracle import bn254 as B
from oracle import ntt as N
from tests.conftest import rand_field


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def test_field_mul_host_path(model, cref):
    rng = np.random.default_rng(1)
    a, b = rand_field(rng, 2000), rand_field(rng, 2000)
    edge = cref.ints_to_mont([0, 1, B.R - 1, 2, B.R - 2], B.R)
    a[:5], b[:5] = edge, edge[::-1]
    out = np.empty_like(a)
    model.model_fr_mul(_p(a), _p(b), _p(out), ctypes.c_size_t(len(a)))
    assert np.array_equal(out, cref.fr_mul(a, b))
    model.model_fp_mul(_p(a), _p(b), _p(out), ctypes.c_size_t(len(a)))
    assert np.array_equal(out, cref.fp_mul(a, b))


def test_field_inverse_host_path(model, cref):
    rng = np.random.default_rng(2)
    a = rand_field(rng, 20)
    out = np.empty_like(a)
    model.model_fr_inv(_p(a), _p(out), ctypes.c_size_t(20))
    assert cref.mont_to_ints(out, B.R) == [pow(x, -1, B.R) for x in cref.mont_to_ints(a, B.R)]
    model.model_fp_inv(_p(a), _p(out), ctypes.c_size_t(20))
    assert cref.mont_to_ints(out, B.P) == [pow(x, -1, B.P) for x in cref.mont_to_ints(a, B.P)]


@pytest.mark.parametrize("log_n,tl,smax", [(1, 12, 12), (3, 12, 12), (5, 3, 3), (6, 3, 2), (7, 4, 3), (9, 4, 2),
                                           (10, 6, 4), (13, 12, 12), (14, 12, 1), (12, 6, 5)])
@pytest.mark.parametrize("inverse", [0, 1])
@pytest.mark.parametrize("dit", [0, 1])
def test_ntt_pass_plan_matches_oracle(model, cref, log_n, tl, smax, inverse, dit):
    rng = np.random.default_rng(100 + log_n)
    n = 1 << log_n
    x = rand_field(rng, n)
    d = N.Domain(n)
    gen, cg = cref.fr(d.generator), cref.fr(d.coset_gen)
    got = x.copy()
    model.model_ntt(_p(got), log_n, inverse, dit, tl, smax, _p(gen))
    want = cref.ntt(x, log_n, bool(inverse), dit, False, gen, cg)
    if inverse:  # the model leaves out the 1/n scaling (the kernel applies it as a post-scale)
        ninv = cref.fr(d.cardinality_inv)
        got = cref.fr_mul(got, np.tile(ninv, (n, 1)))
    assert np.array_equal(got, want)


@pytest.mark.parametrize("n,c", [(1, 0), (7, 0), (300, 0), (300, 5), (2000, 0), (1500, 13)])
def test_msm_model_g1(model, cref, n, c):
    rng = np.random.default_rng(7 + n)
    ks = [int(x) for x in rng.integers(1, 1 << 62, size=n)]
    gen = cref.g1_points_to_array([B.G1_GEN])[0]
    bases = cref.g1_batch_mul(gen, cref.ints_to_mont(ks, B.R))
    bases[n // 2] = 0  # a point at infinity
    scalars = rand_field(rng, n)
    scalars[0] = 0
    if n > 3:
        scalars[1] = cref.fr(B.R - 1)
        bases[3] = bases[2]  # repeated base
    out = np.zeros(12, dtype=np.uint64)
    model.model_msm_g1(_p(scalars), _p(bases), ctypes.c_size_t(n), c, _p(out))
    assert np.array_equal(cref.g1_jac_to_aff(out), cref.g1_jac_to_aff(cref.msm_g1(scalars, bases)))


def test_msm_model_all_equal_bases(model, cref):
    """DummySetup keys: every base is the same point (setup.go:532-546) -> doubling branch."""
    rng = np.random.default_rng(3)
    n = 257
    gen = cref.g1_points_to_array([B.G1_GEN])[0]
    bases = np.tile(gen, (n, 1))
    scalars = rand_field(rng, n)
    scalars[:8] = cref.fr(5)  # equal digits: same bucket gets the same point repeatedly
    out = np.zeros(12, dtype=np.uint64)
    model.model_msm_g1(_p(scalars), _p(bases), ctypes.c_size_t(n), 0, _p(out))
    assert np.array_equal(cref.g1_jac_to_aff(out), cref.g1_jac_to_aff(cref.msm_g1(scalars, bases)))


@pytest.mark.parametrize("n", [1, 65, 400])
def test_msm_model_g2(model, cref, n):
    rng = np.random.default_rng(11 + n)
    ks = [int(x) for x in rng.integers(1, 1 << 62, size=n)]
    gen = cref.g2_points_to_array([B.G2_GEN])[0]
    bases = cref.g2_batch_mul(gen, cref.ints_to_mont(ks, B.R))
    scalars = rand_field(rng, n)
    out = np.zeros(24, dtype=np.uint64)
    model.model_msm_g2(_p(scalars), _p(bases), ctypes.c_size_t(n), 0, _p(out))
    assert np.array_equal(cref.g2_jac_to_aff(out), cref.g2_jac_to_aff(cref.msm_g2(scalars, bases)))
