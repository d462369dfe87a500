"""PlonK building blocks on the device (csrc/poly.cu, through the C ABI) against oracle/plonk.py:
batch inversion, prefix product, evaluation, quotient by (X - z), divideByXMinusOne, KZG commit."""
import random

import numpy as np
import pytest

from oracle import bn254 as B
from oracle import ntt as N
from oracle import plonk as P
from tests import util
from tests.conftest import rand_field

pytestmark = pytest.mark.gpu


def _ints(rng, n, zeros=()):
    xs = [rng.randrange(B.R) for _ in range(n)]
    for z in zeros:
        if z < n:
            xs[z] = 0
    return xs


@pytest.mark.parametrize("n", [1, 2, 31, 32, 33, 1000, 1024, 1025, 40000])
def test_batch_inverse_and_prefix_product(ctx, pkg, cref, n):
    rng = random.Random(n)
    xs = _ints(rng, n, zeros=(0, 5, 33, n - 1) if n > 40 else ())
    a = cref.ints_to_mont(xs, B.R)
    src = pkg.DeviceVector(ctx, n * 32).upload(a)
    dst = pkg.DeviceVector(ctx, n * 32)
    ctx.fr_batch_inverse_dev(src, dst, n)
    assert cref.mont_to_ints(dst.download((n, 4)), B.R) == P.batch_invert(xs)
    ys = [x if x else 1 for x in xs]
    src.upload(cref.ints_to_mont(ys, B.R))
    ctx.fr_prefix_product_dev(src, dst, n)
    assert cref.mont_to_ints(dst.download((n, 4)), B.R) == P.prefix_product(ys)
    ctx.fr_prefix_product_dev(src, src, n)                # in place
    assert cref.mont_to_ints(src.download((n, 4)), B.R) == P.prefix_product(ys)
    src.free(); dst.free()


@pytest.mark.parametrize("n", [1, 2, 32, 33, 1024, 1025, 32 * 32 * 32 + 7, 50000])
def test_poly_eval_and_division(ctx, pkg, cref, n):
    rng = random.Random(100 + n)
    c = _ints(rng, n)
    z = rng.randrange(B.R)
    src = pkg.DeviceVector(ctx, n * 32).upload(cref.ints_to_mont(c, B.R))
    got = ctx.fr_poly_eval_dev(src, n, cref.fr(z))
    assert cref.mont_to_ints(got.reshape(1, 4), B.R)[0] == P.poly_eval(c, z)
    if n > 1:
        q = pkg.DeviceVector(ctx, (n - 1) * 32)
        ctx.fr_poly_div_linear_dev(src, n, cref.fr(z), q)
        want_q, _ = P.poly_div_linear(c, z)
        assert cref.mont_to_ints(q.download((n - 1, 4)), B.R) == want_q
        q.free()
    src.free()


def test_poly_eval_large_matches_cpu_restatement(ctx, pkg, cref):
    """2^20 coefficients: value checked through the identity p(z) = q(z)(z - z0) + p(z0) at a second point."""
    n = 1 << 20
    rng = np.random.default_rng(9)
    c = rand_field(rng, n)
    src = pkg.DeviceVector(ctx, n * 32).upload(c)
    q = pkg.DeviceVector(ctx, (n - 1) * 32)
    z0, z1 = 0x1234567 % B.R, 0x7654321 % B.R
    ctx.fr_poly_div_linear_dev(src, n, cref.fr(z0), q)
    m = lambda v: cref.mont_to_ints(np.asarray(v).reshape(1, 4), B.R)[0]
    p0, p1 = m(ctx.fr_poly_eval_dev(src, n, cref.fr(z0))), m(ctx.fr_poly_eval_dev(src, n, cref.fr(z1)))
    q1 = m(ctx.fr_poly_eval_dev(q, n - 1, cref.fr(z1)))
    assert p1 == (q1 * (z1 - z0) + p0) % B.R
    src.free(); q.free()


@pytest.mark.parametrize("log_small,log_big", [(3, 5), (6, 8), (10, 12), (12, 13), (9, 9)])
def test_divide_by_xn_minus_one(ctx, pkg, cref, log_small, log_big):
    rng = random.Random(log_big)
    d_small, d_big = N.Domain(1 << log_small), N.Domain(1 << log_big)
    r = _ints(rng, 1 << log_big)
    want = P.divide_by_xn_minus_one(list(r), d_small, d_big)
    v = pkg.DeviceVector(ctx, len(r) * 32).upload(cref.ints_to_mont(r, B.R))
    ctx.plonk_divide_by_xn_minus_one_dev(v, log_big, log_small, cref.fr(d_big.generator), cref.fr(d_big.coset_gen))
    assert cref.mont_to_ints(v.download((len(r), 4)), B.R) == want
    v.free()


def test_kzg_commit_matches_msm(ctx, pkg, cref):
    n = 5000
    rng = np.random.default_rng(4)
    srs, _ = util.known_dlog_bases_g1(n, seed=11)
    coeffs = rand_field(rng, n)
    h = ctx.upload_bases(srs, 1)
    cv = pkg.DeviceVector(ctx, n * 32).upload(coeffs)
    got = ctx.kzg_commit_dev(h, cv, n)
    assert np.array_equal(got, cref.g1_jac_to_aff(cref.msm_g1(coeffs, srs)))
    h.free(); cv.free()
