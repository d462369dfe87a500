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


@pytest.mark.parametrize("log_n", [0, 1, 3, 6, 10])
def test_build_ratio_copy_constraint(ctx, pkg, cref, log_n):
    n = 1 << log_n
    rng = random.Random(50 + log_n)
    d = N.Domain(n)
    l, r, o = (_ints(rng, n) for _ in range(3))
    perm = list(range(3 * n))
    rng.shuffle(perm)
    beta, gamma = rng.randrange(B.R), rng.randrange(B.R)
    want = P.build_ratio_copy_constraint([l, r, o], perm, beta, gamma, d)
    dv = [pkg.DeviceVector(ctx, n * 32).upload(cref.ints_to_mont(v, B.R)) for v in (l, r, o)]
    pv = pkg.DeviceVector(ctx, 3 * n * 8).upload(np.array(perm, dtype=np.int64))
    z = pkg.DeviceVector(ctx, n * 32)
    ctx.plonk_build_ratio_dev(dv[0], dv[1], dv[2], pv, cref.fr(beta), cref.fr(gamma), log_n, cref.fr(d.generator), cref.fr(d.coset_gen), z)
    assert cref.mont_to_ints(z.download((n, 4)), B.R) == want
    [v.free() for v in dv + [pv, z]]


def test_build_ratio_closes_for_a_valid_permutation(ctx, pkg, cref):
    """Copy constraints satisfied (wires equal along the permutation cycles): the full product is 1,
    i.e. Z[n-1] * num_{n-1} / den_{n-1} = 1 -- checked on the device output with the oracle's last factor."""
    log_n, n = 8, 256
    rng = random.Random(8)
    d = N.Domain(n)
    perm = list(range(3 * n))
    rng.shuffle(perm)
    vals = [None] * (3 * n)
    for s0 in range(3 * n):                                   # one value per permutation cycle
        if vals[s0] is None:
            v, k = rng.randrange(B.R), s0
            while vals[k] is None:
                vals[k] = v
                k = perm[k]
    l, r, o = vals[:n], vals[n:2 * n], vals[2 * n:]
    beta, gamma = rng.randrange(B.R), rng.randrange(B.R)
    dv = [pkg.DeviceVector(ctx, n * 32).upload(cref.ints_to_mont(v, B.R)) for v in (l, r, o)]
    pv = pkg.DeviceVector(ctx, 3 * n * 8).upload(np.array(perm, dtype=np.int64))
    z = pkg.DeviceVector(ctx, n * 32)
    ctx.plonk_build_ratio_dev(dv[0], dv[1], dv[2], pv, cref.fr(beta), cref.fr(gamma), log_n, cref.fr(d.generator), cref.fr(d.coset_gen), z)
    zz = cref.mont_to_ints(z.download((n, 4)), B.R)
    ident = [pow(d.coset_gen, j, B.R) * pow(d.generator, i, B.R) % B.R for j in range(3) for i in range(n)]
    num = den = 1
    for j, w in enumerate((l, r, o)):
        num = num * (beta * ident[j * n + n - 1] + gamma + w[n - 1]) % B.R
        den = den * (beta * ident[perm[j * n + n - 1]] + gamma + w[n - 1]) % B.R
    assert zz[0] == 1 and zz[n - 1] * num % B.R == den
    [v.free() for v in dv + [pv, z]]


@pytest.mark.parametrize("log_n,log_big,nb_bsb", [(2, 4, 0), (5, 7, 1), (8, 10, 2), (6, 6, 0)])
def test_quotient_rows(ctx, pkg, cref, log_n, log_big, nb_bsb):
    n, big = 1 << log_n, 1 << log_big
    rho = big // n
    rng = random.Random(log_n * 7 + nb_bsb)
    d_small, d_big = N.Domain(n), N.Domain(big)
    x = {k: _ints(rng, n) for k in range(15) if k != P.ID_ZS}
    qci = [_ints(rng, n) for _ in range(2 * nb_bsb)]
    beta, gamma, alpha = (rng.randrange(B.R) for _ in range(3))
    cs = d_big.coset_gen
    bl, br, bo, bz = _ints(rng, 2), _ints(rng, 2), _ints(rng, 2), _ints(rng, 3)
    want = [0] * big
    got_dev = pkg.DeviceVector(ctx, big * 32).upload(np.zeros((big, 4), dtype=np.uint64))
    xd = [None if k == P.ID_ZS else pkg.DeviceVector(ctx, n * 32).upload(cref.ints_to_mont(x[k], B.R)) for k in range(15)]
    qd = [pkg.DeviceVector(ctx, n * 32).upload(cref.ints_to_mont(v, B.R)) for v in qci]
    m = lambda v: cref.ints_to_mont(v, B.R)
    for coset in range(rho):
        P.quotient_rows(x, qci, d_small, log_big, coset, beta, gamma, alpha, cs, bl, br, bo, bz, want)
        ctx.plonk_quotient_rows_dev(xd, qd, log_n, log_big, coset, cref.fr(d_small.generator), cref.fr(beta), cref.fr(gamma),
                                    cref.fr(alpha), cref.fr(cs), m(bl), m(br), m(bo), m(bz), got_dev)
    assert cref.mont_to_ints(got_dev.download((big, 4)), B.R) == want
    [v.free() for v in xd if v is not None]
    [v.free() for v in qd]
    got_dev.free()
