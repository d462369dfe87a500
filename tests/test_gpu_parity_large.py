"""Parity at the sizes BASELINE.json's configs name (2^20 ... 2^24), device vs the threaded C++ restatement
(oracle/cpu_ref.cpp; liboracle377.so for BLS12-377), bit for bit -- not round trips:

  * the four transforms of computeH's recipe (prove.go:369-375,393: FFTInverse DIF, FFT DIT on coset,
    FFTInverse DIF on coset) plus the plain forward DIF, at 2^22 and 2^24;
  * computeH itself (prove.go:353-396) at 2^22 and 2^24, with n_constraints < n;
  * G1 MultiExp (prove.go:201-244) at 2^20 and 2^22 and G2 MultiExp (:290) at 2^20 against the oracle's
    Pippenger, one-shot path and resident folded-window path, points compared after affine normalisation;
  * the top of config 2's transform sweep, 2^26 (two variants);
  * the same on BLS12-377 (config 5): transforms and computeH at 2^22, G1 MSM at 2^20, G2 at 2^18.

The bases are [k_i]G from the device's fixed-base kernel; a slice of them is checked against the oracle's
own batch multiplication so that the MSM inputs are the oracle's too.
"""
import numpy as np
import pytest

from oracle import bn254 as B
from oracle import ntt as N
from tests.conftest import rand_field

pytestmark = pytest.mark.gpu

# (inverse, decimation, on_coset): the three variants computeH uses and the plain forward transform
VARIANTS = [(1, 0, 0), (0, 1, 1), (1, 0, 1), (0, 0, 0)]


@pytest.mark.parametrize("log_n", [22, 24])
def test_ntt_variants_vs_oracle_at_baseline_sizes(ctx, cref, log_n):
    rng = np.random.default_rng(log_n)
    x = rand_field(rng, 1 << log_n)
    d = N.Domain(1 << log_n)
    gen, cg = cref.fr(d.generator), cref.fr(d.coset_gen)
    for inverse, dec, coset in VARIANTS:
        got = ctx.ntt(x, log_n, inverse, dec, coset, gen, cg)
        want = cref.ntt(x, log_n, inverse, dec, coset, gen, cg)
        assert np.array_equal(got, want), (log_n, inverse, dec, coset)
        del got, want


def test_ntt_2p26_vs_oracle(ctx, cref):
    """The top of config 2's sweep (2^16 ... 2^26): three passes on the device; the two transforms computeH is made of."""
    log_n = 26
    rng = np.random.default_rng(26)
    x = rand_field(rng, 1 << log_n)
    d = N.Domain(1 << log_n)
    gen, cg = cref.fr(d.generator), cref.fr(d.coset_gen)
    for inverse, dec, coset in [(1, 0, 0), (0, 1, 1)]:
        got = ctx.ntt(x, log_n, inverse, dec, coset, gen, cg)
        want = cref.ntt(x, log_n, inverse, dec, coset, gen, cg)
        assert np.array_equal(got, want), (inverse, dec, coset)
        del got, want


@pytest.mark.parametrize("log_n", [22, 24])
def test_compute_h_vs_oracle_at_baseline_sizes(ctx, cref, log_n):
    n = 1 << log_n
    rng = np.random.default_rng(100 + log_n)
    nc = n - 3          # zero-padded to the domain like the solver's output (prove.go:363-366)
    a, b, c = rand_field(rng, nc), rand_field(rng, nc), rand_field(rng, nc)
    d = N.Domain(n)
    gen, cg = cref.fr(d.generator), cref.fr(d.coset_gen)
    got = ctx.compute_h(a, b, c, log_n, gen, cg)
    want = cref.compute_h(a, b, c, log_n, gen, cg)
    assert np.array_equal(got, want)


def _edge_cases(scalars, bases, fr_one, fr_minus_one):
    scalars[0] = 0
    scalars[1] = fr_minus_one
    scalars[2] = fr_one
    bases[7] = 0            # infinity base
    bases[9] = bases[8]     # repeated base, same scalar: the P = Q branch
    scalars[9] = scalars[8]


@pytest.mark.parametrize("log_n", [20, 22])
def test_msm_g1_vs_oracle_pippenger_at_baseline_sizes(ctx, cref, log_n):
    n = (1 << log_n) - 1    # ragged: the Z MSM has n - 1 points
    rng = np.random.default_rng(200 + log_n)
    ks = rand_field(rng, n)
    gen = cref.g1_points_to_array([B.G1_GEN])[0]
    bases = ctx.batch_scalar_mul(gen, ks, 1)
    assert np.array_equal(bases[-2048:], cref.g1_batch_mul(gen, ks[-2048:]))
    scalars = rand_field(rng, n)
    _edge_cases(scalars, bases, cref.fr(1), cref.fr(B.R - 1))
    want = cref.g1_jac_to_aff(cref.msm_g1(scalars, bases))
    assert np.array_equal(cref.g1_jac_to_aff(ctx.msm_g1(scalars, bases)), want)          # one-shot path, c <= 17
    h = ctx.upload_bases(bases, 1)                                                      # resident: folded windows, c = 20
    assert h.info()["folded"] == 1
    assert np.array_equal(cref.g1_jac_to_aff(ctx.msm_resident(h, scalars)), want)
    h.free()


def test_msm_g2_vs_oracle_pippenger_2p20(ctx, cref):
    n = 1 << 20
    rng = np.random.default_rng(220)
    ks = rand_field(rng, n)
    gen = cref.g2_points_to_array([B.G2_GEN])[0]
    bases = ctx.batch_scalar_mul(gen, ks, 2)
    assert np.array_equal(bases[-512:], cref.g2_batch_mul(gen, ks[-512:]))
    scalars = rand_field(rng, n)
    _edge_cases(scalars, bases, cref.fr(1), cref.fr(B.R - 1))
    want = cref.g2_jac_to_aff(cref.msm_g2(scalars, bases))
    assert np.array_equal(cref.g2_jac_to_aff(ctx.msm_g2(scalars, bases)), want)
    h = ctx.upload_bases(bases, 2)
    assert h.info()["folded"] == 1
    assert np.array_equal(cref.g2_jac_to_aff(ctx.msm_resident(h, scalars)), want)
    h.free()


# ---- BLS12-377 (config 5) --------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def ctx377(pkg):
    c = pkg.Context(0, curve=pkg.BLS12_377)
    yield c
    c.close()


def test_bls12_377_transforms_and_compute_h_2p22(ctx377, cref):
    from oracle import bls12_377 as B377
    from oracle.suite377 import ntt as N377
    from tests import util377 as U
    o = cref.BLS12_377
    log_n = 22
    n = 1 << log_n
    rng = np.random.default_rng(377)
    d = N377.Domain(n)
    gen, cg = U.fr(d.generator), U.fr(d.coset_gen)
    x = U.rand_fr(rng, n)
    for inverse, dec, coset in VARIANTS:
        got = ctx377.ntt(x, log_n, bool(inverse), dec, bool(coset), gen, cg)
        assert np.array_equal(got, o.ntt(x, log_n, inverse, dec, coset, gen, cg)), (inverse, dec, coset)
    a, b, c = U.rand_fr(rng, n - 5), U.rand_fr(rng, n - 5), U.rand_fr(rng, n - 5)
    assert np.array_equal(ctx377.compute_h(a, b, c, log_n, gen, cg), o.compute_h(a, b, c, log_n, gen, cg))
    assert B377.R.bit_length() == 253


@pytest.mark.parametrize("group,log_n", [(1, 20), (2, 18)])
def test_bls12_377_msm_vs_oracle_pippenger(ctx377, cref, group, log_n):
    from oracle import bls12_377 as B377
    from tests import util377 as U
    o = cref.BLS12_377
    n = (1 << log_n) - 1
    rng = np.random.default_rng(377 + log_n)
    ks = U.rand_fr(rng, n)
    gen = (U.g1_arr([B377.G1_GEN]) if group == 1 else U.g2_arr([B377.G2_GEN]))[0]
    bases = ctx377.batch_scalar_mul(gen, ks, group)
    assert np.array_equal(bases[-256:], (o.g1_batch_mul if group == 1 else o.g2_batch_mul)(gen, ks[-256:]))
    scalars = U.rand_fr(rng, n)
    _edge_cases(scalars, bases, U.fr(1), U.fr(B377.R - 1))
    to_aff = o.g1_jac_to_aff if group == 1 else o.g2_jac_to_aff
    want = to_aff((o.msm_g1 if group == 1 else o.msm_g2)(scalars, bases))
    got = (ctx377.msm_g1 if group == 1 else ctx377.msm_g2)(scalars, bases)
    assert np.array_equal(to_aff(got), want)
    h = ctx377.upload_bases(bases, group)
    assert np.array_equal(to_aff(ctx377.msm_resident(h, scalars)), want)
    h.free()
