"""The batched-affine pre-reduction of the MSM buckets (csrc/msm_ba.cuh; option msm_ba = rounds) against the oracle
and against the plain XYZZ path: G1 and G2, one-shot and resident folded-window tables, with the special cases an
affine addition has to survive -- infinity operands, P + P (DummySetup keys: every base the same point,
setup.go:532-546), P + (-P), empty and single-entry buckets, skewed (witness-like) digits."""
import numpy as np
import pytest

from oracle import bn254 as B
from tests import util
from tests.conftest import rand_field

pytestmark = pytest.mark.gpu


@pytest.fixture()
def ba(ctx):
    def on(rounds):
        ctx.set_option("msm_ba", rounds)
        ctx.set_option("msm_ba_force", 1)
    yield on
    ctx.set_option("msm_ba", 15)     # the default: rounds chosen from the mean bucket load
    ctx.set_option("msm_ba_force", 0)


def _edge(cref, scalars, bases):
    n = len(scalars)
    if n > 40:
        scalars[0] = 0
        scalars[1] = cref.fr(B.R - 1)
        scalars[2] = cref.fr(1)
        bases[7] = 0                 # infinity base
        bases[9] = bases[8]          # repeated base, same scalar: P + P inside a bucket
        scalars[9] = scalars[8]
        bases[11] = bases[10]        # P + (-P): the scalars are r - k and k
        scalars[11] = cref.fr((B.R - util.fr_ints(scalars[10:11])[0]) % B.R)


@pytest.mark.parametrize("rounds", [1, 2, 3, 5])
@pytest.mark.parametrize("n", [1, 2, 33, 1000, 5000, (1 << 15) + 3])
def test_g1_vs_oracle(ctx, cref, ba, rounds, n):
    rng = np.random.default_rng(1000 * rounds + n)
    bases, _ = util.known_dlog_bases_g1(n, seed=n)
    scalars = rand_field(rng, n)
    _edge(cref, scalars, bases)
    want = cref.g1_jac_to_aff(cref.msm_g1(scalars, bases))
    ba(rounds)
    assert np.array_equal(cref.g1_jac_to_aff(ctx.msm_g1(scalars, bases)), want)


@pytest.mark.parametrize("rounds", [1, 3])
@pytest.mark.parametrize("n", [1, 77, 3000, 1 << 13])
def test_g2_vs_oracle(ctx, cref, ba, rounds, n):
    rng = np.random.default_rng(2000 * rounds + n)
    bases, _ = util.known_dlog_bases_g2(n, seed=n)
    scalars = rand_field(rng, n)
    _edge(cref, scalars, bases)
    want = cref.g2_jac_to_aff(cref.msm_g2(scalars, bases))
    ba(rounds)
    assert np.array_equal(cref.g2_jac_to_aff(ctx.msm_g2(scalars, bases)), want)


@pytest.mark.parametrize("rounds", [2, 4])
def test_all_equal_bases(ctx, cref, ba, rounds):
    """Every addition of the first rounds is a doubling or meets an equal-x pair."""
    rng = np.random.default_rng(4)
    n = 3000
    gen = cref.g1_points_to_array([B.G1_GEN])[0]
    bases = np.tile(gen, (n, 1))
    scalars = rand_field(rng, n)
    scalars[:64] = cref.fr(12345)
    want = B.g1_mul(B.G1_GEN, sum(util.fr_ints(scalars)) % B.R)
    ba(rounds)
    assert util.g1_aff_ints(cref.g1_jac_to_aff(ctx.msm_g1(scalars, bases))) == [want[0], want[1]]


@pytest.mark.parametrize("group,n,fold_c,rounds", [(1, 5000, 9, 2), (1, 1 << 15, 13, 3), (2, 3000, 9, 2), (1, (1 << 16) + 7, 0, 3)])
def test_resident_folded_tables(ctx, cref, pkg, ba, group, n, fold_c, rounds):
    rng = np.random.default_rng(n + fold_c)
    if group == 1:
        bases, _ = util.known_dlog_bases_g1(n, seed=3)
        ref, aff = cref.msm_g1, cref.g1_jac_to_aff
    else:
        bases, _ = util.known_dlog_bases_g2(n, seed=3)
        ref, aff = cref.msm_g2, cref.g2_jac_to_aff
    scalars = rand_field(rng, n)
    half = n // 2
    scalars[:half:2] = cref.fr(1)          # witness-like: piles on one bucket of window 0
    scalars[1:half:2] = 0
    want = aff(ref(scalars, bases))
    ctx.set_option("msm_fold_c", fold_c)
    ctx.set_option("msm_precompute", 1 if fold_c else 0)
    try:
        rb = ctx.upload_bases(bases, group)
        plain = aff(ctx.msm_resident(rb, scalars))
        ba(rounds)
        got = aff(ctx.msm_resident(rb, scalars))
        rb.free()
    finally:
        ctx.set_option("msm_fold_c", 0)
        ctx.set_option("msm_precompute", 1)
    assert np.array_equal(plain, want) and np.array_equal(got, want)


def test_known_dlog_2p20(ctx, cref, pkg, ba):
    """2^20 points, the size class the stage is meant for (no force flag needed: 13 x 2^20 entries)."""
    n = 1 << 20
    rng = np.random.default_rng(9)
    ks = rand_field(rng, n)
    gen = cref.g1_points_to_array([B.G1_GEN])[0]
    bases = ctx.batch_scalar_mul(gen, ks, 1)
    scalars = rand_field(rng, n)
    rb = ctx.upload_bases(bases, 1)
    ctx.set_option("msm_ba", 3)
    try:
        got = cref.g1_jac_to_aff(ctx.msm_resident(rb, scalars))
    finally:
        ctx.set_option("msm_ba", 15)
        rb.free()
    dot = sum(a * b for a, b in zip(util.fr_ints(ks), util.fr_ints(scalars))) % B.R
    want = B.g1_mul(B.G1_GEN, dot)
    assert util.g1_aff_ints(got) == [want[0], want[1]]
