"""Bucket accumulation with the running sum in shared memory (msm_accumulate_smem_kernel; option msm_acc_smem, bit 0 = G1,
bit 1 = G2; the default for BN254 G2) against the oracle and against the register kernel: every branch of the mixed
addition -- first point into an empty sum, infinity bases, P + P (DummySetup keys, setup.go:532-546), P + (-P) and the
sum coming back from infinity -- one-shot and resident (folded-table) paths, both curves."""
import numpy as np
import pytest

from oracle import bn254 as B
from tests import util
from tests.conftest import rand_field
from tests.test_gpu_msm_ba import _edge

pytestmark = pytest.mark.gpu


@pytest.fixture()
def smem(ctx):
    def on(mask):
        ctx.set_option("msm_acc_smem", mask)
    yield on
    ctx.set_option("msm_acc_smem", 2)      # the BN254 default


@pytest.mark.parametrize("mask", [0, 1, 2, 3])
@pytest.mark.parametrize("n", [1, 33, 3000, (1 << 14) + 5])
def test_g1_g2_vs_oracle(ctx, cref, smem, mask, n):
    rng = np.random.default_rng(10 * n + mask)
    b1, _ = util.known_dlog_bases_g1(n, seed=n)
    b2, _ = util.known_dlog_bases_g2(n, seed=n)
    s1, s2 = rand_field(rng, n), rand_field(rng, n)
    _edge(cref, s1, b1)
    _edge(cref, s2, b2)
    smem(mask)
    assert np.array_equal(cref.g1_jac_to_aff(ctx.msm_g1(s1, b1)), cref.g1_jac_to_aff(cref.msm_g1(s1, b1)))
    assert np.array_equal(cref.g2_jac_to_aff(ctx.msm_g2(s2, b2)), cref.g2_jac_to_aff(cref.msm_g2(s2, b2)))


@pytest.mark.parametrize("group", [1, 2])
def test_all_equal_bases_and_cancellation(ctx, cref, pkg, smem, group):
    """Every base the same point and scalars drawn from a handful of values: buckets hold runs of P + P; with half
    the scalars negated the sums pass through infinity and start again."""
    n = 4096
    gen = (cref.g1_points_to_array([B.G1_GEN]) if group == 1 else cref.g2_points_to_array([B.G2_GEN]))[0]
    bases = np.tile(gen, (n, 1))
    rng = np.random.default_rng(5)
    vals = [int(v) for v in rng.integers(1, 1 << 40, size=4)]
    ks = [vals[i % 4] if (i // 4) % 2 == 0 else B.R - vals[i % 4] for i in range(n)]
    ks[-1] = 12345
    scalars = util.fr_arr(ks)
    msm = ctx.msm_g1 if group == 1 else ctx.msm_g2
    aff = pkg.g1_jac_to_affine if group == 1 else pkg.g2_jac_to_affine
    smem(0)
    want = aff(msm(scalars, bases))
    smem(3)
    got = aff(msm(scalars, bases))
    assert np.array_equal(got, want)
    exp = B.g1_mul(B.G1_GEN, sum(ks) % B.R) if group == 1 else B.g2_mul(B.G2_GEN, sum(ks) % B.R)
    assert (util.g1_aff_ints(got) == list(exp)) if group == 1 else (util.g2_aff_ints(got) == [exp[0][0], exp[0][1], exp[1][0], exp[1][1]])


@pytest.mark.parametrize("group", [1, 2])
def test_resident_tables_same_result(ctx, cref, pkg, smem, group):
    n = 1 << 15
    bases, _ = (util.known_dlog_bases_g1 if group == 1 else util.known_dlog_bases_g2)(n, seed=3)
    scalars = rand_field(np.random.default_rng(8), n)
    h = ctx.upload_bases(bases, group)
    aff = pkg.g1_jac_to_affine if group == 1 else pkg.g2_jac_to_affine
    out = {}
    for mask in (0, 3):
        smem(mask)
        out[mask] = aff(ctx.msm_resident(h, scalars))
    h.free()
    assert np.array_equal(out[0], out[3])


def test_bls12_377_both_groups(pkg):
    from oracle import bls12_377 as B3
    from tests import util377 as U
    c = pkg.Context(0, curve=pkg.BLS12_377)
    try:
        rng = np.random.default_rng(2)
        for group, n in ((1, 700), (2, 300)):
            pts, ks = U.known_dlog_bases(group, n, seed=group)
            pts[5] = None
            pts[9] = pts[8]
            scalars = U.rand_fr(rng, n)
            scalars[9] = scalars[8]
            arr = U.g1_arr(pts) if group == 1 else U.g2_arr(pts)
            want = (B3.msm_g1 if group == 1 else B3.msm_g2)(pts, U.fr_ints(scalars))
            for mask in (0, 3):
                c.set_option("msm_acc_smem", mask)
                got = (c.msm_g1 if group == 1 else c.msm_g2)(scalars, arr)
                assert (U.g1_jac(got) if group == 1 else U.g2_jac(got)) == want, (group, mask)
    finally:
        c.close()
