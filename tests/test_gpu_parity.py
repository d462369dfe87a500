"""Parity of the CUDA path (through the C ABI, host buffers) with the oracle.  Bit-exact for
field elements; MSM results compared after affine normalisation (Jacobian representatives are
not canonical)."""
import numpy as np
import pytest

from oracle import bn254 as B
from oracle import groth16 as G
from oracle import ntt as N
from tests import util
from tests.conftest import rand_field

pytestmark = pytest.mark.gpu


def test_field_mul(ctx, cref):
    rng = np.random.default_rng(1)
    n = 1 << 18
    a, b = rand_field(rng, n), rand_field(rng, n)
    edge = cref.ints_to_mont([0, 1, B.R - 1, 2, B.R - 2, (1 << 256) % B.R], B.R)
    a[:6], b[:6] = edge, edge[::-1]
    assert np.array_equal(ctx.fr_vec_mul(a, b), cref.fr_mul(a, b))
    edge = cref.ints_to_mont([0, 1, B.P - 1, 2, B.P - 2, (1 << 256) % B.P], B.P)
    a[:6], b[:6] = edge, edge[::-1]
    assert np.array_equal(ctx.fp_vec_mul(a, b), cref.fp_mul(a, b))
    assert ctx.fr_vec_mul(a[:0], b[:0]).shape == (0, 4)


def test_ntt_golden(ctx, cref):
    g = util.golden()["ntt"]
    x, gen, cg = util.fr_arr(g["input"]), cref.fr(g["generator"]), cref.fr(g["coset_gen"])
    for case in g["cases"]:
        got = ctx.ntt(x, g["log_n"], case["inverse"], case["decimation"], case["on_coset"], gen, cg)
        assert util.fr_ints(got) == case["output"], case


@pytest.mark.parametrize("log_n", [0, 1, 2, 5, 11, 12, 13, 16, 18, 20])
def test_ntt_all_variants_vs_oracle(ctx, cref, log_n):
    rng = np.random.default_rng(50 + log_n)
    n = 1 << log_n
    x = rand_field(rng, n)
    d = N.Domain(n)
    gen, cg = cref.fr(d.generator), cref.fr(d.coset_gen)
    for inverse in (0, 1):
        for dec in (0, 1):
            for coset in (0, 1):
                got = ctx.ntt(x, log_n, inverse, dec, coset, gen, cg)
                want = cref.ntt(x, log_n, inverse, dec, coset, gen, cg)
                assert np.array_equal(got, want), (log_n, inverse, dec, coset)


def test_ntt_roundtrip_large(ctx, cref):
    """size-independent property at a BASELINE size: FFTInverse(DIT) o FFT(DIF) = identity."""
    log_n = 22
    rng = np.random.default_rng(9)
    x = rand_field(rng, 1 << log_n)
    d = N.Domain(1 << log_n)
    gen, cg = cref.fr(d.generator), cref.fr(d.coset_gen)
    y = ctx.ntt(x, log_n, 0, 0, 1, gen, cg)
    assert not np.array_equal(x, y)
    z = ctx.ntt(y, log_n, 1, 1, 1, gen, cg)
    assert np.array_equal(x, z)


def test_compute_h_golden_and_oracle(ctx, cref):
    g = util.golden()["compute_h"]
    gen, cg = cref.fr(g["generator"]), cref.fr(g["coset_gen"])
    got = ctx.compute_h(util.fr_arr(g["a"]), util.fr_arr(g["b"]), util.fr_arr(g["c"]), g["log_n"], gen, cg)
    assert util.fr_ints(got) == g["h_bitreversed"]
    for log_n, nc in ((10, 1000), (14, 1 << 14), (17, (1 << 17) - 5)):
        rng = np.random.default_rng(log_n)
        a, b, c = rand_field(rng, nc), rand_field(rng, nc), rand_field(rng, nc)
        d = N.Domain(1 << log_n)
        gen, cg = cref.fr(d.generator), cref.fr(d.coset_gen)
        assert np.array_equal(ctx.compute_h(a, b, c, log_n, gen, cg), cref.compute_h(a, b, c, log_n, gen, cg))


def test_msm_golden(ctx, cref):
    g = util.golden()["msm"]
    s = util.fr_arr(g["scalars"])
    assert util.g1_aff_ints(cref.g1_jac_to_aff(ctx.msm_g1(s, util.g1_arr(g["g1_bases"])))) == g["g1_result"]
    assert util.g2_aff_ints(cref.g2_jac_to_aff(ctx.msm_g2(s, util.g2_arr(g["g2_bases"])))) == g["g2_result"]


@pytest.mark.parametrize("n", [0, 1, 2, 33, 1000, 5000, 1 << 15, (1 << 16) + 3])
def test_msm_g1_vs_oracle(ctx, cref, n):
    rng = np.random.default_rng(70 + n)
    bases, _ = util.known_dlog_bases_g1(max(n, 1), seed=n)
    bases = bases[:n]
    scalars = rand_field(rng, n)
    if n > 40:
        scalars[0] = 0
        scalars[1] = cref.fr(B.R - 1)
        scalars[2] = cref.fr(1)
        bases[7] = 0            # infinity base
        bases[9] = bases[8]     # repeated base
        scalars[9] = scalars[8]
    got = ctx.msm_g1(scalars, bases)
    want = cref.msm_g1(scalars, bases)
    assert np.array_equal(cref.g1_jac_to_aff(got), cref.g1_jac_to_aff(want))


def test_msm_g1_all_equal_bases(ctx, cref):
    """DummySetup (setup.go:532-546): every base is the same point."""
    rng = np.random.default_rng(4)
    n = 3000
    gen = cref.g1_points_to_array([B.G1_GEN])[0]
    bases = np.tile(gen, (n, 1))
    scalars = rand_field(rng, n)
    scalars[:64] = cref.fr(12345)
    want = B.g1_mul(B.G1_GEN, sum(util.fr_ints(scalars)) % B.R)
    assert util.g1_aff_ints(cref.g1_jac_to_aff(ctx.msm_g1(scalars, bases))) == [want[0], want[1]]


def test_msm_g1_witness_like_scalars(ctx, cref):
    """Skewed digits: half the scalars in {0,1}, a quarter below 2^16."""
    rng = np.random.default_rng(6)
    n = 1 << 14
    bases, ks = util.known_dlog_bases_g1(n, seed=3)
    vals = [int(rng.integers(0, 2)) for _ in range(n // 2)] + [int(rng.integers(0, 1 << 16)) for _ in range(n // 4)]
    vals += util.fr_ints(rand_field(rng, n - len(vals)))
    scalars = util.fr_arr(vals)
    want = B.g1_mul(B.G1_GEN, sum(k * s for k, s in zip(ks, vals)) % B.R)
    assert util.g1_aff_ints(cref.g1_jac_to_aff(ctx.msm_g1(scalars, bases))) == [want[0], want[1]]


@pytest.mark.parametrize("n", [1, 77, 3000, 1 << 14])
def test_msm_g2_vs_oracle(ctx, cref, n):
    rng = np.random.default_rng(90 + n)
    bases, _ = util.known_dlog_bases_g2(n, seed=n)
    scalars = rand_field(rng, n)
    if n > 40:
        scalars[0] = 0
        bases[7] = 0
        bases[9] = bases[8]
        scalars[9] = scalars[8]
    got = ctx.msm_g2(scalars, bases)
    want = cref.msm_g2(scalars, bases)
    assert np.array_equal(cref.g2_jac_to_aff(got), cref.g2_jac_to_aff(want))


def test_msm_known_dlog_large(ctx, cref):
    """2^20 points, exact at any size: bases [k_i]G from the device fixed-base kernel (itself
    checked against the oracle on a slice), expected = [sum k_i s_i]G."""
    rng = np.random.default_rng(12)
    n = 1 << 20
    ks = rand_field(rng, n)
    gen = cref.g1_points_to_array([B.G1_GEN])[0]
    bases = ctx.batch_scalar_mul(gen, ks, 1)
    assert np.array_equal(bases[:512], cref.g1_batch_mul(gen, ks[:512]))
    scalars = rand_field(rng, n)
    prod = cref.fr_mul(ks, scalars)  # Montgomery form of k_i * s_i
    total = sum(util.fr_ints(prod)) % B.R
    want = B.g1_mul(B.G1_GEN, total)
    assert util.g1_aff_ints(cref.g1_jac_to_aff(ctx.msm_g1(scalars, bases))) == [want[0], want[1]]


def test_batch_scalar_mul_g2(ctx, cref):
    rng = np.random.default_rng(13)
    ks = rand_field(rng, 300)
    ks[0] = 0
    gen = cref.g2_points_to_array([B.G2_GEN])[0]
    assert np.array_equal(ctx.batch_scalar_mul(gen, ks, 2), cref.g2_batch_mul(gen, ks))


@pytest.mark.parametrize("group,n,fold_c", [(1, 5000, 0), (1, 5000, 9), (1, 1 << 15, 13), (1, (1 << 16) + 7, 0),
                                            (2, 4500, 0), (2, 1 << 13, 11)])
def test_msm_resident_folded_windows(ctx, cref, group, n, fold_c):
    """Resident base sets run the folded-window path (tables T[w][i] = 2^(c w) P_i, one shared
    bucket set); result must equal the oracle's MultiExp, also with fewer scalars than bases."""
    rng = np.random.default_rng(300 + n + fold_c)
    bases, _ = (util.known_dlog_bases_g1 if group == 1 else util.known_dlog_bases_g2)(n, seed=n)
    bases[7] = 0                # infinity base: every multiple stays infinity
    bases[9] = bases[8]         # repeated base
    scalars = rand_field(rng, n)
    scalars[0] = 0
    scalars[1] = cref.fr(B.R - 1)
    scalars[2] = cref.fr(1)
    scalars[9] = scalars[8]
    ctx.set_option("msm_fold_c", fold_c)
    ctx.set_option("msm_precompute", 1)
    try:
        h = ctx.upload_bases(bases, group)
    finally:
        ctx.set_option("msm_fold_c", 0)
    oracle_msm = cref.msm_g1 if group == 1 else cref.msm_g2
    to_aff = cref.g1_jac_to_aff if group == 1 else cref.g2_jac_to_aff
    for m in (n, n - 1234):
        got = ctx.msm_resident(h, scalars[:m])
        assert np.array_equal(to_aff(got), to_aff(oracle_msm(scalars[:m], bases[:m]))), (group, n, fold_c, m)
    h.free()
    # and the plain path on the same inputs
    ctx.set_option("msm_precompute", 0)
    try:
        h = ctx.upload_bases(bases, group)
        got = ctx.msm_resident(h, scalars)
        h.free()
    finally:
        ctx.set_option("msm_precompute", 1)
    assert np.array_equal(to_aff(got), to_aff(oracle_msm(scalars, bases)))


@pytest.mark.parametrize("group,n", [(1, 1 << 14), (2, 1 << 12)])
def test_msm_resident_witness_like_scalars(ctx, cref, group, n):
    """SURVEY 8(d) witness-like scalars (half in {0, 1}, a quarter below 2^16) on the folded path:
    window 0 piles half of the entries into bucket 1 (split into pieces and merged by the warp / CTA
    tiers), higher windows are mostly empty."""
    rng = np.random.default_rng(n + group)
    bases, _ = (util.known_dlog_bases_g1 if group == 1 else util.known_dlog_bases_g2)(n, seed=3)
    kind = rng.integers(0, 4, size=n)
    vals = [int(rng.integers(0, 2)) if k < 2 else int(rng.integers(0, 1 << 16)) if k == 2 else int(rng.integers(0, 1 << 62)) ** 4 % B.R
            for k in kind]
    scalars = cref.ints_to_mont(vals, B.R)
    h = ctx.upload_bases(bases, group)
    assert h.info()["folded"] == 1
    got = ctx.msm_resident(h, scalars)
    h.free()
    want = (cref.msm_g1 if group == 1 else cref.msm_g2)(scalars, bases)
    to_aff = cref.g1_jac_to_aff if group == 1 else cref.g2_jac_to_aff
    assert np.array_equal(to_aff(got), to_aff(want))


def test_domain_cache_trim_keeps_domains_of_resident_keys(ctx, pkg, cref):
    """b200_domain_cache_trim frees the cached NTT domains no key refers to; a resident key's domain survives and the
    key still proves afterwards."""
    from tests.test_gpu_groth16 import build_trapdoor_key
    nb_mul = (1 << 9) - 1
    r1cs, td, pk, vk, pkh, vkpts = build_trapdoor_key(ctx, cref, nb_mul)
    w, a, b, c = G.mul_chain_witness(nb_mul, 2)
    args = [util.fr_arr(x) for x in (a, b, c, w)]
    before = ctx.groth16_prove_core(pkh, *args)[0].as_dict()
    x = rand_field(np.random.default_rng(1), 1 << 11)
    d = N.Domain(1 << 11)
    ctx.ntt(x, 11, 0, 0, 0, cref.fr(d.generator), cref.fr(d.coset_gen))      # caches a 2^11 domain nobody owns
    assert ctx.domain_cache_trim() >= (1 << 10) * 32 * 2
    after = ctx.groth16_prove_core(pkh, *args)[0].as_dict()
    for k in ("ar", "bs1", "krs", "krs2"):      # Jacobian representatives depend on the accumulation order: compare the points
        assert np.array_equal(pkg.g1_jac_to_affine(before[k]), pkg.g1_jac_to_affine(after[k])), k
    assert np.array_equal(pkg.g2_jac_to_affine(before["bs2"]), pkg.g2_jac_to_affine(after["bs2"]))
    assert np.array_equal(ctx.ntt(x, 11, 0, 0, 0, cref.fr(d.generator), cref.fr(d.coset_gen)), cref.ntt(x, 11, 0, 0, 0, cref.fr(d.generator), cref.fr(d.coset_gen)))
    pkh.free()
    ctx.domain_cache_trim()


@pytest.mark.parametrize("log_n,tile_log", [(6, 9), (9, 9), (10, 9), (13, 9), (16, 9), (18, 9), (20, 9), (22, 9), (20, 10), (22, 11), (20, 12), (14, 6)])
def test_ntt_tma_staged_passes_vs_oracle(ctx, cref, log_n, tile_log):
    """Option ntt_tma: tiles staged by cp.async.bulk.tensor (ntt_pass_tma_kernel) -- every variant bit-exact vs the
    oracle, for several tile sizes (the box shapes change with the tile), and computeH with its fused scale tables."""
    rng = np.random.default_rng(500 + log_n + tile_log)
    n = 1 << log_n
    x = rand_field(rng, n)
    d = N.Domain(n)
    gen, cg = cref.fr(d.generator), cref.fr(d.coset_gen)
    ctx.set_option("ntt_tma", 1)
    ctx.set_option("ntt_tile_log", tile_log)
    try:
        for inverse in (0, 1):
            for dec in (0, 1):
                for coset in (0, 1):
                    got = ctx.ntt(x, log_n, inverse, dec, coset, gen, cg)
                    assert np.array_equal(got, cref.ntt(x, log_n, inverse, dec, coset, gen, cg)), (log_n, tile_log, inverse, dec, coset)
        if log_n <= 20:
            a, b, c = rand_field(rng, n - 3), rand_field(rng, n - 3), rand_field(rng, n - 3)
            assert np.array_equal(ctx.compute_h(a, b, c, log_n, gen, cg), cref.compute_h(a, b, c, log_n, gen, cg))
    finally:
        ctx.set_option("ntt_tma", 0)
        ctx.set_option("ntt_tile_log", 9)
