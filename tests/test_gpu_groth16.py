"""Groth16 prove core on the device against the golden proof and the trapdoor checker
(oracle/groth16.py): byte-identical proof under fixed randomness, and acceptance by the
verify equation evaluated in the exponent."""
import numpy as np
import pytest

from oracle import bn254 as B
from oracle import groth16 as G
from tests import util

pytestmark = pytest.mark.gpu


def _prove(ctx, pkg, cref, pkh, vkpts, w, a, b, c, r, s):
    out, tm = ctx.groth16_prove_core(pkh, util.fr_arr(a), util.fr_arr(b), util.fr_arr(c), util.fr_arr(w))
    ar, bs, krs = pkg.groth16_finalize(out, cref.fr(r), cref.fr(s), *vkpts)
    return ar, bs, krs, tm


def test_golden_proof_bytes(ctx, pkg, cref):
    g = util.golden()["groth16"]
    pkj = g["pk"]
    w, a, b, c = G.mul_chain_witness(g["nb_mul"], g["x"])
    d = G.ntt.Domain(len(a))
    pkh = ctx.groth16_pk_upload(log_n=d.log_n, generator=cref.fr(d.generator), coset_gen=cref.fr(d.coset_gen),
                                g1_a=util.g1_arr(pkj["g1_A"]), g1_b=util.g1_arr(pkj["g1_B"]), g1_k=util.g1_arr(pkj["g1_K"]),
                                g1_z=util.g1_arr(pkj["g1_Z"]), g2_b=util.g2_arr(pkj["g2_B"]),
                                infinity_a=pkj["infinity_A"], infinity_b=pkj["infinity_B"], nb_public=2)
    vkpts = [util.g1_arr([pkj[k]])[0] for k in ("g1_alpha", "g1_beta", "g1_delta")] + \
            [util.g2_arr([pkj[k]])[0] for k in ("g2_beta", "g2_delta")]
    ar, bs, krs, _ = _prove(ctx, pkg, cref, pkh, vkpts, w, a, b, c, g["r"], g["s"])
    assert util.g1_aff_ints(ar) == g["ar"] and util.g2_aff_ints(bs) == g["bs"] and util.g1_aff_ints(krs) == g["krs"]
    assert pkg.groth16_proof_bytes(ar, bs, krs).hex() == g["proof_hex"]
    pkh.free()


def build_trapdoor_key(ctx, cref, nb_mul, seed=7, shard=(0, 1)):
    """Setup restatement with a known trapdoor; bases come from the device fixed-base kernel."""
    import random
    rnd = random.Random(seed)
    r1cs = G.mul_chain_circuit(nb_mul)
    td = G.Trapdoor(*[rnd.randrange(1, B.R) for _ in range(5)])
    pk, vk = G.setup(r1cs, td, points=False)
    g1 = cref.g1_points_to_array([B.G1_GEN])[0]
    g2 = cref.g2_points_to_array([B.G2_GEN])[0]
    mul1 = lambda xs: ctx.batch_scalar_mul(g1, util.fr_arr(xs), 1)
    mul2 = lambda xs: ctx.batch_scalar_mul(g2, util.fr_arr(xs), 2)
    arrs = dict(g1_a=mul1(pk.sc["A"]), g1_b=mul1(pk.sc["B"]), g1_k=mul1(pk.sc["K"]), g1_z=mul1(pk.sc["Z"]), g2_b=mul2(pk.sc["B"]))
    vkpts = [mul1([td.alpha])[0], mul1([td.beta])[0], mul1([td.delta])[0], mul2([td.beta])[0], mul2([td.delta])[0]]
    d = pk.domain
    pkh = ctx.groth16_pk_upload(log_n=d.log_n, generator=cref.fr(d.generator), coset_gen=cref.fr(d.coset_gen),
                                infinity_a=pk.infinity_A, infinity_b=pk.infinity_B, nb_public=r1cs.nb_public,
                                shard_index=shard[0], shard_count=shard[1], **arrs)
    return r1cs, td, pk, vk, pkh, vkpts


@pytest.mark.parametrize("nb_mul", [(1 << 10) - 1, (1 << 12) - 1, 3000])
def test_trapdoor_proof_verifies(ctx, pkg, cref, nb_mul):
    r1cs, td, pk, vk, pkh, vkpts = build_trapdoor_key(ctx, cref, nb_mul)
    w, a, b, c = G.mul_chain_witness(nb_mul, 2)
    r, s = 0x1234567890ABCDEF1234567890ABCDEF % B.R, 0xFEDCBA0987654321FEDCBA0987654321 % B.R
    ar, bs, krs, tm = _prove(ctx, pkg, cref, pkh, vkpts, w, a, b, c, r, s)
    dl = G.prove_dlog(pk, td, w, a, b, c, r, s, r1cs.nb_public)
    assert G.trapdoor_check(vk, w[: r1cs.nb_public], dl["ar"], dl["bs"], dl["krs"])
    ear, ekrs, ebs = B.g1_mul(B.G1_GEN, dl["ar"]), B.g1_mul(B.G1_GEN, dl["krs"]), B.g2_mul(B.G2_GEN, dl["bs"])
    assert util.g1_aff_ints(ar) == [ear[0], ear[1]]
    assert util.g1_aff_ints(krs) == [ekrs[0], ekrs[1]]
    assert util.g2_aff_ints(bs) == [ebs[0][0], ebs[0][1], ebs[1][0], ebs[1][1]]
    assert pkg.groth16_proof_bytes(ar, bs, krs) == G.proof_bytes(ear, ebs, ekrs)
    assert tm["total_ms"] > 0
    pkh.free()


def test_sharded_partial_sums_combine(ctx, pkg, cref):
    """MSM sharding by point range (2 shards on one device): host-combined partial sums equal the
    unsharded result."""
    nb_mul = (1 << 10) - 1
    w, a, b, c = G.mul_chain_witness(nb_mul, 2)
    args = [util.fr_arr(x) for x in (a, b, c, w)]
    _, _, _, _, pk_full, _ = build_trapdoor_key(ctx, cref, nb_mul)
    full, _ = ctx.groth16_prove_core(pk_full, *args)
    parts = []
    for i in range(2):
        _, _, _, _, pk_i, _ = build_trapdoor_key(ctx, cref, nb_mul, shard=(i, 2))
        parts.append(ctx.groth16_prove_core(pk_i, *args)[0].as_dict())
        pk_i.free()
    fd = full.as_dict()
    for k in ("ar", "bs1", "krs", "krs2"):
        comb = pkg.g1_jac_add(parts[0][k], parts[1][k])
        assert np.array_equal(pkg.g1_jac_to_affine(comb), pkg.g1_jac_to_affine(fd[k])), k
    comb = pkg.g2_jac_add(parts[0]["bs2"], parts[1]["bs2"])
    assert np.array_equal(pkg.g2_jac_to_affine(comb), pkg.g2_jac_to_affine(fd["bs2"]))
    pk_full.free()


class _LockedEngine:
    """DeviceEngine whose calls into the (non re-entrant) context are serialised: the virtual ranks
    of the tests below share one device and one b200_ctx."""

    def __init__(self, pkg, ctx, lock, log_n, gen, coset, rank, world):
        self.e, self.lock = pkg.sharding.DeviceEngine(ctx, log_n, gen, coset, rank, world), lock

    def phase(self, *a, **k):
        with self.lock:
            self.e.phase(*a, **k)

    def mul_sub(self, *a):
        with self.lock:
            self.e.mul_sub(*a)


def _run_virtual_ranks(pkg, ctx, log_n, gen, coset, world, a, b, c):
    """compute_h_distributed on `world` virtual ranks of one device; the all-to-all goes through the host."""
    import threading
    n_loc = (1 << log_n) // world
    lock, barrier, slots = threading.Lock(), threading.Barrier(world), [None] * world
    res, errs = [None] * world, []

    def pad(x):
        o = np.zeros((1 << log_n, 4), dtype=np.uint64)
        o[: len(x)] = x
        return o
    a, b, c = pad(a), pad(b), pad(c)

    def run(rank):
        try:
            with lock:
                bufs = [pkg.DeviceVector(ctx, n_loc * 32).upload(np.ascontiguousarray(v[rank * n_loc:(rank + 1) * n_loc])) for v in (a, b, c)]
                tmp = pkg.DeviceVector(ctx, n_loc * 32)

            def exchange(send, recv):
                with lock:
                    slots[rank] = send.download((n_loc, 4))
                barrier.wait()
                chunk = n_loc // world
                got = np.concatenate([slots[r][rank * chunk:(rank + 1) * chunk] for r in range(world)])
                barrier.wait()
                with lock:
                    recv.upload(got)

            eng = _LockedEngine(pkg, ctx, lock, log_n, gen, coset, rank, world)
            h = pkg.sharding.compute_h_distributed(eng, exchange, bufs[0], bufs[1], bufs[2], tmp)
            res[rank] = (h, bufs, tmp)
        except Exception as e:   # noqa: BLE001
            errs.append(e)
            barrier.abort()

    ts = [threading.Thread(target=run, args=(r,)) for r in range(world)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errs, errs
    return res


@pytest.mark.parametrize("log_n,world", [(4, 2), (6, 8), (10, 2), (10, 4), (12, 8), (14, 4), (16, 8), (18, 2)])
def test_distributed_compute_h_matches_single_device(ctx, pkg, cref, log_n, world):
    from tests.conftest import rand_field
    n = 1 << log_n
    rng = np.random.default_rng(log_n * 10 + world)
    a, b, c = (rand_field(rng, n - 5) for _ in range(3))
    d = G.ntt.Domain(n)
    gen, coset = cref.fr(d.generator), cref.fr(d.coset_gen)
    want = ctx.compute_h(a, b, c, log_n, gen, coset)
    if log_n <= 14:
        assert np.array_equal(want, cref.compute_h(a, b, c, log_n, gen, coset))
    res = _run_virtual_ranks(pkg, ctx, log_n, gen, coset, world, a, b, c)
    n_loc = n // world
    got = np.concatenate([h.download((n_loc, 4)) for h, _, _ in res])
    assert np.array_equal(got, want)
    for h, bufs, tmp in res:
        [v.free() for v in bufs]
        tmp.free()


def test_distributed_prove_core_two_virtual_ranks(ctx, pkg, cref):
    """Distributed computeH + b200_groth16_msms_dev on 2 virtual ranks: the combined partial sums
    equal the single-device prove core."""
    nb_mul = (1 << 10) - 1
    w, a, b, c = G.mul_chain_witness(nb_mul, 2)
    fa, fb, fc, fw = (util.fr_arr(x) for x in (a, b, c, w))
    r1cs, td, pk, vk, pk_full, _ = build_trapdoor_key(ctx, cref, nb_mul)
    full = ctx.groth16_prove_core(pk_full, fa, fb, fc, fw)[0].as_dict()
    d = pk.domain
    gen, coset = cref.fr(d.generator), cref.fr(d.coset_gen)
    res = _run_virtual_ranks(pkg, ctx, d.log_n, gen, coset, 2, fa, fb, fc)
    wv = pkg.DeviceVector(ctx, fw.nbytes).upload(fw)
    parts = []
    for i in range(2):
        _, _, _, _, pk_i, _ = build_trapdoor_key(ctx, cref, nb_mul, shard=(i, 2))
        parts.append(ctx.groth16_msms_dev(pk_i, wv, res[i][0])[0].as_dict())
        pk_i.free()
    for k in ("ar", "bs1", "krs", "krs2"):
        comb = pkg.g1_jac_add(parts[0][k], parts[1][k])
        assert np.array_equal(pkg.g1_jac_to_affine(comb), pkg.g1_jac_to_affine(full[k])), k
    comb = pkg.g2_jac_add(parts[0]["bs2"], parts[1]["bs2"])
    assert np.array_equal(pkg.g2_jac_to_affine(comb), pkg.g2_jac_to_affine(full["bs2"]))
    pk_full.free()


def test_prove_core_serial_and_overlapped_agree(ctx, pkg, cref):
    """msm_overlap = 0 (MSMs strictly one after the other) and the default (sort of the next MSM on a
    second stream, Bs2 reusing Bs1's sort) give the same five results; so does the device-input entry."""
    nb_mul = 3000
    w, a, b, c = G.mul_chain_witness(nb_mul, 2)
    args = [util.fr_arr(x) for x in (a, b, c, w)]
    _, _, _, _, pkh, _ = build_trapdoor_key(ctx, cref, nb_mul)
    ref = ctx.groth16_prove_core(pkh, *args)[0].as_dict()
    ctx.set_option("msm_overlap", 0)
    try:
        serial = ctx.groth16_prove_core(pkh, *args)[0].as_dict()
    finally:
        ctx.set_option("msm_overlap", 1)
    dv = [pkg.DeviceVector(ctx, x.nbytes).upload(x) for x in args]
    dev = ctx.groth16_prove_core_dev(pkh, dv[0], dv[1], dv[2], len(args[0]), dv[3])[0].as_dict()
    for k in ("ar", "bs1", "krs", "krs2"):
        assert np.array_equal(pkg.g1_jac_to_affine(serial[k]), pkg.g1_jac_to_affine(ref[k])), k
        assert np.array_equal(pkg.g1_jac_to_affine(dev[k]), pkg.g1_jac_to_affine(ref[k])), k
    assert np.array_equal(pkg.g2_jac_to_affine(serial["bs2"]), pkg.g2_jac_to_affine(ref["bs2"]))
    assert np.array_equal(pkg.g2_jac_to_affine(dev["bs2"]), pkg.g2_jac_to_affine(ref["bs2"]))
    [v.free() for v in dv]
    pkh.free()


@pytest.mark.parametrize("h_pos", [0, 1, 2])
def test_prove_core_same_results_wherever_compute_h_is_enqueued(ctx, pkg, cref, h_pos):
    """Option h_pos (groth16.cu): computeH of a host-buffer proof is enqueued after the accumulation of A, Bs1 or Krs.
    The five results do not depend on it, and the timings stay consistent (the stages add up to the total)."""
    nb_mul = 3000
    w, a, b, c = G.mul_chain_witness(nb_mul, 2)
    args = [util.fr_arr(x) for x in (a, b, c, w)]
    _, _, _, _, pkh, _ = build_trapdoor_key(ctx, cref, nb_mul)
    ref = ctx.groth16_prove_core(pkh, *args)[0].as_dict()
    ctx.set_option("h_pos", h_pos)
    try:
        out, tm = ctx.groth16_prove_core(pkh, *args)
    finally:
        ctx.set_option("h_pos", -1)
    got = out.as_dict()
    for k in ("ar", "bs1", "krs", "krs2"):
        assert np.array_equal(pkg.g1_jac_to_affine(got[k]), pkg.g1_jac_to_affine(ref[k])), k
    assert np.array_equal(pkg.g2_jac_to_affine(got["bs2"]), pkg.g2_jac_to_affine(ref["bs2"]))
    parts = sum(tm[k] for k in ("compute_h_ms", "msm_a_ms", "msm_b1_ms", "msm_k_ms", "msm_b2_ms", "msm_z_ms"))
    assert abs(parts - tm["total_ms"]) < 0.25 * tm["total_ms"] + 0.5, (parts, tm)
    with pytest.raises(pkg.B200Error):
        ctx.set_option("h_pos", 3)
    pkh.free()


def test_prove_core_with_persistent_aux_grids(ctx, pkg, cref):
    """The sort's scatter and the round-0 denominator pass as small persistent grids (msm_aux_persistent, msm_host.cuh) --
    the form the prove core uses from 2^22-point MSMs up -- forced here on a small instance (rounds forced on too):
    the five results equal those of the full-grid kernels."""
    nb_mul = (1 << 12) - 1
    w, a, b, c = G.mul_chain_witness(nb_mul, 2)
    args = [util.fr_arr(x) for x in (a, b, c, w)]
    _, _, _, _, pkh, _ = build_trapdoor_key(ctx, cref, nb_mul)
    ref = ctx.groth16_prove_core(pkh, *args)[0].as_dict()
    ctx.set_option("msm_ba", 2)
    ctx.set_option("msm_ba_force", 1)
    try:
        rounds = ctx.groth16_prove_core(pkh, *args)[0].as_dict()
        ctx.set_option("aux_persist_min", 0)
        persistent = ctx.groth16_prove_core(pkh, *args)[0].as_dict()
    finally:
        ctx.set_option("aux_persist_min", -1)
        ctx.set_option("msm_ba_force", 0)
        ctx.set_option("msm_ba", 15)
    for got in (rounds, persistent):
        for k in ("ar", "bs1", "krs", "krs2"):
            assert np.array_equal(pkg.g1_jac_to_affine(got[k]), pkg.g1_jac_to_affine(ref[k])), k
        assert np.array_equal(pkg.g2_jac_to_affine(got["bs2"]), pkg.g2_jac_to_affine(ref["bs2"]))
    pkh.free()
