"""The device PlonK prover (csrc/plonk.cu, through the C ABI) against oracle/plonk_prove.py:
byte-identical proofs under fixed blinding, the committed golden proofs, and acceptance by the oracle's
restatement of plonk.Verify (verify.go:45-294) for proofs made with fresh randomness."""
import json
import os
import random

import numpy as np
import pytest

from oracle import plonk_prove as PP
from tests import plonk_util as PU
from tests import util

pytestmark = pytest.mark.gpu
R = PP.R
HERE = os.path.dirname(os.path.abspath(__file__))


def _case(nb_mul, seed):
    rng = random.Random(seed)
    spr = PP.mul_chain_scs(nb_mul)
    pk = PP.setup(spr, rng.randrange(1, R))
    sol = PP.mul_chain_solution(nb_mul, rng.randrange(2, R))
    blinding = tuple([rng.randrange(R) for _ in range(k)] for k in (2, 2, 2, 3))
    return spr, pk, sol, blinding


def _device_prove(ctx, dpk, spr, sol, blinding):
    L, Rr, O = PP.evaluate_lro(spr, sol)
    bl = None if blinding is None else PU.blinding_array(blinding)
    return ctx.plonk_prove(dpk, util.fr_arr(L), util.fr_arr(Rr), util.fr_arr(O), util.fr_arr(sol[:spr.nb_public]), bl)


# sizeSystem 3 (n = 4, big = 32: rho = 8), 6 / 7 (n = 8, rho = 4), ragged sizes, 2^10 and 2^12 rows
@pytest.mark.parametrize("nb_mul", [1, 2, 4, 5, 13, 100, 1000, 4000])
def test_proof_bytes_identical_to_oracle(ctx, pkg, nb_mul):
    spr, pk, sol, blinding = _case(nb_mul, 500 + nb_mul)
    want, ch = PP.prove(spr, pk, sol, blinding)
    dpk = PU.upload_pk(ctx, pk)
    try:
        got, tm = _device_prove(ctx, dpk, spr, sol, blinding)
    finally:
        dpk.free()
    d = got.as_dict()
    assert util.fr_ints(d["challenges"]) == [ch["gamma"], ch["beta"], ch["alpha"], ch["zeta"], ch["gamma_kzg"]]
    assert util.fr_ints(d["claimed"]) == want.claimed
    assert pkg.plonk_proof_bytes(got) == PP.proof_bytes(want)
    assert tm["total_ms"] > 0


def test_golden_proofs(ctx, pkg):
    with open(os.path.join(HERE, "golden", "plonk_small.json")) as f:
        cases = json.load(f)["cases"]
    for c in cases:
        spr = PP.mul_chain_scs(c["nb_mul"])
        pk = PP.setup(spr, int(c["tau"], 16))
        sol = PP.mul_chain_solution(c["nb_mul"], int(c["x"], 16))
        blinding = tuple([int(v, 16) for v in b] for b in c["blinding"])
        dpk = PU.upload_pk(ctx, pk)
        try:
            got, _ = _device_prove(ctx, dpk, spr, sol, blinding)
        finally:
            dpk.free()
        assert pkg.plonk_proof_bytes(got).hex() == c["proof_bytes"]


def test_fresh_randomness_proofs_verify_and_differ(ctx, pkg):
    spr, pk, sol, _ = _case(300, 77)
    dpk = PU.upload_pk(ctx, pk)
    try:
        p1, _ = _device_prove(ctx, dpk, spr, sol, None)
        p2, _ = _device_prove(ctx, dpk, spr, sol, None)
    finally:
        dpk.free()
    assert pkg.plonk_proof_bytes(p1) != pkg.plonk_proof_bytes(p2)
    for p in (p1, p2):
        assert PP.verify(PU.struct_to_proof(p), pk.vk, sol[:1], pk.tau)
    assert not PP.verify(PU.struct_to_proof(p1), pk.vk, [(sol[0] + 1) % R], pk.tau)


def test_same_key_many_proofs(ctx, pkg):
    """A resident key serves different witnesses (the workspace is reused)."""
    spr, pk, _, blinding = _case(50, 5)
    dpk = PU.upload_pk(ctx, pk)
    try:
        for x in (2, 3, R - 5):
            sol = PP.mul_chain_solution(50, x)
            want, _ = PP.prove(spr, pk, sol, blinding)
            got, _ = _device_prove(ctx, dpk, spr, sol, blinding)
            assert pkg.plonk_proof_bytes(got) == PP.proof_bytes(want)
    finally:
        dpk.free()


def test_bad_arguments(ctx, pkg):
    spr, pk, sol, blinding = _case(5, 1)
    with pytest.raises(pkg.B200Error):
        kzg, lag = PU.srs_arrays(pk)
        f = util.fr_arr
        ctx.plonk_pk_upload(log_n=pk.d0.log_n, log_big=pk.d0.log_n + 1, generator=util.fr_arr([pk.d0.generator])[0],
                            generator_big=util.fr_arr([pk.d1.generator])[0], coset_gen=util.fr_arr([5])[0], ql=f(pk.ql), qr=f(pk.qr),
                            qm=f(pk.qm), qo=f(pk.qo), qk=f(pk.qk), s1=f(pk.s1), s2=f(pk.s2), s3=f(pk.s3),
                            perm=np.array(pk.perm, dtype=np.int64), kzg_g1=kzg, kzg_lagrange_g1=lag,
                            vk_commitments=util.g1_arr(PU.vk_points(pk)), nb_public=1)


def _host_mod():
    from importlib import import_module
    from tests.conftest import load_pkg
    load_pkg()
    return import_module("gnark_fork_b200.plonk_host")


@pytest.mark.parametrize("log_n", [3, 7, 10])
def test_device_setup_mirror_matches_oracle_setup(ctx, pkg, log_n):
    """plonk_host.setup (Setup on the device through the ABI, setup.go:110-161) == the oracle's setup:
    canonical trace, S1..S3, both SRS, the eight vk commitments; then proofs from that key are the oracle's."""
    H = _host_mod()
    n = 1 << log_n
    tau = random.Random(log_n).randrange(1, R)
    trace, lro, nb_vars, nb_pub, L, Rr, O, pub = H.mul_chain(log_n, 5)
    kw = H.setup(ctx, H.Domains(n), trace, lro, nb_vars, nb_pub, tau)
    spr = PP.mul_chain_scs(n - 2)
    pk = PP.setup(spr, tau)
    for k in ("ql", "qr", "qm", "qo", "qk", "s1", "s2", "s3"):
        assert util.fr_ints(kw[k]) == getattr(pk, k), k
    assert kw["perm"].tolist() == pk.perm
    kzg, lag = PU.srs_arrays(pk)
    assert np.array_equal(kw["kzg_g1"], kzg) and np.array_equal(kw["kzg_lagrange_g1"], lag)
    assert np.array_equal(kw["vk_commitments"], util.g1_arr(PU.vk_points(pk)))
    blinding = ([1, 2], [3, 4], [5, 6], [7, 8, 9])
    dpk = ctx.plonk_pk_upload(**kw)
    try:
        got, _ = ctx.plonk_prove(dpk, L, Rr, O, pub, PU.blinding_array(blinding))
    finally:
        dpk.free()
    want, _ = PP.prove(spr, pk, PP.mul_chain_solution(n - 2, 5), blinding)
    assert pkg.plonk_proof_bytes(got) == PP.proof_bytes(want)


@pytest.mark.parametrize("log_n", [16, 22])
def test_large_proof_verifies(ctx, pkg, log_n):
    """A proof at 2^16 and at 2^22 rows (BASELINE.json config 4's size) made from a key the device built, checked by
    the oracle's plonk.Verify restatement (O(1) work: it only needs vk and tau)."""
    H = _host_mod()
    n = 1 << log_n
    tau = random.Random(99).randrange(1, R)
    trace, lro, nb_vars, nb_pub, L, Rr, O, pub = H.mul_chain(log_n, 7)
    d = H.Domains(n)
    kw = H.setup(ctx, d, trace, lro, nb_vars, nb_pub, tau)
    dpk = ctx.plonk_pk_upload(**kw)
    pinned = []
    for a in (L, Rr, O):                                       # the solver's output buffers, page-locked
        b = pkg.pinned_empty(a.shape)
        b[:] = a
        pinned.append(b)
    L, Rr, O = pinned
    try:
        ctx.plonk_prove(dpk, L, Rr, O, pub, None)              # warm-up
        proof, tm = ctx.plonk_prove(dpk, L, Rr, O, pub, None)
    finally:
        dpk.free()
    pts = [util.g1_aff_ints(v) for v in kw["vk_commitments"]]
    pts = [None if q is None else tuple(q) for q in pts]            # [Qk] of a Mul chain is the point at infinity
    vk = {"size": n, "size_inv": pow(n, -1, R), "generator": d.gen, "nb_public": 1, "coset_shift": 5,
          "S": pts[:3], "Ql": pts[3], "Qr": pts[4], "Qm": pts[5], "Qo": pts[6], "Qk": pts[7]}
    public = util.fr_ints(pub)
    assert PP.verify(PU.struct_to_proof(proof), vk, public, tau)
    assert not PP.verify(PU.struct_to_proof(proof), vk, [(public[0] + 1) % R], tau)
    out = os.path.join(os.path.dirname(HERE), "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, "plonk_prove_2p%d.json" % log_n), "w") as f:
            json.dump({"log_n": log_n, "verified_by_oracle": True, "timings_ms": tm}, f, indent=1)
    print("plonk prove 2^%d:" % log_n, tm)


@pytest.mark.parametrize("nb_mul,nb_commit", [(3, 1), (6, 2), (40, 3), (1000, 1)])
def test_bsb22_commitment_circuits(ctx, pkg, nb_mul, nb_commit):
    """api.Commit circuits (prove.go:303-352): Qcp / Pi2 columns through the quotient, the linearised polynomial,
    the transcripts and the batch opening -- proof bytes identical to the oracle, accepted by its verifier."""
    rng = random.Random(900 + nb_mul)
    spr = PP.commit_chain_scs(nb_mul, nb_commit)
    pk = PP.setup(spr, rng.randrange(1, R))
    sol, bsb = PP.commit_chain_solution(spr, pk, rng.randrange(2, R), [(rng.randrange(R), rng.randrange(R)) for _ in range(nb_commit)])
    blinding = tuple([rng.randrange(R) for _ in range(k)] for k in (2, 2, 2, 3))
    want, ch = PP.prove(spr, pk, sol, blinding, bsb)
    dpk = PU.upload_pk(ctx, pk)
    try:
        L, Rr, O = PP.evaluate_lro(spr, sol)
        got, _ = ctx.plonk_prove(dpk, util.fr_arr(L), util.fr_arr(Rr), util.fr_arr(O), util.fr_arr(sol[:1]),
                                 PU.blinding_array(blinding), PU.bsb_arrays(bsb))
        fresh, _ = ctx.plonk_prove(dpk, util.fr_arr(L), util.fr_arr(Rr), util.fr_arr(O), util.fr_arr(sol[:1]), None, PU.bsb_arrays(bsb))
        # the hint's commitment is a KZG commitment against KzgLagrange: the same resident-SRS MSM path
        srs = ctx.upload_bases(PU.srs_arrays(pk)[1], 1)
        vec = pkg.DeviceVector(ctx, pk.d0.cardinality * 32).upload(util.fr_arr(bsb["pi2"][0]))
        assert util.g1_aff_ints(ctx.kzg_commit_dev(srs, vec, pk.d0.cardinality)) == list(bsb["commitments"][0])
        vec.free(); srs.free()
    finally:
        dpk.free()
    assert util.fr_ints(got.as_dict()["challenges"]) == [ch["gamma"], ch["beta"], ch["alpha"], ch["zeta"], ch["gamma_kzg"]]
    assert pkg.plonk_proof_bytes(got) == PP.proof_bytes(want)
    assert PP.verify(PU.struct_to_proof(fresh), pk.vk, sol[:1], pk.tau)


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("nb_mul,nb_commit", [(5, 0), (100, 0), (40, 2)])
def test_sharded_proof_on_virtual_ranks(ctx, pkg, world, nb_mul, nb_commit):
    """The phase ABI with the key sharded over `world` virtual ranks on one GPU (real kernels; the exchanges --
    sum of partial commitments, all-gather of the numerator's coset blocks -- done through the host): the proof is
    byte-identical to the single-GPU one and to the oracle's.  rho = 8 at the smallest size, 4 otherwise."""
    rng = random.Random(7000 + nb_mul)
    if nb_commit:
        spr = PP.commit_chain_scs(nb_mul, nb_commit)
        pk = PP.setup(spr, rng.randrange(1, R))
        sol, bsb = PP.commit_chain_solution(spr, pk, rng.randrange(2, R), [(rng.randrange(R), rng.randrange(R)) for _ in range(nb_commit)])
    else:
        spr, bsb = PP.mul_chain_scs(nb_mul), None
        pk = PP.setup(spr, rng.randrange(1, R))
        sol = PP.mul_chain_solution(nb_mul, rng.randrange(2, R))
    blinding = tuple([rng.randrange(R) for _ in range(k)] for k in (2, 2, 2, 3))
    want, _ = PP.prove(spr, pk, sol, blinding, bsb)
    L, Rr, O = (util.fr_arr(v) for v in PP.evaluate_lro(spr, sol))
    pub, bl, bsb_a = util.fr_arr(sol[:1]), PU.blinding_array(blinding), PU.bsb_arrays(bsb)
    S = pkg.sharding
    keys = [PU.upload_pk(ctx, pk, shard_index=k, shard_count=world) for k in range(world)]
    sess = []
    try:
        sess = [ctx.plonk_session(keys[k], L, Rr, O, pub, bl, bsb_a) for k in range(world)]
        tot = S.sum_partials([s.partials for s in sess])
        tot = S.sum_partials([s.step(1, tot) for s in sess])
        for s in sess:
            s.step(2, tot)
        # exchange of the coset blocks through the host
        ptrs = [s.numerator() for s in sess]
        _, be, lb = ptrs[0]
        rho = 1 << lb
        for k in range(rho):
            owner, blk = k % world, S.coset_block(k, lb)
            src = pkg.DeviceVector.__new__(pkg.DeviceVector)
            src.ctx, src.nbytes, src.ptr = ctx, be * 32, __import__("ctypes").c_void_p(ptrs[owner][0] + blk * be * 32)
            data = src.download((be, 4))
            for r2 in range(world):
                if r2 != owner:
                    dst = pkg.DeviceVector.__new__(pkg.DeviceVector)
                    dst.ctx, dst.nbytes, dst.ptr = ctx, be * 32, __import__("ctypes").c_void_p(ptrs[r2][0] + blk * be * 32)
                    dst.upload(data)
        tot = S.sum_partials([s.step(3) for s in sess])
        tot = S.sum_partials([s.step(4, tot) for s in sess])
        tot = S.sum_partials([s.step(5, tot) for s in sess])
        proofs = [s.finish(tot)[0] for s in sess]
    finally:
        for s in sess:
            s.abandon()
        for k in keys:
            k.free()
    for p in proofs:
        assert pkg.plonk_proof_bytes(p) == PP.proof_bytes(want)


def test_second_session_on_a_key_in_flight_is_refused(ctx, pkg):
    """The per-proof workspace lives in the key: a second session before the first one ends would corrupt both proofs,
    so b200_plonk_session_begin refuses it (and the binding refuses short l, r, o: the C ABI takes no lengths)."""
    rng = random.Random(77)
    spr = PP.mul_chain_scs(30)
    pk = PP.setup(spr, rng.randrange(1, R))
    sol = PP.mul_chain_solution(30, 5)
    L, Rr, O = (util.fr_arr(v) for v in PP.evaluate_lro(spr, sol))
    pub = util.fr_arr(sol[:1])
    bl = PU.blinding_array(([1, 2], [3, 4], [5, 6], [7, 8, 9]))
    dpk = PU.upload_pk(ctx, pk)
    try:
        s1 = ctx.plonk_session(dpk, L, Rr, O, pub, bl)
        with pytest.raises(pkg.B200Error, match="already in flight"):
            ctx.plonk_session(dpk, L, Rr, O, pub, bl)
        with pytest.raises(pkg.B200Error, match="already in flight"):
            ctx.plonk_prove(dpk, L, Rr, O, pub, bl)
        s1.abandon()
        with pytest.raises(pkg.B200Error, match="rows"):
            ctx.plonk_prove(dpk, L[:-1], Rr, O, pub, bl)
        got, _ = ctx.plonk_prove(dpk, L, Rr, O, pub, bl)          # the key is free again
        want, _ = PP.prove(spr, pk, sol, ([1, 2], [3, 4], [5, 6], [7, 8, 9]))
        assert pkg.plonk_proof_bytes(got) == PP.proof_bytes(want)
    finally:
        dpk.free()
