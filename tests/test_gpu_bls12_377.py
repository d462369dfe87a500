"""BLS12-377 (BASELINE.json config 5) through the C ABI of libb200prover_bls12377.so against the BLS12-377 oracle:
12-limb Montgomery Fp, Fr NTT / computeH, G1 and G2 (Fp2 with u^2 = -5) MSM, and the Groth16 prove core with the
fix-ups and Proof.WriteTo -- bit-exact on affine-normalised points and bytes
(reference: backend/groth16/bls12-377/prove.go, the same template as bn254/prove.go)."""
import random

import numpy as np
import pytest

from oracle import bls12_377 as B
from oracle.suite377 import groth16 as G
from oracle.suite377 import ntt as N
from tests import util377 as U

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx377(pkg):
    c = pkg.Context(0, curve=pkg.BLS12_377)
    yield c
    c.close()


def test_library_identity(pkg, ctx377):
    assert ctx377.lib.b200_build_curve() == pkg.BLS12_377 and ctx377.lib.b200_build_fp_words() == 6
    assert b"BLS12-377" in ctx377.lib.b200_version()
    with pytest.raises(pkg.B200Error):      # the BN254 id is refused by this library, with a message
        ctx377._check(ctx377.lib.b200_fr_vec_mul(ctx377.h, pkg.BN254, None, None, None, 0))


def test_field_mul(ctx377):
    rng = np.random.default_rng(1)
    n = 3000
    a, b = U.rand_fr(rng, n), U.rand_fr(rng, n)
    edge = U.fr_arr([0, 1, B.R - 1, 2, B.R - 2])
    a[:5], b[:5] = edge, edge[::-1]
    assert U.fr_ints(ctx377.fr_vec_mul(a, b)) == [x * y % B.R for x, y in zip(U.fr_ints(a), U.fr_ints(b))]
    a, b = U.rand_fp(rng, n), U.rand_fp(rng, n)
    edge = U.fp_arr([0, 1, B.P - 1, 2, B.P - 2])
    a[:5], b[:5] = edge, edge[::-1]
    assert U.fp_ints(ctx377.fp_vec_mul(a, b)) == [x * y % B.P for x, y in zip(U.fp_ints(a), U.fp_ints(b))]


@pytest.mark.parametrize("log_n", [1, 2, 5, 9, 10, 13])
@pytest.mark.parametrize("inverse", [False, True])
@pytest.mark.parametrize("dit", [0, 1])
@pytest.mark.parametrize("coset", [False, True])
def test_ntt_matches_oracle(ctx377, log_n, inverse, dit, coset):
    rng = np.random.default_rng(log_n)
    n = 1 << log_n
    x = U.rand_fr(rng, n)
    d = N.Domain(n)
    got = ctx377.ntt(x, log_n, inverse, dit, coset, U.fr(d.generator), U.fr(d.coset_gen))
    want = U.fr_ints(x)
    (N.fft_inverse if inverse else N.fft)(want, d, dit, coset)
    assert U.fr_ints(got) == want


def test_ntt_round_trip_large(ctx377):
    log_n = 20
    rng = np.random.default_rng(5)
    x = U.rand_fr(rng, 1 << log_n)
    d = N.Domain(1 << log_n)
    g, cg = U.fr(d.generator), U.fr(d.coset_gen)
    y = ctx377.ntt(x, log_n, False, 0, True, g, cg)       # coset FFT, DIF: bit-reversed out
    z = ctx377.ntt(y, log_n, True, 1, True, g, cg)        # inverse, DIT: natural out
    assert np.array_equal(z, x)


def test_compute_h(ctx377):
    rng = np.random.default_rng(9)
    n = 1 << 8
    a, b, c = (U.rand_fr(rng, n - 3) for _ in range(3))
    d = N.Domain(n)
    got = ctx377.compute_h(a, b, c, d.log_n, U.fr(d.generator), U.fr(d.coset_gen))
    pad = lambda v: U.fr_ints(v) + [0] * 3
    want = N.compute_h(pad(a), pad(b), pad(c), d)
    assert U.fr_ints(got) == want


@pytest.mark.parametrize("n", [0, 1, 7, 300, 5000])
def test_msm_g1_matches_oracle(ctx377, n):
    rng = np.random.default_rng(7 + n)
    pts, ks = U.known_dlog_bases(1, n, seed=n + 1)
    scalars = U.rand_fr(rng, n)
    if n > 3:
        pts[n // 2] = None
        pts[3] = pts[2]
        scalars[0] = 0
        scalars[1] = U.fr(B.R - 1)
    got = ctx377.msm_g1(scalars, U.g1_arr(pts)) if n else ctx377.msm_g1(np.zeros((0, 4), np.uint64), np.zeros((0, 12), np.uint64))
    if n <= 300:
        want = B.msm_g1(pts, U.fr_ints(scalars))
    else:
        e = sum(k * s for k, s, p in zip(ks, U.fr_ints(scalars), pts) if p is not None)
        e += (ks[2] - ks[3]) * U.fr_ints(scalars[3:4])[0]          # base 3 was replaced by base 2
        want = B.g1_mul(B.G1_GEN, e % B.R)
    assert U.g1_jac(got) == want


@pytest.mark.parametrize("n", [1, 65, 2000])
def test_msm_g2_matches_oracle(ctx377, n):
    rng = np.random.default_rng(11 + n)
    pts, ks = U.known_dlog_bases(2, n, seed=n + 2)
    scalars = U.rand_fr(rng, n)
    got = ctx377.msm_g2(scalars, U.g2_arr(pts))
    want = B.g2_mul(B.G2_GEN, sum(k * s for k, s in zip(ks, U.fr_ints(scalars))) % B.R)
    assert U.g2_jac(got) == want


def test_msm_resident_folded_tables_large(ctx377):
    """2^17 points through the resident (precomputed-table) path: the result is (sum k_i s_i) G."""
    n = 1 << 17
    rng = np.random.default_rng(3)
    ks = [int(x) for x in rng.integers(1, 1 << 62, size=n)]
    bases = ctx377.batch_scalar_mul(U.g1_arr([B.G1_GEN])[0], U.fr_arr(ks), 1)
    spot = [0, 1, n // 3, n - 1]
    for i in spot:                                       # the device-made bases themselves, spot-checked
        assert U.g1_aff(bases[i]) == B.g1_mul(B.G1_GEN, ks[i])
    scalars = U.rand_fr(rng, n)
    h = ctx377.upload_bases(bases, 1)
    got = ctx377.msm_resident(h, scalars)
    want = B.g1_mul(B.G1_GEN, sum(k * s for k, s in zip(ks, U.fr_ints(scalars))) % B.R)
    assert U.g1_jac(got) == want
    h.free()


def test_host_group_helpers(pkg):
    c = pkg.BLS12_377
    p, q = B.g1_mul(B.G1_GEN, 5), B.g1_mul(B.G1_GEN, 11)
    jac = lambda pt: np.concatenate([U.fp_arr([pt[0], pt[1], 1]).reshape(-1)])
    s = pkg.g1_jac_add(jac(p), jac(q), curve=c)
    assert U.g1_aff(pkg.g1_jac_to_affine(s, curve=c)) == B.g1_mul(B.G1_GEN, 16)
    p2, q2 = B.g2_mul(B.G2_GEN, 5), B.g2_mul(B.G2_GEN, 11)
    jac2 = lambda pt: U.fp_arr([pt[0][0], pt[0][1], pt[1][0], pt[1][1], 1, 0]).reshape(-1)
    s2 = pkg.g2_jac_add(jac2(p2), jac2(q2), curve=c)
    assert U.g2_aff(pkg.g2_jac_to_affine(s2, curve=c)) == B.g2_mul(B.G2_GEN, 16)


def _key(ctx, nb_mul, seed=7, shard=(0, 1)):
    rnd = random.Random(seed)
    r1cs = G.mul_chain_circuit(nb_mul)
    td = G.Trapdoor(*[rnd.randrange(1, B.R) for _ in range(5)])
    pk, vk = G.setup(r1cs, td, points=False)
    g1, g2 = U.g1_arr([B.G1_GEN])[0], U.g2_arr([B.G2_GEN])[0]
    mul1 = lambda xs: ctx.batch_scalar_mul(g1, U.fr_arr(xs), 1)
    mul2 = lambda xs: ctx.batch_scalar_mul(g2, U.fr_arr(xs), 2)
    arrs = dict(g1_a=mul1(pk.sc["A"]), g1_b=mul1(pk.sc["B"]), g1_k=mul1(pk.sc["K"]), g1_z=mul1(pk.sc["Z"]), g2_b=mul2(pk.sc["B"]))
    vkpts = [mul1([td.alpha])[0], mul1([td.beta])[0], mul1([td.delta])[0], mul2([td.beta])[0], mul2([td.delta])[0]]
    d = pk.domain
    pkh = ctx.groth16_pk_upload(log_n=d.log_n, generator=U.fr(d.generator), coset_gen=U.fr(d.coset_gen),
                                infinity_a=pk.infinity_A, infinity_b=pk.infinity_B, nb_public=r1cs.nb_public,
                                shard_index=shard[0], shard_count=shard[1], **arrs)
    return r1cs, td, pk, vk, pkh, vkpts


@pytest.mark.parametrize("nb_mul", [(1 << 8) - 1, 3000])
def test_groth16_trapdoor_proof(ctx377, pkg, nb_mul):
    """prove.go:60-321 on BLS12-377: device prove core + host fix-ups + WriteTo equal the oracle's proof, which
    satisfies the verify equation in the exponent."""
    r1cs, td, pk, vk, pkh, vkpts = _key(ctx377, nb_mul)
    w, a, b, c = G.mul_chain_witness(nb_mul, 3)
    r, s = 0x1234567890ABCDEF1234567890ABCDEF % B.R, 0xFEDCBA0987654321FEDCBA0987654321 % B.R
    out, tm = ctx377.groth16_prove_core(pkh, U.fr_arr(a), U.fr_arr(b), U.fr_arr(c), U.fr_arr(w))
    ar, bs, krs = pkg.groth16_finalize(out, U.fr(r), U.fr(s), *vkpts, curve=pkg.BLS12_377)
    dl = G.prove_dlog(pk, td, w, a, b, c, r, s, r1cs.nb_public)
    assert G.trapdoor_check(vk, w[: r1cs.nb_public], dl["ar"], dl["bs"], dl["krs"])
    ear, ekrs, ebs = B.g1_mul(B.G1_GEN, dl["ar"]), B.g1_mul(B.G1_GEN, dl["krs"]), B.g2_mul(B.G2_GEN, dl["bs"])
    assert U.g1_aff(ar) == ear and U.g1_aff(krs) == ekrs and U.g2_aff(bs) == ebs
    raw = pkg.groth16_proof_bytes(ar, bs, krs, curve=pkg.BLS12_377)
    assert len(raw) == 244 and raw == G.proof_bytes(ear, ebs, ekrs)
    assert tm["total_ms"] > 0
    pkh.free()


@pytest.mark.parametrize("shards", [2, 4])
def test_sharded_partial_sums_combine(ctx377, pkg, shards):
    """MSM sharding by point range on BLS12-377: the host-combined partial sums of the shards (18- and
    36-word Jacobian points through b200_g1_jac_add / b200_g2_jac_add of this library) equal the unsharded results."""
    c = pkg.BLS12_377
    nb_mul = (1 << 9) - 1
    w, a, b, cc = G.mul_chain_witness(nb_mul, 3)
    args = [U.fr_arr(x) for x in (a, b, cc, w)]
    _, _, _, _, pk_full, _ = _key(ctx377, nb_mul)
    full = ctx377.groth16_prove_core(pk_full, *args)[0].as_dict()
    pk_full.free()
    parts = []
    for i in range(shards):
        _, _, _, _, pk_i, _ = _key(ctx377, nb_mul, shard=(i, shards))
        parts.append(ctx377.groth16_prove_core(pk_i, *args)[0].as_dict())
        pk_i.free()
    for k in ("ar", "bs1", "krs", "krs2", "bs2"):
        add, aff = (pkg.g2_jac_add, pkg.g2_jac_to_affine) if k == "bs2" else (pkg.g1_jac_add, pkg.g1_jac_to_affine)
        acc = parts[0][k]
        for p in parts[1:]:
            acc = add(acc, p[k], curve=c)
        assert np.array_equal(aff(acc, curve=c), aff(full[k], curve=c)), k


# ---- PlonK on BLS12-377 (backend/plonk/bls12-377/prove.go: the bn254 template with the curve swapped) -------------
def _plonk377_upload(ctx377, PP, pk):
    mono, lag = PP.srs_scalars(pk)
    gen = U.g1_arr([B.G1_GEN])[0]
    kzg = ctx377.batch_scalar_mul(gen, U.fr_arr(mono), 1)          # [tau^i]G1 and [L_i(tau)]G1 from the fixed-base kernel
    kzg_lag = ctx377.batch_scalar_mul(gen, U.fr_arr(lag), 1)
    vk = pk.vk
    vkp = [vk["S"][0], vk["S"][1], vk["S"][2], vk["Ql"], vk["Qr"], vk["Qm"], vk["Qo"], vk["Qk"]]
    f = U.fr_arr
    return ctx377.plonk_pk_upload(
        log_n=pk.d0.log_n, log_big=pk.d1.log_n, generator=U.fr(pk.d0.generator), generator_big=U.fr(pk.d1.generator),
        coset_gen=U.fr(pk.d0.coset_gen), ql=f(pk.ql), qr=f(pk.qr), qm=f(pk.qm), qo=f(pk.qo), qk=f(pk.qk),
        s1=f(pk.s1), s2=f(pk.s2), s3=f(pk.s3), perm=np.array(pk.perm, dtype=np.int64), kzg_g1=kzg, kzg_lagrange_g1=kzg_lag,
        vk_commitments=U.g1_arr(vkp), nb_public=pk.nb_public, qcp=[f(q) for q in pk.qcp],
        vk_qcp=U.g1_arr(pk.vk["Qcp"]) if pk.qcp else None, commitment_index=[c["index"] for c in pk.commitments])


def _plonk377_struct_to_proof(PP, s):
    d = s.as_dict()
    nb = int(s.nb_bsb22)
    return PP.Proof([U.g1_aff(d["lro"][k]) for k in range(3)], U.g1_aff(d["z"]), [U.g1_aff(d["h"][k]) for k in range(3)],
                    U.g1_aff(d["batched_h"]), U.fr_ints(s.all_claimed()), U.g1_aff(d["zshift_h"]),
                    U.fr_ints(d["zshift_value"].reshape(1, 4))[0], [U.g1_aff(d["bsb22"][k]) for k in range(nb)])


@pytest.mark.parametrize("nb_mul,nb_commit", [(6, 0), (300, 0), (4000, 0), (40, 2)])
def test_plonk_proof_bytes_identical_to_oracle(pkg, ctx377, nb_mul, nb_commit):
    """plonk.Prove of the second library against oracle/plonk_prove.py run on BLS12-377 (oracle/suite377.py): proof bytes
    identical under fixed blinding (696 bytes + 80 per BSB22 commitment), fresh-randomness proofs accepted by the verifier
    restatement, with and without api.Commit columns."""
    from oracle.suite377 import plonk_prove as PP
    rng = random.Random(377 + nb_mul)
    if nb_commit:
        spr = PP.commit_chain_scs(nb_mul, nb_commit)
        pk = PP.setup(spr, rng.randrange(1, B.R))
        sol, bsb = PP.commit_chain_solution(spr, pk, rng.randrange(2, B.R), [(rng.randrange(B.R), rng.randrange(B.R)) for _ in range(nb_commit)])
        bsb_arr = {"pi2": [U.fr_arr(c) for c in bsb["pi2"]], "commitments": U.g1_arr(bsb["commitments"]), "values": U.fr_arr(bsb["values"])}
    else:
        spr = PP.mul_chain_scs(nb_mul)
        pk = PP.setup(spr, rng.randrange(1, B.R))
        sol, bsb, bsb_arr = PP.mul_chain_solution(nb_mul, 5), None, None
    blinding = tuple([rng.randrange(B.R) for _ in range(k)] for k in (2, 2, 2, 3))
    want, ch = PP.prove(spr, pk, sol, blinding, bsb) if bsb else PP.prove(spr, pk, sol, blinding)
    L, Rr, O = (U.fr_arr(v) for v in PP.evaluate_lro(spr, sol))
    pub = U.fr_arr(sol[:1])
    bl = U.fr_arr([x for part in blinding for x in part])
    dpk = _plonk377_upload(ctx377, PP, pk)
    try:
        got, tm = ctx377.plonk_prove(dpk, L, Rr, O, pub, bl, bsb_arr)
        fresh, _ = ctx377.plonk_prove(dpk, L, Rr, O, pub, None, bsb_arr)
    finally:
        dpk.free()
    assert U.fr_ints(got.as_dict()["challenges"]) == [ch["gamma"], ch["beta"], ch["alpha"], ch["zeta"], ch["gamma_kzg"]]
    raw = pkg.plonk_proof_bytes(got)
    assert raw == PP.proof_bytes(want) and len(raw) == 696 + 80 * nb_commit
    assert PP.verify(_plonk377_struct_to_proof(PP, fresh), pk.vk, sol[:1], pk.tau)
    assert pkg.plonk_proof_bytes(fresh) != raw
