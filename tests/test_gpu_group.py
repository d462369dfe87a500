"""The device group (b200_group_*): one process, one call per proof over several ranks.  On the one GPU of the test
box the ranks are virtual (the same device listed several times): the very code that runs on 2/4/8 B200s -- host
threads per rank, point-range shards, computeH distributed with its exchanges fused into the kernels over peer
pointers, partial sums combined in C -- against the single-context prover and the oracle."""
import random

import numpy as np
import pytest

from oracle import bn254 as B
from oracle import groth16 as G
from tests import util
from tests.test_gpu_groth16_setup import device_setup, r1cs_to_csr, wire_classes
from tests.test_oracle import commitment_chain_circuit, commitment_chain_witness

pytestmark = pytest.mark.gpu


def _same(pkg, got, want):
    for k in ("ar", "bs1", "krs", "krs2"):
        assert np.array_equal(pkg.g1_jac_to_affine(got[k]), pkg.g1_jac_to_affine(want[k])), k
    assert np.array_equal(pkg.g2_jac_to_affine(got["bs2"]), pkg.g2_jac_to_affine(want["bs2"]))


def _group_setup(grp, cref, r1cs, td, **kw):
    d = G.ntt.Domain(r1cs.nb_constraints)
    mats, coeffs = r1cs_to_csr(r1cs)
    toxic = dict(tau=cref.fr(td.t), alpha=cref.fr(td.alpha), beta=cref.fr(td.beta), gamma=cref.fr(td.gamma), delta=cref.fr(td.delta))
    wc = wire_classes(r1cs) if r1cs.commitments else None
    return grp.groth16_setup(log_n=d.log_n, generator=cref.fr(d.generator), coset_gen=cref.fr(d.coset_gen),
                             nb_wires=r1cs.nb_wires, nb_public=r1cs.nb_public, matrices=mats, coeffs=coeffs, toxic=toxic,
                             g1_gen=cref.g1_points_to_array([B.G1_GEN])[0], g2_gen=cref.g2_points_to_array([B.G2_GEN])[0],
                             wire_class=wc, nb_commitments=len(r1cs.commitments), **kw)


@pytest.mark.parametrize("ranks,log_n", [(2, 6), (2, 12), (4, 10), (8, 6), (8, 13), (3, 9), (1, 8), (4, 16)])
def test_group_prove_matches_single_context_and_oracle(ctx, pkg, cref, ranks, log_n):
    nb_mul = (1 << log_n) - 1
    r1cs = G.mul_chain_circuit(nb_mul)
    rnd = random.Random(ranks * 100 + log_n)
    td = G.Trapdoor(*[rnd.randrange(1, B.R) for _ in range(5)])
    w, a, b, c = G.mul_chain_witness(nb_mul, 3)
    fa, fb, fc, fw = (util.fr_arr(x) for x in (a, b, c, w))
    pk1, res1 = device_setup(ctx, cref, r1cs, td, want_points=True)
    want = ctx.groth16_prove_core(pk1, fa, fb, fc, fw)[0].as_dict()
    pk1.free()
    with pkg.Group([0] * ranks) as grp:
        gpk, res = _group_setup(grp, cref, r1cs, td, want_points=True)
        info = gpk.info(ranks - 1)
        assert info["distributed_h"] == (ranks > 1 and ranks & (ranks - 1) == 0 and 2 * (ranks.bit_length() - 1) <= log_n)
        for k in ("g1_a", "g1_b", "g1_k", "g1_z", "g2_b", "vk_k", "g1_alpha", "g2_gamma", "infinity_a", "infinity_b"):
            assert np.array_equal(res[k], res1[k]), k      # the ranks filled their ranges of the point arrays
        out, tm = grp.groth16_prove_core(gpk, fa, fb, fc, fw)
        _same(pkg, out.as_dict(), want)
        out2, _ = grp.groth16_prove_core(gpk, fa[:-3], fb[:-3], fc[:-3], fw)     # a second proof on the same buffers, ragged n_constraints
        want2 = None
        assert tm["total_ms"] > 0
        # the same key uploaded from host arrays (b200_group_groth16_pk_upload), inputs resident on the devices
        gpk2 = grp.groth16_pk_upload(log_n=log_n, generator=cref.fr(G.ntt.Domain(1 << log_n).generator), coset_gen=cref.fr(5),
                                     g1_a=res["g1_a"], g1_b=res["g1_b"], g1_k=res["g1_k"], g1_z=res["g1_z"], g2_b=res["g2_b"],
                                     infinity_a=res["infinity_a"], infinity_b=res["infinity_b"], nb_public=2)
        if info["distributed_h"]:
            n_loc = (1 << log_n) // ranks
            c0 = grp.ctx(0)
            blocks = [[pkg.DeviceVector(c0, n_loc * 32).upload(np.ascontiguousarray(v[r * n_loc:(r + 1) * n_loc])) for r in range(ranks)]
                      for v in (fa, fb, fc)]
            wv = [pkg.DeviceVector(c0, fw.nbytes).upload(fw) for _ in range(ranks)]
            out3, _ = grp.groth16_prove_core_dev(gpk2, blocks[0], blocks[1], blocks[2], wv)
            _same(pkg, out3.as_dict(), want)
            [v.free() for vs in blocks for v in vs]
            [v.free() for v in wv]
        else:
            out3, _ = grp.groth16_prove_core(gpk2, fa, fb, fc, fw)
            _same(pkg, out3.as_dict(), want)
        gpk2.free()
        gpk.free()
    # the ragged proof against the single context on the same truncated inputs
    pk1, _ = device_setup(ctx, cref, r1cs, td)
    want2 = ctx.groth16_prove_core(pk1, fa[:-3], fb[:-3], fc[:-3], fw)[0].as_dict()
    pk1.free()
    _same(pkg, out2.as_dict(), want2)
    # and the oracle: proof bytes under fixed r, s
    pk, vk = G.setup(r1cs, td, points=False)
    r, s = rnd.randrange(B.R), rnd.randrange(B.R)
    dl = G.prove_dlog(pk, td, w, a, b, c, r, s, 2)
    ar, bs, krs = pkg.groth16_finalize(out, cref.fr(r), cref.fr(s), res["g1_alpha"], res["g1_beta"], res["g1_delta"], res["g2_beta"], res["g2_delta"])
    assert pkg.groth16_proof_bytes(ar, bs, krs) == G.proof_bytes(B.g1_mul(B.G1_GEN, dl["ar"]), B.g2_mul(B.G2_GEN, dl["bs"]), B.g1_mul(B.G1_GEN, dl["krs"]))


def test_group_commitment_circuit_and_errors(ctx, pkg, cref):
    nb_mul = 250
    r1cs = commitment_chain_circuit(nb_mul, committed=(2, 9, 100), two_commitments=True)
    rnd = random.Random(5)
    td = G.Trapdoor(*[rnd.randrange(1, B.R) for _ in range(5)])
    pk, vk = G.setup(r1cs, td, points=False)
    w, a, b, c = commitment_chain_witness(r1cs, nb_mul)
    remove = sorted(set(vk["commit_wires"]) | {i for ws in vk["committed"] for i in ws})
    k_remove = np.array([i - 2 for i in remove], dtype=np.uint32)
    fa, fb, fc, fw = (util.fr_arr(x) for x in (a, b, c, w))
    with pkg.Group([0, 0, 0, 0]) as grp:
        gpk, res = _group_setup(grp, cref, r1cs, td)
        out, _ = grp.groth16_prove_core(gpk, fa, fb, fc, fw, k_remove=k_remove)
        dl = G.prove_dlog(pk, td, w, a, b, c, 0, 0, 2, to_remove=remove)
        od = out.as_dict()
        for name, e in (("ar", dl["msm"]["A"]), ("bs1", dl["msm"]["B"]), ("krs", dl["msm"]["K"]), ("krs2", dl["msm"]["Z"])):
            p = B.g1_mul(B.G1_GEN, e)
            assert util.g1_aff_ints(pkg.g1_jac_to_affine(od[name])) == [p[0], p[1]], name
        with pytest.raises(pkg.B200Error, match="rank"):       # every rank refuses; the group reports one, nobody hangs
            grp.groth16_prove_core(gpk, fa, fb, fc, fw, k_remove=k_remove[:-1])
        out2, _ = grp.groth16_prove_core(gpk, fa, fb, fc, fw, k_remove=k_remove)     # and the group still works
        _same(pkg, out2.as_dict(), od)
        gpk.free()
    with pytest.raises(pkg.B200Error):
        pkg.Group([0, 99])


@pytest.mark.parametrize("ranks", [2, 4, 8, 3])
@pytest.mark.parametrize("nb_mul,nb_commit", [(5, 0), (100, 0), (40, 2)])
def test_group_plonk_prove_matches_oracle(pkg, ranks, nb_mul, nb_commit):
    """b200_group_plonk_prove on virtual ranks: commitments sharded by SRS point range and summed in C, quotient by
    coset with peer copies of the numerator blocks -- proof bytes identical to the oracle's under fixed blinding (rho = 8
    at the smallest size, 4 otherwise; 3 ranks: an uneven split), fresh blinding drawn once for all ranks verifies."""
    from oracle import plonk_prove as PP
    from tests import plonk_util as PU
    rng = random.Random(8000 + nb_mul)
    if nb_commit:
        spr = PP.commit_chain_scs(nb_mul, nb_commit)
        pk = PP.setup(spr, rng.randrange(1, B.R))
        sol, bsb = PP.commit_chain_solution(spr, pk, rng.randrange(2, B.R), [(rng.randrange(B.R), rng.randrange(B.R)) for _ in range(nb_commit)])
    else:
        spr, bsb = PP.mul_chain_scs(nb_mul), None
        pk = PP.setup(spr, rng.randrange(1, B.R))
        sol = PP.mul_chain_solution(nb_mul, rng.randrange(2, B.R))
    blinding = tuple([rng.randrange(B.R) for _ in range(k)] for k in (2, 2, 2, 3))
    want, _ = PP.prove(spr, pk, sol, blinding, bsb)
    L, Rr, O = (util.fr_arr(v) for v in PP.evaluate_lro(spr, sol))
    pub, bl, bsb_a = util.fr_arr(sol[:1]), PU.blinding_array(blinding), PU.bsb_arrays(bsb)
    with pkg.Group([0] * ranks) as grp:
        gpk = PU.upload_pk(grp, pk)
        try:
            got, tm = grp.plonk_prove(gpk, L, Rr, O, pub, bl, bsb_a)
            fresh, _ = grp.plonk_prove(gpk, L, Rr, O, pub, None, bsb_a)
        finally:
            gpk.free()
    assert pkg.plonk_proof_bytes(got) == PP.proof_bytes(want)
    assert PP.verify(PU.struct_to_proof(fresh), pk.vk, sol[:1], pk.tau) and pkg.plonk_proof_bytes(fresh) != pkg.plonk_proof_bytes(got)
    assert tm["total_ms"] > 0
