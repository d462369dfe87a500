"""b200_groth16_setup (Setup on the device, setup.go:85-337), api.Commit circuits through the device prover
(prove.go:238-242: `k_remove`), and BASELINE.json config 1: a real 2^20-constraint refCircuit proof
(groth16_test.go:120-156) that verifies -- all against oracle/groth16.py and oracle/cpu_ref.cpp."""
import random

import numpy as np
import pytest

from oracle import bn254 as B
from oracle import groth16 as G
from tests import util
from tests.test_oracle import commitment_chain_circuit, commitment_chain_witness

pytestmark = pytest.mark.gpu


def r1cs_to_csr(r1cs):
    """oracle R1CS -> the ABI's CSR matrices over (coefficient id, wire id) terms + the coefficient table."""
    coeff_id, coeffs = {}, []
    mats = []
    for side in range(3):
        row_ptr, terms = [0], []
        for cons in r1cs.constraints:
            for k, wire in cons[side]:
                k %= B.R
                if k not in coeff_id:
                    coeff_id[k] = len(coeffs)
                    coeffs.append(k)
                terms.append((coeff_id[k], wire))
            row_ptr.append(len(terms))
        mats.append((np.array(row_ptr, dtype=np.uint64), np.array(terms, dtype=np.uint32).reshape(-1, 2)))
    return mats, util.fr_arr(coeffs)


def wire_classes(r1cs):
    wc = np.zeros(r1cs.nb_wires, dtype=np.uint8)
    for j, (cw, ws) in enumerate(r1cs.commitments):
        wc[cw] = 1
        wc[list(ws)] = 2 + j
    return wc


def device_setup(ctx, cref, r1cs, td, **kw):
    d = G.ntt.Domain(r1cs.nb_constraints)
    mats, coeffs = r1cs_to_csr(r1cs)
    toxic = dict(tau=cref.fr(td.t), alpha=cref.fr(td.alpha), beta=cref.fr(td.beta), gamma=cref.fr(td.gamma), delta=cref.fr(td.delta))
    g1 = cref.g1_points_to_array([B.G1_GEN])[0]
    g2 = cref.g2_points_to_array([B.G2_GEN])[0]
    wc = wire_classes(r1cs) if r1cs.commitments else None
    return ctx.groth16_setup(log_n=d.log_n, generator=cref.fr(d.generator), coset_gen=cref.fr(d.coset_gen),
                             nb_wires=r1cs.nb_wires, nb_public=r1cs.nb_public, matrices=mats, coeffs=coeffs, toxic=toxic,
                             g1_gen=g1, g2_gen=g2, wire_class=wc, nb_commitments=len(r1cs.commitments), **kw)


def random_r1cs(rnd, nb_constraints, nb_wires, nb_public):
    """Linear combinations with several terms, repeated wires, coefficients 1, -1, 2 and random ones, a wire that
    appears in no L (infinity in A) and one that appears in no R."""
    def lc():
        return [(rnd.choice([1, B.R - 1, 2, rnd.randrange(B.R)]), rnd.randrange(2, nb_wires)) for _ in range(rnd.randrange(1, 5))]
    cons = [(lc(), lc(), lc()) for _ in range(nb_constraints)]
    cons[0] = ([(1, 0)], [(5, 1)], [(1, 1), (B.R - 1, 1)])      # wires 0 / 1: only L / only R (and an O that cancels to zero)
    return G.R1CS(nb_public=nb_public, nb_secret=1, nb_internal=nb_wires - nb_public - 1, constraints=cons)


@pytest.mark.parametrize("kind", ["chain15", "chain1000", "random300", "commit"])
def test_setup_matches_oracle_setup(ctx, cref, kind):
    rnd = random.Random(len(kind))
    if kind.startswith("chain"):
        r1cs = G.mul_chain_circuit(int(kind[5:]))
    elif kind == "random300":
        r1cs = random_r1cs(rnd, 300, 190, 3)
    else:
        r1cs = commitment_chain_circuit(29, two_commitments=True)
    td = G.Trapdoor(*[rnd.randrange(1, B.R) for _ in range(5)])
    pk, vk = G.setup(r1cs, td, points=False)
    pkh, res = device_setup(ctx, cref, r1cs, td, want_points=True)
    g1 = cref.g1_points_to_array([B.G1_GEN])[0]
    g2 = cref.g2_points_to_array([B.G2_GEN])[0]
    mul1 = lambda xs: cref.g1_batch_mul(g1, util.fr_arr(xs)) if len(xs) else np.zeros((0, 8), dtype=np.uint64)
    mul2 = lambda xs: cref.g2_batch_mul(g2, util.fr_arr(xs)) if len(xs) else np.zeros((0, 16), dtype=np.uint64)
    assert [bool(x) for x in res["infinity_a"]] == pk.infinity_A and [bool(x) for x in res["infinity_b"]] == pk.infinity_B
    assert (res["n_a"], res["n_b"], res["n_k"], res["n_z"]) == (len(pk.sc["A"]), len(pk.sc["B"]), len(pk.sc["K"]), len(pk.sc["Z"]))
    assert np.array_equal(res["g1_a"], mul1(pk.sc["A"])) and np.array_equal(res["g1_b"], mul1(pk.sc["B"]))
    assert np.array_equal(res["g1_k"], mul1(pk.sc["K"])) and np.array_equal(res["g1_z"], mul1(pk.sc["Z"]))
    assert np.array_equal(res["g2_b"], mul2(pk.sc["B"]))
    assert np.array_equal(res["vk_k"], mul1(vk["vkK"]))
    for j, basis in enumerate(pk.sc["ckK"]):
        assert np.array_equal(res["ck_k"][j], mul1(basis)), j
    assert np.array_equal(np.stack([res["g1_alpha"], res["g1_beta"], res["g1_delta"]]), mul1([td.alpha, td.beta, td.delta]))
    assert np.array_equal(np.stack([res["g2_beta"], res["g2_delta"], res["g2_gamma"]]), mul2([td.beta, td.delta, td.gamma]))
    pkh.free()


def test_setup_rejects_bad_input(ctx, pkg, cref):
    r1cs = G.mul_chain_circuit(15)
    td = G.Trapdoor(*[5, 6, 7, 8, 9])
    d = G.ntt.Domain(16)
    with pytest.raises(pkg.B200Error, match="tau lies in the evaluation domain"):
        device_setup(ctx, cref, r1cs, G.Trapdoor(d.generator, 6, 7, 8, 9))
    bad = G.mul_chain_circuit(15)
    bad.constraints[3] = ([(1, 999)], [(1, 2)], [(1, 3)])
    with pytest.raises(pkg.B200Error, match="out of range"):
        device_setup(ctx, cref, bad, td)


def _finalize(pkg, cref, out, r, s, res):
    return pkg.groth16_finalize(out, cref.fr(r), cref.fr(s), res["g1_alpha"], res["g1_beta"], res["g1_delta"], res["g2_beta"], res["g2_delta"])


@pytest.mark.parametrize("two", [False, True])
def test_commitment_circuit_k_remove(ctx, pkg, cref, two):
    """api.Commit circuits (backend/groth16/bn254/commitment_test.go; prove.go:83-110,238-242): the K MultiExp runs on
    W[nb_public:] minus the private committed wires and the commitment wires (`k_remove`); the proof then satisfies
    the verify equation with the Pedersen commitments in kSum (verify.go:117-119) and equals the oracle's byte for byte."""
    nb_mul = 200
    r1cs = commitment_chain_circuit(nb_mul, committed=(2, 5, 150), two_commitments=two)
    rnd = random.Random(17)
    td = G.Trapdoor(*[rnd.randrange(1, B.R) for _ in range(5)])
    pk, vk = G.setup(r1cs, td, points=False)
    pkh, res = device_setup(ctx, cref, r1cs, td)
    w, a, b, c = commitment_chain_witness(r1cs, nb_mul)
    remove = sorted(set(vk["commit_wires"]) | {i for ws in vk["committed"] for i in ws})
    k_remove = np.array([i - r1cs.nb_public for i in remove], dtype=np.uint32)
    r, s = rnd.randrange(B.R), rnd.randrange(B.R)
    fa, fb, fc, fw = (util.fr_arr(x) for x in (a, b, c, w))
    out, _ = ctx.groth16_prove_core(pkh, fa, fb, fc, fw, k_remove=k_remove)
    ar, bs, krs = _finalize(pkg, cref, out, r, s, res)
    dl = G.prove_dlog(pk, td, w, a, b, c, r, s, r1cs.nb_public, to_remove=remove)
    assert G.trapdoor_check(vk, w[:2], dl["ar"], dl["bs"], dl["krs"], [w[i] for i in vk["commit_wires"]], G.commitment_dlogs(pk, vk, w))
    ear, ekrs, ebs = B.g1_mul(B.G1_GEN, dl["ar"]), B.g1_mul(B.G1_GEN, dl["krs"]), B.g2_mul(B.G2_GEN, dl["bs"])
    assert util.g1_aff_ints(ar) == [ear[0], ear[1]] and util.g1_aff_ints(krs) == [ekrs[0], ekrs[1]]
    assert pkg.groth16_proof_bytes(ar, bs, krs) == G.proof_bytes(ear, ebs, ekrs)
    # the MSM outputs one by one, and the serial schedule
    want = dict(ar=dl["msm"]["A"], bs1=dl["msm"]["B"], krs=dl["msm"]["K"], krs2=dl["msm"]["Z"])
    od = out.as_dict()
    for name, e in want.items():
        p = B.g1_mul(B.G1_GEN, e)
        assert util.g1_aff_ints(pkg.g1_jac_to_affine(od[name])) == (None if p is None else [p[0], p[1]]), name
    ctx.set_option("msm_overlap", 0)
    try:
        out2, _ = ctx.groth16_prove_core(pkh, fa, fb, fc, fw, k_remove=k_remove)
    finally:
        ctx.set_option("msm_overlap", 1)
    assert np.array_equal(pkg.g1_jac_to_affine(out2.as_dict()["krs"]), pkg.g1_jac_to_affine(od["krs"]))
    # a k_remove that does not match the key is refused
    with pytest.raises(pkg.B200Error):
        ctx.groth16_prove_core(pkh, fa, fb, fc, fw, k_remove=k_remove[:-1])
    with pytest.raises(pkg.B200Error):
        ctx.groth16_prove_core(pkh, fa, fb, fc, fw, k_remove=k_remove[::-1].copy())
    pkh.free()


def test_config1_refcircuit_2p20_proof_verifies(ctx, pkg, cref):
    """BASELINE.json config 1: Groth16 Prove of the 2^20-constraint api.Mul-chain circuit with its SATISFYING witness,
    key from b200_groth16_setup.  Checked three ways: (1) every MSM result and the proof bytes equal the oracle's
    (oracle Setup scalars, h from the C++ computeH, fixed r, s); (2) the oracle's verify equation holds for those
    discrete logs; (3) the product-side checker (groth16_host, no h involved: A(tau) B(tau) - C(tau) = H(tau) Z(tau))
    agrees with the oracle on all five discrete logs."""
    H = pkg.groth16_host
    log_n = 20
    nb_mul = (1 << log_n) - 1
    fld = H.Field(pkg.BN254)
    rnd = random.Random(2020)
    td = G.Trapdoor(*[rnd.randrange(1, B.R) for _ in range(5)])
    toxic = dict(tau=td.t, alpha=td.alpha, beta=td.beta, gamma=td.gamma, delta=td.delta)
    g1 = cref.g1_points_to_array([B.G1_GEN])[0]
    g2 = cref.g2_points_to_array([B.G2_GEN])[0]
    cs, pkh, res = H.setup_mul_chain(ctx, fld, log_n, toxic, g1, g2)
    w, a, b, c = H.mul_chain_witness(ctx, fld, log_n, 2)
    # the witness generator against the oracle's
    ow, oa, ob, oc = G.mul_chain_witness(nb_mul, 2)
    assert util.fr_ints(w[:5]) + util.fr_ints(w[-3:]) == ow[:5] + ow[-3:] and util.fr_ints(c[-2:]) == oc[-2:]
    r, s = rnd.randrange(B.R), rnd.randrange(B.R)
    out, tm = ctx.groth16_prove_core(pkh, a, b, c, w)
    ar, bs, krs = _finalize(pkg, cref, out, r, s, res)
    # (1) oracle: Setup scalars (big-int), h from the threaded C++ computeH on the same a, b, c
    r1cs = G.mul_chain_circuit(nb_mul)
    opk, ovk = G.setup(r1cs, td, points=False)
    assert [bool(x) for x in res["infinity_a"]] == opk.infinity_A and [bool(x) for x in res["infinity_b"]] == opk.infinity_B
    d = opk.domain
    h = util.fr_ints(cref.compute_h(a, b, c, log_n, cref.fr(d.generator), cref.fr(d.coset_gen)))
    dl = G.prove_dlog(opk, td, ow, oa, ob, oc, r, s, 2, h=h)
    assert G.trapdoor_check(ovk, ow[:2], dl["ar"], dl["bs"], dl["krs"])                     # (2)
    ear, ekrs, ebs = B.g1_mul(B.G1_GEN, dl["ar"]), B.g1_mul(B.G1_GEN, dl["krs"]), B.g2_mul(B.G2_GEN, dl["bs"])
    assert util.g1_aff_ints(ar) == [ear[0], ear[1]] and util.g1_aff_ints(krs) == [ekrs[0], ekrs[1]]
    assert util.g2_aff_ints(bs) == [ebs[0][0], ebs[0][1], ebs[1][0], ebs[1][1]]
    assert pkg.groth16_proof_bytes(ar, bs, krs) == G.proof_bytes(ear, ebs, ekrs)
    # (3) the product-side checker bench.py uses at 2^24
    exp = H.mul_chain_expected_dlogs(ctx, fld, log_n, toxic, w, a, b, c)
    assert (exp["ar"], exp["bs1"], exp["krs"], exp["krs2"]) == (dl["msm"]["A"], dl["msm"]["B"], dl["msm"]["K"], dl["msm"]["Z"])
    assert all(H.points_match(ctx, pkg, fld, g1, g2, out.as_dict(), exp).values())
    proof = H.expected_proof_dlogs(fld, toxic, exp, r, s)
    assert (proof["ar"], proof["bs"], proof["krs"]) == (dl["ar"], dl["bs"], dl["krs"]) and H.verify_equation(fld, toxic, exp, proof)
    # a witness that does not satisfy the circuit must not pass the product-side checker
    bad = c.copy()
    bad[12345] = fld.limbs(7)
    exp_bad = H.mul_chain_expected_dlogs(ctx, fld, log_n, toxic, w, a, b, bad)
    out_bad, _ = ctx.groth16_prove_core(pkh, a, b, bad, w)
    assert not all(H.points_match(ctx, pkg, fld, g1, g2, out_bad.as_dict(), exp_bad).values())
    pkh.free()
