"""Pins the oracle: in-tree constants of the reference, the committed golden vectors (computed
from the big-int definitions, tests/golden/make_golden.py), and C++ restatement == Python
restatement at sizes Python can reach.  No GPU."""
import numpy as np
import pytest

from oracle import bn254 as B
from oracle import groth16 as G
from oracle import ntt as N
from tests import util
from tests.conftest import rand_field


def test_constants_pinned_in_reference_tree():
    # backend/groth16/bn254/solidity.go:41-42
    assert B.P == 0x30644e72e131a029b85045b68181585d97816a916871ca8d3c208c16d87cfd47
    assert B.R == 0x30644e72e131a029b85045b68181585d2833e84879b9709143e1f593f0000001
    # backend/plonk/bn254/solidity.go:25-27,56
    assert B.R == 21888242871839275222246405745257275088548364400416034343698204186575808495617
    assert B.P == 21888242871839275222246405745257275088696311157297823662689037894645226208583
    assert B.FR_MULT_GEN == 5
    # backend/groth16/bn254/solidity.go:52-54,57-58 (fractions in Fp pin the G2 twist 3/(9+u))
    assert pow(2, -1, B.P) == 0x183227397098d014dc2822db40c0ac2ecbc0b548b438e5469e10460b6c3e7ea4
    assert 27 * pow(82, -1, B.P) % B.P == 0x2b149d40ceb8aaae81be18991be06ac3b5b4c5e559dbefa33267e6dc24a138e5
    assert 3 * pow(82, -1, B.P) % B.P == 0x2fcd3ac2a640a154eb23960892a85a68f031ca0c8344b23a577dcf1052b9e775
    assert B.G2_B == (27 * pow(82, -1, B.P) % B.P, (-3 * pow(82, -1, B.P)) % B.P)
    assert (B.P + 1) // 4 == 0xC19139CB84C680A6E14116DA060561765E05AA45A1C72A34F082305B61F3F52
    # BN parameter t (solidity.go:36-40): p, r polynomials in t
    t = 4965661367192848881
    assert B.P == 36 * t**4 + 36 * t**3 + 24 * t**2 + 6 * t + 1
    assert B.R == 36 * t**4 + 36 * t**3 + 18 * t**2 + 6 * t + 1


def test_recalled_constants_self_check():
    assert pow(B.FR_ROOT_2_28, 1 << 28, B.R) == 1 and pow(B.FR_ROOT_2_28, 1 << 27, B.R) == B.R - 1
    assert (B.R - 1) % (1 << 28) == 0 and (B.R - 1) % (1 << 29) != 0
    assert B.g1_on_curve(B.G1_GEN) and B.g2_on_curve(B.G2_GEN)
    assert B.g1_mul(B.G1_GEN, B.R - 1) == B.g1_neg(B.G1_GEN)
    assert B._aff_mul(B._Fp2, B.G2_GEN, B.R) is None  # G2 generator has order r


def test_python_ntt_vs_naive_dft():
    d = N.Domain(16)
    x = [pow(7, i + 1, B.R) for i in range(16)]
    y = list(x)
    N.fft(y, d, N.DIF)
    N.bit_reverse(y)
    assert y == N.naive_dft(x, d.generator)
    N.bit_reverse(y)
    N.fft_inverse(y, d, N.DIT)
    assert y == x


def test_cref_ntt_golden(cref):
    g = util.golden()["ntt"]
    x, gen, cg = util.fr_arr(g["input"]), cref.fr(g["generator"]), cref.fr(g["coset_gen"])
    for case in g["cases"]:
        got = cref.ntt(x, g["log_n"], case["inverse"], case["decimation"], case["on_coset"], gen, cg)
        assert util.fr_ints(got) == case["output"], case


def test_python_ntt_golden():
    g = util.golden()["ntt"]
    d = N.Domain(1 << g["log_n"], g["generator"], g["coset_gen"])
    for case in g["cases"]:
        v = list(g["input"])
        (N.fft_inverse if case["inverse"] else N.fft)(v, d, case["decimation"], bool(case["on_coset"]))
        assert v == case["output"], case


def test_compute_h_golden(cref):
    g = util.golden()["compute_h"]
    gen, cg = cref.fr(g["generator"]), cref.fr(g["coset_gen"])
    got = cref.compute_h(util.fr_arr(g["a"]), util.fr_arr(g["b"]), util.fr_arr(g["c"]), g["log_n"], gen, cg)
    assert util.fr_ints(got) == g["h_bitreversed"]
    d = N.Domain(1 << g["log_n"], g["generator"], g["coset_gen"])
    assert N.compute_h(g["a"], g["b"], g["c"], d) == g["h_bitreversed"]


def test_msm_golden(cref):
    g = util.golden()["msm"]
    s = util.fr_arr(g["scalars"])
    assert util.g1_aff_ints(cref.g1_jac_to_aff(cref.msm_g1(s, util.g1_arr(g["g1_bases"])))) == g["g1_result"]
    assert util.g2_aff_ints(cref.g2_jac_to_aff(cref.msm_g2(s, util.g2_arr(g["g2_bases"])))) == g["g2_result"]


@pytest.mark.parametrize("log_n", [6, 10])
def test_cref_vs_python_ntt(cref, log_n):
    rng = np.random.default_rng(log_n)
    n = 1 << log_n
    x = rand_field(rng, n)
    xi = util.fr_ints(x)
    d = N.Domain(n)
    gen, cg = cref.fr(d.generator), cref.fr(d.coset_gen)
    for inverse in (0, 1):
        for dec in (0, 1):
            for coset in (0, 1):
                v = list(xi)
                (N.fft_inverse if inverse else N.fft)(v, d, dec, bool(coset))
                assert util.fr_ints(cref.ntt(x, log_n, inverse, dec, coset, gen, cg)) == v


def test_cref_msm_vs_python_definition(cref):
    rng = np.random.default_rng(5)
    n = 40
    bases, ks = util.known_dlog_bases_g1(n)
    scalars = rand_field(rng, n)
    si = util.fr_ints(scalars)
    want = B.g1_mul(B.G1_GEN, sum(k * s for k, s in zip(ks, si)) % B.R)
    assert util.g1_aff_ints(cref.g1_jac_to_aff(cref.msm_g1(scalars, bases))) == [want[0], want[1]]
    bases2, ks2 = util.known_dlog_bases_g2(n)
    want2 = B.g2_mul(B.G2_GEN, sum(k * s for k, s in zip(ks2, si)) % B.R)
    assert util.g2_aff_ints(cref.g2_jac_to_aff(cref.msm_g2(scalars, bases2))) == [want2[0][0], want2[0][1], want2[1][0], want2[1][1]]


def test_groth16_golden_trapdoor_and_bytes():
    g = util.golden()["groth16"]
    r1cs = G.mul_chain_circuit(g["nb_mul"])
    w, a, b, c = G.mul_chain_witness(g["nb_mul"], g["x"])
    assert G.solve(r1cs, w) == (a, b, c)
    td = G.Trapdoor(*g["trapdoor"])
    pk, vk = G.setup(r1cs, td, points=False)
    dl = G.prove_dlog(pk, td, w, a, b, c, g["r"], g["s"], r1cs.nb_public)
    assert G.trapdoor_check(vk, w[: r1cs.nb_public], dl["ar"], dl["bs"], dl["krs"])
    assert not G.trapdoor_check(vk, w[: r1cs.nb_public], dl["ar"], dl["bs"], (dl["krs"] + 1) % B.R)
    ar, krs = B.g1_mul(B.G1_GEN, dl["ar"]), B.g1_mul(B.G1_GEN, dl["krs"])
    bs = B.g2_mul(B.G2_GEN, dl["bs"])
    assert [ar[0], ar[1]] == g["ar"] and [krs[0], krs[1]] == g["krs"]
    assert G.proof_bytes(ar, bs, krs).hex() == g["proof_hex"]


def test_cref_bls12_377_build_vs_python_oracle(cref):
    """liboracle377.so (cpu_ref.cpp with the BLS12-377 constants, 6-limb Fp, u^2 = -5) against the big-int
    restatement: field products, fixed-base multiples, G1/G2 MultiExp, all eight transforms, computeH."""
    import random

    from oracle import bls12_377 as B3
    from oracle.suite377 import ntt as N3
    from tests import util377 as U
    o = cref.BLS12_377
    rnd = random.Random(1)
    xs = [0, 1, B3.P - 1] + [rnd.randrange(B3.P) for _ in range(40)]
    ys = [B3.P - 1, B3.P - 2, B3.P - 1] + [rnd.randrange(B3.P) for _ in range(40)]
    assert U.fp_ints(o.fp_mul(U.fp_arr(xs), U.fp_arr(ys))) == [x * y % B3.P for x, y in zip(xs, ys)]
    ks = [0, 1, B3.R - 1] + [rnd.randrange(B3.R) for _ in range(13)]
    g1, g2 = U.g1_arr([B3.G1_GEN])[0], U.g2_arr([B3.G2_GEN])[0]
    p1, p2 = o.g1_batch_mul(g1, U.fr_arr(ks)), o.g2_batch_mul(g2, U.fr_arr(ks))
    assert [U.g1_aff(p) for p in p1] == [B3.g1_mul(B3.G1_GEN, k) for k in ks]
    assert [U.g2_aff(p) for p in p2] == [B3.g2_mul(B3.G2_GEN, k) for k in ks]
    sc = [rnd.randrange(B3.R) for _ in ks]
    tot = sum(a * b for a, b in zip(sc, ks)) % B3.R
    assert U.g1_jac(o.msm_g1(U.fr_arr(sc), p1)) == B3.g1_mul(B3.G1_GEN, tot)
    assert U.g2_jac(o.msm_g2(U.fr_arr(sc), p2)) == B3.g2_mul(B3.G2_GEN, tot)
    d = N3.Domain(64)
    v = [rnd.randrange(B3.R) for _ in range(64)]
    gen, cg = U.fr(d.generator), U.fr(d.coset_gen)
    for inv in (0, 1):
        for dec in (0, 1):
            for cos in (0, 1):
                want = list(v)
                (N3.fft_inverse if inv else N3.fft)(want, d, dec, bool(cos))
                assert U.fr_ints(o.ntt(U.fr_arr(v), 6, inv, dec, cos, gen, cg)) == want, (inv, dec, cos)
    a, b, c = ([rnd.randrange(B3.R) for _ in range(61)] for _ in range(3))
    pad = lambda t: t + [0] * 3
    assert U.fr_ints(o.compute_h(U.fr_arr(a), U.fr_arr(b), U.fr_arr(c), 6, gen, cg)) == N3.compute_h(pad(a), pad(b), pad(c), d)


def commitment_chain_circuit(nb_mul, committed=(2, 5), two_commitments=False):
    """refCircuit plus api.Commit (commitment_test.go shapes): a commitment to some private wires, its value a new
    internal wire Cw constrained by AssertIsDifferent(Cw, 0), i.e. Cw * Iw = ONE with the inverse Iw as another wire."""
    r1cs = G.mul_chain_circuit(nb_mul)
    base = r1cs.nb_wires
    cons = list(r1cs.constraints)
    groups = [list(committed)] if not two_commitments else [list(committed), [3, 4, 7]]
    commitments = []
    for j, ws in enumerate(groups):
        cw, iw = base + 2 * j, base + 2 * j + 1
        cons.append(([(1, cw)], [(1, iw)], [(1, 0)]))
        commitments.append((cw, sorted(ws)))
    return G.R1CS(nb_public=2, nb_secret=1, nb_internal=nb_mul + 2 * len(groups), constraints=cons, commitments=commitments)


def commitment_chain_witness(r1cs, nb_mul, x=2):
    w, _, _, _ = G.mul_chain_witness(nb_mul, x)
    for j, _ in enumerate(r1cs.commitments):
        h = pow(7 + j, 12345, B.R)              # stands for the hash-to-field value the solver hint injects (prove.go:98-109)
        w += [h, pow(h, -1, B.R)]
    a, b, c = G.solve(r1cs, w)
    assert all(x * y % B.R == z for x, y, z in zip(a, b, c))
    return w, a, b, c


@pytest.mark.parametrize("two", [False, True])
def test_groth16_commitment_circuit_trapdoor_identity(two):
    """Setup / Prove with api.Commit wires (setup.go:155-200, prove.go:238-242, verify.go:75-119): the K MultiExp skips
    the committed and commitment wires and the verify equation closes with the Pedersen commitments added to kSum."""
    import random
    nb_mul = 13
    r1cs = commitment_chain_circuit(nb_mul, two_commitments=two)
    rnd = random.Random(3)
    td = G.Trapdoor(*[rnd.randrange(1, B.R) for _ in range(5)])
    pk, vk = G.setup(r1cs, td, points=False)
    w, a, b, c = commitment_chain_witness(r1cs, nb_mul)
    remove = sorted(set(vk["commit_wires"]) | {i for ws in vk["committed"] for i in ws})
    r, s = rnd.randrange(B.R), rnd.randrange(B.R)
    dl = G.prove_dlog(pk, td, w, a, b, c, r, s, r1cs.nb_public, to_remove=remove)
    assert len(dl["scalars"]["K"]) == len(pk.sc["K"]) == r1cs.nb_wires - 2 - len(remove)
    com = G.commitment_dlogs(pk, vk, w)
    assert G.trapdoor_check(vk, w[:2], dl["ar"], dl["bs"], dl["krs"], [w[i] for i in vk["commit_wires"]], com)
    assert not G.trapdoor_check(vk, w[:2], dl["ar"], dl["bs"], dl["krs"])       # without the commitments it must fail


def test_plonk_oracle_on_bls12_377_proves_and_verifies():
    """oracle/plonk_prove.py executed on BLS12-377 (oracle/suite377.py): a proof of the SCS Mul chain is accepted by the
    verifier restatement, a wrong public input is not, and Proof.WriteTo has the 48-byte-coordinate size."""
    import random

    from oracle.suite377 import curve as B3
    from oracle.suite377 import plonk_prove as PP3
    rng = random.Random(1)
    spr = PP3.mul_chain_scs(6)
    pk = PP3.setup(spr, rng.randrange(1, B3.R))
    sol = PP3.mul_chain_solution(6, 5)
    proof, _ = PP3.prove(spr, pk, sol, ([1, 2], [3, 4], [5, 6], [7, 8, 9]))
    assert PP3.verify(proof, pk.vk, sol[:1], pk.tau)
    assert not PP3.verify(proof, pk.vk, [(sol[0] + 1) % B3.R], pk.tau)
    assert len(PP3.proof_bytes(proof)) == 9 * 48 + 8 + 32 * 8
