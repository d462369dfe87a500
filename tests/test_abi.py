"""The C-ABI library: loads, exports every symbol include/b200prover.h declares, refuses to run
without a device, and its host-side group / Groth16 fix-up / marshalling entry points agree with
the golden vectors.  No compute call touches a GPU here."""
import ctypes
import os
import re

import numpy as np
import pytest

from oracle import groth16 as G
from oracle import ntt as N
from tests import util
from tests.conftest import ROOT


def header_symbols():
    src = open(os.path.join(ROOT, "include", "b200prover.h")).read()
    src = src[src.index("/* END per-curve symbol names */"):]
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(pkg):
    lib = ctypes.CDLL(pkg.library_path())
    names = header_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in b200prover.h but not exported"
    pkg.load_library()  # binds every prototype the Python mirror uses


def test_bls12_377_library_exports_and_identity(pkg):
    """The second library (make CURVE=bls12_377) exports the same ABI and says which curve it was built for."""
    lib = ctypes.CDLL(pkg.library_path(pkg.BLS12_377))
    lib254 = ctypes.CDLL(pkg.library_path())
    for n in header_symbols():
        renamed = "b200_bls12377_" + n[5:]
        assert hasattr(lib, renamed), f"{n} declared in b200prover.h but not exported (as {renamed}) by the BLS12-377 build"
        # the two libraries share no exported name: both can be linked into one binary (gnark links every curve)
        assert not hasattr(lib, n) and not hasattr(lib254, renamed)
    lib377, lib254 = pkg.load_library(pkg.BLS12_377), pkg.load_library()
    assert (lib377.b200_build_curve(), lib377.b200_build_fp_words()) == (pkg.BLS12_377, 6)
    assert (lib254.b200_build_curve(), lib254.b200_build_fp_words()) == (pkg.BN254, 4)
    assert ctypes.sizeof(pkg.api._MSM_OUT[pkg.BLS12_377]) == 8 * (4 * 18 + 36)


def test_bls12_377_host_fixups_and_bytes_match_oracle(pkg):
    """b200_groth16_finalize + b200_groth16_proof_bytes of the BLS12-377 build (pure host code) on MSM results made
    by the oracle: prove.go:178-321 and marshal.go:41-66 with 48-byte coordinates and three flag bits."""
    from oracle import bls12_377 as B
    from oracle.suite377 import groth16 as G3
    from tests import util377 as U
    import random
    rnd = random.Random(5)
    c = pkg.BLS12_377
    k = [rnd.randrange(1, B.R) for _ in range(12)]
    g1 = lambda e: B.g1_mul(B.G1_GEN, e)
    g2 = lambda e: B.g2_mul(B.G2_GEN, e)
    jac1 = lambda pt: U.fp_arr([pt[0], pt[1], 1]).reshape(-1)
    jac2 = lambda pt: U.fp_arr([pt[0][0], pt[0][1], pt[1][0], pt[1][1], 1, 0]).reshape(-1)
    out = pkg.api._MSM_OUT[c]()
    for name, e in (("ar", k[0]), ("bs1", k[1]), ("krs", k[2]), ("krs2", k[3])):
        getattr(out, name)[:] = [int(v) for v in jac1(g1(e))]
    out.bs2[:] = [int(v) for v in jac2(g2(k[1]))]
    r, s, alpha, beta, delta = k[4:9]
    ar, bs, krs = pkg.groth16_finalize(out, U.fr(r), U.fr(s), U.g1_arr([g1(alpha)])[0], U.g1_arr([g1(beta)])[0],
                                       U.g1_arr([g1(delta)])[0], U.g2_arr([g2(beta)])[0], U.g2_arr([g2(delta)])[0], curve=c)
    e_ar = (k[0] + alpha + r * delta) % B.R
    e_bs = (k[1] + beta + s * delta) % B.R
    e_krs = (k[2] + k[3] - r * s * delta + s * e_ar + r * e_bs) % B.R
    assert U.g1_aff(ar) == g1(e_ar) and U.g2_aff(bs) == g2(e_bs) and U.g1_aff(krs) == g1(e_krs)
    raw = pkg.groth16_proof_bytes(ar, bs, krs, curve=c)
    assert len(raw) == 244 and raw == G3.proof_bytes(g1(e_ar), g2(e_bs), g1(e_krs))
    inf = np.zeros(12, dtype=np.uint64)
    raw = pkg.groth16_proof_bytes(inf, np.zeros(24, dtype=np.uint64), inf, curve=c)
    assert raw == G3.proof_bytes(None, None, None)


def test_no_cpu_fallback_without_device(pkg):
    lib = pkg.load_library()
    if lib.b200_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(pkg.B200Error):
        pkg.Context(0)


def test_host_group_ops(pkg, cref):
    rng = np.random.default_rng(2)
    bases, _ = util.known_dlog_bases_g1(4)
    j = [cref.msm_g1(util.fr_arr([int(rng.integers(1, 1 << 60))]), bases[i:i + 1]) for i in range(4)]
    s01 = pkg.g1_jac_add(j[0], j[1])
    want = cref.msm_g1(util.fr_arr([1, 1]), np.stack([cref.g1_jac_to_aff(j[0]), cref.g1_jac_to_aff(j[1])]))
    assert np.array_equal(pkg.g1_jac_to_affine(s01), cref.g1_jac_to_aff(want))
    # doubling and cancellation branches
    dbl = pkg.g1_jac_add(j[2], j[2])
    want = cref.msm_g1(util.fr_arr([2]), cref.g1_jac_to_aff(j[2]).reshape(1, 8))
    assert np.array_equal(pkg.g1_jac_to_affine(dbl), cref.g1_jac_to_aff(want))
    inf = np.zeros(12, dtype=np.uint64)
    assert np.array_equal(pkg.g1_jac_to_affine(pkg.g1_jac_add(j[3], inf)), cref.g1_jac_to_aff(j[3]))
    assert not pkg.g1_jac_to_affine(inf).any()
    b2, _ = util.known_dlog_bases_g2(2)
    k = [cref.msm_g2(util.fr_arr([3 + i]), b2[i:i + 1]) for i in range(2)]
    want2 = cref.msm_g2(util.fr_arr([3, 4]), b2)
    assert np.array_equal(pkg.g2_jac_to_affine(pkg.g2_jac_add(k[0], k[1])), cref.g2_jac_to_aff(want2))


def test_groth16_finalize_and_bytes_match_golden(pkg, cref):
    g = util.golden()["groth16"]
    pkj = g["pk"]
    w, a, b, c = G.mul_chain_witness(g["nb_mul"], g["x"])
    d = N.Domain(len(a))
    h = N.compute_h(a, b, c, d)
    wa = [x for x, inf in zip(w, pkj["infinity_A"]) if not inf]
    wb = [x for x, inf in zip(w, pkj["infinity_B"]) if not inf]
    out = pkg.Groth16MsmOut()
    def put(field, arr):
        for i, v in enumerate(arr):
            field[i] = int(v)
    put(out.ar, cref.msm_g1(util.fr_arr(wa), util.g1_arr(pkj["g1_A"])))
    put(out.bs1, cref.msm_g1(util.fr_arr(wb), util.g1_arr(pkj["g1_B"])))
    put(out.krs, cref.msm_g1(util.fr_arr(w[2:]), util.g1_arr(pkj["g1_K"])))
    put(out.krs2, cref.msm_g1(util.fr_arr(h[: d.cardinality - 1]), util.g1_arr(pkj["g1_Z"])))
    put(out.bs2, cref.msm_g2(util.fr_arr(wb), util.g2_arr(pkj["g2_B"])))
    ar, bs, krs = pkg.groth16_finalize(out, cref.fr(g["r"]), cref.fr(g["s"]),
                                       util.g1_arr([pkj["g1_alpha"]])[0], util.g1_arr([pkj["g1_beta"]])[0],
                                       util.g1_arr([pkj["g1_delta"]])[0], util.g2_arr([pkj["g2_beta"]])[0],
                                       util.g2_arr([pkj["g2_delta"]])[0])
    assert util.g1_aff_ints(ar) == g["ar"]
    assert util.g2_aff_ints(bs) == g["bs"]
    assert util.g1_aff_ints(krs) == g["krs"]
    assert pkg.groth16_proof_bytes(ar, bs, krs).hex() == g["proof_hex"]
