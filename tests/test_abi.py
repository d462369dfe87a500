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
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(pkg):
    lib = ctypes.CDLL(pkg.library_path())
    names = header_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in b200prover.h but not exported"
    pkg.load_library()  # binds every prototype the Python mirror uses


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
