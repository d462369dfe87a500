"""A pure-C translation unit drives the whole Groth16 path through include/b200prover.h (tests/abi_c/smoke.c, gcc
-std=c11): the header is valid C, the two libraries link into ONE binary (per-curve symbol names), and on the GPU the
C caller's proofs equal the oracle's byte for byte on both curves."""
import os
import random
import struct
import subprocess

import numpy as np
import pytest

from tests.conftest import ROOT

HERE = os.path.join(ROOT, "tests", "abi_c")
LIBDIR = os.path.join(ROOT, "gnark-fork_b200")


def build_binary():
    exe = os.path.join(HERE, "smoke_both")
    cc = ["gcc", "-std=c11", "-O1", "-Wall", "-Wextra", "-Werror", "-I" + os.path.join(ROOT, "include")]
    objs = []
    for name, flags in (("smoke_bn254.o", []), ("smoke_bls12377.o", ["-DB200_CURVE_BLS12_377", "-DSMOKE_MAIN=smoke_bls12377"])):
        o = os.path.join(HERE, name)
        subprocess.check_call(cc + flags + ["-c", os.path.join(HERE, "smoke.c"), "-o", o])
        objs.append(o)
    o = os.path.join(HERE, "main.o")
    subprocess.check_call(cc + ["-c", os.path.join(HERE, "main.c"), "-o", o])
    subprocess.check_call(["gcc", "-o", exe, o] + objs + ["-L" + LIBDIR, "-lb200prover", "-lb200prover_bls12377",
                                                         "-Wl,-rpath," + LIBDIR, "-Wl,-rpath,/usr/local/cuda/lib64"])
    return exe


def test_c11_caller_compiles_and_links_both_libraries():
    exe = build_binary()
    out = subprocess.run(["nm", "-u", exe], capture_output=True, text=True, check=True).stdout
    assert " b200_groth16_setup" in out and " b200_bls12377_groth16_setup" in out


def _input_file(path, curve):
    """refCircuit at 2^8 with a seeded toxic waste; returns the proof bytes the oracle expects."""
    if curve == "bn254":
        from oracle import bn254 as B
        from oracle import groth16 as G
        from oracle import cref
        fr = lambda xs: cref.ints_to_mont([x % B.R for x in xs], B.R)
        fp = lambda xs: cref.ints_to_mont(xs, B.P)
    else:
        from oracle import bls12_377 as B
        from oracle.suite377 import groth16 as G
        from tests import util377 as U
        fr, fp = U.fr_arr, U.fp_arr
    log_n = 8
    nb_mul = (1 << log_n) - 1
    rnd = random.Random(99)
    td = G.Trapdoor(*[rnd.randrange(1, B.R) for _ in range(5)])
    r, s = rnd.randrange(B.R), rnd.randrange(B.R)
    r1cs = G.mul_chain_circuit(nb_mul)
    pk, vk = G.setup(r1cs, td, points=False)
    w, a, b, c = G.mul_chain_witness(nb_mul, 2)
    d = pk.domain
    g2 = B.G2_GEN
    blob = struct.pack("<I", log_n) + fr([d.generator, d.coset_gen, td.t, td.alpha, td.beta, td.gamma, td.delta, r, s]).tobytes()
    blob += fp([B.G1_GEN[0], B.G1_GEN[1]]).tobytes() + fp([g2[0][0], g2[0][1], g2[1][0], g2[1][1]]).tobytes()
    blob += fr(w).tobytes() + fr(a).tobytes() + fr(b).tobytes() + fr(c).tobytes()
    open(path, "wb").write(blob)
    dl = G.prove_dlog(pk, td, w, a, b, c, r, s, 2)
    assert G.trapdoor_check(vk, w[:2], dl["ar"], dl["bs"], dl["krs"])
    return G.proof_bytes(B.g1_mul(B.G1_GEN, dl["ar"]), B.g2_mul(B.G2_GEN, dl["bs"]), B.g1_mul(B.G1_GEN, dl["krs"]))


@pytest.mark.gpu
@pytest.mark.parametrize("curve", ["bn254", "bls12377"])
def test_c_caller_proof_matches_oracle(tmp_path, curve):
    exe = build_binary()
    path = str(tmp_path / f"{curve}.bin")
    want = _input_file(path, curve)
    res = subprocess.run([exe, curve, path], capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stderr
    assert res.stdout.strip() == want.hex()
