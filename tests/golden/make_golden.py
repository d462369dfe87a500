"""Generates tests/golden/bn254_small.json from the big-int DEFINITIONS in oracle/*.py
(naive O(n^2) DFT, double-and-add MSM, trapdoor Groth16) -- no fast algorithm is involved, so the
file pins both the C++ oracle and the CUDA path to independently computed values.

The reference itself holds no golden vector for this path and cannot run here (no Go, gnark-crypto
not vendored: SURVEY.md 8(c)), so these are "definition" vectors, not gnark outputs.
Run from the repo root:  python tests/golden/make_golden.py
"""
import json
import os
import random
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import bn254 as B  # noqa: E402
from oracle import groth16 as G  # noqa: E402
from oracle import ntt as N  # noqa: E402

rnd = random.Random(0xB200)
out = {}

# --- NTT: every (inverse, decimation, coset) variant at n = 16, from the naive DFT ---------
log_n, n = 4, 16
d = N.Domain(n)
x = [rnd.randrange(B.R) for _ in range(n)]
out["ntt"] = {"log_n": log_n, "generator": d.generator, "coset_gen": d.coset_gen, "input": x, "cases": []}
br = [N.bitrev(i, log_n) for i in range(n)]
for inverse in (0, 1):
    for dec in (N.DIF, N.DIT):
        for coset in (0, 1):
            # natural-order semantics, then place according to the decimation's ordering contract
            src = list(x) if dec == N.DIF else [x[br[i]] for i in range(n)]  # DIT reads bit-reversed input
            if not inverse:
                v = [src[i] * pow(d.coset_gen, i, B.R) % B.R for i in range(n)] if coset else src
                y = N.naive_dft(v, d.generator)
            else:
                y = N.naive_dft(src, d.generator_inv)
                y = [y[i] * d.cardinality_inv % B.R for i in range(n)]
                if coset:
                    y = [y[i] * pow(d.coset_gen_inv, i, B.R) % B.R for i in range(n)]
            res = [y[br[i]] for i in range(n)] if dec == N.DIF else y  # DIF leaves bit-reversed output
            out["ntt"]["cases"].append({"inverse": inverse, "decimation": dec, "on_coset": coset, "output": res})

# --- computeH at n = 8 from the polynomial definition ---------------------------------------
nb_mul = 6
r1cs = G.mul_chain_circuit(nb_mul)
w, a, b, c = G.mul_chain_witness(nb_mul, 3)
dom = N.Domain(r1cs.nb_constraints)
m = dom.cardinality
pad = lambda v: list(v) + [0] * (m - len(v))
coef = lambda v: [t * dom.cardinality_inv % B.R for t in N.naive_dft(pad(v), dom.generator_inv)]
ca, cb, cc = coef(a), coef(b), coef(c)
ev = lambda co, pt: sum(k * pow(pt, i, B.R) for i, k in enumerate(co)) % B.R
g = dom.coset_gen
den = pow((pow(g, m, B.R) - 1) % B.R, -1, B.R)
hv = [(ev(ca, g * pow(dom.generator, i, B.R)) * ev(cb, g * pow(dom.generator, i, B.R)) - ev(cc, g * pow(dom.generator, i, B.R))) * den % B.R
      for i in range(m)]
hc = [t * dom.cardinality_inv % B.R for t in N.naive_dft(hv, dom.generator_inv)]
hc = [hc[i] * pow(dom.coset_gen_inv, i, B.R) % B.R for i in range(m)]
lg = m.bit_length() - 1
out["compute_h"] = {"log_n": lg, "a": a, "b": b, "c": c, "generator": dom.generator, "coset_gen": g,
                    "h_bitreversed": [hc[N.bitrev(i, lg)] for i in range(m)]}

# --- MSM G1 / G2, n = 12, with infinity, zero scalar, r-1 and a repeated base ---------------
nm = 12
ks = [rnd.randrange(1, B.R) for _ in range(nm)]
p1 = [B.g1_mul(B.G1_GEN, k) for k in ks]
p2 = [B.g2_mul(B.G2_GEN, k) for k in ks]
p1[5] = None; p2[5] = None
p1[7] = p1[6]; p2[7] = p2[6]
sc = [rnd.randrange(B.R) for _ in range(nm)]
sc[0] = 0; sc[1] = B.R - 1; sc[2] = 1
enc1 = lambda p: None if p is None else [p[0], p[1]]
enc2 = lambda p: None if p is None else [p[0][0], p[0][1], p[1][0], p[1][1]]
out["msm"] = {"scalars": sc, "g1_bases": [enc1(p) for p in p1], "g2_bases": [enc2(p) for p in p2],
              "g1_result": enc1(B.msm_g1(p1, sc)), "g2_result": enc2(B.msm_g2(p2, sc))}

# --- Groth16 (trapdoor setup, fixed r, s), proof points and bytes ---------------------------
td = G.Trapdoor(*[rnd.randrange(1, B.R) for _ in range(5)])
pk, vk = G.setup(r1cs, td)
r_, s_ = rnd.randrange(B.R), rnd.randrange(B.R)
ar, bs, krs = G.prove_points(pk, w, a, b, c, r_, s_, r1cs.nb_public)
dl = G.prove_dlog(pk, td, w, a, b, c, r_, s_, r1cs.nb_public)
assert G.trapdoor_check(vk, w[: r1cs.nb_public], dl["ar"], dl["bs"], dl["krs"])
assert ar == B.g1_mul(B.G1_GEN, dl["ar"]) and krs == B.g1_mul(B.G1_GEN, dl["krs"]) and bs == B.g2_mul(B.G2_GEN, dl["bs"])
out["groth16"] = {
    "nb_mul": nb_mul, "x": 3, "trapdoor": [td.t, td.alpha, td.beta, td.gamma, td.delta], "r": r_, "s": s_,
    "ar": enc1(ar), "bs": enc2(bs), "krs": enc1(krs), "proof_hex": G.proof_bytes(ar, bs, krs).hex(),
    "pk": {"g1_alpha": enc1(pk.g1_alpha), "g1_beta": enc1(pk.g1_beta), "g1_delta": enc1(pk.g1_delta),
           "g2_beta": enc2(pk.g2_beta), "g2_delta": enc2(pk.g2_delta),
           "g1_A": [enc1(p) for p in pk.g1_A], "g1_B": [enc1(p) for p in pk.g1_B], "g1_K": [enc1(p) for p in pk.g1_K],
           "g1_Z": [enc1(p) for p in pk.g1_Z], "g2_B": [enc2(p) for p in pk.g2_B],
           "infinity_A": [int(v) for v in pk.infinity_A], "infinity_B": [int(v) for v in pk.infinity_B]},
}

path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "bn254_small.json")
with open(path, "w") as f:
    json.dump(out, f, separators=(",", ":"))
print("wrote", path, os.path.getsize(path), "bytes")
