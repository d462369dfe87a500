"""Generates tests/golden/plonk_small.json from the big-int PlonK oracle (oracle/plonk_prove.py): a Mul-chain
SCS circuit, a known-tau setup, fixed blinding coefficients -> the proof, its 552 serialised bytes and the
challenges.  Run from the repo root:  python tests/golden/make_plonk_golden.py"""
import json
import os
import random
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle import plonk_prove as PP  # noqa: E402

R = PP.R
cases = []
for nb_mul, seed in ((2, 11), (6, 12), (29, 13), (100, 14)):
    rng = random.Random(seed)
    spr = PP.mul_chain_scs(nb_mul)
    tau = rng.randrange(1, R)
    x = rng.randrange(2, R)
    pk = PP.setup(spr, tau)
    sol = PP.mul_chain_solution(nb_mul, x)
    blinding = tuple([rng.randrange(R) for _ in range(k)] for k in (2, 2, 2, 3))
    proof, ch = PP.prove(spr, pk, sol, blinding)
    assert PP.verify(proof, pk.vk, sol[:1], tau)
    cases.append({
        "nb_mul": nb_mul, "tau": hex(tau), "x": hex(x), "blinding": [[hex(v) for v in b] for b in blinding],
        "log_n": pk.d0.log_n, "log_big": pk.d1.log_n,
        "challenges": {k: hex(v) for k, v in ch.items()},
        "claimed": [hex(v) for v in proof.claimed], "zshift_value": hex(proof.zshift_value),
        "proof_bytes": PP.proof_bytes(proof).hex(),
    })
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "plonk_small.json")
with open(out, "w") as f:
    json.dump({"generator": "tests/golden/make_plonk_golden.py", "cases": cases}, f, indent=1)
print("wrote", out)
