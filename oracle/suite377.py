"""CPU oracle (TEST INFRASTRUCTURE ONLY) -- the NTT and Groth16 restatements instantiated on BLS12-377.

`oracle/ntt.py` and `oracle/groth16.py` are written against the names a curve module exports (P, R, G1_GEN,
msm_g1, ...) and import them relatively from `.bn254`.  The BLS12-377 instances are the SAME source files
executed a second time inside a synthetic package whose `bn254` submodule is `oracle/bls12_377.py`: one
restatement of the algorithm (prove.go / setup.go are generated from one template for every curve in the
reference, `internal/generator/backend/template/zkpschemes/groth16/`), two curves.

    from oracle.suite377 import curve, ntt, groth16

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline leg may import this module.
"""
import importlib.util
import os
import sys
import types

from . import bls12_377 as curve

_PKG = "oracle._on_bls12_377"
_HERE = os.path.dirname(os.path.abspath(__file__))


def _instantiate(name: str):
    full = f"{_PKG}.{name}"
    if full in sys.modules:
        return sys.modules[full]
    spec = importlib.util.spec_from_file_location(full, os.path.join(_HERE, name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[full] = mod
    spec.loader.exec_module(mod)
    return mod


if _PKG not in sys.modules:
    _pkg = types.ModuleType(_PKG)
    _pkg.__path__ = []          # a package, so the relative imports of ntt.py / groth16.py resolve inside it
    sys.modules[_PKG] = _pkg
    sys.modules[_PKG + ".bn254"] = curve
    _pkg.bn254 = curve

ntt = _instantiate("ntt")
sys.modules[_PKG].ntt = ntt
groth16 = _instantiate("groth16")
sys.modules[_PKG].groth16 = groth16
# PlonK (backend/plonk/bls12-377/prove.go is the bn254 template with the curve swapped): the same two oracle files
plonk = _instantiate("plonk")
sys.modules[_PKG].plonk = plonk
plonk_prove = _instantiate("plonk_prove")
