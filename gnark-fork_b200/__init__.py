"""b200 prover: host-side mirror (Python, ctypes) of the C ABI in include/b200prover.h.

The product is libb200prover.so (hand-written sm_100a CUDA behind a C ABI); this package only
loads it and mirrors, for callers without Go, the names of the reference interface it slots
under (groth16.Prove -> backend/groth16/bn254/prove.go; icicle device surface ->
backend/groth16/bn254/icicle/icicle.go).  There is no CPU fallback: importing works anywhere,
but creating a Context without the library or without a CUDA device raises B200Error.
"""
from .api import (  # noqa: F401
    B200Error, Context, Bases, DeviceVector, Groth16ProvingKey, Groth16MsmOut,
    g1_jac_add, g2_jac_add, g1_jac_to_affine, g2_jac_to_affine, groth16_finalize, groth16_proof_bytes,
    pinned_empty, library_path, load_library, DIF, DIT, BN254, BLS12_377, fp_words,
    Group, GroupGroth16ProvingKey, PlonkProof, PlonkProvingKey, PlonkSession, plonk_proof_bytes, plonk_derive_challenges,
)
from . import sharding  # noqa: F401,E402
from . import groth16_host  # noqa: F401,E402
from . import pk_format, plonk_pk_format  # noqa: F401,E402
from . import api  # noqa: F401,E402
