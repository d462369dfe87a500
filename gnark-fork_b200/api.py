"""ctypes binding of include/b200prover.h (numpy arrays of u64 limbs in, numpy arrays out).

Layouts (identical to gnark-crypto's in-memory ones, see the header): Fr/Fp element = 4 u64
Montgomery limbs; G1 affine (n, 8); G1 Jacobian (12,); G2 affine (n, 16); G2 Jacobian (24,).

BLS12-377 (`curve=BLS12_377`) binds the second library, libb200prover_bls12377.so: Fr is still 4 u64, Fp is
6 u64 (`fp_words(curve)`), so G1 affine (n, 12), G1 Jacobian (18,), G2 affine (n, 24), G2 Jacobian (36,).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

BN254 = 0
BLS12_377 = 1
DIF = 0
DIT = 1
PLONK_MAX_BSB22 = 8      # B200_PLONK_MAX_BSB22

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIBS: dict = {}
_LIB_FILES = {BN254: ("B200PROVER_LIB", "libb200prover.so"), BLS12_377: ("B200PROVER_LIB_BLS12_377", "libb200prover_bls12377.so")}
_FP_WORDS = {BN254: 4, BLS12_377: 6}


def fp_words(curve: int = BN254) -> int:
    """u64 words of one base-field element (B200_FP_WORDS of the header)."""
    return _FP_WORDS[curve]


class B200Error(RuntimeError):
    pass


def library_path(curve: int = BN254) -> str:
    env, name = _LIB_FILES[curve]
    return os.environ.get(env, os.path.join(_HERE, name))


class _Timings(C.Structure):
    _fields_ = [(k, C.c_float) for k in
                ("h2d_ms", "compute_h_ms", "msm_a_ms", "msm_b1_ms", "msm_k_ms", "msm_z_ms", "msm_b2_ms", "total_ms")]


class _MsmSetInfo(C.Structure):
    """b200_msm_set_info."""
    _fields_ = [("n", C.c_uint64), ("device_bytes", C.c_uint64), ("c", C.c_int32), ("windows", C.c_int32),
                ("folded", C.c_int32), ("ba_rounds", C.c_int32)]

    def as_dict(self):
        return {"n": int(self.n), "device_bytes": int(self.device_bytes), "c": int(self.c),
                "windows": int(self.windows), "folded": int(self.folded), "ba_rounds": int(self.ba_rounds)}


class _PlonkRowsDesc(C.Structure):
    """b200_plonk_rows_desc."""
    _fields_ = [("x", C.c_void_p * 15), ("qci", C.c_void_p * 16), ("nb_bsb", C.c_uint), ("log_n", C.c_uint),
                ("log_big", C.c_uint), ("coset_index", C.c_uint), ("generator_small", C.c_void_p),
                ("beta", C.c_uint64 * 4), ("gamma", C.c_uint64 * 4), ("alpha", C.c_uint64 * 4), ("cs", C.c_uint64 * 4),
                ("bl", (C.c_uint64 * 4) * 2), ("br", (C.c_uint64 * 4) * 2), ("bo", (C.c_uint64 * 4) * 2), ("bz", (C.c_uint64 * 4) * 3)]


class _PlonkPkDesc(C.Structure):
    """b200_plonk_pk_desc."""
    _fields_ = [("log_n", C.c_uint), ("log_big", C.c_uint), ("generator", C.c_void_p), ("generator_big", C.c_void_p),
                ("coset_gen", C.c_void_p), ("ql", C.c_void_p), ("qr", C.c_void_p), ("qm", C.c_void_p), ("qo", C.c_void_p),
                ("qk", C.c_void_p), ("s1", C.c_void_p), ("s2", C.c_void_p), ("s3", C.c_void_p), ("perm", C.c_void_p),
                ("kzg_g1", C.c_void_p), ("kzg_lagrange_g1", C.c_void_p), ("vk_commitments", C.c_void_p),
                ("nb_public", C.c_size_t), ("nb_bsb22", C.c_size_t), ("qcp", C.c_void_p), ("vk_qcp", C.c_void_p),
                ("commitment_index", C.c_void_p), ("raw_file", C.c_int), ("shard_index", C.c_int), ("shard_count", C.c_int)]


class _PlonkBsb22(C.Structure):
    """b200_plonk_bsb22."""
    _fields_ = [("pi2", C.c_void_p), ("commitments", C.c_void_p), ("values", C.c_void_p)]


def _plonk_proof_struct(fw):
    pw = 2 * fw        # u64 words of an affine G1 point

    class _Proof(C.Structure):
        """b200_plonk_proof: plonk.Proof (prove.go:96-114), points affine, scalars Montgomery."""
        _fields_ = [("lro", (C.c_uint64 * pw) * 3), ("z", C.c_uint64 * pw), ("h", (C.c_uint64 * pw) * 3),
                    ("batched_h", C.c_uint64 * pw), ("claimed", (C.c_uint64 * 4) * 7), ("zshift_h", C.c_uint64 * pw),
                    ("zshift_value", C.c_uint64 * 4), ("challenges", (C.c_uint64 * 4) * 5), ("nb_bsb22", C.c_uint64),
                    ("claimed_qcp", (C.c_uint64 * 4) * 8), ("bsb22", (C.c_uint64 * pw) * 8)]
        FP_WORDS = fw

        def as_dict(self):
            return {k: np.array(getattr(self, k), dtype=np.uint64) for k, _ in self._fields_}

        def all_claimed(self) -> np.ndarray:
            """BatchedProof.ClaimedValues: the 7 fixed ones then Qcp_i(zeta), (7 + nb_bsb22, 4)."""
            d = self.as_dict()
            return np.concatenate([d["claimed"], d["claimed_qcp"][:int(self.nb_bsb22)]])
    return _Proof


PlonkProof = _plonk_proof_struct(4)             # BN254
PlonkProof377 = _plonk_proof_struct(6)          # the struct of libb200prover_bls12377.so
_PLONK_PROOF = {BN254: PlonkProof, BLS12_377: PlonkProof377}


class _PlonkTimings(C.Structure):
    _fields_ = [(k, C.c_float) for k in ("h2d_ms", "commit_lro_ms", "ratio_ms", "commit_z_ms", "quotient_ms", "commit_h_ms",
                                          "open_z_ms", "linearize_ms", "batch_open_ms", "total_ms")]


class Groth16MsmOut(C.Structure):
    """b200_groth16_out: the five MultiExp results before the fix-ups (prove.go:201-290)."""
    _fields_ = [("ar", C.c_uint64 * 12), ("bs1", C.c_uint64 * 12), ("krs", C.c_uint64 * 12),
                ("krs2", C.c_uint64 * 12), ("bs2", C.c_uint64 * 24)]

    def as_dict(self):
        return {k: np.array(getattr(self, k), dtype=np.uint64) for k, _ in self._fields_}


class Groth16MsmOut377(C.Structure):
    """b200_groth16_out of the BLS12-377 library (B200_FP_WORDS = 6)."""
    _fields_ = [("ar", C.c_uint64 * 18), ("bs1", C.c_uint64 * 18), ("krs", C.c_uint64 * 18),
                ("krs2", C.c_uint64 * 18), ("bs2", C.c_uint64 * 36)]
    as_dict = Groth16MsmOut.as_dict


_MSM_OUT = {BN254: Groth16MsmOut, BLS12_377: Groth16MsmOut377}


class _PkDesc(C.Structure):
    _fields_ = [
        ("log_n", C.c_uint), ("generator", C.c_void_p), ("coset_gen", C.c_void_p),
        ("g1_a", C.c_void_p), ("n_a", C.c_size_t), ("g1_b", C.c_void_p), ("n_b", C.c_size_t),
        ("g1_k", C.c_void_p), ("n_k", C.c_size_t), ("g1_z", C.c_void_p), ("n_z", C.c_size_t),
        ("g2_b", C.c_void_p), ("n_g2b", C.c_size_t),
        ("infinity_a", C.c_void_p), ("infinity_b", C.c_void_p),
        ("n_wires", C.c_size_t), ("nb_public", C.c_size_t),
        ("shard_index", C.c_int), ("shard_count", C.c_int), ("presliced", C.c_int), ("raw_points", C.c_int),
    ]


class _R1csMatrix(C.Structure):
    """b200_r1cs_matrix."""
    _fields_ = [("row_ptr", C.c_void_p), ("terms", C.c_void_p)]


def _setup_structs(fw):
    class _SetupDesc(C.Structure):
        """b200_groth16_setup_desc."""
        _fields_ = [("log_n", C.c_uint), ("generator", C.c_void_p), ("coset_gen", C.c_void_p),
                    ("nb_constraints", C.c_size_t), ("nb_wires", C.c_size_t), ("nb_public", C.c_size_t),
                    ("l", _R1csMatrix), ("r", _R1csMatrix), ("o", _R1csMatrix),
                    ("coeffs", C.c_void_p), ("nb_coeffs", C.c_size_t),
                    ("wire_class", C.c_void_p), ("nb_commitments", C.c_size_t),
                    ("tau", C.c_uint64 * 4), ("alpha", C.c_uint64 * 4), ("beta", C.c_uint64 * 4),
                    ("gamma", C.c_uint64 * 4), ("delta", C.c_uint64 * 4),
                    ("g1_gen", C.c_void_p), ("g2_gen", C.c_void_p), ("shard_index", C.c_int), ("shard_count", C.c_int)]

    class _SetupOut(C.Structure):
        """b200_groth16_setup_out of a library with B200_FP_WORDS = fw."""
        _fields_ = [("g1_alpha", C.c_uint64 * (2 * fw)), ("g1_beta", C.c_uint64 * (2 * fw)), ("g1_delta", C.c_uint64 * (2 * fw)),
                    ("g2_beta", C.c_uint64 * (4 * fw)), ("g2_delta", C.c_uint64 * (4 * fw)), ("g2_gamma", C.c_uint64 * (4 * fw)),
                    ("n_a", C.c_size_t), ("n_b", C.c_size_t), ("n_k", C.c_size_t), ("n_z", C.c_size_t), ("n_vk_k", C.c_size_t),
                    ("infinity_a", C.c_void_p), ("infinity_b", C.c_void_p), ("vk_k", C.c_void_p), ("ck_k", C.c_void_p),
                    ("g1_a", C.c_void_p), ("g1_b", C.c_void_p), ("g1_k", C.c_void_p), ("g1_z", C.c_void_p), ("g2_b", C.c_void_p)]
    return _SetupDesc, _SetupOut


_SETUP = {BN254: _setup_structs(4), BLS12_377: _setup_structs(6)}


PROF_SLOTS = ("msm_sort", "msm_acc_g1", "msm_acc_g2", "msm_reduce", "ntt_pass", "pointwise", "copy_in", "_7")


class _Profile(C.Structure):
    _fields_ = [("ms", C.c_double * 8), ("launches", C.c_uint64 * 8), ("units", C.c_double * 8),
                ("kernel_launches", C.c_uint64)]


_SYM_PREFIX = {BN254: "b200_", BLS12_377: "b200_bls12377_"}


class _Lib:
    """The loaded library under the header's names: the BLS12-377 build exports every entry point b200_x as
    b200_bls12377_x (include/b200prover.h, per-curve symbol names) so that both libraries link into one binary."""

    def __init__(self, cdll, prefix):
        self._cdll, self._prefix = cdll, prefix

    def __getattr__(self, name):
        fn = getattr(self._cdll, self._prefix + name[5:] if name.startswith("b200_") else name)
        setattr(self, name, fn)
        return fn


def load_library(curve: int = BN254):
    """Load libb200prover.so (or the BLS12-377 build); fails loudly (no fallback of any kind)."""
    if curve in _LIBS:
        return _LIBS[curve]
    if curve not in _LIB_FILES:
        raise B200Error(f"unknown curve id {curve}")
    path = library_path(curve)
    if not os.path.exists(path):
        raise B200Error(f"{path} not found: build it with `make -C gnark-fork_b200/csrc"
                        f"{' CURVE=bls12_377' if curve == BLS12_377 else ''}` "
                        "(or __graft_entry__.build()); there is no CPU fallback")
    lib = _Lib(C.CDLL(path), _SYM_PREFIX[curve])
    Groth16MsmOut = _MSM_OUT[curve]     # the structs of THIS library's B200_FP_WORDS
    PlonkProof = _PLONK_PROOF[curve]
    vp, sz, u, i = C.c_void_p, C.c_size_t, C.c_uint, C.c_int
    sig = {
        "b200_device_count": (i, []),
        "b200_init": (i, [i, C.POINTER(vp)]),
        "b200_destroy": (None, [vp]),
        "b200_last_error": (C.c_char_p, [vp]),
        "b200_version": (C.c_char_p, []),
        "b200_build_curve": (i, []),
        "b200_build_fp_words": (i, []),
        "b200_set_option": (i, [vp, C.c_char_p, C.c_long]),
        "b200_host_alloc": (vp, [sz]),
        "b200_host_free": (None, [vp]),
        "b200_dev_alloc": (i, [vp, sz, C.POINTER(vp)]),
        "b200_dev_free": (i, [vp, vp]),
        "b200_h2d": (i, [vp, vp, vp, sz]),
        "b200_d2h": (i, [vp, vp, vp, sz]),
        "b200_sync": (i, [vp]),
        "b200_fr_vec_mul": (i, [vp, i, vp, vp, vp, sz]),
        "b200_fp_vec_mul": (i, [vp, i, vp, vp, vp, sz]),
        "b200_imad_peak": (i, [vp, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double)]),
        "b200_ntt": (i, [vp, i, vp, u, i, i, i, vp, vp]),
        "b200_ntt_dev": (i, [vp, i, vp, u, i, i, i, vp, vp, C.POINTER(C.c_float)]),
        "b200_groth16_compute_h": (i, [vp, i, vp, vp, vp, sz, u, vp, vp, vp]),
        "b200_msm_g1": (i, [vp, i, vp, vp, sz, vp]),
        "b200_msm_g2": (i, [vp, i, vp, vp, sz, vp]),
        "b200_bases_upload": (i, [vp, i, i, vp, sz, C.POINTER(vp)]),
        "b200_bases_upload_raw": (i, [vp, i, i, vp, sz, C.POINTER(vp)]),
        "b200_points_from_raw": (i, [vp, i, i, vp, sz, vp]),
        "b200_bases_free": (i, [vp, vp]),
        "b200_bases_info": (i, [vp, vp, C.POINTER(_MsmSetInfo)]),
        "b200_msm_resident": (i, [vp, vp, vp, sz, vp]),
        "b200_msm_dev": (i, [vp, vp, vp, sz, vp, C.POINTER(C.c_float)]),
        "b200_g1_jac_add": (i, [vp, vp, vp]),
        "b200_g2_jac_add": (i, [vp, vp, vp]),
        "b200_g1_jac_to_affine": (i, [vp, vp]),
        "b200_g2_jac_to_affine": (i, [vp, vp]),
        "b200_batch_scalar_mul": (i, [vp, i, i, vp, vp, sz, vp]),
        "b200_kzg_to_lagrange_g1": (i, [vp, i, vp, C.c_uint, vp, vp]),
        "b200_groth16_pk_upload": (i, [vp, i, C.POINTER(_PkDesc), C.POINTER(vp)]),
        "b200_groth16_pk_free": (i, [vp, vp]),
        "b200_groth16_setup": (i, [vp, i, C.POINTER(_SETUP[curve][0]), C.POINTER(vp), C.POINTER(_SETUP[curve][1])]),
        "b200_group_init": (i, [C.POINTER(C.c_int), i, C.POINTER(vp)]),
        "b200_group_destroy": (None, [vp]),
        "b200_group_last_error": (C.c_char_p, [vp]),
        "b200_group_size": (i, [vp]),
        "b200_group_ctx": (vp, [vp, i]),
        "b200_group_set_option": (i, [vp, C.c_char_p, C.c_long]),
        "b200_group_groth16_pk_upload": (i, [vp, i, C.POINTER(_PkDesc), C.POINTER(vp)]),
        "b200_group_groth16_setup": (i, [vp, i, C.POINTER(_SETUP[curve][0]), C.POINTER(vp), C.POINTER(_SETUP[curve][1])]),
        "b200_group_groth16_pk_free": (i, [vp, vp]),
        "b200_group_groth16_pk_info": (i, [vp, vp, i, C.POINTER(_MsmSetInfo), C.POINTER(i)]),
        "b200_group_plonk_pk_upload": (i, [vp, i, C.POINTER(_PlonkPkDesc), C.POINTER(vp)]),
        "b200_group_plonk_pk_free": (i, [vp, vp]),
        "b200_group_plonk_prove": (i, [vp, vp, vp, vp, vp, vp, vp, C.POINTER(_PlonkBsb22), C.POINTER(PlonkProof), C.POINTER(_PlonkTimings)]),
        "b200_group_groth16_prove_core": (i, [vp, vp, vp, vp, vp, sz, vp, vp, sz, C.POINTER(Groth16MsmOut), C.POINTER(_Timings)]),
        "b200_group_groth16_prove_core_dev": (i, [vp, vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp),
                                                  C.POINTER(Groth16MsmOut), C.POINTER(_Timings)]),
        "b200_groth16_pk_info": (i, [vp, vp, C.POINTER(_MsmSetInfo)]),
        "b200_groth16_pk_wire_range": (i, [vp, vp, C.POINTER(sz), C.POINTER(sz)]),
        "b200_groth16_prove_core": (i, [vp, vp, vp, vp, vp, sz, vp, vp, sz, C.POINTER(Groth16MsmOut), C.POINTER(_Timings)]),
        "b200_groth16_prove_core_dev": (i, [vp, vp, vp, vp, vp, sz, vp, C.POINTER(Groth16MsmOut), C.POINTER(_Timings)]),
        "b200_groth16_msms_dev": (i, [vp, vp, vp, vp, C.POINTER(Groth16MsmOut), C.POINTER(_Timings)]),
        "b200_ntt_dist_phase": (i, [vp, i, u, vp, vp, vp, vp, u, u, i, i, i]),
        "b200_fr_mul_sub_dev": (i, [vp, i, vp, vp, vp, sz]),
        "b200_use_stream": (i, [vp, vp]),
        "b200_domain_cache_trim": (i, [vp, C.POINTER(sz)]),
        "b200_fr_batch_inverse_dev": (i, [vp, i, vp, vp, sz]),
        "b200_fr_prefix_product_dev": (i, [vp, i, vp, vp, sz]),
        "b200_fr_prefix_product_static": (i, [vp, i, vp, vp, sz]),
        "b200_fr_poly_eval_dev": (i, [vp, i, vp, sz, vp, vp]),
        "b200_fr_poly_div_linear_dev": (i, [vp, i, vp, sz, vp, vp]),
        "b200_plonk_divide_by_xn_minus_one_dev": (i, [vp, i, vp, u, u, vp, vp]),
        "b200_kzg_commit_dev": (i, [vp, vp, vp, sz, vp]),
        "b200_plonk_build_ratio_dev": (i, [vp, i, vp, vp, vp, vp, vp, vp, u, vp, vp, vp]),
        "b200_plonk_quotient_rows_dev": (i, [vp, i, C.POINTER(_PlonkRowsDesc), vp]),
        "b200_plonk_pk_upload": (i, [vp, i, C.POINTER(_PlonkPkDesc), C.POINTER(vp)]),
        "b200_plonk_pk_free": (i, [vp, vp]),
        "b200_plonk_prove": (i, [vp, vp, vp, vp, vp, vp, vp, C.POINTER(PlonkProof), C.POINTER(_PlonkTimings)]),
        "b200_plonk_prove_bsb22": (i, [vp, vp, vp, vp, vp, vp, vp, C.POINTER(_PlonkBsb22), C.POINTER(PlonkProof), C.POINTER(_PlonkTimings)]),
        "b200_plonk_session_begin": (i, [vp, vp, vp, vp, vp, vp, vp, C.POINTER(_PlonkBsb22), C.POINTER(vp), vp]),
        "b200_plonk_session_step": (i, [vp, vp, i, vp, vp]),
        "b200_plonk_session_numerator": (i, [vp, vp, C.POINTER(vp), C.POINTER(sz), C.POINTER(u)]),
        "b200_plonk_session_finish": (i, [vp, vp, vp, C.POINTER(PlonkProof), C.POINTER(_PlonkTimings)]),
        "b200_plonk_derive_challenges": (i, [vp, vp, vp, sz, C.POINTER(PlonkProof), vp]),
        "b200_plonk_proof_bytes": (i, [C.POINTER(PlonkProof), vp, sz]),
        "b200_profile_start": (i, [vp]),
        "b200_profile_read": (i, [vp, C.POINTER(_Profile)]),
        "b200_groth16_finalize": (i, [C.POINTER(Groth16MsmOut), vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
        "b200_groth16_proof_bytes": (i, [vp, vp, vp, vp, sz]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)  # AttributeError if the library lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    if lib.b200_build_curve() != curve or lib.b200_build_fp_words() != _FP_WORDS[curve]:
        raise B200Error(f"{path} was built for curve {lib.b200_build_curve()} / {lib.b200_build_fp_words()} Fp words, wanted curve {curve}")
    _LIBS[curve] = lib
    return lib


EXPORTED_SYMBOLS = None  # filled lazily by tests from the header


def _u64(a, shape=None) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.uint64)
    if shape is not None:
        a = a.reshape(shape)
    return a


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def pinned_empty(shape, dtype=np.uint64) -> np.ndarray:
    """numpy array over page-locked memory from b200_host_alloc (prove.go:119-125 buffers)."""
    lib = load_library()
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    p = lib.b200_host_alloc(max(n, 1))
    if not p:
        raise B200Error("b200_host_alloc failed")
    buf = (C.c_uint8 * max(n, 1)).from_address(p)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    _PINNED[arr.ctypes.data] = p
    return arr


_PINNED: dict = {}


# ---- host-side group helpers (no ctx) -----------------------------------------------------
def _host_call(name, *arrays, out_words, curve=BN254):
    lib = load_library(curve)
    out = np.zeros(out_words, dtype=np.uint64)
    st = getattr(lib, name)(*[_ptr(_u64(a)) for a in arrays], _ptr(out))
    if st != 0:
        raise B200Error(f"{name} failed with status {st}")
    return out


def g1_jac_add(p, q, curve=BN254):
    return _host_call("b200_g1_jac_add", p, q, out_words=3 * fp_words(curve), curve=curve)


def g2_jac_add(p, q, curve=BN254):
    return _host_call("b200_g2_jac_add", p, q, out_words=6 * fp_words(curve), curve=curve)


def g1_jac_to_affine(p, curve=BN254):
    return _host_call("b200_g1_jac_to_affine", p, out_words=2 * fp_words(curve), curve=curve)


def g2_jac_to_affine(p, curve=BN254):
    return _host_call("b200_g2_jac_to_affine", p, out_words=4 * fp_words(curve), curve=curve)


def groth16_finalize(msm: Groth16MsmOut, r, s, g1_alpha, g1_beta, g1_delta, g2_beta, g2_delta, curve=BN254):
    """Fix-ups of prove.go:178-321 on the five MSM results -> affine (Ar, Bs, Krs)."""
    lib = load_library(curve)
    fw = fp_words(curve)
    ar, bs, krs = np.zeros(2 * fw, np.uint64), np.zeros(4 * fw, np.uint64), np.zeros(2 * fw, np.uint64)
    ins = [_u64(x) for x in (r, s, g1_alpha, g1_beta, g1_delta, g2_beta, g2_delta)]
    st = lib.b200_groth16_finalize(C.byref(msm), *[_ptr(x) for x in ins], _ptr(ar), _ptr(bs), _ptr(krs))
    if st != 0:
        raise B200Error(f"b200_groth16_finalize failed with status {st}")
    return ar, bs, krs


def groth16_proof_bytes(ar, bs, krs, curve=BN254) -> bytes:
    """Proof.WriteTo for a proof without commitments (marshal.go:41-66): 164 bytes on BN254, 244 on BLS12-377."""
    lib = load_library(curve)
    cap = 5 * 8 * fp_words(curve) + 4
    out = np.zeros(cap, dtype=np.uint8)
    n = lib.b200_groth16_proof_bytes(_ptr(_u64(ar)), _ptr(_u64(bs)), _ptr(_u64(krs)), _ptr(out), cap)
    if n < 0:
        raise B200Error(f"b200_groth16_proof_bytes failed with status {n}")
    return out[:n].tobytes()


def _proof_curve(proof) -> int:
    return BLS12_377 if getattr(proof, "FP_WORDS", 4) == 6 else BN254


def plonk_proof_bytes(proof: PlonkProof) -> bytes:
    """Proof.WriteTo (plonk/bn254/marshal.go:41-66), 552 bytes for a BN254 proof without BSB22 commitments."""
    lib = load_library(_proof_curve(proof))
    out = np.zeros(9 * 48 + 8 + 32 * 16 + 48 * 8, dtype=np.uint8)
    n = lib.b200_plonk_proof_bytes(C.byref(proof), _ptr(out), out.shape[0])
    if n < 0:
        raise B200Error(f"b200_plonk_proof_bytes failed with status {n}")
    return out[:n].tobytes()


def plonk_derive_challenges(vk_commitments, public_witness, proof: PlonkProof, vk_qcp=None) -> np.ndarray:
    """gamma, beta, alpha, zeta (4 x 4 u64 Montgomery) of the proof's commitments (verify.go:296-360)."""
    curve = _proof_curve(proof)
    lib = load_library(curve)
    vk = _u64(vk_commitments, (8, 2 * fp_words(curve)))
    pw = _u64(public_witness, (-1, 4))
    qcp = None if vk_qcp is None else _u64(vk_qcp, (-1, 2 * fp_words(curve)))
    out = np.zeros((4, 4), dtype=np.uint64)
    st = lib.b200_plonk_derive_challenges(_ptr(vk), _ptr(qcp), _ptr(pw) if pw.shape[0] else None, pw.shape[0], C.byref(proof), _ptr(out))
    if st != 0:
        raise B200Error(f"b200_plonk_derive_challenges failed with status {st}")
    return out


# ---- handles ------------------------------------------------------------------------------
def _dev_ptr(x):
    """Device address of a DeviceVector, a torch CUDA tensor, or a raw integer."""
    if hasattr(x, "data_ptr"):
        return C.c_void_p(x.data_ptr())
    if isinstance(x, np.ndarray):          # host memory: the library copies with cudaMemcpyDefault
        return C.c_void_p(x.ctypes.data)
    p = getattr(x, "ptr", x)
    return p if isinstance(p, C.c_void_p) else C.c_void_p(int(p))


DIST_PACK, DIST_DIF_LOCAL, DIST_DIF_CROSS, DIST_DIT_LOCAL, DIST_DIT_CROSS = range(5)
SCALE_NONE, SCALE_H_A, SCALE_H_BC, SCALE_COSETINV_BR, SCALE_NINV = range(5)


class DeviceVector:
    """A raw device buffer (iciclegnark.OnDeviceData analogue, icicle.go:228,248)."""

    def __init__(self, ctx: "Context", nbytes: int):
        self.ctx, self.nbytes = ctx, nbytes
        p = C.c_void_p()
        ctx._check(ctx.lib.b200_dev_alloc(ctx.h, nbytes, C.byref(p)))
        self.ptr = p

    def upload(self, host: np.ndarray):
        host = np.ascontiguousarray(host)
        assert host.nbytes <= self.nbytes
        self.ctx._check(self.ctx.lib.b200_h2d(self.ctx.h, self.ptr, _ptr(host), host.nbytes))
        return self

    def download(self, shape, dtype=np.uint64) -> np.ndarray:
        out = np.empty(shape, dtype=dtype)
        assert out.nbytes <= self.nbytes
        self.ctx._check(self.ctx.lib.b200_d2h(self.ctx.h, _ptr(out), self.ptr, out.nbytes))
        return out

    def free(self):
        if self.ptr:
            self.ctx.lib.b200_dev_free(self.ctx.h, self.ptr)
            self.ptr = None


class Bases:
    """Resident MSM base set (CopyPointsToDevice, icicle.go:88-126)."""

    def __init__(self, ctx, handle, n, group):
        self.ctx, self.h, self.n, self.group = ctx, handle, n, group

    def info(self):
        o = _MsmSetInfo()
        self.ctx._check(self.ctx.lib.b200_bases_info(self.ctx.h, self.h, C.byref(o)))
        return o.as_dict()

    def free(self):
        if self.h:
            self.ctx.lib.b200_bases_free(self.ctx.h, self.h)
            self.h = None


class Groth16ProvingKey:
    """Device-resident groth16_bn254.ProvingKey (setup.go:35-58; icicle/provingkey.go:10-28)."""

    def __init__(self, ctx, handle, keep):
        self.ctx, self.h, self._keep = ctx, handle, keep

    def info(self):
        """{'a','b','k','z','g2b'} -> layout and Pippenger plan of each resident base set."""
        arr = (_MsmSetInfo * 5)()
        self.ctx._check(self.ctx.lib.b200_groth16_pk_info(self.ctx.h, self.h, arr))
        return {k: arr[j].as_dict() for j, k in enumerate(("a", "b", "k", "z", "g2b"))}

    def wire_range(self):
        """[lo, hi) of W this shard reads; only that slice is copied per proof."""
        lo, hi = C.c_size_t(), C.c_size_t()
        self.ctx._check(self.ctx.lib.b200_groth16_pk_wire_range(self.ctx.h, self.h, C.byref(lo), C.byref(hi)))
        return int(lo.value), int(hi.value)

    def free(self):
        if self.h:
            self.ctx.lib.b200_groth16_pk_free(self.ctx.h, self.h)
            self.h = None


class PlonkProvingKey:
    """Device-resident plonk_bn254.ProvingKey (setup.go:79-107): trace, coset evaluations, both SRS."""

    def __init__(self, ctx, handle, nb_public, nb_bsb22=0, shard_index=0, shard_count=1):
        self.ctx, self.h, self.nb_public, self.nb_bsb22 = ctx, handle, nb_public, nb_bsb22
        self.shard_index, self.shard_count = shard_index, max(1, shard_count)

    def free(self):
        if self.h:
            self.ctx.lib.b200_plonk_pk_free(self.ctx.h, self.h)
            self.h = None


class PlonkSession:
    """One proof as phases (b200_plonk_session_*): each phase returns this rank's PARTIAL commitments ((k, 12) G1Jac)
    and takes the TOTALS of the previous phase's partials over all ranks.  Sequence: begin (3 partials) ->
    step(1, totals3) -> step(2, totals1) -> [all-gather numerator blocks] -> step(3) -> step(4, totals3) ->
    step(5, totals2) -> finish(totals1)."""
    N_OUT = {1: 1, 2: 0, 3: 3, 4: 2, 5: 1}

    def __init__(self, ctx, handle, partials):
        self.ctx, self.h, self.partials = ctx, handle, partials

    def step(self, phase: int, totals=None) -> np.ndarray:
        jw = 3 * self.ctx.fw
        t = None if totals is None else _u64(totals, (-1, jw))
        out = np.zeros((max(1, self.N_OUT[phase]), jw), dtype=np.uint64)
        self.ctx._check(self.ctx.lib.b200_plonk_session_step(self.ctx.h, self.h, phase, _ptr(t), _ptr(out) if self.N_OUT[phase] else None))
        self.partials = out[:self.N_OUT[phase]]
        return self.partials

    def numerator(self):
        """(device address, elements per block, log2 of the block count) of the quotient numerator buffer."""
        p, be, lb = C.c_void_p(), C.c_size_t(), C.c_uint()
        self.ctx._check(self.ctx.lib.b200_plonk_session_numerator(self.ctx.h, self.h, C.byref(p), C.byref(be), C.byref(lb)))
        return int(p.value), int(be.value), int(lb.value)

    def finish(self, totals):
        proof, tm = _PLONK_PROOF[self.ctx.curve](), _PlonkTimings()
        t = _u64(totals, (-1, 3 * self.ctx.fw))
        h, self.h = self.h, None
        self.ctx._check(self.ctx.lib.b200_plonk_session_finish(self.ctx.h, h, _ptr(t), C.byref(proof), C.byref(tm)))
        return proof, {k: getattr(tm, k) for k, _ in tm._fields_}

    def abandon(self):
        if self.h:
            self.ctx.lib.b200_plonk_session_finish(self.ctx.h, self.h, None, None, None)
            self.h = None


class Context:
    """One b200_ctx = one device. Not re-entrant."""

    def __init__(self, device: int = 0, curve: int = BN254):
        self.lib = load_library(curve)
        self.curve = curve
        self.fw = fp_words(curve)        # u64 words per Fp element: 4 (BN254) or 6 (BLS12-377)
        h = C.c_void_p()
        st = self.lib.b200_init(device, C.byref(h))
        if st != 0:
            msg = self.lib.b200_last_error(None)
            raise B200Error(f"b200_init(device={device}) failed ({st}): {msg.decode() if msg else ''}; "
                            "a CUDA device is required, there is no CPU fallback")
        self.h = h
        self.device = device

    def close(self):
        if self.h:
            self.lib.b200_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, st: int):
        if st != 0:
            msg = self.lib.b200_last_error(self.h)
            raise B200Error(f"status {st}: {msg.decode() if msg else ''}")

    def sync(self):
        self._check(self.lib.b200_sync(self.h))

    def domain_cache_trim(self) -> int:
        """Free the cached NTT domains no resident key refers to; returns the bytes released."""
        n = C.c_size_t()
        self._check(self.lib.b200_domain_cache_trim(self.h, C.byref(n)))
        return int(n.value)

    def use_stream(self, cuda_stream: int | None):
        """Order all further work of this context on the caller's CUDA stream (an integer
        cudaStream_t, e.g. torch.cuda.current_stream().cuda_stream); None: the context's own."""
        if cuda_stream is not None and int(cuda_stream) == 0:
            raise ValueError("the legacy default stream (handle 0) cannot be shared: make a torch.cuda.Stream current first")
        self._check(self.lib.b200_use_stream(self.h, None if cuda_stream is None else C.c_void_p(int(cuda_stream))))

    def set_option(self, key: str, value: int):
        self._check(self.lib.b200_set_option(self.h, key.encode(), int(value)))

    # -- element-wise / roofline
    def fr_vec_mul(self, a, b):
        a, b = _u64(a, (-1, 4)), _u64(b, (-1, 4))
        out = np.empty_like(a)
        self._check(self.lib.b200_fr_vec_mul(self.h, self.curve, _ptr(a), _ptr(b), _ptr(out), a.shape[0]))
        return out

    def fp_vec_mul(self, a, b):
        a, b = _u64(a, (-1, self.fw)), _u64(b, (-1, self.fw))
        out = np.empty_like(a)
        self._check(self.lib.b200_fp_vec_mul(self.h, self.curve, _ptr(a), _ptr(b), _ptr(out), a.shape[0]))
        return out

    def imad_peak(self):
        m, p, f = C.c_double(), C.c_double(), C.c_double()
        self._check(self.lib.b200_imad_peak(self.h, C.byref(m), C.byref(p), C.byref(f)))
        return {"mac32_per_s": m.value, "imad32_per_s": p.value, "sm_mhz": f.value}

    # -- NTT (fft.Domain.FFT / FFTInverse)
    def ntt(self, data, log_n, inverse, decimation, on_coset, generator, coset_gen):
        out = _u64(data, (-1, 4)).copy()
        assert out.shape[0] == 1 << log_n
        self._check(self.lib.b200_ntt(self.h, self.curve, _ptr(out), log_n, int(inverse), int(decimation),
                                      int(on_coset), _ptr(_u64(generator)), _ptr(_u64(coset_gen))))
        return out

    def ntt_dev(self, vec: DeviceVector, log_n, inverse, decimation, on_coset, generator, coset_gen) -> float:
        ms = C.c_float()
        self._check(self.lib.b200_ntt_dev(self.h, self.curve, vec.ptr, log_n, int(inverse), int(decimation),
                                          int(on_coset), _ptr(_u64(generator)), _ptr(_u64(coset_gen)), C.byref(ms)))
        return ms.value

    def compute_h(self, a, b, c, log_n, generator, coset_gen):
        a, b, c = (_u64(x, (-1, 4)) for x in (a, b, c))
        out = np.empty((1 << log_n, 4), dtype=np.uint64)
        self._check(self.lib.b200_groth16_compute_h(self.h, self.curve, _ptr(a), _ptr(b), _ptr(c), a.shape[0], log_n,
                                                    _ptr(_u64(generator)), _ptr(_u64(coset_gen)), _ptr(out)))
        return out

    # -- MSM (G1Jac.MultiExp / G2Jac.MultiExp)
    def msm_g1(self, scalars, bases):
        s, b = _u64(scalars, (-1, 4)), _u64(bases, (-1, 2 * self.fw))
        out = np.zeros(3 * self.fw, dtype=np.uint64)
        self._check(self.lib.b200_msm_g1(self.h, self.curve, _ptr(s), _ptr(b), s.shape[0], _ptr(out)))
        return out

    def msm_g2(self, scalars, bases):
        s, b = _u64(scalars, (-1, 4)), _u64(bases, (-1, 4 * self.fw))
        out = np.zeros(6 * self.fw, dtype=np.uint64)
        self._check(self.lib.b200_msm_g2(self.h, self.curve, _ptr(s), _ptr(b), s.shape[0], _ptr(out)))
        return out

    def upload_bases(self, bases, group: int) -> Bases:
        b = _u64(bases, (-1, (2 if group == 1 else 4) * self.fw))
        h = C.c_void_p()
        self._check(self.lib.b200_bases_upload(self.h, self.curve, group, _ptr(b), b.shape[0], C.byref(h)))
        return Bases(self, h, b.shape[0], group)

    def _raw(self, raw, n, group):
        a = np.frombuffer(raw, dtype=np.uint8)
        want = n * (2 if group == 1 else 4) * 8 * self.fw
        if a.shape[0] != want:
            raise ValueError(f"{n} raw group-{group} points are {want} bytes, got {a.shape[0]}")
        return a

    def points_from_raw(self, raw, n: int, group: int) -> np.ndarray:
        """gnark-crypto RawEncoding point bytes (a section of a WriteRawTo file) -> (n, 2|4 x fw) affine limbs."""
        a = self._raw(raw, n, group)
        out = np.zeros((n, (2 if group == 1 else 4) * self.fw), dtype=np.uint64)
        self._check(self.lib.b200_points_from_raw(self.h, self.curve, group, _ptr(a) if n else None, n, _ptr(out) if n else None))
        return out

    def upload_bases_raw(self, raw, n: int, group: int) -> Bases:
        a = self._raw(raw, n, group)
        h = C.c_void_p()
        self._check(self.lib.b200_bases_upload_raw(self.h, self.curve, group, _ptr(a) if n else None, n, C.byref(h)))
        return Bases(self, h, n, group)

    def msm_resident(self, bases: Bases, scalars):
        s = _u64(scalars, (-1, 4))
        out = np.zeros((3 if bases.group == 1 else 6) * self.fw, dtype=np.uint64)
        self._check(self.lib.b200_msm_resident(self.h, bases.h, _ptr(s), s.shape[0], _ptr(out)))
        return out

    def msm_dev(self, bases: Bases, scalars: DeviceVector, n: int):
        out = np.zeros((3 if bases.group == 1 else 6) * self.fw, dtype=np.uint64)
        ms = C.c_float()
        self._check(self.lib.b200_msm_dev(self.h, bases.h, scalars.ptr, n, _ptr(out), C.byref(ms)))
        return out, ms.value

    def batch_scalar_mul(self, base, scalars, group: int):
        s = _u64(scalars, (-1, 4))
        w = (2 if group == 1 else 4) * self.fw
        out = np.zeros((s.shape[0], w), dtype=np.uint64)
        self._check(self.lib.b200_batch_scalar_mul(self.h, self.curve, group, _ptr(_u64(base)), _ptr(s), s.shape[0], _ptr(out)))
        return out

    def kzg_to_lagrange_g1(self, srs_g1, generator):
        """kzg.ToLagrangeG1 (plonk setup.go:134): [L_i(tau)]G1 from the monomial SRS [tau^j]G1, j < n = 2^k, without tau."""
        pts = _u64(srs_g1, (-1, 2 * self.fw))
        n = pts.shape[0]
        if n == 0 or n & (n - 1):
            raise B200Error("ToLagrangeG1 takes a power-of-two number of points")
        out = np.zeros_like(pts)
        self._check(self.lib.b200_kzg_to_lagrange_g1(self.h, self.curve, _ptr(pts), n.bit_length() - 1, _ptr(_u64(generator)), _ptr(out)))
        return out

    # -- Groth16
    def groth16_pk_upload(self, *, log_n, generator, coset_gen, g1_a, g1_b, g1_k, g1_z, g2_b, infinity_a,
                          infinity_b, nb_public, shard_index=0, shard_count=1, presliced=False,
                          full_counts=None, raw_points=False, _invoke=None) -> Groth16ProvingKey:
        """presliced=True: the base arrays hold only this shard's point range; full_counts =
        dict(a=, b=, k=, z=) then gives the full lengths of pk.G1.A/B/K/Z."""
        if raw_points:      # bytes as ProvingKey.WriteRawTo stores them; converted on the device
            pts = lambda x, w: np.frombuffer(x, dtype=np.uint8).reshape(-1, 8 * w * self.fw)
        else:
            pts = lambda x, w: _u64(x, (-1, w * self.fw))
        keep = dict(gen=_u64(generator), cg=_u64(coset_gen), a=pts(g1_a, 2), b=pts(g1_b, 2),
                    k=pts(g1_k, 2), z=pts(g1_z, 2), g2b=pts(g2_b, 4),
                    ia=np.ascontiguousarray(infinity_a, dtype=np.uint8), ib=np.ascontiguousarray(infinity_b, dtype=np.uint8))
        d = _PkDesc()
        d.log_n = log_n
        d.generator, d.coset_gen = _ptr(keep["gen"]), _ptr(keep["cg"])
        d.g1_a, d.n_a = _ptr(keep["a"]), keep["a"].shape[0]
        d.g1_b, d.n_b = _ptr(keep["b"]), keep["b"].shape[0]
        d.g1_k, d.n_k = _ptr(keep["k"]), keep["k"].shape[0]
        d.g1_z, d.n_z = _ptr(keep["z"]), keep["z"].shape[0]
        d.g2_b, d.n_g2b = _ptr(keep["g2b"]), keep["g2b"].shape[0]
        d.infinity_a, d.infinity_b = _ptr(keep["ia"]), _ptr(keep["ib"])
        d.n_wires, d.nb_public = keep["ia"].shape[0], nb_public
        d.shard_index, d.shard_count, d.presliced = shard_index, shard_count, int(bool(presliced))
        d.raw_points = int(bool(raw_points))
        if presliced:
            d.n_a, d.n_b, d.n_k, d.n_z = (full_counts[k] for k in ("a", "b", "k", "z"))
            d.n_g2b = d.n_b
        h = C.c_void_p()
        if _invoke is not None:
            return _invoke(d, h)
        self._check(self.lib.b200_groth16_pk_upload(self.h, self.curve, C.byref(d), C.byref(h)))
        return Groth16ProvingKey(self, h, None)

    def groth16_setup(self, *, log_n, generator, coset_gen, nb_wires, nb_public, matrices, coeffs, toxic, g1_gen, g2_gen,
                      wire_class=None, nb_commitments=0, shard_index=0, shard_count=1, want_points=False, _invoke=None):
        """b200_groth16_setup: R1CS + toxic waste -> resident proving key (setup.go:85-337).  matrices = three
        (row_ptr u64[nb_constraints + 1], terms u32[nnz, 2] = (coefficient id, wire id)) for L, R, O; toxic = dict of
        Montgomery Fr limbs tau, alpha, beta, gamma, delta.  Returns (Groth16ProvingKey, dict): vk-side points,
        infinity bitmaps, vk.G1.K, the Pedersen bases, and (want_points) the proving key's point arrays."""
        Desc, Out = _SETUP[self.curve]
        fw = self.fw
        d, o = Desc(), Out()
        keep = [_u64(generator), _u64(coset_gen), _u64(coeffs, (-1, 4)), _u64(g1_gen), _u64(g2_gen)]
        d.log_n, d.generator, d.coset_gen = log_n, _ptr(keep[0]), _ptr(keep[1])
        nb_constraints = len(matrices[0][0]) - 1
        d.nb_constraints, d.nb_wires, d.nb_public = nb_constraints, nb_wires, nb_public
        for name, (rp, tm) in zip(("l", "r", "o"), matrices):
            rp = np.ascontiguousarray(rp, dtype=np.uint64)
            tm = np.ascontiguousarray(tm, dtype=np.uint32).reshape(-1, 2)
            assert rp.shape[0] == nb_constraints + 1 and int(rp[-1]) == tm.shape[0]
            keep += [rp, tm]
            m = getattr(d, name)
            m.row_ptr, m.terms = _ptr(rp), _ptr(tm)
        d.coeffs, d.nb_coeffs = _ptr(keep[2]), keep[2].shape[0]
        wc = None
        if wire_class is not None:
            wc = np.ascontiguousarray(wire_class, dtype=np.uint8)
            assert wc.shape[0] == nb_wires
            d.wire_class = _ptr(wc)
        d.nb_commitments = nb_commitments
        for k in ("tau", "alpha", "beta", "gamma", "delta"):
            getattr(d, k)[:] = [int(v) for v in _u64(toxic[k], (4,))]
        d.g1_gen, d.g2_gen = _ptr(keep[3]), _ptr(keep[4])
        d.shard_index, d.shard_count = shard_index, shard_count
        res = dict(infinity_a=np.zeros(nb_wires, dtype=np.uint8), infinity_b=np.zeros(nb_wires, dtype=np.uint8),
                   vk_k=np.zeros((nb_public + nb_commitments, 2 * fw), dtype=np.uint64))
        o.infinity_a, o.infinity_b, o.vk_k = _ptr(res["infinity_a"]), _ptr(res["infinity_b"]), _ptr(res["vk_k"])
        ck = []
        if nb_commitments:
            for j in range(nb_commitments):
                ck.append(np.zeros((int((wc == 2 + j).sum()), 2 * fw), dtype=np.uint64))
            arr = (C.c_void_p * nb_commitments)(*[c.ctypes.data for c in ck])
            keep.append(arr)
            o.ck_k = C.cast(arr, C.c_void_p)
        res["ck_k"] = ck
        if want_points:
            n = 1 << log_n
            res.update(g1_a=np.zeros((nb_wires, 2 * fw), dtype=np.uint64), g1_b=np.zeros((nb_wires, 2 * fw), dtype=np.uint64),
                       g1_k=np.zeros((nb_wires, 2 * fw), dtype=np.uint64), g1_z=np.zeros((max(n - 1, 1), 2 * fw), dtype=np.uint64),
                       g2_b=np.zeros((nb_wires, 4 * fw), dtype=np.uint64))
            for k in ("g1_a", "g1_b", "g1_k", "g1_z", "g2_b"):
                setattr(o, k, _ptr(res[k]))
        h = C.c_void_p()
        if _invoke is not None:         # the group's entry point (Group.groth16_setup) on the same descriptor
            wrap = _invoke(d, h, o)
        else:
            self._check(self.lib.b200_groth16_setup(self.h, self.curve, C.byref(d), C.byref(h), C.byref(o)))
            wrap = Groth16ProvingKey(self, h, None)
        for k in ("g1_alpha", "g1_beta", "g1_delta", "g2_beta", "g2_delta", "g2_gamma"):
            res[k] = np.array(getattr(o, k), dtype=np.uint64)
        res.update(n_a=int(o.n_a), n_b=int(o.n_b), n_k=int(o.n_k), n_z=int(o.n_z), n_vk_k=int(o.n_vk_k))
        if want_points:
            res["g1_a"], res["g1_b"], res["g2_b"] = res["g1_a"][:o.n_a], res["g1_b"][:o.n_b], res["g2_b"][:o.n_b]
            res["g1_k"], res["g1_z"] = res["g1_k"][:o.n_k], res["g1_z"][:o.n_z]
        return wrap, res

    def groth16_prove_core(self, pk: Groth16ProvingKey, a, b, c, w, k_remove=None):
        """computeH + the five MSMs (prove.go:139-301). Returns (Groth16MsmOut, timings dict)."""
        a, b, c, w = (_u64(x, (-1, 4)) for x in (a, b, c, w))
        out, tm = _MSM_OUT[self.curve](), _Timings()
        kr = None if k_remove is None or len(k_remove) == 0 else np.ascontiguousarray(k_remove, dtype=np.uint32)
        self._check(self.lib.b200_groth16_prove_core(self.h, pk.h, _ptr(a), _ptr(b), _ptr(c), a.shape[0], _ptr(w),
                                                     _ptr(kr), 0 if kr is None else kr.shape[0],
                                                     C.byref(out), C.byref(tm)))
        return out, {k: getattr(tm, k) for k, _ in _Timings._fields_}

    def groth16_prove_core_dev(self, pk: Groth16ProvingKey, a: DeviceVector, b: DeviceVector, c: DeviceVector,
                               n_constraints: int, w: DeviceVector):
        """Same with a, b, c, W already resident on the device (copied D2D into the workspace)."""
        out, tm = _MSM_OUT[self.curve](), _Timings()
        self._check(self.lib.b200_groth16_prove_core_dev(self.h, pk.h, a.ptr, b.ptr, c.ptr, n_constraints, w.ptr,
                                                         C.byref(out), C.byref(tm)))
        return out, {k: getattr(tm, k) for k, _ in _Timings._fields_}

    def groth16_msms_dev(self, pk: Groth16ProvingKey, w, h):
        """The five MSMs with this shard's range of h supplied by the distributed computeH.
        w, h: DeviceVector or any object with an integer/ctypes `.ptr` (e.g. a torch tensor's data_ptr)."""
        out, tm = _MSM_OUT[self.curve](), _Timings()
        self._check(self.lib.b200_groth16_msms_dev(self.h, pk.h, _dev_ptr(w), _dev_ptr(h), C.byref(out), C.byref(tm)))
        return out, {k: getattr(tm, k) for k, _ in _Timings._fields_}

    # -- distributed computeH phases (include/b200prover.h: b200_dist_phase / b200_dist_scale)
    def ntt_dist_phase(self, log_n, generator, coset_gen, data, tmp, log_ranks, rank, phase, inverse, scale=0):
        self._check(self.lib.b200_ntt_dist_phase(self.h, self.curve, log_n, _ptr(_u64(generator)), _ptr(_u64(coset_gen)),
                                                 _dev_ptr(data), _dev_ptr(tmp), log_ranks, rank, int(phase),
                                                 int(inverse), int(scale)))

    def fr_mul_sub_dev(self, a, b, c, n):
        self._check(self.lib.b200_fr_mul_sub_dev(self.h, self.curve, _dev_ptr(a), _dev_ptr(b), _dev_ptr(c), n))

    # -- PlonK building blocks on device vectors (plonk/bn254/prove.go)
    def fr_batch_inverse_dev(self, src, dst, n):
        self._check(self.lib.b200_fr_batch_inverse_dev(self.h, self.curve, _dev_ptr(src), _dev_ptr(dst), n))

    def fr_prefix_product_dev(self, src, dst, n):
        self._check(self.lib.b200_fr_prefix_product_dev(self.h, self.curve, _dev_ptr(src), _dev_ptr(dst), n))

    def fr_prefix_product_static(self, src, dst, n):
        """b200_fr_prefix_product_static: no allocation, scratch in a static __device__ array (n <= 2^22); enqueue only."""
        self._check(self.lib.b200_fr_prefix_product_static(self.h, self.curve, _dev_ptr(src), _dev_ptr(dst), n))

    def fr_poly_eval_dev(self, coeffs, n, x) -> np.ndarray:
        out = np.empty(4, dtype=np.uint64)
        self._check(self.lib.b200_fr_poly_eval_dev(self.h, self.curve, _dev_ptr(coeffs), n, _ptr(_u64(x)), _ptr(out)))
        return out

    def fr_poly_div_linear_dev(self, coeffs, n, z, quotient):
        self._check(self.lib.b200_fr_poly_div_linear_dev(self.h, self.curve, _dev_ptr(coeffs), n, _ptr(_u64(z)), _dev_ptr(quotient)))

    def plonk_divide_by_xn_minus_one_dev(self, vec, log_big, log_small, generator_big, coset_gen):
        self._check(self.lib.b200_plonk_divide_by_xn_minus_one_dev(self.h, self.curve, _dev_ptr(vec), log_big, log_small,
                                                                   _ptr(_u64(generator_big)), _ptr(_u64(coset_gen))))

    def plonk_build_ratio_dev(self, l, r, o, perm, beta, gamma, log_n, generator, coset_gen, z_out):
        """Z of the copy constraints (prove.go:610-621); perm: device int64[3 n]."""
        self._check(self.lib.b200_plonk_build_ratio_dev(self.h, self.curve, _dev_ptr(l), _dev_ptr(r), _dev_ptr(o), _dev_ptr(perm),
                                                        _ptr(_u64(beta)), _ptr(_u64(gamma)), log_n, _ptr(_u64(generator)),
                                                        _ptr(_u64(coset_gen)), _dev_ptr(z_out)))

    def plonk_quotient_rows_dev(self, x, qci, log_n, log_big, coset_index, generator_small, beta, gamma, alpha, cs,
                                bl, br, bo, bz, cres):
        """One coset of computeNumerator's row evaluation (prove.go:837-1044).  x: 15 device vectors
        indexed by prove.go's ids (x[4] may be None); qci: list of device vectors [Qcp_0, Pi_0, ...]."""
        d = _PlonkRowsDesc()
        for k in range(15):
            d.x[k] = None if x[k] is None else _dev_ptr(x[k]).value
        for k, v in enumerate(qci):
            d.qci[k] = _dev_ptr(v).value
        d.nb_bsb, d.log_n, d.log_big, d.coset_index = len(qci) // 2, log_n, log_big, coset_index
        gs = _u64(generator_small)
        d.generator_small = gs.ctypes.data
        for name, val in (("beta", beta), ("gamma", gamma), ("alpha", alpha), ("cs", cs)):
            getattr(d, name)[:] = [int(v) for v in _u64(val).reshape(4)]
        for name, val in (("bl", bl), ("br", br), ("bo", bo), ("bz", bz)):
            arr = _u64(val).reshape(-1, 4)
            for j in range(arr.shape[0]):
                getattr(d, name)[j][:] = [int(v) for v in arr[j]]
        self._check(self.lib.b200_plonk_quotient_rows_dev(self.h, self.curve, C.byref(d), _dev_ptr(cres)))

    def kzg_commit_dev(self, srs: "Bases", coeffs, n) -> np.ndarray:
        out = np.empty(2 * self.fw, dtype=np.uint64)
        self._check(self.lib.b200_kzg_commit_dev(self.h, srs.h, _dev_ptr(coeffs), n, _ptr(out)))
        return out

    # -- PlonK prover
    def plonk_pk_upload(self, *, log_n, log_big, generator, generator_big, coset_gen, ql, qr, qm, qo, qk, s1, s2, s3, perm,
                        kzg_g1, kzg_lagrange_g1, vk_commitments, nb_public, qcp=(), vk_qcp=None, commitment_index=(),
                        shard_index=0, shard_count=1, raw_file=False, _invoke=None) -> PlonkProvingKey:
        """qcp: list of pk.trace.Qcp columns (canonical), vk_qcp: their commitments, commitment_index: CommitmentInfo[i].CommitmentIndex.
        raw_file=True: the arrays are byte sections of a ProvingKey.WriteRawTo file (plonk_pk_format.py), converted on the device."""
        n = 1 << log_n
        pb = 16 * self.fw      # bytes of a raw / in-memory affine G1 point
        if raw_file:
            def sec(x, nbytes):
                a = np.frombuffer(x, dtype=np.uint8)
                if a.shape[0] != nbytes:
                    raise B200Error(f"key file section holds {a.shape[0]} bytes, expected {nbytes}")
                return a
            keep = {k: sec(v, 32 * n) for k, v in dict(ql=ql, qr=qr, qm=qm, qo=qo, qk=qk, s1=s1, s2=s2, s3=s3).items()}
            keep["perm"] = sec(perm, 24 * n)
            keep["kzg_g1"], keep["kzg_lagrange_g1"], keep["vk_commitments"] = sec(kzg_g1, (n + 3) * pb), sec(kzg_lagrange_g1, n * pb), sec(vk_commitments, 8 * pb)
        else:
            keep = {k: _u64(v, (n, 4)) for k, v in dict(ql=ql, qr=qr, qm=qm, qo=qo, qk=qk, s1=s1, s2=s2, s3=s3).items()}
            keep["perm"] = np.ascontiguousarray(perm, dtype=np.int64).reshape(3 * n)
            keep["kzg_g1"] = _u64(kzg_g1, (n + 3, 2 * self.fw))
            keep["kzg_lagrange_g1"] = _u64(kzg_lagrange_g1, (n, 2 * self.fw))
            keep["vk_commitments"] = _u64(vk_commitments, (8, 2 * self.fw))
        keep["generator"], keep["generator_big"], keep["coset_gen"] = _u64(generator), _u64(generator_big), _u64(coset_gen)
        d = _PlonkPkDesc()
        d.log_n, d.log_big, d.nb_public, d.nb_bsb22 = log_n, log_big, nb_public, len(qcp)
        d.shard_index, d.shard_count, d.raw_file = shard_index, shard_count, int(bool(raw_file))
        for k, v in keep.items():
            setattr(d, k, v.ctypes.data)
        if len(qcp):
            cols = [np.frombuffer(q, dtype=np.uint8) if raw_file else _u64(q, (n, 4)) for q in qcp]
            ptrs = (C.c_void_p * len(cols))(*[c.ctypes.data for c in cols])
            vq = np.frombuffer(vk_qcp, dtype=np.uint8) if raw_file else _u64(vk_qcp, (len(cols), 2 * self.fw))
            ci = np.ascontiguousarray(commitment_index, dtype=np.uint64)
            d.qcp, d.vk_qcp, d.commitment_index = C.addressof(ptrs), vq.ctypes.data, ci.ctypes.data
        h = C.c_void_p()
        if _invoke is not None:
            key = _invoke(d, h, nb_public, len(qcp))
            key.n = n
            return key
        self._check(self.lib.b200_plonk_pk_upload(self.h, self.curve, C.byref(d), C.byref(h)))
        key = PlonkProvingKey(self, h, nb_public, len(qcp), shard_index, shard_count)
        key.n = n          # the C ABI takes no lengths: the binding checks them (out-of-bounds host reads otherwise)
        return key

    def _plonk_bsb(self, pk, bsb):
        if not pk.nb_bsb22:
            return None, None
        cols = [_u64(c, (-1, 4)) for c in bsb["pi2"]]
        assert len(cols) == pk.nb_bsb22
        if any(c.shape[0] != getattr(pk, "n", c.shape[0]) for c in cols):
            raise B200Error("every Pi2 column must hold n rows")
        ptrs = (C.c_void_p * len(cols))(*[c.ctypes.data for c in cols])
        com, val = _u64(bsb["commitments"], (len(cols), 2 * self.fw)), _u64(bsb["values"], (len(cols), 4))
        return _PlonkBsb22(C.addressof(ptrs), com.ctypes.data, val.ctypes.data), (cols, ptrs, com, val)

    @staticmethod
    def _plonk_check_lengths(pk, l, r, o, pw):
        n = getattr(pk, "n", l.shape[0])
        if not (l.shape[0] == r.shape[0] == o.shape[0] == n):
            raise B200Error(f"l, r, o must hold n = {n} rows each (got {l.shape[0]}, {r.shape[0]}, {o.shape[0]})")
        if pw.shape[0] != pk.nb_public:
            raise B200Error(f"public_witness must hold nb_public = {pk.nb_public} elements (got {pw.shape[0]})")

    def plonk_session(self, pk: PlonkProvingKey, l, r, o, public_witness, blinding, bsb=None) -> "PlonkSession":
        """Phase 0 of a (possibly sharded) proof; see PlonkSession."""
        l, r, o = _u64(l, (-1, 4)), _u64(r, (-1, 4)), _u64(o, (-1, 4))
        pw = _u64(public_witness, (-1, 4))
        self._plonk_check_lengths(pk, l, r, o, pw)
        bl = None if blinding is None else _u64(blinding, (9, 4))
        cb, keep = self._plonk_bsb(pk, bsb)
        h = C.c_void_p()
        partials = np.zeros((3, 3 * self.fw), dtype=np.uint64)
        self._check(self.lib.b200_plonk_session_begin(self.h, pk.h, _ptr(l), _ptr(r), _ptr(o), _ptr(pw) if pw.shape[0] else None,
                                                      _ptr(bl), None if cb is None else C.byref(cb), C.byref(h), _ptr(partials)))
        return PlonkSession(self, h, partials)

    def plonk_prove(self, pk: PlonkProvingKey, l, r, o, public_witness, blinding=None, bsb=None):
        """plonk_bn254.Prove after the solver (prove.go:116-180).  blinding: (9, 4) Montgomery coefficients
        bl0 bl1 br0 br1 bo0 bo1 bz0 bz1 bz2, or None for fresh randomness.  bsb (api.Commit circuits): dict with
        'pi2' (list of (n, 4) Lagrange columns), 'commitments' ((k, 8) affine), 'values' ((k, 4)) from the solver's
        bsb22Hint (prove.go:303-352).  Returns (PlonkProof, timings dict)."""
        l, r, o = _u64(l, (-1, 4)), _u64(r, (-1, 4)), _u64(o, (-1, 4))
        pw = _u64(public_witness, (-1, 4))
        self._plonk_check_lengths(pk, l, r, o, pw)
        bl = None if blinding is None else _u64(blinding, (9, 4))
        proof, tm = _PLONK_PROOF[self.curve](), _PlonkTimings()
        cb, keep = self._plonk_bsb(pk, bsb)
        self._check(self.lib.b200_plonk_prove_bsb22(self.h, pk.h, _ptr(l), _ptr(r), _ptr(o), _ptr(pw) if pw.shape[0] else None,
                                                    _ptr(bl), None if cb is None else C.byref(cb), C.byref(proof), C.byref(tm)))
        return proof, {k: getattr(tm, k) for k, _ in tm._fields_}

    # -- per-kernel-class device times (CUDA events on the launching stream)
    def profile_start(self):
        self._check(self.lib.b200_profile_start(self.h))

    def profile_read(self):
        p = _Profile()
        self._check(self.lib.b200_profile_read(self.h, C.byref(p)))
        out = {name: {"ms": p.ms[i], "launches": int(p.launches[i]), "units": p.units[i]}
               for i, name in enumerate(PROF_SLOTS) if not name.startswith("_")}
        out["kernel_launches"] = int(p.kernel_launches)
        return out


# ---------------------------------------------------------------------------------------------------------------
# the device group: multi-GPU in one process (include/b200prover.h: b200_group_*)
# ---------------------------------------------------------------------------------------------------------------
class _BorrowedContext(Context):
    """A rank's b200_ctx, owned by the group (never destroyed from here)."""

    def __init__(self, lib, curve, handle, device):
        self.lib, self.curve, self.fw, self.h, self.device = lib, curve, fp_words(curve), handle, device

    def close(self):
        self.h = None


class GroupGroth16ProvingKey:
    def __init__(self, group, handle):
        self.group, self.h = group, handle

    def info(self, rank=0):
        arr, dist = (_MsmSetInfo * 5)(), C.c_int()
        self.group._check(self.group.lib.b200_group_groth16_pk_info(self.group.h, self.h, rank, arr, C.byref(dist)))
        d = {k: arr[j].as_dict() for j, k in enumerate(("a", "b", "k", "z", "g2b"))}
        d["distributed_h"] = bool(dist.value)
        return d

    def free(self):
        if self.h:
            self.group.lib.b200_group_groth16_pk_free(self.group.h, self.h)
            self.h = None


class Group:
    """b200_group: one process, one call per proof over all the listed devices (a device may be listed several
    times: virtual ranks).  Not re-entrant."""

    def __init__(self, devices, curve: int = BN254):
        self.lib, self.curve, self.fw = load_library(curve), curve, fp_words(curve)
        self.devices = [int(d) for d in devices]
        arr = (C.c_int * len(self.devices))(*self.devices)
        h = C.c_void_p()
        st = self.lib.b200_group_init(arr, len(self.devices), C.byref(h))
        if st != 0:
            raise B200Error(f"b200_group_init({self.devices}) failed ({st}); CUDA devices are required, there is no CPU fallback")
        self.h = h
        self.size = int(self.lib.b200_group_size(h))

    def _check(self, st):
        if st != 0:
            msg = self.lib.b200_group_last_error(self.h)
            raise B200Error(f"status {st}: {msg.decode() if msg else ''}")

    def ctx(self, rank=0) -> Context:
        return _BorrowedContext(self.lib, self.curve, C.c_void_p(self.lib.b200_group_ctx(self.h, rank)), self.devices[rank])

    def set_option(self, key, value):
        self._check(self.lib.b200_group_set_option(self.h, key.encode(), int(value)))

    def close(self):
        if self.h:
            self.lib.b200_group_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def groth16_pk_upload(self, **kw) -> GroupGroth16ProvingKey:
        def invoke(d, h):
            self._check(self.lib.b200_group_groth16_pk_upload(self.h, self.curve, C.byref(d), C.byref(h)))
            return GroupGroth16ProvingKey(self, h)
        return Context.groth16_pk_upload(self.ctx(0), _invoke=invoke, **kw)

    def groth16_setup(self, **kw):
        def invoke(d, h, o):
            self._check(self.lib.b200_group_groth16_setup(self.h, self.curve, C.byref(d), C.byref(h), C.byref(o)))
            return GroupGroth16ProvingKey(self, h)
        return Context.groth16_setup(self.ctx(0), _invoke=invoke, **kw)

    def groth16_prove_core(self, pk: GroupGroth16ProvingKey, a, b, c, w, k_remove=None):
        """One proof over the whole group from host buffers: (combined Groth16MsmOut, per-stage max timings)."""
        a, b, c, w = (_u64(x, (-1, 4)) for x in (a, b, c, w))
        out, tm = _MSM_OUT[self.curve](), _Timings()
        kr = None if k_remove is None or len(k_remove) == 0 else np.ascontiguousarray(k_remove, dtype=np.uint32)
        self._check(self.lib.b200_group_groth16_prove_core(self.h, pk.h, _ptr(a), _ptr(b), _ptr(c), a.shape[0], _ptr(w),
                                                           _ptr(kr), 0 if kr is None else kr.shape[0], C.byref(out), C.byref(tm)))
        return out, {k: getattr(tm, k) for k, _ in _Timings._fields_}

    def groth16_prove_core_dev(self, pk: GroupGroth16ProvingKey, a_blocks, b_blocks, c_blocks, w_full):
        """Inputs resident on the devices: per rank its block of a, b, c and the full W (lists of DeviceVector)."""
        arrs = [(C.c_void_p * self.size)(*[_dev_ptr(x).value for x in v]) for v in (a_blocks, b_blocks, c_blocks, w_full)]
        out, tm = _MSM_OUT[self.curve](), _Timings()
        self._check(self.lib.b200_group_groth16_prove_core_dev(self.h, pk.h, *[C.cast(x, C.POINTER(C.c_void_p)) for x in arrs],
                                                               C.byref(out), C.byref(tm)))
        return out, {k: getattr(tm, k) for k, _ in _Timings._fields_}


class GroupPlonkProvingKey:
    def __init__(self, group, handle, nb_public, nb_bsb22):
        self.group, self.h, self.nb_public, self.nb_bsb22 = group, handle, nb_public, nb_bsb22

    def free(self):
        if self.h:
            self.group.lib.b200_group_plonk_pk_free(self.group.h, self.h)
            self.h = None


def _group_plonk_pk_upload(self, **kw) -> GroupPlonkProvingKey:
    """b200_group_plonk_pk_upload: every rank keeps its shard of the key (SRS point ranges, its cosets)."""
    def invoke(d, h, nb_public, nb_bsb):
        self._check(self.lib.b200_group_plonk_pk_upload(self.h, self.curve, C.byref(d), C.byref(h)))
        return GroupPlonkProvingKey(self, h, nb_public, nb_bsb)
    return Context.plonk_pk_upload(self.ctx(0), _invoke=invoke, **kw)


def _group_plonk_prove(self, pk: GroupPlonkProvingKey, l, r, o, public_witness, blinding=None, bsb=None):
    """b200_group_plonk_prove: one proof over the whole group; (PlonkProof, rank 0's timings)."""
    c0 = self.ctx(0)
    l, r, o = _u64(l, (-1, 4)), _u64(r, (-1, 4)), _u64(o, (-1, 4))
    pw = _u64(public_witness, (-1, 4))
    Context._plonk_check_lengths(pk, l, r, o, pw)
    bl = None if blinding is None else _u64(blinding, (9, 4))
    proof, tm = _PLONK_PROOF[self.curve](), _PlonkTimings()
    cb, keep = c0._plonk_bsb(pk, bsb)
    self._check(self.lib.b200_group_plonk_prove(self.h, pk.h, _ptr(l), _ptr(r), _ptr(o), _ptr(pw) if pw.shape[0] else None,
                                                _ptr(bl), None if cb is None else C.byref(cb), C.byref(proof), C.byref(tm)))
    return proof, {k: getattr(tm, k) for k, _ in tm._fields_}


Group.plonk_pk_upload = _group_plonk_pk_upload
Group.plonk_prove = _group_plonk_prove
