"""Host-side mirror of plonk_bn254.Setup for callers without Go (tests, benches): the stage BEFORE the hot
path, written over the C ABI so that a 2^22-gate key is built in seconds on the device.

Follows /root/reference/backend/plonk/bn254/setup.go:
  Setup                          :110-161     (domains, SRS slices, KzgLagrange, trace, permutation, commitments)
  BuildTrace                     :176-231     (the caller passes the Lagrange columns)
  buildPermutation               :304-358     (vectorised: a cyclic shift inside every variable's position list)
  computePermutationPolynomials  :360-420
  commitTrace                    :233-273
The SRS is test-style (powers of a known tau on G1, test/kzg_srs.go semantics) and is produced with the
fixed-base kernel (b200_batch_scalar_mul); KzgLagrange = [L_i(tau)]G1 where L(tau) = F^-1 (tau^j).
No oracle import: this is product-side host code."""
from __future__ import annotations

import numpy as np

from .api import Context, DeviceVector, DIF

R_MOD = 0x30644E72E131A029B85045B68181585D2833E84879B9709143E1F593F0000001
P_MOD = 0x30644E72E131A029B85045B68181585D97816A916871CA8D3C208C16D87CFD47
FR_ROOT_2_28 = 19103219067921713944291392827692070036145651957329286315305642004821462161904
FR_MULT_GEN = 5
MONT = 1 << 256
G1_GEN = np.frombuffer(((1 * MONT) % P_MOD).to_bytes(32, "little") + ((2 * MONT) % P_MOD).to_bytes(32, "little"), dtype=np.uint64).copy()


def fr_limbs(x: int) -> np.ndarray:
    return np.frombuffer(((x % R_MOD) * MONT % R_MOD).to_bytes(32, "little"), dtype=np.uint64).copy()


def to_mont_array(xs) -> np.ndarray:
    buf = b"".join(((x * MONT) % R_MOD).to_bytes(32, "little") for x in xs)
    return np.frombuffer(buf, dtype=np.uint64).reshape(-1, 4).copy()


def from_mont_array(a) -> list:
    rinv = pow(MONT, -1, R_MOD)
    b = np.ascontiguousarray(a, dtype=np.uint64).reshape(-1, 4).tobytes()
    return [(int.from_bytes(b[i:i + 32], "little") * rinv) % R_MOD for i in range(0, len(b), 32)]


def const_array(x: int, n: int) -> np.ndarray:
    return np.tile(fr_limbs(x), (n, 1))


def bitrev_indices(log_n: int) -> np.ndarray:
    idx = np.arange(1 << log_n, dtype=np.uint64)
    out = np.zeros_like(idx)
    for b in range(log_n):
        out |= ((idx >> np.uint64(b)) & np.uint64(1)) << np.uint64(log_n - 1 - b)
    return out.astype(np.int64)


def domain_generator(log_n: int) -> int:
    return pow(FR_ROOT_2_28, 1 << (28 - log_n), R_MOD)


class Domains:
    """pk.Domain[0], pk.Domain[1] (initDomains, setup.go:275-290)."""

    def __init__(self, size_system: int):
        self.log_n = max(1, (size_system - 1).bit_length())
        big = 8 * size_system if size_system < 6 else 4 * size_system
        self.log_big = (big - 1).bit_length()
        self.n, self.big = 1 << self.log_n, 1 << self.log_big
        self.gen, self.gen_big = domain_generator(self.log_n), domain_generator(self.log_big)


def build_permutation(lro: np.ndarray, nb_variables: int) -> np.ndarray:
    """buildPermutation (setup.go:304-358): position i of l||r||o is sent to the previous position holding
    the same variable, the first one to the last."""
    lro = np.asarray(lro, dtype=np.int64)
    order = np.argsort(lro, kind="stable")                 # positions grouped by variable, ascending inside
    v = lro[order]
    first = np.ones(len(v), dtype=bool)
    first[1:] = v[1:] != v[:-1]
    prev = np.empty_like(order)
    prev[1:] = order[:-1]
    # the first position of a group takes the group's last position
    starts = np.flatnonzero(first)
    ends = np.append(starts[1:], len(v)) - 1
    prev[starts] = order[ends]
    perm = np.empty_like(order)
    perm[order] = prev
    return perm


def _to_regular_canonical(ctx: Context, lagrange: np.ndarray, d: Domains, br: np.ndarray) -> np.ndarray:
    """ToCanonical(&Domain[0]).ToRegular(): FFTInverse(DIF) then the bit-reversal permutation."""
    out = ctx.ntt(lagrange, d.log_n, True, DIF, False, fr_limbs(d.gen), fr_limbs(FR_MULT_GEN))
    return out[br]


LAGRANGE_FROM_POINTS_MAX_LOG = 16      # b200_kzg_to_lagrange_g1: n log n / 2 scalar multiplications (0.06 s at 2^16, ~4.5 s at 2^22)


def setup(ctx: Context, d: Domains, trace_lagrange: dict, lro: np.ndarray, nb_variables: int, nb_public: int, tau: int) -> dict:
    """Returns the keyword arguments of Context.plonk_pk_upload plus 'vk' (the 8 commitments, (8, 8) affine)
    and 'perm'.  trace_lagrange: ql qr qm qo qk as (n, 4) Montgomery arrays in Lagrange / regular form."""
    n, log_n = d.n, d.log_n
    br = bitrev_indices(log_n)
    gen_l, cg_l = fr_limbs(d.gen), fr_limbs(FR_MULT_GEN)
    # SRS: tau^i (n + 3) and L_i(tau) (n)
    pw, acc = [], 1
    for _ in range(n + 3):
        pw.append(acc)
        acc = acc * tau % R_MOD
    mono = to_mont_array(pw)
    kzg = ctx.batch_scalar_mul(G1_GEN, mono, 1)
    if log_n <= LAGRANGE_FROM_POINTS_MAX_LOG:
        # setup.go:134 as the reference does it: from the SRS POINTS (kzg.ToLagrangeG1), tau not used
        kzg_lag = ctx.kzg_to_lagrange_g1(kzg[:n], gen_l)
    else:
        # large test keys: L(tau) = F^-1 (tau^j) on the scalar side, then the fixed-base kernel (needs tau; same points --
        # tests/test_gpu_kzg_lagrange.py compares the two routes)
        lag = ctx.ntt(mono[:n], log_n, True, DIF, False, gen_l, cg_l)[br]
        kzg_lag = ctx.batch_scalar_mul(G1_GEN, lag, 1)
    # support of the permutation: w^i | g w^i | g^2 w^i  (getSupportPermutation, setup.go:398-415)
    e1 = np.zeros((n, 4), dtype=np.uint64)
    e1[1] = fr_limbs(1)
    omega = ctx.ntt(e1, log_n, False, DIF, False, gen_l, cg_l)[br]           # evaluations of X = w^i
    support = np.concatenate([omega, ctx.fr_vec_mul(omega, const_array(FR_MULT_GEN, n)),
                              ctx.fr_vec_mul(omega, const_array(FR_MULT_GEN * FR_MULT_GEN, n))])
    perm = build_permutation(lro, nb_variables)
    s_lag = [support[perm[k * n:(k + 1) * n]] for k in range(3)]
    out = {"log_n": log_n, "log_big": d.log_big, "generator": gen_l, "generator_big": fr_limbs(d.gen_big), "coset_gen": cg_l,
           "perm": perm, "kzg_g1": kzg, "kzg_lagrange_g1": kzg_lag, "nb_public": nb_public}
    for k in ("ql", "qr", "qm", "qo", "qk"):
        out[k] = _to_regular_canonical(ctx, trace_lagrange[k], d, br)
    for k, v in zip(("s1", "s2", "s3"), s_lag):
        out[k] = _to_regular_canonical(ctx, v, d, br)
    # commitTrace (setup.go:233-273): S1 S2 S3 Ql Qr Qm Qo Qk against pk.Kzg
    srs = ctx.upload_bases(kzg, 1)
    vec = DeviceVector(ctx, n * 32)
    vk = np.zeros((8, 8), dtype=np.uint64)
    for j, k in enumerate(("s1", "s2", "s3", "ql", "qr", "qm", "qo", "qk")):
        vec.upload(out[k])
        vk[j] = ctx.kzg_commit_dev(srs, vec, n)
    vec.free()
    srs.free()
    out["vk_commitments"] = vk
    return out


def mul_chain(log_n: int, x: int = 3):
    """The SCS Mul chain of config 4 (backend/plonk/plonk_test.go:248-260 shape) filling 2^log_n rows exactly:
    one public input Y, nb_mul = 2^log_n - 2 constraints x <- x*x, one AssertIsEqual(x, Y).
    Returns (trace_lagrange, lro ids, nb_variables, nb_public, L, R, O, public) with field arrays in Montgomery form."""
    n = 1 << log_n
    nb_mul = n - 2
    vals = [0, x % R_MOD]
    cur = x % R_MOD
    for _ in range(nb_mul):
        cur = cur * cur % R_MOD
        vals.append(cur)
    vals[0] = cur
    v = to_mont_array(vals)                                  # [public Y | secret X | internal X_1..X_nb_mul]
    ids = np.arange(1, nb_mul + 1, dtype=np.int64)           # xa = xb of constraint j: wire 1 + j
    la = np.concatenate([[0], ids, [nb_mul + 1]])            # row 0: placeholder of public 0; last: assert
    lb = np.concatenate([[0], ids, [0]])
    lc = np.concatenate([[0], ids + 1, [0]])
    lro = np.concatenate([la, lb, lc]).astype(np.int64)
    one, minus = fr_limbs(1), fr_limbs(R_MOD - 1)
    z = np.zeros((n, 4), dtype=np.uint64)
    ql, qr, qm, qo, qk = z.copy(), z.copy(), z.copy(), z.copy(), z.copy()
    ql[0] = minus
    qm[1:n - 1] = one
    qo[1:n - 1] = minus
    ql[n - 1] = one
    qr[n - 1] = minus
    trace = {"ql": ql, "qr": qr, "qm": qm, "qo": qo, "qk": qk}
    L, Rr, O = v[la], v[lb], v[lc]
    return trace, lro, nb_mul + 2, 1, L, Rr, O, v[:1].copy()


# ---------------------------------------------------------------------------------------------
# plonk_bn254.Verify (verify.go:45-294) for proofs made against a test SRS whose tau is known: the two KZG
# checks e(C - [v]G1, G2) == e(H, [tau - z]G2) become C - [v]G1 == [tau - z]H in G1 (no pairing needed).
# Host-side big-int code, O(nb_public) field work + ~25 scalar multiplications.
# ---------------------------------------------------------------------------------------------
def _g1_add(p, q):
    if p is None:
        return q
    if q is None:
        return p
    (x1, y1), (x2, y2) = p, q
    if x1 == x2:
        if (y1 + y2) % P_MOD == 0:
            return None
        lam = 3 * x1 * x1 * pow(2 * y1, -1, P_MOD) % P_MOD
    else:
        lam = (y2 - y1) * pow(x2 - x1, -1, P_MOD) % P_MOD
    x3 = (lam * lam - x1 - x2) % P_MOD
    return x3, (lam * (x1 - x3) - y1) % P_MOD


def _g1_mul(p, k):
    k %= R_MOD
    acc = None
    while k:
        if k & 1:
            acc = _g1_add(acc, p)
        p = _g1_add(p, p)
        k >>= 1
    return acc


def _g1_neg(p):
    return None if p is None else (p[0], (-p[1]) % P_MOD)


def g1_from_limbs(a):
    """(8,) u64 affine Montgomery -> (x, y) ints or None for infinity."""
    rinv = pow(MONT, -1, P_MOD)
    b = np.ascontiguousarray(a, dtype=np.uint64).tobytes()
    x = int.from_bytes(b[:32], "little") * rinv % P_MOD
    y = int.from_bytes(b[32:], "little") * rinv % P_MOD
    return None if x == 0 and y == 0 else (x, y)


def _g1_raw(p) -> bytes:
    return bytes([0x40]) + bytes(63) if p is None else p[0].to_bytes(32, "big") + p[1].to_bytes(32, "big")


def _challenge(name: bytes, prev: bytes, *chunks: bytes) -> bytes:
    import hashlib
    return hashlib.sha256(name + prev + b"".join(chunks)).digest()


def verify_with_trapdoor(proof, vk_commitments, public, d: Domains, tau: int) -> bool:
    """proof: api.PlonkProof; vk_commitments: (8, 8) S1 S2 S3 Ql Qr Qm Qo Qk; public: list of ints."""
    pd = proof.as_dict()
    pt = g1_from_limbs
    lro, z_com, h_com = [pt(pd["lro"][k]) for k in range(3)], pt(pd["z"]), [pt(pd["h"][k]) for k in range(3)]
    batched_h, zs_h = pt(pd["batched_h"]), pt(pd["zshift_h"])
    claimed = from_mont_array(pd["claimed"])
    zu = from_mont_array(pd["zshift_value"].reshape(1, 4))[0]
    vk = [pt(v) for v in vk_commitments]
    s_com, ql, qr, qm, qo, qk = vk[:3], vk[3], vk[4], vk[5], vk[6], vk[7]
    fb = lambda x: (x % R_MOD).to_bytes(32, "big")
    g = _challenge(b"gamma", b"", *[_g1_raw(p) for p in vk], *[fb(w) for w in public], *[_g1_raw(p) for p in lro])
    b = _challenge(b"beta", g)
    a = _challenge(b"alpha", b, _g1_raw(z_com))
    zt = _challenge(b"zeta", a, *[_g1_raw(p) for p in h_com])
    gamma, beta, alpha, zeta = (int.from_bytes(v, "big") % R_MOD for v in (g, b, a, zt))
    n, R = d.n, R_MOD
    zzeta = (pow(zeta, n, R) - 1) % R
    lag1 = zzeta * pow((zeta - 1) % R, -1, R) % R * pow(n, -1, R) % R
    pi, lag, wp = 0, lag1, 1
    for i, w in enumerate(public):                                   # PI = sum L_i(zeta) w_i, verify.go:103-125
        pi = (pi + lag * w) % R
        if i + 1 != len(public):
            lag = lag * d.gen % R * ((zeta - wp) % R) % R
            wp = wp * d.gen % R
            lag = lag * pow((zeta - wp) % R, -1, R) % R
    cq, lin_z, l, r, o, s1, s2 = claimed
    t = (s1 * beta + l + gamma) * ((s2 * beta + r + gamma) % R) % R * ((o + gamma) % R) % R * alpha % R * zu % R
    asl = lag1 * alpha % R * alpha % R
    if (lin_z + pi + t - asl) % R * pow(zzeta, -1, R) % R != cq:
        return False
    zm2 = pow(zeta, n + 2, R)
    folded_h = _g1_add(_g1_mul(_g1_add(_g1_mul(h_com[2], zm2), h_com[1]), zm2), h_com[0])
    cs = FR_MULT_GEN
    u1 = zu * beta % R * ((beta * s1 + l + gamma) % R) % R * ((beta * s2 + r + gamma) % R) % R * alpha % R
    u2 = (beta * zeta + l + gamma) * ((beta * zeta * cs + r + gamma) % R) % R * ((beta * zeta * cs * cs + o + gamma) % R) % R
    u2 = ((R - u2) * alpha + asl) % R
    lin = None
    for p, s in zip((ql, qr, qm, qo, qk, s_com[2], z_com), (l, r, l * r % R, o, 1, u1, u2)):
        lin = _g1_add(lin, _g1_mul(p, s))
    digests = [folded_h, lin, lro[0], lro[1], lro[2], s_com[0], s_com[1]]
    gk = int.from_bytes(_challenge(b"gamma", b"", fb(zeta), *[_g1_raw(p) for p in digests], *[fb(v) for v in claimed], fb(zu)), "big") % R
    fd, fv, acc = None, 0, 1
    for dg, v in zip(digests, claimed):
        fd = _g1_add(fd, _g1_mul(dg, acc))
        fv = (fv + acc * v) % R
        acc = acc * gk % R
    gen = (1, 2)

    def kzg_ok(c, h, v, z):
        return _g1_add(c, _g1_neg(_g1_mul(gen, v))) == _g1_mul(h, (tau - z) % R)
    return kzg_ok(fd, batched_h, fv, zeta) and kzg_ok(z_com, zs_h, zu, zeta * d.gen % R)
