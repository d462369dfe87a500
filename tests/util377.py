"""BLS12-377 limb marshalling for the parity tests (pure Python / numpy; oracle/cref.py's C helpers are BN254-only).

Fr element = 4 u64 Montgomery limbs (R = 2^256), Fp element = 6 u64 (R = 2^384); G1 affine = 12 u64,
G1 Jacobian = 18, G2 affine = X.A0 X.A1 Y.A0 Y.A1 = 24, G2 Jacobian = 36; (0, 0) / Z = 0 is infinity.
"""
import numpy as np

from oracle import bls12_377 as B

_M64 = (1 << 64) - 1
_RFR_INV = pow(B.MONT_R_FR, -1, B.R)
_RFP_INV = pow(B.MONT_R_FP, -1, B.P)


def _limbs(x, n):
    return [(x >> (64 * i)) & _M64 for i in range(n)]


def _int(row):
    return sum(int(v) << (64 * i) for i, v in enumerate(row))


def fr_arr(xs):
    return np.array([_limbs(x % B.R * B.MONT_R_FR % B.R, 4) for x in xs], dtype=np.uint64).reshape(-1, 4)


def fr(x):
    return fr_arr([x])[0]


def fr_ints(a):
    return [_int(row) * _RFR_INV % B.R for row in np.asarray(a, dtype=np.uint64).reshape(-1, 4)]


def fp_arr(xs):
    return np.array([_limbs(x % B.P * B.MONT_R_FP % B.P, 6) for x in xs], dtype=np.uint64).reshape(-1, 6)


def fp_ints(a):
    return [_int(row) * _RFP_INV % B.P for row in np.asarray(a, dtype=np.uint64).reshape(-1, 6)]


def g1_arr(pts):
    out = np.zeros((len(pts), 12), dtype=np.uint64)
    for i, p in enumerate(pts):
        if p is not None:
            out[i] = fp_arr([p[0], p[1]]).reshape(12)
    return out


def g2_arr(pts):
    out = np.zeros((len(pts), 24), dtype=np.uint64)
    for i, p in enumerate(pts):
        if p is not None:
            (x0, x1), (y0, y1) = p
            out[i] = fp_arr([x0, x1, y0, y1]).reshape(24)
    return out


def g1_aff(a12):
    x, y = fp_ints(np.asarray(a12).reshape(2, 6))
    return None if x == 0 and y == 0 else (x, y)


def g2_aff(a24):
    x0, x1, y0, y1 = fp_ints(np.asarray(a24).reshape(4, 6))
    return None if (x0, x1, y0, y1) == (0, 0, 0, 0) else ((x0, x1), (y0, y1))


def g1_jac(a18):
    X, Y, Z = fp_ints(np.asarray(a18).reshape(3, 6))
    return B.jac_to_affine_g1(X, Y, Z)


def g2_jac(a36):
    v = fp_ints(np.asarray(a36).reshape(6, 6))
    return B.jac_to_affine_g2((v[0], v[1]), (v[2], v[3]), (v[4], v[5]))


def rand_fr(rng, n):
    """n pseudo-random reduced Fr elements, taken as Montgomery limbs."""
    top = B.R >> 192
    a = rng.integers(0, 1 << 63, size=(n, 4), dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, size=(n, 4), dtype=np.uint64)
    a[:, 3] = rng.integers(0, top, size=n, dtype=np.uint64)
    return a


def rand_fp(rng, n):
    top = B.P >> 320
    a = rng.integers(0, 1 << 63, size=(n, 6), dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, size=(n, 6), dtype=np.uint64)
    a[:, 5] = rng.integers(0, top, size=n, dtype=np.uint64)
    return a


def known_dlog_bases(group, n, seed=1):
    """n points k_i * G with known k_i, one affine addition each: P_0 = k_0 G, P_{i+1} = P_i + T[j], T[j] = t_j G."""
    rng = np.random.default_rng(seed)
    gen, mul, add = (B.G1_GEN, B.g1_mul, B.g1_add) if group == 1 else (B.G2_GEN, B.g2_mul, B.g2_add)
    ts = [int(x) for x in rng.integers(1, 1 << 62, size=8)]
    T = [mul(gen, t) for t in ts]
    k = int(rng.integers(1, 1 << 62))
    p = mul(gen, k)
    pts, ks = [], []
    for i in range(n):
        pts.append(p)
        ks.append(k)
        j = int(rng.integers(0, 8))
        p, k = add(p, T[j]), k + ts[j]
    return pts, ks
