"""CPU oracle (TEST INFRASTRUCTURE ONLY) -- BLS12-377 field and curve restatement (BASELINE.json config 5).

Same role and same exported names as `oracle/bn254.py`, for the second curve gnark's Groth16 / PlonK
backends are instantiated on (`/root/reference/backend/groth16/bls12-377/prove.go`; the arithmetic lives in
gnark-crypto's ecc/bls12-377, a go.mod dependency that is not vendored, see bn254.py's header).

Constants pinned on the reference tree:
  * p, r                 std/math/emulated/emparams/emparams.go:123,137 (BLS12377Fp / BLS12377Fr moduli)
  * seed x0              std/algebra/native/sw_bls12377/pairing.go:29 (9586122913090633729);
                         r = x0^4 - x0^2 + 1 and p = (x0-1)^2 r / 3 + x0 are re-derived below and asserted
  * Fp2 = Fp[u]/(u^2+5)  std/algebra/native/fields_bls12377/e12.go:79 (non-residue -5)
  * G1: y^2 = x^3 + 1    (BLS12-377 curve equation; generator recalled from gnark-crypto and self-checked:
                         on the curve and of order r)
  * G2: y^2 = x^3 + 1/u  (D-twist; the order of E'(Fp2) is derived from the trace and asserted divisible by r)
The G2 generator is NOT gnark-crypto's: it is derived here by clearing the cofactor of the first curve point
with a small x.  Every device entry point takes its bases as inputs, so the choice of generator is immaterial
to the product; it only seeds test fixtures.

PARITY UNPINNED (as for bn254.py): no golden MSM / FFT / proof vectors exist in the reference tree.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline leg may import this module.
"""
from __future__ import annotations

from math import isqrt

from .bn254 import _aff_add, _aff_neg, _aff_mul  # field-generic affine formulas (a = 0 curves)

X0 = 9586122913090633729
R = X0 ** 4 - X0 ** 2 + 1
P = ((X0 - 1) ** 2 * R) // 3 + X0
assert P == 0x01AE3A4617C510EAC63B05C06CA1493B1A22D9F300F5138F1EF3622FBA094800170B5D44300000008508C00000000001
assert R == 0x12AB655E9A2CA55660B44D1E5C37B00159AA76FED00000010A11800000000001

FP_LIMBS64 = 6
FR_LIMBS64 = 4
MONT_R_FP = 1 << 384
MONT_R_FR = 1 << 256

# Fr: 2-adicity 47; multiplicative generator 22 (gnark-crypto fr.FrMultiplicativeGen for BLS12-377; checked
# below to be a non-residue of every prime order dividing r-1 that matters for the FFT: order 2^47 exactly)
FR_MULT_GEN = 22
FR_TWO_ADICITY = 47
assert (R - 1) % (1 << 47) == 0 and ((R - 1) >> 47) & 1
FR_ROOT_OF_UNITY = pow(FR_MULT_GEN, (R - 1) >> FR_TWO_ADICITY, R)
assert pow(FR_ROOT_OF_UNITY, 1 << (FR_TWO_ADICITY - 1), R) == R - 1   # order is exactly 2^47

FP2_BETA = -5          # u^2 = -5

# point compression flags in the top three bits of byte 0: smallest y, largest y, infinity
# (recalled from gnark-crypto ecc/bls12-377/marshal.go, the ZCash BLS12-381 convention; source absent)
COMPRESS_FLAGS = (0b100 << 5, 0b101 << 5, 0b110 << 5)

G1_B = 1
G1_GEN = (
    81937999373150964239938255573465948239988671502647976594219695644855304257327692006745978603320413799295628339695,
    241266749859715473739788878240585681733927191168601896383759122102112907357779751001206799952863815012735208165030,
)


def to_mont(x: int, mod: int) -> int:
    return (x * (MONT_R_FP if mod == P else MONT_R_FR)) % mod


def from_mont(x: int, mod: int) -> int:
    return (x * pow(MONT_R_FP if mod == P else MONT_R_FR, -1, mod)) % mod


# ----------------------------------------------------------------------------------------
# Fp2 = Fp[u] / (u^2 + 5)
# ----------------------------------------------------------------------------------------
def fp2_add(a, b):
    return ((a[0] + b[0]) % P, (a[1] + b[1]) % P)


def fp2_sub(a, b):
    return ((a[0] - b[0]) % P, (a[1] - b[1]) % P)


def fp2_mul(a, b):
    return ((a[0] * b[0] + FP2_BETA * a[1] * b[1]) % P, (a[0] * b[1] + a[1] * b[0]) % P)


def fp2_sqr(a):
    return fp2_mul(a, a)


def fp2_neg(a):
    return ((-a[0]) % P, (-a[1]) % P)


def fp2_inv(a):
    d = pow(a[0] * a[0] - FP2_BETA * a[1] * a[1], -1, P)
    return (a[0] * d % P, (-a[1]) * d % P)


def fp2_scalar(a, k):
    return (a[0] * k % P, a[1] * k % P)


def fp2_pow(a, e: int):
    acc = (1, 0)
    while e:
        if e & 1:
            acc = fp2_mul(acc, a)
        a = fp2_sqr(a)
        e >>= 1
    return acc


def fp2_sqrt(a):
    """A square root of a in Fp2, or None.  p = 1 mod 4 here, so the complex method through the norm is used:
    with n = sqrt(a0^2 + 5 a1^2) in Fp, x0^2 = (a0 + n)/2 (or (a0 - n)/2), x1 = a1 / (2 x0)."""
    if a == (0, 0):
        return (0, 0)
    norm = (a[0] * a[0] - FP2_BETA * a[1] * a[1]) % P
    n = fp_sqrt(norm)
    if n is None:
        return None
    inv2 = pow(2, -1, P)
    for cand in ((a[0] + n) * inv2 % P, (a[0] - n) * inv2 % P):
        x0 = fp_sqrt(cand)
        if x0 is None or x0 == 0:
            continue
        x1 = a[1] * pow(2 * x0, -1, P) % P
        if fp2_sqr((x0, x1)) == (a[0] % P, a[1] % P):
            return (x0, x1)
    if a[1] % P == 0:                      # a in Fp and a non-residue there: sqrt is purely imaginary
        x1 = fp_sqrt(a[0] * pow(FP2_BETA, -1, P) % P)
        if x1 is not None:
            return (0, x1)
    return None


def fp_sqrt(a: int):
    """Tonelli-Shanks in Fp (p - 1 = 2^46 * odd)."""
    a %= P
    if a == 0:
        return 0
    if pow(a, (P - 1) // 2, P) != 1:
        return None
    s, q = 0, P - 1
    while q % 2 == 0:
        s, q = s + 1, q // 2
    z = 2
    while pow(z, (P - 1) // 2, P) == 1:
        z += 1
    m, c, t, x = s, pow(z, q, P), pow(a, q, P), pow(a, (q + 1) // 2, P)
    while t != 1:
        i, t2 = 0, t
        while t2 != 1:
            t2, i = t2 * t2 % P, i + 1
        b = pow(c, 1 << (m - i - 1), P)
        m, c = i, b * b % P
        t, x = t * c % P, x * b % P
    return x


G2_B = fp2_inv((0, 1))     # 1/u = (0, -1/5): y^2 = x^3 + 1/u
assert G2_B == (0, 155198655607781456406391640216936120121836107652948796323930557600032281009004493664981332883744016074664192874906)


class _Fp:
    zero = 0
    one = 1

    @staticmethod
    def add(a, b): return (a + b) % P
    @staticmethod
    def sub(a, b): return (a - b) % P
    @staticmethod
    def mul(a, b): return (a * b) % P
    @staticmethod
    def inv(a): return pow(a, -1, P)
    @staticmethod
    def neg(a): return (-a) % P
    @staticmethod
    def is_zero(a): return a % P == 0


class _Fp2:
    zero = (0, 0)
    one = (1, 0)
    add = staticmethod(fp2_add)
    sub = staticmethod(fp2_sub)
    mul = staticmethod(fp2_mul)
    inv = staticmethod(fp2_inv)
    neg = staticmethod(fp2_neg)

    @staticmethod
    def is_zero(a): return a[0] % P == 0 and a[1] % P == 0


def g1_add(p, q): return _aff_add(_Fp, p, q)
def g1_neg(p): return _aff_neg(_Fp, p)
def g1_mul(p, k): return _aff_mul(_Fp, p, k % R)
def g2_add(p, q): return _aff_add(_Fp2, p, q)
def g2_neg(p): return _aff_neg(_Fp2, p)
def g2_mul(p, k): return _aff_mul(_Fp2, p, k % R)


def g1_on_curve(p) -> bool:
    if p is None:
        return True
    x, y = p
    return (y * y - x * x * x - G1_B) % P == 0


def g2_on_curve(p) -> bool:
    if p is None:
        return True
    x, y = p
    return fp2_sub(fp2_sqr(y), fp2_add(fp2_mul(fp2_sqr(x), x), G2_B)) == (0, 0)


# order of the twist E'(Fp2): with t = x0 + 1 the trace over Fp, t2 = t^2 - 2p the trace over Fp2 and
# 3 f^2 = 4 p^2 - t2^2, the sextic twists have orders p^2 + 1 - (+-3f +- t2)/2; the D-twist above is the
# one whose order r divides.
_T = X0 + 1
_T2 = _T * _T - 2 * P
_F = isqrt((4 * P * P - _T2 * _T2) // 3)
assert 3 * _F * _F == 4 * P * P - _T2 * _T2
G2_ORDER = P * P + 1 - (-3 * _F + _T2) // 2
assert G2_ORDER % R == 0
G2_COFACTOR = G2_ORDER // R


def _derive_g2_gen():
    x = (1, 0)
    while True:
        y = fp2_sqrt(fp2_add(fp2_mul(fp2_sqr(x), x), G2_B))
        if y is not None:
            g = _aff_mul(_Fp2, (x, y), G2_COFACTOR)
            if g is not None:
                return g
        x = (x[0] + 1, x[1] + 1)


G2_GEN = _derive_g2_gen()


def jac_to_affine_g1(X, Y, Z):
    if Z % P == 0:
        return None
    zi = pow(Z, -1, P)
    return (X * zi * zi % P, Y * zi * zi * zi % P)


def jac_to_affine_g2(X, Y, Z):
    if _Fp2.is_zero(Z):
        return None
    zi = fp2_inv(Z)
    zi2 = fp2_sqr(zi)
    return (fp2_mul(X, zi2), fp2_mul(Y, fp2_mul(zi2, zi)))


def msm_g1(points, scalars):
    """Definition of G1Jac.MultiExp (reference call sites: backend/groth16/bls12-377/prove.go, same lines as bn254)."""
    acc = None
    for pt, k in zip(points, scalars):
        if pt is not None and k % R:
            acc = g1_add(acc, g1_mul(pt, k))
    return acc


def msm_g2(points, scalars):
    acc = None
    for pt, k in zip(points, scalars):
        if pt is not None and k % R:
            acc = g2_add(acc, g2_mul(pt, k))
    return acc


# ----------------------------------------------------------------------------------------
# limbs <-> ints (fp.Element = 6 x u64, fr.Element = 4 x u64, little-endian Montgomery)
# ----------------------------------------------------------------------------------------
def int_to_limbs(x: int, n: int) -> list[int]:
    return [(x >> (64 * i)) & 0xFFFFFFFFFFFFFFFF for i in range(n)]


def limbs_to_int(l) -> int:
    return sum(int(v) << (64 * i) for i, v in enumerate(l))
