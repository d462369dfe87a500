"""CPU oracle (TEST INFRASTRUCTURE ONLY) -- BN254 field, curve, NTT and MSM restatement.

Python big-int restatement of the arithmetic that gnark's BN254 Groth16 prover
(`/root/reference/backend/groth16/bn254/prove.go`) delegates to gnark-crypto
v0.12.2-0.20231117165148-e77308824822 (`/root/reference/go.mod:10`), a module
that is NOT vendored in the reference tree.  Every function below restates the
*published mathematical definition* of the gnark-crypto routine the reference
calls, and cites the reference call site it has to agree with.

PARITY UNPINNED: the reference holds no golden vector for MSM / FFT / proof
bytes, and neither Go nor gnark-crypto is available in
this environment.  What *is* pinned against in-tree constants is checked in
`tests/test_oracle_constants.py` (moduli, coset shift, Fp fractions, G2 twist).

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline leg may
import this module.  Product code under `gnark-fork_b200/` must never do so.
"""
from __future__ import annotations

# --- moduli (reference: backend/groth16/bn254/solidity.go:41-42) -------------------------
P = 0x30644E72E131A029B85045B68181585D97816A916871CA8D3C208C16D87CFD47  # base field Fp
R = 0x30644E72E131A029B85045B68181585D2833E84879B9709143E1F593F0000001  # scalar field Fr

LIMBS64 = 4
MONT_R = 1 << 256  # Montgomery radix: 2^(64*limbs) (gnark-crypto fr/fp element.go)

# Fr: multiplicative generator (reference: backend/plonk/bn254/solidity.go:56, VK_COSET_SHIFT)
FR_MULT_GEN = 5
# Fr: 2-adicity and the 2^28-th root of unity gnark-crypto's fft.NewDomain starts from
# (recalled from gnark-crypto ecc/bn254/fr/fft/domain.go; self-checked in tests: order is
# exactly 2^28).  The product never hard-codes it: the generator is an INPUT of the C ABI.
FR_TWO_ADICITY = 28
FR_ROOT_2_28 = 19103219067921713944291392827692070036145651957329286315305642004821462161904
FR_ROOT_OF_UNITY = FR_ROOT_2_28   # curve-neutral name used by ntt.py (order 2^FR_TWO_ADICITY)

# G1: y^2 = x^3 + 3, generator (1, 2)
G1_B = 3
G1_GEN = (1, 2)
# G2 over Fp2 = Fp[u]/(u^2+1): y^2 = x^3 + 3/(9+u)
#   (reference: backend/groth16/bn254/solidity.go:52-58 pins 27/82 and 3/82)
G2_GEN = (
    (10857046999023057135944570762232829481370756359578518086990519993285655852781,
     11559732032986387107991004021392285783925812861821192530917403151452391805634),
    (8495653923123431417604973247489272438418190587263600148770280649306958101930,
     4082367875863433681332203403145435568316851327593401208105741076214120093531),
)


# ----------------------------------------------------------------------------------------
# limbs <-> ints   (layout of fr.Element / fp.Element: 4 x u64 little-endian, Montgomery)
# ----------------------------------------------------------------------------------------
def to_mont(x: int, mod: int) -> int:
    return (x * MONT_R) % mod


def from_mont(x: int, mod: int) -> int:
    return (x * pow(MONT_R, -1, mod)) % mod


def int_to_limbs(x: int) -> list[int]:
    return [(x >> (64 * i)) & 0xFFFFFFFFFFFFFFFF for i in range(LIMBS64)]


def limbs_to_int(l) -> int:
    return sum(int(v) << (64 * i) for i, v in enumerate(l))


# ----------------------------------------------------------------------------------------
# Fp2
# ----------------------------------------------------------------------------------------
def fp2_add(a, b):
    return ((a[0] + b[0]) % P, (a[1] + b[1]) % P)


def fp2_sub(a, b):
    return ((a[0] - b[0]) % P, (a[1] - b[1]) % P)


def fp2_mul(a, b):
    return ((a[0] * b[0] - a[1] * b[1]) % P, (a[0] * b[1] + a[1] * b[0]) % P)


def fp2_sqr(a):
    return fp2_mul(a, a)


def fp2_neg(a):
    return ((-a[0]) % P, (-a[1]) % P)


def fp2_inv(a):
    d = pow(a[0] * a[0] + a[1] * a[1], -1, P)
    return (a[0] * d % P, (-a[1]) * d % P)


def fp2_scalar(a, k):
    return (a[0] * k % P, a[1] * k % P)


G2_B = fp2_mul((3, 0), fp2_inv((9, 1)))  # 3/(9+u)


# ----------------------------------------------------------------------------------------
# generic short-Weierstrass (a = 0) arithmetic on affine points; None = infinity
# field ops are passed in so the same code serves G1 (ints mod P) and G2 (Fp2 tuples)
# ----------------------------------------------------------------------------------------
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


def _aff_add(F, p, q):
    if p is None:
        return q
    if q is None:
        return p
    x1, y1 = p
    x2, y2 = q
    if x1 == x2:
        if F.is_zero(F.add(y1, y2)):
            return None
        lam = F.mul(F.mul(F.add(F.add(x1, x1), x1), x1), F.inv(F.add(y1, y1)))
    else:
        lam = F.mul(F.sub(y2, y1), F.inv(F.sub(x2, x1)))
    x3 = F.sub(F.sub(F.mul(lam, lam), x1), x2)
    y3 = F.sub(F.mul(lam, F.sub(x1, x3)), y1)
    return (x3, y3)


def _aff_neg(F, p):
    return None if p is None else (p[0], F.neg(p[1]))


def _aff_mul(F, p, k: int):
    if k < 0:
        return _aff_mul(F, _aff_neg(F, p), -k)
    acc = None
    while k:
        if k & 1:
            acc = _aff_add(F, acc, p)
        p = _aff_add(F, p, p)
        k >>= 1
    return acc


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


def jac_to_affine_g1(X, Y, Z):
    """G1Affine.FromJacobian (reference use: backend/groth16/bn254/prove.go:221,276)."""
    if Z % P == 0:
        return None
    zi = pow(Z, -1, P)
    return (X * zi * zi % P, Y * zi * zi * zi % P)


def jac_to_affine_g2(X, Y, Z):
    """G2Affine.FromJacobian (reference use: backend/groth16/bn254/prove.go:299)."""
    if _Fp2.is_zero(Z):
        return None
    zi = fp2_inv(Z)
    zi2 = fp2_sqr(zi)
    return (fp2_mul(X, zi2), fp2_mul(Y, fp2_mul(zi2, zi)))


def msm_g1(points, scalars):
    """Definition of G1Jac.MultiExp: sum_i scalars[i]*points[i]
    (reference call sites: backend/groth16/bn254/prove.go:201,214,234,244). Affine result."""
    acc = None
    for pt, k in zip(points, scalars):
        if pt is not None and k % R:
            acc = g1_add(acc, g1_mul(pt, k))
    return acc


def msm_g2(points, scalars):
    """Definition of G2Jac.MultiExp (reference call site: prove.go:290). Affine result."""
    acc = None
    for pt, k in zip(points, scalars):
        if pt is not None and k % R:
            acc = g2_add(acc, g2_mul(pt, k))
    return acc
