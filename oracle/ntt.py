"""CPU oracle (TEST INFRASTRUCTURE ONLY) -- Fr NTT in gnark-crypto's fft.Domain conventions
and the Groth16 quotient `computeH`.

Restates gnark-crypto `ecc/bn254/fr/fft` (module absent from /root/reference, pinned in
/root/reference/go.mod:10) from its published algorithm; the in-tree structural witness of
that algorithm is `/root/reference/backend/groth16/bn254/mpcsetup/lagrange.go:143-179`
(DIF butterfly (a,b)->(a+b,(a-b)*w), twiddles[stage][i] = w^(i*2^stage), bit-reversed
output), and the call sites are `/root/reference/backend/groth16/bn254/prove.go:369-393`.

PARITY UNPINNED (no golden vectors in the reference, no Go here).  The NTT of a vector over
a given (generator, coset shift) is mathematically unique and stored fully reduced, so the
only convention risk is the choice of generator -- which the product takes as an input.
"""
from __future__ import annotations

from .bn254 import R, FR_MULT_GEN, FR_ROOT_OF_UNITY, FR_TWO_ADICITY

DIF = 0
DIT = 1


def bitrev(i: int, log_n: int) -> int:
    r = 0
    for _ in range(log_n):
        r = (r << 1) | (i & 1)
        i >>= 1
    return r


def bit_reverse(a: list) -> None:
    """fft.BitReverse; same permutation as setup.go:678-688 applies to pk.G1.Z."""
    n = len(a)
    log_n = n.bit_length() - 1
    for i in range(n):
        j = bitrev(i, log_n)
        if j > i:
            a[i], a[j] = a[j], a[i]


class Domain:
    """fft.Domain: fields used by the reference backends (SURVEY 8(a) census):
    Cardinality, CardinalityInv, Generator, GeneratorInv, FrMultiplicativeGen (+Inv)."""

    def __init__(self, m: int, generator: int | None = None, coset_gen: int = FR_MULT_GEN):
        log_n = max(0, (m - 1).bit_length())  # ecc.NextPowerOfTwo
        self.log_n = log_n
        self.cardinality = 1 << log_n
        if generator is None:
            assert log_n <= FR_TWO_ADICITY
            generator = pow(FR_ROOT_OF_UNITY, 1 << (FR_TWO_ADICITY - log_n), R)
        self.generator = generator
        self.generator_inv = pow(generator, -1, R)
        self.cardinality_inv = pow(self.cardinality, -1, R)
        self.coset_gen = coset_gen
        self.coset_gen_inv = pow(coset_gen, -1, R)


def _dif_inplace(a: list[int], w: int) -> None:
    """In-place radix-2 DIF: natural order in, bit-reversed out. Stage s works on blocks of
    size n>>s with twiddle w^(j * 2^s) on the lower half (lagrange.go:143-179 structure)."""
    n = len(a)
    m = n
    ws = w
    while m >= 2:
        h = m >> 1
        tw = [1] * h
        for j in range(1, h):
            tw[j] = tw[j - 1] * ws % R
        for base in range(0, n, m):
            for j in range(h):
                u = a[base + j]
                v = a[base + j + h]
                a[base + j] = (u + v) % R
                a[base + j + h] = (u - v) * tw[j] % R
        ws = ws * ws % R
        m = h


def _dit_inplace(a: list[int], w: int) -> None:
    """In-place radix-2 DIT: bit-reversed order in, natural out."""
    n = len(a)
    log_n = n.bit_length() - 1
    m = 2
    for s in range(log_n):
        h = m >> 1
        ws = pow(w, n // m, R)
        tw = [1] * h
        for j in range(1, h):
            tw[j] = tw[j - 1] * ws % R
        for base in range(0, n, m):
            for j in range(h):
                u = a[base + j]
                v = a[base + j + h] * tw[j] % R
                a[base + j] = (u + v) % R
                a[base + j + h] = (u - v) % R
        m <<= 1


def fft(a: list[int], d: Domain, decimation: int, on_coset: bool = False) -> None:
    """Domain.FFT(a, decimation, [OnCoset]) -- call sites prove.go:373-375.
    With OnCoset the input is first scaled by g^i; if decimation is DIT the input is taken to
    be in bit-reversed order so element i is scaled by g^bitrev(i)."""
    n = len(a)
    assert n == d.cardinality
    if on_coset:
        for i in range(n):
            e = i if decimation == DIF else bitrev(i, d.log_n)
            a[i] = a[i] * pow(d.coset_gen, e, R) % R
    if decimation == DIF:
        _dif_inplace(a, d.generator)
    else:
        _dit_inplace(a, d.generator)


def fft_inverse(a: list[int], d: Domain, decimation: int, on_coset: bool = False) -> None:
    """Domain.FFTInverse(a, decimation, [OnCoset]) -- call sites prove.go:369-371,393.
    Transform with the inverse generator, scale by 1/n and, with OnCoset, by g^-i where i is
    the *natural* index of the output slot (bit-reversed slot index when decimation is DIF)."""
    n = len(a)
    assert n == d.cardinality
    if decimation == DIF:
        _dif_inplace(a, d.generator_inv)
    else:
        _dit_inplace(a, d.generator_inv)
    for i in range(n):
        s = d.cardinality_inv
        if on_coset:
            e = bitrev(i, d.log_n) if decimation == DIF else i
            s = s * pow(d.coset_gen_inv, e, R) % R
        a[i] = a[i] * s % R


def naive_dft(a: list[int], w: int) -> list[int]:
    """O(n^2) definition, used to pin the fast transforms on small sizes."""
    n = len(a)
    return [sum(a[j] * pow(w, i * j, R) for j in range(n)) % R for i in range(n)]


def compute_h(a: list[int], b: list[int], c: list[int], d: Domain) -> list[int]:
    """computeH, /root/reference/backend/groth16/bn254/prove.go:353-396.
    Output is in BIT-REVERSED order (final FFTInverse is DIF with no reordering)."""
    n = d.cardinality
    a = list(a) + [0] * (n - len(a))
    b = list(b) + [0] * (n - len(b))
    c = list(c) + [0] * (n - len(c))
    for v in (a, b, c):
        fft_inverse(v, d, DIF)
    for v in (a, b, c):
        fft(v, d, DIT, on_coset=True)
    den = pow((pow(d.coset_gen, n, R) - 1) % R, -1, R)
    for i in range(n):
        a[i] = (a[i] * b[i] - c[i]) * den % R
    fft_inverse(a, d, DIF, on_coset=True)
    return a
