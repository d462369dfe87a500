"""CPU checks of oracle/plonk.py against definitions (no GPU): the fast forms agree with the naive ones."""
import random

from oracle import bn254 as B
from oracle import ntt as N
from oracle import plonk as P


def test_poly_div_linear_identity():
    rnd = random.Random(1)
    for n in (1, 2, 5, 33, 100):
        c = [rnd.randrange(B.R) for _ in range(n)]
        z = rnd.randrange(B.R)
        q, rem = P.poly_div_linear(c, z)
        assert rem == P.poly_eval(c, z)
        # q(X)(X - z) + rem == p(X), coefficient by coefficient
        back = [0] * n
        for i, qi in enumerate(q):
            back[i + 1] = (back[i + 1] + qi) % B.R
            back[i] = (back[i] - qi * z) % B.R
        back[0] = (back[0] + rem) % B.R
        assert back == c


def test_batch_invert_and_prefix_product():
    xs = [3, 0, B.R - 1, 7, 0, 12345]
    inv = P.batch_invert(xs)
    assert [x * y % B.R for x, y in zip(xs, inv)] == [1, 0, 1, 1, 0, 1]
    assert P.prefix_product([2, 3, 5]) == [2, 6, 30]


def test_divide_by_xn_minus_one_is_exact_division():
    """h = t / (X^n - 1) for a t that is a multiple of X^n - 1: evaluate t on the big coset in the
    LagrangeCoset / BitReverse form, divide, compare with h's coefficients."""
    rnd = random.Random(2)
    d_small, d_big = N.Domain(8), N.Domain(32)
    n = 8
    h = [rnd.randrange(B.R) for _ in range(20)]               # deg(h) + n < 32
    t = [0] * 32
    for i, hi in enumerate(h):                                 # t = h * (X^n - 1)
        t[i + n] = (t[i + n] + hi) % B.R
        t[i] = (t[i] - hi) % B.R
    ev = list(t)
    N.fft(ev, d_big, N.DIF, on_coset=True)                     # canonical regular -> LagrangeCoset bit-reversed
    got = P.divide_by_xn_minus_one(ev, d_small, d_big)
    assert got == h + [0] * 12


def test_to_lagrange_g1_restatement_matches_the_definition():
    """oracle/plonk_prove.to_lagrange_g1 (kzg.ToLagrangeG1 as setup.go:134 calls it: a DIF over group elements, no tau)
    against the definition out[i] = [L_i(tau)]G1 with L(tau) = F^-1 (tau^j) from the scalar transform."""
    from oracle import bn254 as B
    from oracle import ntt as N
    from oracle import plonk_prove as PP
    for n, tau in ((1, 7), (2, 11), (8, 123456789)):
        d = N.Domain(n)
        mono = [pow(tau, j, B.R) for j in range(n)]
        lag = list(mono)
        N.fft_inverse(lag, d, N.DIF)
        N.bit_reverse(lag)
        got = PP.to_lagrange_g1([B.g1_mul(B.G1_GEN, m) for m in mono], d)
        assert got == [B.g1_mul(B.G1_GEN, x) for x in lag]
        assert sum(lag) % B.R == 1      # partition of unity
