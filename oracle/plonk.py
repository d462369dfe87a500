"""CPU restatement of the PlonK prover's vector building blocks (TEST INFRASTRUCTURE -- PARITY UNPINNED:
the reference is Go and gnark-crypto is not vendored; nothing here has been run against it).

Definitions followed:
  batch_invert            fr.BatchInvert semantics (zero stays zero), used at
                          /root/reference/backend/plonk/bn254/prove.go:1272
  prefix_product          the running product of iop.BuildRatioCopyConstraint (gnark-crypto, absent;
                          call site prove.go:610-621)
  poly_eval               Polynomial.Evaluate in canonical basis (prove.go:723,1127,1300-1305)
  poly_div_linear         the quotient of kzg.Open (gnark-crypto kzg, absent; call sites prove.go:646,823)
  divide_by_xn_minus_one  divideByXMinusOne + evaluateXnMinusOneDomainBigCoset, prove.go:1223-1276
"""
from . import bn254 as B
from . import ntt as N

R = B.R


def batch_invert(xs):
    return [pow(x, R - 2, R) if x % R else 0 for x in xs]


def prefix_product(xs):
    out, acc = [], 1
    for x in xs:
        acc = acc * x % R
        out.append(acc)
    return out


def poly_eval(coeffs, x):
    acc = 0
    for c in reversed(coeffs):
        acc = (acc * x + c) % R
    return acc


def poly_div_linear(coeffs, z):
    """(quotient, remainder) of p(X) / (X - z)."""
    n = len(coeffs)
    q, acc = [0] * max(n - 1, 0), 0
    for i in range(n - 1, -1, -1):
        acc = (acc * z + coeffs[i]) % R
        if i >= 1:
            q[i - 1] = acc
    return q, acc


def xn_minus_one_inverse_big_coset(d_small: N.Domain, d_big: N.Domain):
    """evaluateXnMinusOneDomainBigCoset (prove.go:1253-1276)."""
    rho = d_big.cardinality // d_small.cardinality
    n = d_small.cardinality
    res = [pow(d_big.coset_gen, n, R)]
    t = pow(d_big.generator, n, R)
    for _ in range(1, rho):
        res.append(res[-1] * t % R)
    return batch_invert([(v - 1) % R for v in res])


def divide_by_xn_minus_one(r, d_small: N.Domain, d_big: N.Domain):
    """divideByXMinusOne (prove.go:1223-1250): r in LagrangeCoset / BitReverse on the big domain;
    returns the canonical, regular-layout coefficients."""
    inv = xn_minus_one_inverse_big_coset(d_small, d_big)
    rho = len(inv)
    m = len(r)
    out = [r[i] * inv[N.bitrev(i, d_big.log_n) % rho] % R for i in range(m)]
    N.fft_inverse(out, d_big, N.DIT, on_coset=True)     # LagrangeCoset + BitReverse -> ToCanonical = FFTInverse(DIT, OnCoset)
    return out
