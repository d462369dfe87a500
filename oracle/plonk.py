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


def build_ratio_copy_constraint(entries, perm, beta, gamma, d: N.Domain):
    """iop.BuildRatioCopyConstraint (gnark-crypto iop/ratios.go, absent -- restated from its call site
    prove.go:610-621 and the published algorithm): entries = [l, r, o] in Lagrange / Regular form,
    perm = pk.trace.S (3 n indices).  Returns Z in Lagrange / Regular form."""
    n = d.cardinality
    g = d.coset_gen
    ident = []
    for j in range(len(entries)):
        ident += [pow(g, j, R) * pow(d.generator, i, R) % R for i in range(n)]
    num, den = [1] * n, [1] * n
    for i in range(n - 1):
        a = b = 1
        for j, w in enumerate(entries):
            a = a * ((beta * ident[j * n + i] + gamma + w[i]) % R) % R
            b = b * ((beta * ident[perm[j * n + i]] + gamma + w[i]) % R) % R
        num[i + 1], den[i + 1] = a, b
    num, den = prefix_product(num), prefix_product(den)
    return [x * y % R for x, y in zip(num, batch_invert(den))]


# prove.go:60-77
ID_L, ID_R, ID_O, ID_Z, ID_ZS, ID_QL, ID_QR, ID_QM, ID_QO, ID_QK, ID_S1, ID_S2, ID_S3, ID_ID, ID_LONE, ID_QCI = range(16)


def quotient_rows(x, qci, d_small: N.Domain, log_big, coset_index, beta, gamma, alpha, cs, bl, br, bo, bz, cres):
    """One coset of computeNumerator (prove.go:837-1044): allConstraints over the rows of the small
    domain, scattered into the bit-reversed big buffer `cres` (list, modified in place).
    x: dict id -> list (Lagrange / Regular on the current coset), bl..bz: blinding coefficients as
    scaled for this coset; ZS = Z shifted by one row (prove.go:945-946)."""
    n = d_small.cardinality
    rho = (1 << log_big) // n
    css = cs * cs % R
    ev = lambda p, t: sum(c * pow(t, k, R) for k, c in enumerate(p)) % R
    for i in range(n):
        wi, wi1 = pow(d_small.generator, i, R), pow(d_small.generator, (i + 1) % n, R)
        L = (x[ID_L][i] + ev(bl, wi)) % R
        Rr = (x[ID_R][i] + ev(br, wi)) % R
        O = (x[ID_O][i] + ev(bo, wi)) % R
        Z = (x[ID_Z][i] + ev(bz, wi)) % R
        ZS = (x[ID_Z][(i + 1) % n] + ev(bz, wi1)) % R
        gate = (x[ID_QL][i] * L + x[ID_QR][i] * Rr + x[ID_QM][i] * L * Rr + x[ID_QO][i] * O + x[ID_QK][i]) % R
        for k in range(len(qci) // 2):
            gate = (gate + qci[2 * k][i] * qci[2 * k + 1][i]) % R
        idv = x[ID_ID][i]
        r = (gamma + L + idv) * (idv * cs + Rr + gamma) % R * (idv * css + O + gamma) % R * Z % R
        l = (x[ID_S1][i] * beta + L + gamma) * (x[ID_S2][i] * beta + Rr + gamma) % R * (x[ID_S3][i] * beta + O + gamma) % R * ZS % R
        c = (Z - 1) * x[ID_LONE][i] % R
        c = ((c * alpha + (l - r)) % R * alpha + gate) % R
        cres[N.bitrev(rho * i + coset_index, log_big)] = c
    return cres
