"""Host-side helpers around the Groth16 device path for callers without Go (tests, bench.py): the reference
circuit of BASELINE.json configs 1 and 3 as the arrays the C ABI takes, its witness, and the pairing-free
acceptance check of a proof made with a key whose toxic waste is known.

Follows (under /root/reference):
  refCircuit                backend/groth16/groth16_test.go:120-156   (X <- X*X chain, AssertIsEqual(X_N, Y))
  R1C shapes                frontend/cs/r1cs/api.go:198-232 (Mul: L = x, R = x, O = new), api_assertions.go:31-36
  solver outputs a, b, c    constraint/bn254/solver.go:116-121
  Verify                    backend/groth16/bn254/verify.go:65-137, read in the exponent (every element of a key made
                            by b200_groth16_setup from a known tau, alpha, beta, gamma, delta has a known discrete log)
No oracle import: this is product-side host code; field vectors go through the library (b200_fr_vec_mul,
b200_ntt_dev, b200_fr_poly_eval_dev, b200_batch_scalar_mul)."""
from __future__ import annotations

import numpy as np

from .api import BLS12_377, BN254, Context, DeviceVector, DIT
from .plonk_host import bitrev_indices

_CURVES = {
    BN254: dict(r=0x30644E72E131A029B85045B68181585D2833E84879B9709143E1F593F0000001,
                root=19103219067921713944291392827692070036145651957329286315305642004821462161904, adicity=28, coset=5),
    BLS12_377: dict(r=0x12AB655E9A2CA55660B44D1E5C37B00159AA76FED00000010A11800000000001, root=None, adicity=47, coset=22),
}
_CURVES[BLS12_377]["root"] = pow(22, (_CURVES[BLS12_377]["r"] - 1) >> 47, _CURVES[BLS12_377]["r"])
MONT = 1 << 256


class Field:
    """Fr of a curve: limb conversions and the fft.Domain constants the ABI takes as inputs."""

    def __init__(self, curve: int = BN254):
        c = _CURVES[curve]
        self.curve, self.r, self.root, self.adicity, self.coset = curve, c["r"], c["root"], c["adicity"], c["coset"]
        self.r2_raw = np.frombuffer((MONT * MONT % self.r).to_bytes(32, "little"), dtype=np.uint64).copy()

    def limbs(self, x: int) -> np.ndarray:
        return np.frombuffer(((x % self.r) * MONT % self.r).to_bytes(32, "little"), dtype=np.uint64).copy()

    def value(self, limbs) -> int:
        return int.from_bytes(np.ascontiguousarray(limbs, dtype=np.uint64).tobytes(), "little") * pow(MONT, -1, self.r) % self.r

    def generator(self, log_n: int) -> int:
        return pow(self.root, 1 << (self.adicity - log_n), self.r)

    def to_mont(self, ctx: Context, canonical: np.ndarray) -> np.ndarray:
        """(n, 4) canonical little-endian limbs -> Montgomery form: one Montgomery product by R^2 on the device."""
        return ctx.fr_vec_mul(canonical, np.broadcast_to(self.r2_raw, canonical.shape).copy())

    def mont_sum(self, arr) -> int:
        """Montgomery form of the sum of Montgomery-form elements (exact, vectorised)."""
        h = np.ascontiguousarray(arr).view(np.uint32).reshape(-1, 8)
        return sum(int(h[:, j].sum(dtype=np.uint64)) << (32 * j) for j in range(8)) % self.r


# ---------------------------------------------------------------------------------------------------------------
# refCircuit as CSR matrices over constraint.Term {CID, VID}
# ---------------------------------------------------------------------------------------------------------------
def mul_chain_r1cs(log_n: int) -> dict:
    """nb_mul = 2^log_n - 1 multiplications: wires 0 = ONE, 1 = Y (public), 2 = X (secret), 3.. = X_1..X_N; constraint
    j < N: X_j * X_j = X_{j+1}; constraint N: ONE * X_N = Y.  One coefficient (1) with id 0."""
    m = 1 << log_n
    nb_mul = m - 1
    row_ptr = np.arange(m + 1, dtype=np.uint64)
    prev = np.arange(2, 2 + nb_mul, dtype=np.uint32)              # X_j of constraint j
    terms = lambda wires: np.stack([np.zeros(m, dtype=np.uint32), wires.astype(np.uint32)], axis=1)
    l = terms(np.concatenate([prev, [0]]))
    r = terms(np.concatenate([prev, [2 + nb_mul]]))
    o = terms(np.concatenate([prev + 1, [1]]))
    return dict(log_n=log_n, nb_constraints=m, nb_wires=m + 2, nb_public=2, nb_mul=nb_mul,
                matrices=[(row_ptr, l), (row_ptr, r), (row_ptr, o)])


def mul_chain_witness(ctx: Context, fld: Field, log_n: int, x: int = 2, alloc=None):
    """(w, a, b, c) of the chain in Montgomery form; alloc(shape) -> array (e.g. page-locked) for the four outputs."""
    m = 1 << log_n
    nb_mul = m - 1
    r = fld.r
    buf = bytearray(32 * (nb_mul + 1))
    cur = x % r
    buf[0:32] = cur.to_bytes(32, "little")
    for j in range(1, nb_mul + 1):
        cur = cur * cur % r
        buf[32 * j:32 * j + 32] = cur.to_bytes(32, "little")
    chain = fld.to_mont(ctx, np.frombuffer(bytes(buf), dtype=np.uint64).reshape(-1, 4))      # X, X_1 .. X_N
    alloc = alloc or (lambda shape: np.empty(shape, dtype=np.uint64))
    w, a, b, c = alloc((m + 2, 4)), alloc((m, 4)), alloc((m, 4)), alloc((m, 4))
    one = fld.limbs(1)
    w[0], w[1], w[2:] = one, chain[nb_mul], chain
    a[:nb_mul], a[nb_mul] = chain[:nb_mul], one
    b[:nb_mul], b[nb_mul] = chain[:nb_mul], chain[nb_mul]
    c[:nb_mul], c[nb_mul] = chain[1:], chain[nb_mul]
    return w, a, b, c


def sample_toxic(fld: Field, rng: np.random.Generator) -> dict:
    """tau, alpha, beta, gamma, delta as integers (sampleToxicWaste, setup.go:435-470; test-style: seeded)."""
    return {k: int.from_bytes(rng.bytes(40), "little") % (fld.r - 1) + 1 for k in ("tau", "alpha", "beta", "gamma", "delta")}


def setup_mul_chain(ctx: Context, fld: Field, log_n: int, toxic: dict, g1_gen, g2_gen, **kw):
    cs = mul_chain_r1cs(log_n)
    pk, res = ctx.groth16_setup(log_n=log_n, generator=fld.limbs(fld.generator(log_n)), coset_gen=fld.limbs(fld.coset),
                                nb_wires=cs["nb_wires"], nb_public=cs["nb_public"], matrices=cs["matrices"],
                                coeffs=fld.limbs(1).reshape(1, 4), toxic={k: fld.limbs(v) for k, v in toxic.items()},
                                g1_gen=g1_gen, g2_gen=g2_gen, **kw)
    return cs, pk, res


# ---------------------------------------------------------------------------------------------------------------
# acceptance in the exponent
# ---------------------------------------------------------------------------------------------------------------
def poly_at_tau(ctx: Context, fld: Field, evals: np.ndarray, log_n: int, tau: int, br=None) -> int:
    """p(tau) for the polynomial with p(w^j) = evals[j] (zero beyond len(evals)): FFTInverse + Horner, both on the device."""
    n = 1 << log_n
    br = bitrev_indices(log_n) if br is None else br
    full = np.zeros((n, 4), dtype=np.uint64)
    full[:evals.shape[0]] = evals
    v = DeviceVector(ctx, n * 32).upload(np.ascontiguousarray(full[br]))
    ctx.ntt_dev(v, log_n, 1, DIT, 0, fld.limbs(fld.generator(log_n)), fld.limbs(fld.coset))      # bit-reversed in, natural out
    val = ctx.fr_poly_eval_dev(v, n, fld.limbs(tau))
    v.free()
    return fld.value(val)


def lagrange_at_tau(fld: Field, log_n: int, j: int, tau: int) -> int:
    """L_j(tau) = (tau^n - 1) / n * w^j / (tau - w^j)   (setup.go:373-382)."""
    n, r = 1 << log_n, fld.r
    wj = pow(fld.generator(log_n), j, r)
    return (pow(tau, n, r) - 1) * pow(n, -1, r) % r * wj % r * pow((tau - wj) % r, -1, r) % r


def mul_chain_expected_dlogs(ctx: Context, fld: Field, log_n: int, toxic: dict, w, a, b, c) -> dict:
    """Discrete logs of the five MSM results a correct prover returns for the chain's key: <A, w> = A(tau) with
    A = the polynomial interpolating the solver's a (prove.go:214), likewise B; <K, w_private> = (beta A_priv + alpha
    B_priv + C_priv)(tau) / delta (:244); <Z, h> = (A B - C)(tau) / delta (:234) -- h never enters: this is the
    statement 'the witness satisfies the R1CS', independent of the device's computeH."""
    r, tau = fld.r, toxic["tau"]
    br = bitrev_indices(log_n)
    At, Bt, Ct = (poly_at_tau(ctx, fld, v, log_n, tau, br) for v in (a, b, c))
    m = 1 << log_n
    last = lagrange_at_tau(fld, log_n, m - 1, tau)            # the constraint ONE * X_N = Y
    w0, w1 = fld.value(w[0]), fld.value(w[1])
    a_pub, b_pub, c_pub = w0 * last % r, 0, w1 * last % r       # ONE sits in L, Y in O of the last constraint only
    dinv = pow(toxic["delta"], -1, r)
    k = (toxic["beta"] * (At - a_pub) + toxic["alpha"] * (Bt - b_pub) + (Ct - c_pub)) % r * dinv % r
    z = (At * Bt - Ct) % r * dinv % r
    vk_k = [toxic["beta"] * last % r * pow(toxic["gamma"], -1, r) % r, last * pow(toxic["gamma"], -1, r) % r]   # ONE, Y
    return dict(ar=At, bs1=Bt, bs2=Bt, krs=k, krs2=z, vk_k=vk_k, public=[w0, w1])


def expected_proof_dlogs(fld: Field, toxic: dict, msm: dict, r_blind: int, s_blind: int) -> dict:
    """prove.go:206-207,219-221,248-271,294-299 in the exponent."""
    q = fld.r
    ar = (toxic["alpha"] + msm["ar"] + r_blind * toxic["delta"]) % q
    bs = (toxic["beta"] + msm["bs1"] + s_blind * toxic["delta"]) % q
    krs = (msm["krs"] + msm["krs2"] + s_blind * ar + r_blind * bs - r_blind * s_blind % q * toxic["delta"]) % q
    return dict(ar=ar, bs=bs, krs=krs)


def verify_equation(fld: Field, toxic: dict, msm: dict, proof: dict) -> bool:
    """verify.go:65-137 in the exponent: e(Ar, Bs) = e(alpha, beta) e(sum_pub w_i K_i, gamma) e(Krs, delta)."""
    q = fld.r
    ksum = sum(x * k for x, k in zip(msm["public"], msm["vk_k"])) % q
    return (proof["ar"] * proof["bs"] - toxic["alpha"] * toxic["beta"] - ksum * toxic["gamma"] - proof["krs"] * toxic["delta"]) % q == 0


def points_match(ctx: Context, pkg, fld: Field, g1_gen, g2_gen, out: dict, want: dict) -> dict:
    """{name: bool}: each Jacobian MSM result equals [want[name]]G (G2 for bs2), compared after affine normalisation."""
    res = {}
    for name, group in (("ar", 1), ("bs1", 1), ("krs", 1), ("krs2", 1), ("bs2", 2)):
        exp = ctx.batch_scalar_mul(g1_gen if group == 1 else g2_gen, fld.limbs(want[name]).reshape(1, 4), group)[0]
        got = (pkg.g1_jac_to_affine if group == 1 else pkg.g2_jac_to_affine)(out[name], curve=fld.curve)
        res[name] = bool(np.array_equal(got, exp))
    return res
