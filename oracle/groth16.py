"""CPU oracle (TEST INFRASTRUCTURE ONLY) -- BN254 Groth16 Setup/Prove restatement with a
KNOWN trapdoor, and the pairing-free acceptance check built on it.

Follows (file:line under /root/reference):
  * circuit shape      backend/groth16/groth16_test.go:120-156 (refCircuit: X <- X*X chain,
                       then AssertIsEqual = R1C(ONE, X_N, Y), frontend/cs/r1cs/api_assertions.go:31-36)
  * Setup              backend/groth16/bn254/setup.go:85-337 (+ setupABC :352-433)
  * Prove              backend/groth16/bn254/prove.go:63-322 (fix-ups :192-301)
  * Verify equation    backend/groth16/bn254/verify.go:65-137
  * proof wire format  backend/groth16/bn254/marshal.go:41-66  (order Ar, Bs, Krs)

PARITY UNPINNED vs gnark itself (no Go toolchain, gnark-crypto not vendored; the reference
tests hold no golden proof).  Acceptance is proven by `trapdoor_check`: with tau, alpha,
beta, gamma, delta known every proof element has a known discrete log, so
e(Ar,Bs) == e(alpha,beta) e(sum_pub, gamma) e(Krs, delta) becomes an identity in Fr.
"""
from __future__ import annotations

from dataclasses import dataclass, field

from . import bn254 as C
from . import ntt
from .bn254 import R, P


# ----------------------------------------------------------------------------------------
# R1CS (just enough of constraint/bn254/system.go for linear-combination constraints)
# ----------------------------------------------------------------------------------------
@dataclass
class R1CS:
    nb_public: int  # includes the ONE wire (gnark: GetNbPublicVariables)
    nb_secret: int
    nb_internal: int
    constraints: list  # [(L, R, O)], each a list of (coeff:int, wire:int)
    # api.Commit circuits (constraint.Groth16Commitments): [(commitment wire, [private committed wires, ascending])]
    commitments: list = field(default_factory=list)

    @property
    def nb_wires(self):
        return self.nb_public + self.nb_secret + self.nb_internal

    @property
    def nb_constraints(self):
        return len(self.constraints)


def mul_chain_circuit(nb_mul: int) -> R1CS:
    """refCircuit (groth16_test.go:120-132). Wires: 0=ONE, 1=Y(public), 2=X(secret),
    3..2+nb_mul = X_1..X_N. Constraint j<N: X_{j} * X_{j} = X_{j+1}; last: ONE * X_N = Y."""
    cons = []
    prev = 2
    for j in range(nb_mul):
        new = 3 + j
        cons.append(([(1, prev)], [(1, prev)], [(1, new)]))
        prev = new
    cons.append(([(1, 0)], [(1, prev)], [(1, 1)]))
    return R1CS(nb_public=2, nb_secret=1, nb_internal=nb_mul, constraints=cons)


def mul_chain_witness(nb_mul: int, x: int = 2):
    """Full wire vector W and the per-constraint A,B,C the solver emits
    (constraint/bn254/solver.go:116-121: a=L.w, b=R.w, c=O.w)."""
    w = [1, 0, x % R]
    cur = x % R
    a, b, c = [], [], []
    for _ in range(nb_mul):
        nxt = cur * cur % R
        a.append(cur); b.append(cur); c.append(nxt)
        w.append(nxt)
        cur = nxt
    w[1] = cur
    a.append(1); b.append(cur); c.append(cur)
    return w, a, b, c


def solve(r1cs: R1CS, w: list[int]):
    """Evaluate a, b, c for a fully assigned wire vector (no hint solving: test circuits are
    generated together with their witness)."""
    def lc(t):
        return sum(k * w[i] for k, i in t) % R
    a = [lc(L) for L, _, _ in r1cs.constraints]
    b = [lc(Rr) for _, Rr, _ in r1cs.constraints]
    c = [lc(O) for _, _, O in r1cs.constraints]
    return a, b, c


# ----------------------------------------------------------------------------------------
# Setup with a known trapdoor (setup.go:85-337)
# ----------------------------------------------------------------------------------------
@dataclass
class Trapdoor:
    t: int
    alpha: int
    beta: int
    gamma: int
    delta: int


@dataclass
class ProvingKey:
    domain: ntt.Domain
    g1_alpha: tuple
    g1_beta: tuple
    g1_delta: tuple
    g1_A: list
    g1_B: list
    g1_Z: list
    g1_K: list
    g2_beta: tuple
    g2_delta: tuple
    g2_B: list
    infinity_A: list
    infinity_B: list
    nb_infinity_A: int = 0
    nb_infinity_B: int = 0
    # scalars behind the bases (only a trapdoor setup has these; used by trapdoor_check)
    sc: dict = field(default_factory=dict)


class _FixedBase:
    """k*G by adding precomputed 2^j*G (test-size only)."""

    def __init__(self, add, gen):
        self.add = add
        self.tab = [gen]
        for _ in range(255):
            self.tab.append(add(self.tab[-1], self.tab[-1]))

    def mul(self, k: int):
        k %= R
        acc = None
        j = 0
        while k:
            if k & 1:
                acc = self.add(acc, self.tab[j])
            k >>= 1
            j += 1
        return acc


_FB1 = None
_FB2 = None


def g1_gen_mul(k):
    global _FB1
    if _FB1 is None:
        _FB1 = _FixedBase(C.g1_add, C.G1_GEN)
    return _FB1.mul(k)


def g2_gen_mul(k):
    global _FB2
    if _FB2 is None:
        _FB2 = _FixedBase(C.g2_add, C.G2_GEN)
    return _FB2.mul(k)


def setup_abc(r1cs: R1CS, d: ntt.Domain, t: int):
    """setupABC (setup.go:352-433): A_i(t), B_i(t), C_i(t) via Lagrange basis at t."""
    n = d.cardinality
    A = [0] * r1cs.nb_wires
    B = [0] * r1cs.nb_wires
    Cc = [0] * r1cs.nb_wires
    zt = (pow(t, n, R) - 1) % R
    # [t - w^j] and its inverses in one batch inversion, as setup.go:362-371 does (fr.BatchInvert)
    ws, wi = [], 1
    for _ in r1cs.constraints:
        ws.append(wi)
        wi = wi * d.generator % R
    pre, acc = [], 1
    for wj in ws:
        pre.append(acc)
        acc = acc * ((t - wj) % R) % R
    inv = pow(acc, -1, R)
    tinv = [0] * len(ws)
    for j in range(len(ws) - 1, -1, -1):
        tinv[j] = inv * pre[j] % R
        inv = inv * ((t - ws[j]) % R) % R
    scale = zt * d.cardinality_inv % R
    for j, (L, Rr, O) in enumerate(r1cs.constraints):
        # L_j(t) = (t^n - 1)/n * w^j / (t - w^j)
        wi = ws[j]
        lag = scale * wi % R * tinv[j] % R
        for k, i in L:
            A[i] = (A[i] + k * lag) % R
        for k, i in Rr:
            B[i] = (B[i] + k * lag) % R
        for k, i in O:
            Cc[i] = (Cc[i] + k * lag) % R
        wi = wi * d.generator % R
    return A, B, Cc


def setup(r1cs: R1CS, td: Trapdoor, points: bool = True) -> tuple[ProvingKey, dict]:
    """Setup (setup.go:85-337) without commitments. Returns (pk, vk-dict)."""
    d = ntt.Domain(r1cs.nb_constraints)
    n = d.cardinality
    A, B, Cc = setup_abc(r1cs, d, td.t)
    gamma_inv = pow(td.gamma, -1, R)
    delta_inv = pow(td.delta, -1, R)
    nb_pub = r1cs.nb_public
    vkK = []
    pkK = []
    # setup.go:155-200: commitment wires join vk.K (after the public wires, in wire order); private wires committed to
    # by commitment j go to its Pedersen basis ckK[j]; both are divided by gamma.  Everything else: pk.K, divided by delta.
    commit_wires = {cw for cw, _ in r1cs.commitments}
    committed_by = {w: j for j, (_, ws_) in enumerate(r1cs.commitments) for w in ws_}
    ckK = [[] for _ in r1cs.commitments]
    for i in range(r1cs.nb_wires):
        k = (A[i] * td.beta + B[i] * td.alpha + Cc[i]) % R
        if i < nb_pub or i in commit_wires:
            vkK.append(k * gamma_inv % R)
        elif i in committed_by:
            ckK[committed_by[i]].append(k * gamma_inv % R)
        else:
            pkK.append(k * delta_inv % R)
    zdt = (pow(td.t, n, R) - 1) * delta_inv % R
    Z = []
    for _ in range(n):
        Z.append(zdt)
        zdt = zdt * td.t % R
    inf_a = [x == 0 for x in A]
    inf_b = [x == 0 for x in B]
    A_f = [x for x in A if x]
    B_f = [x for x in B if x]
    ntt.bit_reverse(Z)          # setup.go:265
    Z = Z[: n - 1]              # setup.go:266-267
    mul1 = g1_gen_mul if points else (lambda k: None)
    mul2 = g2_gen_mul if points else (lambda k: None)
    pk = ProvingKey(
        domain=d,
        g1_alpha=mul1(td.alpha), g1_beta=mul1(td.beta), g1_delta=mul1(td.delta),
        g1_A=[mul1(x) for x in A_f], g1_B=[mul1(x) for x in B_f],
        g1_Z=[mul1(x) for x in Z], g1_K=[mul1(x) for x in pkK],
        g2_beta=mul2(td.beta), g2_delta=mul2(td.delta), g2_B=[mul2(x) for x in B_f],
        infinity_A=inf_a, infinity_B=inf_b,
        nb_infinity_A=sum(inf_a), nb_infinity_B=sum(inf_b),
        sc=dict(A=A_f, B=B_f, Z=Z, K=pkK, vkK=vkK, ckK=ckK),
    )
    vk = dict(vkK=vkK, alpha=td.alpha, beta=td.beta, gamma=td.gamma, delta=td.delta,
              commit_wires=sorted(commit_wires), committed=[list(ws_) for _, ws_ in r1cs.commitments])
    return pk, vk


# ----------------------------------------------------------------------------------------
# Prove (prove.go:63-322) -- the MSMs are done on dlogs when `dlog=True` (exact at any size)
# ----------------------------------------------------------------------------------------
def filter_wires(w, infinity):
    """prove.go:154-175: drop wires whose base is the point at infinity."""
    return [x for x, inf in zip(w, infinity) if not inf]


def prove_scalars(pk: ProvingKey, w, a, b, c, r: int, s: int, nb_public: int, to_remove=(), h=None):
    """The five MSM scalar vectors of prove.go:201,214,234,244,290 in MSM order.  to_remove: wire ids of the
    private committed wires and the commitment wires, dropped from the K scalars (prove.go:238-242); h: computeH's
    output when the caller already has it (e.g. from the C++ restatement at sizes Python cannot reach)."""
    if h is None:
        h = ntt.compute_h(a, b, c, pk.domain)
    n = pk.domain.cardinality
    drop = set(to_remove)
    return dict(
        A=filter_wires(w, pk.infinity_A),
        B=filter_wires(w, pk.infinity_B),
        K=[x for i, x in enumerate(w) if i >= nb_public and i not in drop],
        Z=h[: n - 1],
    )


def prove_dlog(pk: ProvingKey, td: Trapdoor, w, a, b, c, r: int, s: int, nb_public: int, to_remove=(), h=None):
    """Discrete logs (w.r.t. the G1/G2 generators) of the proof elements Prove must output:
    ar = alpha + <A,w> + r*delta; bs = beta + <B,w> + s*delta;
    krs = <K,w_priv> + <Z,h> + s*ar + r*bs - r*s*delta   (prove.go:206-207,219-221,248-271,294-299)."""
    sc = prove_scalars(pk, w, a, b, c, r, s, nb_public, to_remove, h)
    dot = lambda x, y: sum(p * q for p, q in zip(x, y)) % R
    msm = dict(A=dot(pk.sc["A"], sc["A"]), B=dot(pk.sc["B"], sc["B"]),
               K=dot(pk.sc["K"], sc["K"]), Z=dot(pk.sc["Z"], sc["Z"]))
    ar = (td.alpha + msm["A"] + r * td.delta) % R
    bs = (td.beta + msm["B"] + s * td.delta) % R
    krs = (msm["K"] + msm["Z"] + s * ar + r * bs - r * s % R * td.delta) % R
    return dict(ar=ar, bs=bs, krs=krs, msm=msm, scalars=sc)


def commitment_dlogs(pk: ProvingKey, vk: dict, w) -> list:
    """Discrete logs of proof.Commitments[j] = pk.CommitmentKeys[j].Commit(private committed values) (prove.go:95):
    a Pedersen commitment is the MultiExp of the values against the basis ckK[j]."""
    return [sum(w[i] * k for i, k in zip(ws_, pk.sc["ckK"][j])) % R for j, ws_ in enumerate(vk["committed"])]


def trapdoor_check(vk: dict, public_w, ar: int, bs: int, krs: int, commitment_values=(), commitments=()) -> bool:
    """verify.go:65-137 in the exponent: ar*bs == alpha*beta + ksum*gamma + krs*delta,
    ksum = sum_{i public} w_i * vkK_i (verify.go:111-115; public_w includes the ONE wire) -- for api.Commit circuits
    the public witness is extended by the commitment wires' values (:75-99) and every proof.Commitments[j] is added
    to ksum (:117-119).  (The Pedersen proof of knowledge, :101-107, is a host-side check outside this path.)"""
    ksum = (sum(x * k for x, k in zip(list(public_w) + list(commitment_values), vk["vkK"])) + sum(commitments)) % R
    return (ar * bs - vk["alpha"] * vk["beta"] - ksum * vk["gamma"] - krs * vk["delta"]) % R == 0


def prove_points(pk: ProvingKey, w, a, b, c, r: int, s: int, nb_public: int):
    """prove.go:192-301 on actual points (naive MSM; test sizes only). Returns affine
    (Ar, Bs, Krs)."""
    sc = prove_scalars(pk, w, a, b, c, r, s, nb_public)
    kr = (-(r * s)) % R
    d_r, d_s, d_kr = (C.g1_mul(pk.g1_delta, k) for k in (r, s, kr))   # prove.go:192
    ar = C.msm_g1(pk.g1_A, sc["A"])
    ar = C.g1_add(C.g1_add(ar, pk.g1_alpha), d_r)                     # :219-220
    bs1 = C.msm_g1(pk.g1_B, sc["B"])
    bs1 = C.g1_add(C.g1_add(bs1, pk.g1_beta), d_s)                    # :206-207
    krs = C.msm_g1(pk.g1_K, sc["K"])
    krs2 = C.msm_g1(pk.g1_Z, sc["Z"])
    krs = C.g1_add(krs, d_kr)                                         # :248
    krs = C.g1_add(krs, krs2)                                         # :257
    krs = C.g1_add(krs, C.g1_mul(ar, s))                              # :263-264
    krs = C.g1_add(krs, C.g1_mul(bs1, r))                             # :270-271
    bs2 = C.msm_g2(pk.g2_B, sc["B"])
    bs2 = C.g2_add(C.g2_add(bs2, C.g2_mul(pk.g2_delta, s)), pk.g2_beta)   # :294-297
    return ar, bs2, krs


# ----------------------------------------------------------------------------------------
# Proof.WriteTo (marshal.go:41-66): compressed Ar | Bs | Krs | u32 len(Commitments) | Pok
# gnark-crypto point compression (recalled, source absent): 32-byte big-endian X with the two
# top bits of byte 0 = 0b10 (y is the lexicographically smallest), 0b11 (largest), 0b01
# (infinity); G2 = X.A1 | X.A0, 64 bytes, Y order decided on A1 then A0.
# ----------------------------------------------------------------------------------------
# BLS12-377 (oracle/bls12_377.py exports COMPRESS_FLAGS): 48-byte coordinates and the three-bit flags
# gnark-crypto uses on its BLS curves, 0b100 / 0b101 / 0b110 in the top three bits (recalled as well).
M_SMALLEST, M_LARGEST, M_INFINITY = getattr(C, "COMPRESS_FLAGS", (0b10 << 6, 0b11 << 6, 0b01 << 6))
FP_BYTES = (P.bit_length() + 7) // 8


def _lex_largest(y: int) -> bool:
    return y > (P - 1) // 2


def g1_compress(p) -> bytes:
    if p is None:
        return bytes([M_INFINITY]) + bytes(FP_BYTES - 1)
    b = bytearray(p[0].to_bytes(FP_BYTES, "big"))
    b[0] |= M_LARGEST if _lex_largest(p[1]) else M_SMALLEST
    return bytes(b)


def g2_compress(p) -> bytes:
    if p is None:
        return bytes([M_INFINITY]) + bytes(2 * FP_BYTES - 1)
    (x0, x1), (y0, y1) = p
    largest = _lex_largest(y1) if y1 else _lex_largest(y0)
    b = bytearray(x1.to_bytes(FP_BYTES, "big") + x0.to_bytes(FP_BYTES, "big"))
    b[0] |= M_LARGEST if largest else M_SMALLEST
    return bytes(b)


def proof_bytes(ar, bs, krs) -> bytes:
    """Proof.WriteTo for a proof without commitments: Ar, Bs, Krs, u32 0, Pok = infinity."""
    return g1_compress(ar) + g2_compress(bs) + g1_compress(krs) + (0).to_bytes(4, "big") + g1_compress(None)
