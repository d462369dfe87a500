"""CPU restatement of the BN254 PlonK prover and verifier (TEST INFRASTRUCTURE -- the product never
imports this; PARITY UNPINNED against gnark-crypto itself: the reference is Go, gnark-crypto is not
vendored and there is no Go toolchain here).

What is followed, statement by statement:
  Setup / BuildTrace / buildPermutation / computePermutationPolynomials / commitTrace
        /root/reference/backend/plonk/bn254/setup.go:110-161,176-231,275-420
  evaluateLROSmallDomain (the solver's L, R, O)
        /root/reference/constraint/bn254/system.go:221-265
  Prove and every stage of it                /root/reference/backend/plonk/bn254/prove.go:116-1389
  Verify, bindPublicData, deriveRandomness   /root/reference/backend/plonk/bn254/verify.go:45-360
  Proof.WriteTo                              /root/reference/backend/plonk/bn254/marshal.go:41-66

gnark-crypto pieces that are absent and restated from their call sites plus an IN-TREE pin:
  fiat-shamir transcript (sha256(name || previous challenge, NOT reduced || bindings)), the KZG
  folding challenge (sha256("gamma" || zeta || digests || claimed values || Z(w zeta))) and the
  fold sum gamma^i p_i -- all three are restated by the reference's own Solidity verifier:
  /root/reference/backend/plonk/bn254/solidity.go:407-462 (gamma), :469-483 (beta), :492-517 (alpha),
  :524-538 (zeta), :980-1032 (KZG gamma), :906-961 (fold).  tests/test_plonk_prover_oracle.py checks the
  transcript here against those byte layouts.
  kzg.Commit / Open / BatchOpenSinglePoint: commitment = MSM against the SRS; here the SRS comes from a
  KNOWN tau, so a commitment is [p(tau)]G1 (one scalar multiplication) and an opening check is
  C - [v]G1 == [tau - z]H, pairing-free and equivalent to kzg.Verify accepting.

BSB22 commitments (api.Commit circuits, prove.go:303-352, frontend/cs/scs/api.go:623-655) are restated too: Qcp / Pi2
columns, the commitment hint (kzg.Commit against KzgLagrange + hash_to_field: RFC 9380 expand_message_xmd over SHA-256
with DST "BSB22-Plonk", gnark-crypto fr/hash_to_field -- absent, unpinned) and the verifier's extra PI term.
"""
import hashlib
from dataclasses import dataclass, field

from . import bn254 as B
from . import ntt as N
from . import plonk as PB
from .groth16 import g1_gen_mul, g1_compress

R = B.R


# ---------------------------------------------------------------------------------------------
# constraint system (the subset of cs.SparseR1CS the prover reads)
# ---------------------------------------------------------------------------------------------
@dataclass
class SparseR1CS:
    nb_public: int
    nb_secret: int
    nb_internal: int
    # one tuple per constraint: (qL, qR, qM, qO, qC, xa, xb, xc)   qL*a + qR*b + qM*a*b + qO*c + qC = 0
    constraints: list = field(default_factory=list)
    # constraint.PlonkCommitments: [{"committed": [constraint indices], "index": constraint index}]
    commitments: list = field(default_factory=list)

    @property
    def nb_variables(self):
        return self.nb_public + self.nb_secret + self.nb_internal


def mul_chain_scs(nb_mul: int) -> SparseR1CS:
    """The SCS form of refCircuit (backend/plonk/plonk_test.go:248-260 shape): public Y, secret X,
    nb_mul times x <- x*x, then AssertIsEqual(x, Y).  Wires: [public | secret | internal]."""
    spr = SparseR1CS(1, 1, nb_mul)
    cur = 1
    for j in range(nb_mul):
        new = 2 + j
        spr.constraints.append((0, 0, 1, R - 1, 0, cur, cur, new))
        cur = new
    spr.constraints.append((1, R - 1, 0, 0, 0, cur, 0, 0))
    return spr


def commit_chain_scs(nb_mul: int, nb_commit: int = 1) -> SparseR1CS:
    """A Mul chain that also calls api.Commit (frontend/cs/scs/api.go:623-655) nb_commit times on (X, x_1), multiplies
    the chain's end by every commitment value and asserts the product equals the public Y.
    Wires: 0 = Y, 1 = X, then internal: chain, then per commitment (c_k, product)."""
    spr = SparseR1CS(1, 1, 0)
    nxt = 2
    cur = 1
    chain = []
    for _ in range(nb_mul):
        spr.constraints.append((0, 0, 1, R - 1, 0, cur, cur, nxt))
        cur = nxt
        chain.append(cur)
        nxt += 1
    for k in range(nb_commit):
        committed = []
        for v in (1, chain[0]):                                   # - v + Qcp * Pi2 = 0   (api.go:630-636)
            committed.append(len(spr.constraints))
            spr.constraints.append((R - 1, 0, 0, 0, 0, v, 0, 0))
        c = nxt
        nxt += 1
        index = len(spr.constraints)
        spr.constraints.append((R - 1, 0, 0, 0, 0, c, 0, 0))      # - c + qk = 0, qk injected by the prover (api.go:646-649)
        spr.commitments.append({"committed": committed, "index": index})
        spr.constraints.append((0, 0, 1, R - 1, 0, c, cur, nxt))  # use the commitment
        cur = nxt
        nxt += 1
    spr.constraints.append((1, R - 1, 0, 0, 0, cur, 0, 0))
    spr.nb_internal = nxt - 2
    return spr


def expand_message_xmd(msg: bytes, dst: bytes, length: int) -> bytes:
    """RFC 9380 5.3.1 with SHA-256."""
    ell = (length + 31) // 32
    dst_prime = dst + bytes([len(dst)])
    b0 = hashlib.sha256(bytes(64) + msg + length.to_bytes(2, "big") + b"\x00" + dst_prime).digest()
    bi = hashlib.sha256(b0 + b"\x01" + dst_prime).digest()
    out = bi
    for i in range(2, ell + 1):
        bi = hashlib.sha256(bytes(a ^ b for a, b in zip(b0, bi)) + bytes([i]) + dst_prime).digest()
        out += bi
    return out[:length]


def hash_commitment(point) -> int:
    """bsb22Hint's tail (prove.go:341-350): hash_to_field("BSB22-Plonk") of the commitment's Marshal()."""
    return int.from_bytes(expand_message_xmd(g1_raw(point), b"BSB22-Plonk", 48), "big") % R


def commit_chain_solution(spr: SparseR1CS, pk, x: int, bsb_blinding):
    """Solver + bsb22Hint (prove.go:303-352) for commit_chain_scs.  bsb_blinding[k] = the two values SetRandom
    draws for commitment k (prove.go:329,332).  Returns (solution, bsb dict for prove())."""
    vals = [0] * spr.nb_variables
    vals[1] = x % R
    n = pk.d0.cardinality
    off = spr.nb_public
    _, lag = srs_scalars(pk)
    bsb = {"pi2": [], "commitments": [], "values": []}
    depth = 0
    for j, c in enumerate(spr.constraints):
        info = spr.commitments[depth] if depth < len(spr.commitments) else None
        if info is not None and j == info["index"]:
            cv = [0] * n
            for row in info["committed"]:
                cv[off + row] = vals[spr.constraints[row][5]]
            cv[off + info["index"]] = bsb_blinding[depth][0]
            cv[off + len(spr.constraints) - 1] = bsb_blinding[depth][1]
            com = g1_gen_mul(sum(a * b for a, b in zip(cv, lag)) % R)
            h = hash_commitment(com)
            vals[c[5]] = h
            bsb["pi2"].append(cv)
            bsb["commitments"].append(com)
            bsb["values"].append(h)
            depth += 1
        elif c[2] == 1:                                            # a * b = c
            vals[c[7]] = vals[c[5]] * vals[c[6]] % R
        elif c[0] == 1 and c[1] == R - 1:                          # the final AssertIsEqual fixes Y
            vals[0] = vals[c[5]]
    return vals, bsb


def mul_chain_solution(nb_mul: int, x: int = 2):
    """values of [public | secret | internal] for mul_chain_scs."""
    vals = [0, x]
    cur = x
    for _ in range(nb_mul):
        cur = cur * cur % R
        vals.append(cur)
    vals[0] = cur
    return vals


def evaluate_lro(spr: SparseR1CS, solution):
    """evaluateLROSmallDomain, system.go:221-265."""
    s = N.Domain(len(spr.constraints) + spr.nb_public).cardinality
    s0 = solution[0]
    l, r, o = [s0] * s, [s0] * s, [s0] * s
    for i in range(spr.nb_public):
        l[i] = solution[i]
    off = spr.nb_public
    for j, c in enumerate(spr.constraints):
        l[off + j], r[off + j], o[off + j] = solution[c[5]], solution[c[6]], solution[c[7]]
    return l, r, o


# ---------------------------------------------------------------------------------------------
# setup
# ---------------------------------------------------------------------------------------------
@dataclass
class ProvingKey:
    d0: N.Domain
    d1: N.Domain
    nb_public: int
    # trace, canonical / regular (setup.go:233-247)
    ql: list
    qr: list
    qm: list
    qo: list
    qk: list            # incomplete: zero at the public rows
    s1: list
    s2: list
    s3: list
    perm: list          # trace.S
    qk_lagrange: list   # what completeQk (prove.go:397-423) gets back by ToLagrange
    vk: dict
    tau: int
    qcp: list = field(default_factory=list)      # trace.Qcp, canonical
    commitments: list = field(default_factory=list)


def build_trace(spr: SparseR1CS, size: int):
    """BuildTrace, setup.go:176-231 (Lagrange / regular)."""
    ql, qr, qm, qo, qk = ([0] * size for _ in range(5))
    for i in range(spr.nb_public):
        ql[i] = R - 1
    off = spr.nb_public
    for j, c in enumerate(spr.constraints):
        ql[off + j], qr[off + j], qm[off + j], qo[off + j], qk[off + j] = c[0], c[1], c[2], c[3], c[4]
    return ql, qr, qm, qo, qk


def build_permutation(spr: SparseR1CS, size: int):
    """buildPermutation, setup.go:304-358."""
    lro = [0] * (3 * size)
    for i in range(spr.nb_public):
        lro[i] = i
    off = spr.nb_public
    for j, c in enumerate(spr.constraints):
        lro[off + j], lro[size + off + j], lro[2 * size + off + j] = c[5], c[6], c[7]
    perm = [-1] * (3 * size)
    cycle = [-1] * spr.nb_variables
    for i, v in enumerate(lro):
        if cycle[v] != -1:
            perm[i] = cycle[v]
        cycle[v] = i
    for i in range(3 * size):
        if perm[i] == -1:
            perm[i] = cycle[lro[i]]
    return perm


def support_permutation(d: N.Domain):
    """getSupportPermutation, setup.go:398-415: <g> || u<g> || u^2<g>."""
    n = d.cardinality
    res = [0] * (3 * n)
    res[0], res[n], res[2 * n] = 1, d.coset_gen, d.coset_gen * d.coset_gen % R
    for i in range(1, n):
        for k in range(3):
            res[k * n + i] = res[k * n + i - 1] * d.generator % R
    return res


def _to_canonical(lagrange, d):
    a = list(lagrange)
    N.fft_inverse(a, d, N.DIF)
    N.bit_reverse(a)
    return a


def commit_tau(coeffs, tau):
    """kzg.Commit against the SRS of a known tau: [p(tau)]G1."""
    return g1_gen_mul(PB.poly_eval(coeffs, tau))


def setup(spr: SparseR1CS, tau: int) -> ProvingKey:
    """Setup, setup.go:110-161."""
    size_system = len(spr.constraints) + spr.nb_public
    d0 = N.Domain(size_system)
    d1 = N.Domain(8 * size_system if size_system < 6 else 4 * size_system)       # initDomains, setup.go:275-290
    n = d0.cardinality
    lag = build_trace(spr, n)
    perm = build_permutation(spr, n)
    sup = support_permutation(d0)
    s_lag = [[sup[perm[k * n + i]] for i in range(n)] for k in range(3)]
    ql, qr, qm, qo, qk = (_to_canonical(v, d0) for v in lag)
    s1, s2, s3 = (_to_canonical(v, d0) for v in s_lag)
    qcp = []
    for info in spr.commitments:                                   # BuildTrace, setup.go:222-229
        col = [0] * n
        for row in info["committed"]:
            col[spr.nb_public + row] = 1
        qcp.append(_to_canonical(col, d0))
    vk = {
        "Qcp": [commit_tau(q, tau) for q in qcp], "commitment_indexes": [info["index"] for info in spr.commitments],
        "size": n, "size_inv": pow(n, -1, R), "generator": d0.generator, "nb_public": spr.nb_public,
        "coset_shift": d0.coset_gen,
        "S": [commit_tau(s1, tau), commit_tau(s2, tau), commit_tau(s3, tau)],
        "Ql": commit_tau(ql, tau), "Qr": commit_tau(qr, tau), "Qm": commit_tau(qm, tau),
        "Qo": commit_tau(qo, tau), "Qk": commit_tau(qk, tau),
    }
    return ProvingKey(d0, d1, spr.nb_public, ql, qr, qm, qo, qk, s1, s2, s3, perm, list(lag[4]), vk, tau, qcp, list(spr.commitments))


def srs_scalars(pk: ProvingKey):
    """Discrete logs of the two SRS the device needs: pk.Kzg.G1[:n+3] = tau^i and
    pk.KzgLagrange.G1[:n] = L_i(tau) (setup.go:132-137, kzg.ToLagrangeG1)."""
    n, t = pk.d0.cardinality, pk.tau
    mono = [pow(t, i, R) for i in range(n + 3)]
    lag = mono[:n]
    N.fft_inverse(lag, pk.d0, N.DIF)       # L(tau) = F^-1 (tau^j): the inverse DFT matrix is symmetric
    N.bit_reverse(lag)
    return mono, lag


def to_lagrange_g1(points, d: N.Domain, curve=B):
    """kzg.ToLagrangeG1 as setup.go:134 calls it (gnark-crypto kzg, restated; module absent): copy the monomial SRS
    [tau^j]G1, radix-2 DIF over the GROUP with the domain's inverse twiddles (a' = a + b, b' = [w^-k](a - b)), bit-reverse,
    scale every point by 1/n.  Works on points only -- tau is not an input.  points: affine tuples or None (infinity)."""
    n = len(points)
    assert n == d.cardinality
    a = list(points)
    m, ws = n, d.generator_inv
    while m >= 2:
        h = m >> 1
        tw = [1] * h
        for j in range(1, h):
            tw[j] = tw[j - 1] * ws % R
        for base in range(0, n, m):
            for j in range(h):
                u, v = a[base + j], a[base + j + h]
                a[base + j] = curve.g1_add(u, v)
                diff = curve.g1_add(u, curve.g1_neg(v))
                a[base + j + h] = diff if j == 0 else curve.g1_mul(diff, tw[j])
        ws = ws * ws % R
        m = h
    N.bit_reverse(a)
    return [curve.g1_mul(p, d.cardinality_inv) for p in a]


# ---------------------------------------------------------------------------------------------
# Fiat-Shamir (gnark-crypto fiat-shamir.Transcript; byte layout pinned by solidity.go, see header)
# ---------------------------------------------------------------------------------------------
def g1_raw(p) -> bytes:
    """G1Affine.RawBytes / Marshal: X || Y big-endian, infinity = 0x40 then zeros."""
    fb = (B.P.bit_length() + 7) // 8       # 32 on BN254, 48 on BLS12-377 (oracle/suite377.py runs this file on that curve)
    if p is None:
        return bytes([0x40]) + bytes(2 * fb - 1)
    return p[0].to_bytes(fb, "big") + p[1].to_bytes(fb, "big")


def fr_bytes(x: int) -> bytes:
    return (x % R).to_bytes(32, "big")


class Transcript:
    def __init__(self, *names):
        self.names = list(names)
        self.bind = {k: [] for k in names}
        self.value = {}

    def Bind(self, name, data: bytes):
        assert name not in self.value
        self.bind[name].append(bytes(data))

    def ComputeChallenge(self, name) -> bytes:
        if name in self.value:
            return self.value[name]
        h = hashlib.sha256()
        h.update(name.encode())
        pos = self.names.index(name)
        if pos:
            h.update(self.value[self.names[pos - 1]])
        for b in self.bind[name]:
            h.update(b)
        self.value[name] = h.digest()
        return self.value[name]


def bind_public_data(fs, name, vk, public):
    """verify.go:296-340."""
    for p in (vk["S"][0], vk["S"][1], vk["S"][2], vk["Ql"], vk["Qr"], vk["Qm"], vk["Qo"], vk["Qk"]):
        fs.Bind(name, g1_raw(p))
    for p in vk.get("Qcp", []):
        fs.Bind(name, g1_raw(p))
    for w in public:
        fs.Bind(name, fr_bytes(w))


def derive_randomness(fs, name, *points) -> int:
    """verify.go:342-360."""
    for p in points:
        fs.Bind(name, g1_raw(p))
    return int.from_bytes(fs.ComputeChallenge(name), "big") % R


def derive_gamma_kzg(point, digests, claimed, data) -> int:
    """kzg.deriveGamma (layout: solidity.go:980-1032)."""
    fs = Transcript("gamma")
    fs.Bind("gamma", fr_bytes(point))
    for dg in digests:
        fs.Bind("gamma", g1_raw(dg))
    for v in claimed:
        fs.Bind("gamma", fr_bytes(v))
    for dt in data:
        fs.Bind("gamma", dt)
    return int.from_bytes(fs.ComputeChallenge("gamma"), "big") % R


# ---------------------------------------------------------------------------------------------
# prover
# ---------------------------------------------------------------------------------------------
@dataclass
class Proof:
    LRO: list
    Z: tuple
    H: list
    batched_H: tuple
    claimed: list          # BatchedProof.ClaimedValues (7)
    zshift_H: tuple
    zshift_value: int
    bsb22: list = field(default_factory=list)      # Bsb22Commitments


def _commit_blinded_lagrange(p_lagrange, bp, pk: ProvingKey):
    """commitToPolyAndBlinding (prove.go:492-502) + commitBlindingFactor (:1159-1172):
    [p](Lagrange SRS) + [b(tau) (tau^n - 1)]."""
    n, t = pk.d0.cardinality, pk.tau
    canon = _to_canonical(p_lagrange, pk.d0)
    v = (PB.poly_eval(canon, t) + PB.poly_eval(bp, t) * (pow(t, n, R) - 1)) % R
    return g1_gen_mul(v)


def _blinded_coefficients(p_canon, bp):
    """getBlindedCoefficients, prove.go:1148-1156: p + b (X^n - 1)."""
    cp = list(p_canon) + list(bp)
    for i, b in enumerate(bp):
        cp[i] = (cp[i] - b) % R
    return cp


def _kzg_open(p, point, tau):
    q, v = PB.poly_div_linear(p, point)
    return commit_tau(q, tau), v


def prove(spr: SparseR1CS, pk: ProvingKey, solution, blinding, bsb=None):
    """Prove, prove.go:116-180.  blinding = (bl[2], br[2], bo[2], bz[3]) -- getRandomPolynomial's
    coefficients, injected (the reference draws them from crypto/rand, prove.go:295-301).
    bsb: {"pi2": Lagrange columns, "commitments": points, "values": hashed commitments} from the solver's
    bsb22Hint for api.Commit circuits.  Returns (Proof, challenges dict)."""
    bsb = bsb or {"pi2": [], "commitments": [], "values": []}
    nb_bsb = len(bsb["pi2"])
    assert nb_bsb == len(pk.qcp)
    d0, d1, tau = pk.d0, pk.d1, pk.tau
    n, big = d0.cardinality, d1.cardinality
    rho = big // n
    bl, br, bo, bz = (list(v) for v in blinding)
    public = solution[:spr.nb_public]
    vk = pk.vk

    # solveConstraints + commitToLRO (prove.go:365-392, 425-452)
    L, Rr, O = evaluate_lro(spr, solution)
    lro_com = [_commit_blinded_lagrange(L, bl, pk), _commit_blinded_lagrange(Rr, br, pk), _commit_blinded_lagrange(O, bo, pk)]

    # completeQk (prove.go:397-423)
    qk_lag = list(pk.qk_lagrange)
    qk_lag[:spr.nb_public] = public
    for info, val in zip(pk.commitments, bsb["values"]):
        qk_lag[spr.nb_public + info["index"]] = val

    # deriveGammaAndBeta (prove.go:455-485)
    fs = Transcript("gamma", "beta", "alpha", "zeta")
    bind_public_data(fs, "gamma", vk, public)
    gamma = derive_randomness(fs, "gamma", *lro_com)
    beta = int.from_bytes(fs.ComputeChallenge("beta"), "big") % R

    # buildRatioCopyConstraint (prove.go:597-631)
    Z = PB.build_ratio_copy_constraint([L, Rr, O], pk.perm, beta, gamma, d0)
    z_com = _commit_blinded_lagrange(Z, bz, pk)
    alpha = derive_randomness(fs, "alpha", *bsb["commitments"], z_com)     # deriveAlpha, prove.go:504-512

    # evaluateConstraints -> computeNumerator (prove.go:521-595, 837-1079), literally: every polynomial
    # goes ToCanonical -> scaled -> ToLagrange on each of the rho cosets
    x = {
        PB.ID_L: list(L), PB.ID_R: list(Rr), PB.ID_O: list(O), PB.ID_Z: list(Z),
        PB.ID_QL: None, PB.ID_QR: None, PB.ID_QM: None, PB.ID_QO: None, PB.ID_QK: qk_lag,
        PB.ID_S1: None, PB.ID_S2: None, PB.ID_S3: None,
    }
    lone = [0] * n
    lone[0] = 1
    x[PB.ID_LONE] = lone
    ident = [0] * n
    ident[1] = beta
    canon = {}
    for k, v in x.items():
        canon[k] = _to_canonical(v, d0) if v is not None else None
    canon[PB.ID_QL], canon[PB.ID_QR], canon[PB.ID_QM], canon[PB.ID_QO] = pk.ql, pk.qr, pk.qm, pk.qo
    canon[PB.ID_S1], canon[PB.ID_S2], canon[PB.ID_S3] = pk.s1, pk.s2, pk.s3
    canon[PB.ID_ID] = ident
    pi2_canon = [_to_canonical(v, d0) for v in bsb["pi2"]]
    for k in range(nb_bsb):                                # x[id_Qci + 2k] = Qcp_k, x[id_Qci + 2k + 1] = Pi2_k
        canon[PB.ID_QCI + 2 * k] = pk.qcp[k]
        canon[PB.ID_QCI + 2 * k + 1] = pi2_canon[k]
    cres = [0] * big
    shifters = [d1.coset_gen] + [d1.generator] * (rho - 1)
    coset = 1
    bps = [list(bl), list(br), list(bo), list(bz)]
    for i in range(rho):
        coset = coset * shifters[i] % R
        tmp = (pow(coset, n, R) - 1) % R
        for q in bps:                                   # prove.go:985-993
            acc = tmp
            for j in range(len(q)):
                q[j] = q[j] * acc % R
                acc = acc * shifters[i] % R
        ev = {}
        for k, c in canon.items():                      # batchApply block, prove.go:1000-1017
            a = [c[j] * pow(coset, j, R) % R for j in range(n)]
            N.fft(a, d0, N.DIF)
            N.bit_reverse(a)
            ev[k] = a
        PB.quotient_rows(ev, [ev[PB.ID_QCI + k] for k in range(2 * nb_bsb)], d0, d1.log_n, i, beta, gamma, alpha, d1.coset_gen, bps[0], bps[1], bps[2], bps[3], cres)
        inv = pow(tmp, -1, R)
        for q in bps:
            for j in range(len(q)):
                q[j] = q[j] * inv % R

    # divideByXMinusOne (prove.go:1223-1250), commitToQuotient (:1198-1218)
    h = PB.divide_by_xn_minus_one(cres, d0, d1)
    h1, h2, h3 = h[:n + 2], h[n + 2:2 * (n + 2)], h[2 * (n + 2):3 * (n + 2)]
    assert all(v == 0 for v in h[3 * (n + 2):]), "quotient degree"
    h_com = [commit_tau(h1, tau), commit_tau(h2, tau), commit_tau(h3, tau)]
    zeta = derive_randomness(fs, "zeta", *h_com)

    # openZ (prove.go:634-650)
    zeta_shifted = zeta * vk["generator"] % R
    blinded_z = _blinded_coefficients(canon[PB.ID_Z], bz)
    zs_H, zu = _kzg_open(blinded_z, zeta_shifted, tau)

    # foldH (prove.go:670-705)
    zn2 = pow(zeta, n + 2, R)
    folded_h = [((h3[i] * zn2 + h2[i]) * zn2 + h1[i]) % R for i in range(n + 2)]
    folded_h_digest = commit_tau(folded_h, tau)

    # computeLinearizedPolynomial (prove.go:708-777, 1289-1389)
    zn1 = (pow(zeta, n, R) - 1) % R
    blz = (PB.poly_eval(canon[PB.ID_L], zeta) + PB.poly_eval(bl, zeta) * zn1) % R       # evaluateBlinded
    brz = (PB.poly_eval(canon[PB.ID_R], zeta) + PB.poly_eval(br, zeta) * zn1) % R
    boz = (PB.poly_eval(canon[PB.ID_O], zeta) + PB.poly_eval(bo, zeta) * zn1) % R
    s1z, s2z = PB.poly_eval(pk.s1, zeta), PB.poly_eval(pk.s2, zeta)
    qcpz = [PB.poly_eval(q, zeta) for q in pk.qcp]
    rl = brz * blz % R
    c1 = (s1z * beta + blz + gamma) * ((s2z * beta + brz + gamma) % R) % R * zu % R * beta % R
    u = vk["coset_shift"]
    c2 = (beta * zeta + blz + gamma) * ((beta * u * zeta + brz + gamma) % R) % R * ((beta * u * u % R * zeta + boz + gamma) % R) % R
    c2 = (R - c2) % R
    lag = zn1 * pow((zeta - 1) % R, -1, R) % R * alpha % R * alpha % R * d0.cardinality_inv % R
    lin = []
    for i in range(len(blinded_z)):
        t = blinded_z[i] * c2 % R
        if i < n:
            t = (t + pk.s3[i] * c1) % R
        t = t * alpha % R
        if i < n:
            t = (t + pk.qm[i] * rl + pk.ql[i] * blz + pk.qr[i] * brz + pk.qo[i] * boz + pk.qk[i]) % R
            for j in range(nb_bsb):
                t = (t + pi2_canon[j][i] * qcpz[j]) % R
        lin.append((t + blinded_z[i] * lag) % R)
    lin_digest = commit_tau(lin, tau)

    # batchOpening (prove.go:779-833) -> kzg.BatchOpenSinglePoint
    polys = [folded_h, lin, _blinded_coefficients(canon[PB.ID_L], bl), _blinded_coefficients(canon[PB.ID_R], br),
             _blinded_coefficients(canon[PB.ID_O], bo), pk.s1, pk.s2] + list(pk.qcp)
    digests = [folded_h_digest, lin_digest, lro_com[0], lro_com[1], lro_com[2], vk["S"][0], vk["S"][1]] + list(vk["Qcp"])
    claimed = [PB.poly_eval(p, zeta) for p in polys]
    g = derive_gamma_kzg(zeta, digests, claimed, [fr_bytes(zu)])
    width = max(len(p) for p in polys)
    folded = [0] * width
    acc = 1
    for p in polys:
        for i, c in enumerate(p):
            folded[i] = (folded[i] + acc * c) % R
        acc = acc * g % R
    q, _ = PB.poly_div_linear(folded, zeta)
    batched_H = commit_tau(q, tau)

    proof = Proof(lro_com, z_com, h_com, batched_H, claimed, zs_H, zu, list(bsb["commitments"]))
    return proof, {"gamma": gamma, "beta": beta, "alpha": alpha, "zeta": zeta, "gamma_kzg": g}


# ---------------------------------------------------------------------------------------------
# verifier (verify.go:45-294); the two KZG checks use the known tau instead of pairings
# ---------------------------------------------------------------------------------------------
def _kzg_check(digest, H, value, point, tau) -> bool:
    """e(C - [v]G1, G2) == e(H, [tau - z]G2)   <=>   C - [v]G1 == [tau - z]H."""
    lhs = B.g1_add(digest, B.g1_neg(g1_gen_mul(value)))
    rhs = B.g1_mul(H, (tau - point) % R) if H is not None else None
    return lhs == rhs


def verify(proof: Proof, vk: dict, public, tau: int) -> bool:
    fs = Transcript("gamma", "beta", "alpha", "zeta")
    bind_public_data(fs, "gamma", vk, public)
    gamma = derive_randomness(fs, "gamma", *proof.LRO)
    beta = derive_randomness(fs, "beta")
    alpha = derive_randomness(fs, "alpha", *proof.bsb22, proof.Z)
    zeta = derive_randomness(fs, "zeta", *proof.H)
    n = vk["size"]
    zeta_m = pow(zeta, n, R)
    zzeta = (zeta_m - 1) % R
    # PI = sum L_i(zeta) w_i   (verify.go:103-125)
    lagrange_one = zzeta * pow((zeta - 1) % R, -1, R) % R * vk["size_inv"] % R
    pi, lagrange, wpow = 0, lagrange_one, 1
    den = (zeta - 1) % R
    for i, w in enumerate(public):
        pi = (pi + lagrange * w) % R
        if i + 1 != len(public):
            lagrange = lagrange * vk["generator"] % R * den % R
            wpow = wpow * vk["generator"] % R
            den = (zeta - wpow) % R
            lagrange = lagrange * pow(den, -1, R) % R
    for k, ci in enumerate(vk.get("commitment_indexes", [])):          # verify.go:136-156
        hashed = hash_commitment(proof.bsb22[k])
        wpi = pow(vk["generator"], vk["nb_public"] + ci, R)
        lag_i = (zeta - 1) * wpi % R * pow((zeta - wpi) % R, -1, R) % R * lagrange_one % R
        pi = (pi + lag_i * hashed) % R
    zu = proof.zshift_value
    cq, lin_z, l, r, o, s1, s2 = proof.claimed[:7]
    _s1 = (s1 * beta + l + gamma) * ((s2 * beta + r + gamma) % R) % R * ((o + gamma) % R) % R * alpha % R * zu % R
    asl = lagrange_one * alpha % R * alpha % R
    t = (lin_z + pi + _s1 - asl) % R
    t = t * pow(zzeta, -1, R) % R
    if t != cq:
        return False
    zm2 = pow(zeta, n + 2, R)
    folded_h = B.g1_add(B.g1_mul(B.g1_add(B.g1_mul(proof.H[2], zm2), proof.H[1]), zm2), proof.H[0])
    rl = l * r % R
    u1 = zu * beta % R * ((beta * s1 + l + gamma) % R) % R * ((beta * s2 + r + gamma) % R) % R * alpha % R
    cs = vk["coset_shift"]
    u2 = (beta * zeta + l + gamma) * ((beta * zeta % R * cs + r + gamma) % R) % R * ((beta * zeta % R * cs % R * cs + o + gamma) % R) % R
    u2 = ((R - u2) * alpha + asl) % R
    points = list(proof.bsb22) + [vk["Ql"], vk["Qr"], vk["Qm"], vk["Qo"], vk["Qk"], vk["S"][2], proof.Z]
    scalars = list(proof.claimed[7:]) + [l, r, rl, o, 1, u1, u2]
    lin_digest = None
    for p, s in zip(points, scalars):
        lin_digest = B.g1_add(lin_digest, B.g1_mul(p, s))
    digests = [folded_h, lin_digest, proof.LRO[0], proof.LRO[1], proof.LRO[2], vk["S"][0], vk["S"][1]] + list(vk.get("Qcp", []))
    if len(proof.bsb22) != len(vk.get("Qcp", [])) or len(proof.claimed) != len(digests):
        return False
    g = derive_gamma_kzg(zeta, digests, proof.claimed, [fr_bytes(zu)])        # kzg.FoldProof
    fd, fv, acc = None, 0, 1
    for dg, v in zip(digests, proof.claimed):
        fd = B.g1_add(fd, B.g1_mul(dg, acc))
        fv = (fv + acc * v) % R
        acc = acc * g % R
    return (_kzg_check(fd, proof.batched_H, fv, zeta, tau)
            and _kzg_check(proof.Z, proof.zshift_H, zu, zeta * vk["generator"] % R, tau))


def proof_bytes(p: Proof) -> bytes:
    """Proof.WriteTo, marshal.go:41-66 (compressed points; []fr.Element = u32 BE length + elements)."""
    out = b"".join(g1_compress(q) for q in (*p.LRO, p.Z, *p.H, p.batched_H))
    out += len(p.claimed).to_bytes(4, "big") + b"".join(fr_bytes(v) for v in p.claimed)
    out += g1_compress(p.zshift_H) + fr_bytes(p.zshift_value)
    out += len(p.bsb22).to_bytes(4, "big") + b"".join(g1_compress(q) for q in p.bsb22)
    return out
