"""Groth16 ProvingKey on-disk format (WriteRawTo) -> device, the step before the path.

Mirror of `ProvingKey.readFrom` (/root/reference/backend/groth16/bn254/marshal.go:319-374; identical for bls12-377):
the file is walked once on the host -- only lengths and offsets are read -- and every point slice goes to the device
as the bytes it has on disk; `b200_bases_upload_raw` turns big-endian canonical coordinates into Montgomery limbs in
one kernel (no per-point host work, no curve or subgroup checks: `UnsafeReadFrom` semantics, marshal.go:326-330).

Layout (encodings of gnark-crypto's raw Encoder as cited in include/b200prover.h, b200_points_from_raw):
    Domain: u64 Cardinality | CardinalityInv | Generator | GeneratorInv | FrMultiplicativeGen | FrMultiplicativeGenInv
    G1.Alpha G1.Beta G1.Delta | [A] [B] [Z] [K] (u32 length + points) | G2.Beta G2.Delta | [G2.B]
    u64 nbWires | u64 NbInfinityA | u64 NbInfinityB | InfinityA (nbWires bytes) | InfinityB | u32 nbCommitmentKeys
Keys with Pedersen commitment keys (api.Commit circuits) are rejected: their section is not parsed.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

from . import api


class PkFormatError(ValueError):
    pass


@dataclass
class RawProvingKey:
    curve: int
    cardinality: int
    generator: int                  # canonical integers (the ABI wants Montgomery limbs: see fr_limbs)
    coset_gen: int
    points: dict = field(default_factory=dict)     # 'alpha','beta','delta','beta2','delta2' -> (offset, 1)
    slices: dict = field(default_factory=dict)     # 'a','b','z','k','g2b' -> (offset of first point, count)
    nb_wires: int = 0
    nb_infinity_a: int = 0
    nb_infinity_b: int = 0
    infinity_a: np.ndarray | None = None
    infinity_b: np.ndarray | None = None
    size: int = 0
    commitment_keys: list = field(default_factory=list)   # per Pedersen key: ((basis offset, n), (basisExpSigma offset, n))
    domain_header_bytes: int = 168                        # which fft.Domain layout matched (168, or 169 with a trailing flag byte)


_R_MOD = {api.BN254: 0x30644E72E131A029B85045B68181585D2833E84879B9709143E1F593F0000001,
          api.BLS12_377: 0x12AB655E9A2CA55660B44D1E5C37B00159AA76FED00000010A11800000000001}


def fr_limbs(x: int, curve: int) -> np.ndarray:
    return np.frombuffer(((x << 256) % _R_MOD[curve]).to_bytes(32, "little"), dtype=np.uint64).copy()


def parse(buf, curve: int = api.BN254) -> RawProvingKey:
    """Offsets and lengths of every section of a WriteRawTo file; raises PkFormatError on truncation or inconsistency.
    The fft.Domain header is the 168 bytes gnark-crypto v0.12 writes; later versions append one `withPrecompute` byte:
    when the 168-byte layout does not validate the 169-byte one is tried and `domain_header_bytes` says which matched
    (the layout is recalled, not pinned against the module -- a wrong guess is a loud PkFormatError, never a misparse:
    every section length is cross-checked against the infinity bitmaps)."""
    first = None
    for extra in (0, 1):
        try:
            pk = _parse(buf, curve, extra)
            pk.domain_header_bytes = 168 + extra
            return pk
        except PkFormatError as e:
            first = first or e
    raise first


def _parse(buf, curve: int, domain_extra: int) -> RawProvingKey:
    mv = memoryview(buf)
    fb = 8 * api.fp_words(curve)
    g1, g2 = 2 * fb, 4 * fb
    pos = 0

    def take(n):
        nonlocal pos
        if pos + n > len(mv):
            raise PkFormatError(f"truncated proving key: need {n} bytes at offset {pos}, file has {len(mv)}")
        pos += n
        return pos - n

    def u(n):
        o = take(n)
        return int.from_bytes(mv[o:o + n], "big")

    card = u(8)
    if card == 0 or card & (card - 1):
        raise PkFormatError(f"Domain.Cardinality = {card} is not a power of two")
    _card_inv, gen, _gen_inv, coset, _coset_inv = (u(32) for _ in range(5))
    r = _R_MOD[curve]
    if pow(gen, card, r) != 1 or (card > 1 and pow(gen, card // 2, r) != r - 1):
        raise PkFormatError("Domain.Generator is not a primitive root of unity of order Cardinality (wrong curve?)")
    take(domain_extra)
    pk = RawProvingKey(curve, card, gen, coset)
    for name in ("alpha", "beta", "delta"):
        pk.points[name] = (take(g1), 1)
    for name in ("a", "b", "z", "k"):
        n = u(4)
        pk.slices[name] = (take(n * g1), n)
    for name in ("beta2", "delta2"):
        pk.points[name] = (take(g2), 1)
    n = u(4)
    pk.slices["g2b"] = (take(n * g2), n)
    pk.nb_wires, pk.nb_infinity_a, pk.nb_infinity_b = u(8), u(8), u(8)
    oa, ob = take(pk.nb_wires), take(pk.nb_wires)
    pk.infinity_a = np.frombuffer(mv[oa:oa + pk.nb_wires], dtype=np.uint8).copy()
    pk.infinity_b = np.frombuffer(mv[ob:ob + pk.nb_wires], dtype=np.uint8).copy()
    if int(pk.infinity_a.sum()) != pk.nb_infinity_a or int(pk.infinity_b.sum()) != pk.nb_infinity_b:
        raise PkFormatError("InfinityA/B bitmaps disagree with NbInfinityA/B")
    if pk.slices["a"][1] != pk.nb_wires - pk.nb_infinity_a or pk.slices["b"][1] != pk.nb_wires - pk.nb_infinity_b \
            or pk.slices["g2b"][1] != pk.slices["b"][1]:
        raise PkFormatError("len(G1.A), len(G1.B), len(G2.B) disagree with the infinity bitmaps (setup.go:231-262)")
    # pk.CommitmentKeys (marshal.go:276-293,364): per key the Pedersen basis and basisExpSigma as two raw G1 slices.
    # They stay host-side data (prove.go:95,134: Commit / BatchProve run on the CPU in the reference and here); the
    # loader records where they are so the glue can hand them to pedersen without decoding the rest of the file.
    nb_keys = u(4)
    if nb_keys > 4096:
        raise PkFormatError(f"implausible number of commitment keys: {nb_keys}")
    pk.commitment_keys = []
    for _ in range(nb_keys):
        nb = u(4)
        basis = (take(nb * g1), nb)
        ns = u(4)
        if ns != nb:
            raise PkFormatError("a Pedersen key's basis and basisExpSigma differ in length")
        pk.commitment_keys.append((basis, (take(ns * g1), ns)))
    pk.size = pos
    return pk


def upload(ctx: "api.Context", buf, nb_public: int, shard_index: int = 0, shard_count: int = 1):
    """Parse `buf` (bytes / mmap of a WriteRawTo file) and make the key resident: returns (Groth16ProvingKey,
    RawProvingKey, vk_points) with vk_points = [alpha, beta, delta (G1), beta, delta (G2)] as affine Montgomery limbs,
    the arguments of groth16_finalize.  With shard_count > 1 only this rank's point ranges are converted and kept."""
    pk = parse(buf, ctx.curve)
    mv = memoryview(buf)
    fb = 8 * ctx.fw
    log_n = pk.cardinality.bit_length() - 1

    def rng(name):
        off, n = pk.slices[name]
        if name == "z" and shard_count > 1 and shard_count & (shard_count - 1) == 0 and shard_count <= pk.cardinality:
            blk = pk.cardinality // shard_count          # G1.Z follows the block layout of the distributed computeH
            return off, min(n, blk * shard_index), min(n, blk * (shard_index + 1))
        return off, n * shard_index // shard_count, n * (shard_index + 1) // shard_count

    arrays = {}
    for name, group in (("a", 1), ("b", 1), ("k", 1), ("z", 1), ("g2b", 2)):
        off, lo, hi = rng(name)
        stride = (2 if group == 1 else 4) * fb
        arrays[name] = mv[off + lo * stride: off + hi * stride]      # file bytes, no copy, no host-side decoding
    vk = [ctx.points_from_raw(mv[pk.points[k][0]: pk.points[k][0] + (2 if g == 1 else 4) * fb], 1, g)[0]
          for k, g in (("alpha", 1), ("beta", 1), ("delta", 1), ("beta2", 2), ("delta2", 2))]
    full = {k: pk.slices[k][1] for k in ("a", "b", "k", "z")}
    h = ctx.groth16_pk_upload(log_n=log_n, generator=fr_limbs(pk.generator, ctx.curve), coset_gen=fr_limbs(pk.coset_gen, ctx.curve),
                              g1_a=arrays["a"], g1_b=arrays["b"], g1_k=arrays["k"], g1_z=arrays["z"], g2_b=arrays["g2b"],
                              infinity_a=pk.infinity_a, infinity_b=pk.infinity_b, nb_public=nb_public,
                              shard_index=shard_index, shard_count=shard_count, presliced=shard_count > 1, full_counts=full,
                              raw_points=True)
    return h, pk, vk
