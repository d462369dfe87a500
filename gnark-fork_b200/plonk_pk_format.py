"""Reader of the PlonK proving key's on-disk form: ProvingKey.WriteRawTo,
/root/reference/backend/plonk/bn254/marshal.go:101-180 (+ VerifyingKey.writeTo :314-350), decoded by readFrom :186-307
element by element on the CPU.  Here the file is walked once for its section offsets and the sections are handed to
the device as they are (b200_plonk_pk_upload with raw_file = 1): scalars, permutation and points are converted by
kernels (fr_from_be_kernel, i64_from_be_kernel, points_from_raw_kernel) after one DMA each.

Element encodings are gnark-crypto's and recalled (module absent; see oracle/plonk_raw.py): PARITY UNPINNED.  The one
variable-size unknown, vk.Kzg.Lines (a fixed-size array of precomputed pairing lines the prover never reads), is not
sized from memory: its end is found by the Domain[0] header that follows it (Size u64 || SizeInv, both known from the
start of the file).  Inconsistent lengths raise PkFormatError; nothing is guessed silently."""
from __future__ import annotations

from dataclasses import dataclass, field

from . import api
from .pk_format import PkFormatError, _R_MOD, fr_limbs


@dataclass
class RawPlonkKey:
    curve: int
    size: int                       # vk.Size = Domain[0].Cardinality
    nb_public: int
    generator: int
    coset_shift: int
    generator_big: int = 0
    size_big: int = 0
    sections: dict = field(default_factory=dict)    # name -> (offset, byte length)
    qcp: list = field(default_factory=list)         # [(offset, byte length)] of trace.Qcp[i]
    nb_qcp: int = 0
    commitment_indexes: list = field(default_factory=list)
    lines_bytes: int = 0
    file_bytes: int = 0


def parse(buf, curve: int = api.BN254) -> RawPlonkKey:
    mv = memoryview(buf)
    fb = 8 * api.fp_words(curve)
    g1, g2 = 2 * fb, 4 * fb
    r = _R_MOD[curve]
    pos = 0

    def take(n):
        nonlocal pos
        if n < 0 or pos + n > len(mv):
            raise PkFormatError(f"truncated PlonK proving key: need {n} bytes at offset {pos}, file has {len(mv)}")
        pos += n
        return pos - n

    def u(n):
        o = take(n)
        return int.from_bytes(mv[o:o + n], "big")

    size = u(8)
    if size == 0 or size & (size - 1):
        raise PkFormatError(f"vk.Size = {size} is not a power of two")
    size_inv, gen = u(32), u(32)
    nb_public = u(8)
    coset = u(32)
    if size_inv * size % r != 1 or pow(gen, size, r) != 1 or (size > 1 and pow(gen, size // 2, r) != r - 1):
        raise PkFormatError("vk.SizeInv / vk.Generator do not match vk.Size (wrong curve?)")
    pk = RawPlonkKey(curve, size, nb_public, gen, coset)
    pk.sections["vk_commitments"] = (take(8 * g1), 8 * g1)
    pk.nb_qcp = u(4)
    if pk.nb_qcp > api.PLONK_MAX_BSB22:
        raise PkFormatError(f"{pk.nb_qcp} BSB22 commitments: more than the prover supports")
    pk.sections["vk_qcp"] = (take(pk.nb_qcp * g1), pk.nb_qcp * g1)
    take(g1 + 2 * g2)                                   # vk.Kzg.G1, G2[0], G2[1]: the verifier's
    # vk.Kzg.Lines, then CommitmentConstraintIndexes (u32 k, k x u64), then Domain[0] = Size u64 || SizeInv ...
    sig = size.to_bytes(8, "big") + size_inv.to_bytes(32, "big")
    tail = 4 + 8 * pk.nb_qcp
    at = bytes(mv[pos:pos + (1 << 20)]).find(sig, tail)
    while at >= 0:
        cand = pos + at - tail
        if int.from_bytes(mv[cand:cand + 4], "big") == pk.nb_qcp:
            break
        at = bytes(mv[pos:pos + (1 << 20)]).find(sig, at + 1)
    if at < 0:
        raise PkFormatError("Domain[0] header not found behind the verifying key")
    pk.lines_bytes = at - tail
    take(pk.lines_bytes)
    if u(4) != pk.nb_qcp:
        raise PkFormatError("len(CommitmentConstraintIndexes) != len(Qcp)")
    pk.commitment_indexes = [u(8) for _ in range(pk.nb_qcp)]

    def domain(expect_card=None):
        card = u(8)
        _inv, g, _ginv, cs, _csinv = (u(32) for _ in range(5))
        if card == 0 or card & (card - 1) or (expect_card and card != expect_card):
            raise PkFormatError(f"bad fft.Domain cardinality {card}")
        if pow(g, card, r) != 1 or pow(g, card // 2, r) != r - 1:
            raise PkFormatError("fft.Domain generator is not a primitive root of its cardinality")
        return card, g, cs
    # both domain headers may carry the one-byte flag later gnark-crypto versions append (see pk_format.parse)
    start = pos
    for extra in (0, 1):
        pos = start
        try:
            _, g0, _ = domain(size)
            take(extra)
            pk.size_big, pk.generator_big, _ = domain()
            take(extra)
            n_srs = u(4)
            if n_srs != size + 3:
                raise PkFormatError(f"len(pk.Kzg.G1) = {n_srs}, expected Size + 3")
            break
        except PkFormatError:
            if extra:
                raise
    if g0 != gen:
        raise PkFormatError("Domain[0].Generator differs from vk.Generator")
    pk.sections["kzg_g1"] = (take(n_srs * g1), n_srs * g1)
    if u(4) != size:
        raise PkFormatError("len(pk.KzgLagrange.G1) != Size")
    pk.sections["kzg_lagrange_g1"] = (take(size * g1), size * g1)

    def fr_vec(name):
        if u(4) != size:
            raise PkFormatError(f"len(trace.{name}) != Size")
        return take(32 * size), 32 * size
    for name in ("ql", "qr", "qm", "qo", "qk"):
        pk.sections[name] = fr_vec(name)
    if u(4) != pk.nb_qcp:
        raise PkFormatError("len(trace.Qcp) != len(vk.Qcp)")
    pk.qcp = [fr_vec("Qcp") for _ in range(pk.nb_qcp)]
    for name in ("s1", "s2", "s3"):
        pk.sections[name] = fr_vec(name)
    pk.sections["perm"] = (take(24 * size), 24 * size)
    pk.file_bytes = pos
    return pk


def upload(prover, buf, shard_index: int = 0, shard_count: int = 1):
    """prover: a Context or a Group.  Parse `buf` (bytes / mmap of a WriteRawTo file) and make the key resident:
    returns (PlonkProvingKey or GroupPlonkProvingKey, RawPlonkKey)."""
    pk = parse(buf, prover.curve)
    mv = memoryview(buf)
    sec = lambda name: mv[pk.sections[name][0]: pk.sections[name][0] + pk.sections[name][1]]
    kw = dict(log_n=pk.size.bit_length() - 1, log_big=pk.size_big.bit_length() - 1, generator=fr_limbs(pk.generator, pk.curve),
              generator_big=fr_limbs(pk.generator_big, pk.curve), coset_gen=fr_limbs(pk.coset_shift, pk.curve),
              nb_public=pk.nb_public, qcp=[mv[o:o + n] for o, n in pk.qcp], vk_qcp=sec("vk_qcp") if pk.nb_qcp else None,
              commitment_index=pk.commitment_indexes, raw_file=True,
              **{k: sec(k) for k in ("ql", "qr", "qm", "qo", "qk", "s1", "s2", "s3", "perm", "kzg_g1", "kzg_lagrange_g1", "vk_commitments")})
    if isinstance(prover, api.Context):
        kw.update(shard_index=shard_index, shard_count=shard_count)
    return prover.plonk_pk_upload(**kw), pk
