"""CPU oracle (TEST INFRASTRUCTURE ONLY) -- plonk ProvingKey.WriteRawTo restated (the on-disk format before the path).

Follows `/root/reference/backend/plonk/bn254/marshal.go:101-180` (ProvingKey.writeTo, withCompression = false) and
`:314-350` (VerifyingKey.writeTo) field by field:
    Vk: Size u64 | SizeInv | Generator | NbPublicVariables u64 | CosetShift | S[0..2] Ql Qr Qm Qo Qk (8 raw G1) |
        Qcp (G1 slice) | Kzg.G1 | Kzg.G2[0] Kzg.G2[1] (raw G2) | Kzg.Lines (fixed-size array) | CommitmentConstraintIndexes ([]u64)
    Domain[0] | Domain[1] | Kzg.G1 (G1 slice, n + 3) | KzgLagrange.G1 (G1 slice, n) |
    Ql Qr Qm Qo Qk ([]fr.Element) | Qcp ([][]fr.Element) | S1 S2 S3 ([]fr.Element) | S (3n int64, no length prefix)
Element encodings are gnark-crypto's (module absent, recalled -- see oracle/pk_raw.py): fr.Element = 32-byte big-endian
canonical; slices = u32 big-endian length + elements; raw points as in pk_raw.  Kzg.Lines (the verifier's precomputed
pairing lines) is a fixed-size array the prover never reads: written here as `lines_bytes` opaque bytes, and the reader
finds its end by the Domain[0] header that follows.  PARITY UNPINNED.  Only `tests/` may import this module.
"""
from __future__ import annotations

from . import pk_raw


def fr_be(x: int, r_mod: int) -> bytes:
    return (x % r_mod).to_bytes(32, "big")


def fr_slice(xs, r_mod: int) -> bytes:
    return len(xs).to_bytes(4, "big") + b"".join(fr_be(x, r_mod) for x in xs)


def write_raw(pk, curve, srs_points, lagrange_points, g2_points=(None, None), lines_bytes: int = 33280) -> bytes:
    """pk: oracle plonk_prove.ProvingKey; srs_points / lagrange_points: pk.Kzg.G1 (n + 3) and pk.KzgLagrange.G1 (n) as
    affine points; g2_points: vk.Kzg.G2 (not read by the prover)."""
    fb = (curve.P.bit_length() + 7) // 8
    R = curve.R
    vk = pk.vk
    n = pk.d0.cardinality
    out = [n.to_bytes(8, "big"), fr_be(pow(n, -1, R), R), fr_be(pk.d0.generator, R), int(pk.nb_public).to_bytes(8, "big"),
           fr_be(pk.d0.coset_gen, R)]
    out += [pk_raw.g1_raw(p, fb) for p in (vk["S"][0], vk["S"][1], vk["S"][2], vk["Ql"], vk["Qr"], vk["Qm"], vk["Qo"], vk["Qk"])]
    out.append(pk_raw.slice_raw(list(vk.get("Qcp", [])), pk_raw.g1_raw, fb))
    out.append(pk_raw.g1_raw(srs_points[0], fb))
    out += [pk_raw.g2_raw(p, fb) for p in g2_points]
    out.append(bytes((i * 37 + 11) & 0xFF for i in range(lines_bytes)))            # Kzg.Lines: opaque to the prover
    idx = [c["index"] for c in pk.commitments]
    out.append(len(idx).to_bytes(4, "big") + b"".join(int(i).to_bytes(8, "big") for i in idx))
    out += [pk_raw.domain_bytes(pk.d0, R), pk_raw.domain_bytes(pk.d1, R)]
    out += [pk_raw.slice_raw(list(srs_points), pk_raw.g1_raw, fb), pk_raw.slice_raw(list(lagrange_points), pk_raw.g1_raw, fb)]
    out += [fr_slice(v, R) for v in (pk.ql, pk.qr, pk.qm, pk.qo, pk.qk)]
    out.append(len(pk.qcp).to_bytes(4, "big") + b"".join(fr_slice(q, R) for q in pk.qcp))
    out += [fr_slice(v, R) for v in (pk.s1, pk.s2, pk.s3)]
    out.append(b"".join(int(v).to_bytes(8, "big", signed=True) for v in pk.perm))
    return b"".join(out)
