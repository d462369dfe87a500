"""CPU oracle (TEST INFRASTRUCTURE ONLY) -- ProvingKey.WriteRawTo restated (the on-disk format before the path).

Follows `/root/reference/backend/groth16/bn254/marshal.go:246-304` (writeTo with raw = true), field by field:
    Domain.WriteTo | G1.Alpha, G1.Beta, G1.Delta | G1.A, G1.B, G1.Z, G1.K | G2.Beta, G2.Delta | G2.B |
    nbWires, NbInfinityA, NbInfinityB (u64) | InfinityA, InfinityB | u32 len(CommitmentKeys) [| keys]
The element encodings are gnark-crypto's (`curve.NewEncoder(w, curve.RawEncoding())`, module absent, recalled; what the
reference tree itself corroborates is cited):
  * point, raw: X || Y big-endian CANONICAL bytes (G2: X.A1 || X.A0 || Y.A1 || Y.A0); infinity = first byte 0x40,
    the rest zero; flag bits 00 otherwise (bellman_test.go:101 sizes a slice as SizeOfG1AffineUncompressed + 4)
  * point slice: u32 big-endian length, then the points ("uint32(len(Kvk)),[Kvk]1", marshal.go:149)
  * u64 / u32: big-endian; []bool: one byte per entry, no length prefix -- encoding/binary's default, which is why
    readFrom sizes the slices from nbWires before decoding them (marshal.go:342-350)
  * fft.Domain.WriteTo (gnark-crypto v0.12): Cardinality u64, then CardinalityInv, Generator, GeneratorInv,
    FrMultiplicativeGen, FrMultiplicativeGenInv as 32-byte big-endian canonical fr.Element
PARITY UNPINNED: no serialized proving key exists in the reference tree.

Only `tests/` may import this module.
"""
from __future__ import annotations


def fp_bytes(x: int, nbytes: int) -> bytes:
    return x.to_bytes(nbytes, "big")


def g1_raw(p, fb: int) -> bytes:
    if p is None:
        return bytes([0x40]) + bytes(2 * fb - 1)
    return fp_bytes(p[0], fb) + fp_bytes(p[1], fb)


def g2_raw(p, fb: int) -> bytes:
    if p is None:
        return bytes([0x40]) + bytes(4 * fb - 1)
    (x0, x1), (y0, y1) = p
    return fp_bytes(x1, fb) + fp_bytes(x0, fb) + fp_bytes(y1, fb) + fp_bytes(y0, fb)


def slice_raw(pts, enc, fb: int) -> bytes:
    return len(pts).to_bytes(4, "big") + b"".join(enc(p, fb) for p in pts)


def domain_bytes(d, r_mod: int) -> bytes:
    fr = lambda x: (x % r_mod).to_bytes(32, "big")
    return d.cardinality.to_bytes(8, "big") + fr(d.cardinality_inv) + fr(d.generator) + fr(d.generator_inv) + \
        fr(d.coset_gen) + fr(d.coset_gen_inv)


def write_raw(pk, curve, commitment_keys=(), domain_trailer: bytes = b"") -> bytes:
    """pk: oracle groth16.ProvingKey with points (setup(..., points=True)); curve: the oracle curve module.
    commitment_keys: [(basis points, basisExpSigma points)] -- pedersen.ProvingKey.WriteRawTo (marshal.go:276-293) writes
    the two point slices of a key one after the other (gnark-crypto fr/pedersen, recalled).  domain_trailer: bytes a
    later gnark-crypto appends to Domain.WriteTo (a one-byte `withPrecompute` flag); the pinned commit writes none."""
    fb = (curve.P.bit_length() + 7) // 8
    out = [domain_bytes(pk.domain, curve.R) + domain_trailer]
    out += [g1_raw(p, fb) for p in (pk.g1_alpha, pk.g1_beta, pk.g1_delta)]
    out += [slice_raw(s, g1_raw, fb) for s in (pk.g1_A, pk.g1_B, pk.g1_Z, pk.g1_K)]
    out += [g2_raw(p, fb) for p in (pk.g2_beta, pk.g2_delta)]
    out.append(slice_raw(pk.g2_B, g2_raw, fb))
    nb_wires = len(pk.infinity_A)
    out += [nb_wires.to_bytes(8, "big"), int(pk.nb_infinity_A).to_bytes(8, "big"), int(pk.nb_infinity_B).to_bytes(8, "big")]
    out += [bytes(1 if b else 0 for b in pk.infinity_A), bytes(1 if b else 0 for b in pk.infinity_B)]
    out.append(len(commitment_keys).to_bytes(4, "big"))
    for basis, sigma in commitment_keys:
        out += [slice_raw(basis, g1_raw, fb), slice_raw(sigma, g1_raw, fb)]
    return b"".join(out)
