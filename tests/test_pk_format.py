"""The proving key's on-disk format in front of the path: ProvingKey.WriteRawTo as restated
by oracle/pk_raw.py (marshal.go:246-304) against the product's reader gnark-fork_b200/pk_format.py (readFrom,
marshal.go:319-374) and the device conversion of raw point slices (b200_points_from_raw, raw_points key upload)."""
import random

import numpy as np
import pytest

from oracle import bn254 as B
from oracle import groth16 as G
from oracle import pk_raw
from tests import util


def _key(nb_mul, seed=3):
    rnd = random.Random(seed)
    r1cs = G.mul_chain_circuit(nb_mul)
    td = G.Trapdoor(*[rnd.randrange(1, B.R) for _ in range(5)])
    pk, vk = G.setup(r1cs, td, points=True)
    return r1cs, td, pk, vk


@pytest.fixture(scope="module")
def small_key():
    return _key(13)


def _fmt(pkg):
    import importlib
    return importlib.import_module("gnark_fork_b200.pk_format")


def test_parser_walks_the_writer_layout(pkg, small_key):
    r1cs, td, pk, vk = small_key
    raw = pk_raw.write_raw(pk, B)
    F = _fmt(pkg)
    p = F.parse(raw, pkg.BN254)
    assert p.size == len(raw) and p.cardinality == pk.domain.cardinality
    assert p.generator == pk.domain.generator and p.coset_gen == pk.domain.coset_gen == 5
    assert [p.slices[k][1] for k in ("a", "b", "z", "k", "g2b")] == [len(pk.g1_A), len(pk.g1_B), len(pk.g1_Z), len(pk.g1_K), len(pk.g2_B)]
    assert p.nb_wires == r1cs.nb_wires and list(p.infinity_a) == [int(x) for x in pk.infinity_A]
    assert list(p.infinity_b) == [int(x) for x in pk.infinity_B]
    # sections hold what the writer put there: first point of A, last of G2.B, delta
    off, n = p.slices["a"]
    assert raw[off:off + 64] == pk_raw.g1_raw(pk.g1_A[0], 32)
    off, n = p.slices["g2b"]
    assert raw[off + 128 * (n - 1): off + 128 * n] == pk_raw.g2_raw(pk.g2_B[-1], 32)
    assert raw[p.points["delta"][0]:][:64] == pk_raw.g1_raw(pk.g1_delta, 32)
    # layout constants: Domain header 8 + 5 x 32 bytes, then three G1 points
    assert p.points["alpha"][0] == 168 and p.slices["a"][0] == 168 + 3 * 64 + 4


def test_parser_rejects_damaged_files(pkg, small_key):
    _, _, pk, _ = small_key
    raw = bytearray(pk_raw.write_raw(pk, B))
    F = _fmt(pkg)
    with pytest.raises(F.PkFormatError):
        F.parse(bytes(raw[:-9]), pkg.BN254)                 # truncated
    bad = bytearray(raw); bad[7] = 3                          # cardinality not a power of two
    with pytest.raises(F.PkFormatError):
        F.parse(bytes(bad), pkg.BN254)
    bad = bytearray(raw); bad[-1] = 1                         # says a commitment key follows, and none does: truncated
    with pytest.raises(F.PkFormatError):
        F.parse(bytes(bad), pkg.BN254)
    p = F.parse(bytes(raw), pkg.BN254)
    bad = bytearray(raw); bad[p.size - 4 - 1] ^= 1            # flip the last InfinityB entry: counts disagree
    with pytest.raises(F.PkFormatError):
        F.parse(bytes(bad), pkg.BN254)
    with pytest.raises(F.PkFormatError):                      # a BN254 file read as BLS12-377
        F.parse(bytes(raw), pkg.BLS12_377)


def test_parser_commitment_keys_and_domain_header_variants(pkg, small_key):
    """pk.CommitmentKeys (marshal.go:276-293,364): the Pedersen keys' two point slices are located, not refused; and the
    fft.Domain header is accepted with or without the one-byte flag later gnark-crypto versions append."""
    _, _, pk, _ = small_key
    F = _fmt(pkg)
    basis = [B.g1_mul(B.G1_GEN, k) for k in (3, 5, 7)]
    sigma = [B.g1_mul(B.G1_GEN, 11 * k) for k in (3, 5, 7)]
    keys = [(basis, sigma), (basis[:1], sigma[:1])]
    for trailer in (b"", b"\x01"):
        raw = pk_raw.write_raw(pk, B, commitment_keys=keys, domain_trailer=trailer)
        p = F.parse(raw, pkg.BN254)
        assert p.size == len(raw) and p.domain_header_bytes == 168 + len(trailer)
        assert [(b[1], s[1]) for b, s in p.commitment_keys] == [(3, 3), (1, 1)]
        (boff, _), (soff, _) = p.commitment_keys[0]
        assert raw[boff:boff + 64] == pk_raw.g1_raw(basis[0], 32) and raw[soff + 128: soff + 192] == pk_raw.g1_raw(sigma[2], 32)
        assert p.slices["a"][1] == len(pk.g1_A)
    bad = bytearray(pk_raw.write_raw(pk, B, commitment_keys=keys))
    bad[p.commitment_keys[0][1][0] - 1 - len(b"\x01")] ^= 2     # basisExpSigma length differs from the basis length
    with pytest.raises(F.PkFormatError):
        F.parse(bytes(bad), pkg.BN254)


@pytest.mark.gpu
def test_points_from_raw_matches_oracle(ctx, pkg, cref):
    rng = np.random.default_rng(4)
    pts = [B.g1_mul(B.G1_GEN, int(k)) for k in rng.integers(1, 1 << 62, size=40)]
    pts[7] = None
    raw = b"".join(pk_raw.g1_raw(p, 32) for p in pts)
    got = ctx.points_from_raw(raw, len(pts), 1)
    assert np.array_equal(got, util.g1_arr([None if p is None else list(p) for p in pts]))
    q = [B.g2_mul(B.G2_GEN, int(k)) for k in rng.integers(1, 1 << 62, size=9)]
    q[0] = None
    raw2 = b"".join(pk_raw.g2_raw(p, 32) for p in q)
    got2 = ctx.points_from_raw(raw2, len(q), 2)
    want2 = util.g2_arr([None if p is None else [p[0][0], p[0][1], p[1][0], p[1][1]] for p in q])
    assert np.array_equal(got2, want2)
    assert ctx.points_from_raw(b"", 0, 1).shape == (0, 8)
    # a compressed-form flag, a coordinate >= p, a dirty infinity: all refused
    for bad in (bytes([0x80]) + raw[1:64], (B.P).to_bytes(32, "big") + raw[32:64], bytes([0x40, 1]) + bytes(62)):
        with pytest.raises(pkg.B200Error):
            ctx.points_from_raw(bad, 1, 1)


@pytest.mark.gpu
def test_resident_bases_from_raw(ctx, pkg, cref):
    n = 5000
    bases, ks = util.known_dlog_bases_g1(n, seed=9)
    ints = cref.g1_array_to_points(bases)
    raw = b"".join(pk_raw.g1_raw(p, 32) for p in ints)
    scalars = util.fr_arr([int(x) for x in np.random.default_rng(1).integers(1, 1 << 62, size=n)])
    h_raw, h_mem = ctx.upload_bases_raw(raw, n, 1), ctx.upload_bases(bases, 1)
    assert np.array_equal(pkg.g1_jac_to_affine(ctx.msm_resident(h_raw, scalars)), pkg.g1_jac_to_affine(ctx.msm_resident(h_mem, scalars)))
    h_raw.free(); h_mem.free()


@pytest.mark.gpu
@pytest.mark.parametrize("shards", [1, 2])
def test_key_file_to_proof(ctx, pkg, cref, shards):
    """WriteRawTo bytes -> pk_format.upload (sections DMA'd as they are, converted on the device) -> the prove core gives
    the oracle's proof; with 2 shards the partial sums combine to the same proof."""
    nb_mul = 300
    r1cs, td, pk, vk = _key(nb_mul, seed=8)
    raw = pk_raw.write_raw(pk, B)
    F = _fmt(pkg)
    w, a, b, c = G.mul_chain_witness(nb_mul, 2)
    r, s = 0x1234567890ABCDEF % B.R, 0xFEDCBA0987654321 % B.R
    outs = []
    for i in range(shards):
        h, parsed, vkp = F.upload(ctx, raw, r1cs.nb_public, i, shards)
        out, _ = ctx.groth16_prove_core(h, util.fr_arr(a), util.fr_arr(b), util.fr_arr(c), util.fr_arr(w))
        outs.append(out)
        h.free()
    out = outs[0]
    if shards == 2:
        d0, d1 = outs[0].as_dict(), outs[1].as_dict()
        for k in ("ar", "bs1", "krs", "krs2"):
            getattr(out, k)[:] = [int(v) for v in pkg.g1_jac_add(d0[k], d1[k])]
        out.bs2[:] = [int(v) for v in pkg.g2_jac_add(d0["bs2"], d1["bs2"])]
    ar, bs, krs = pkg.groth16_finalize(out, cref.fr(r), cref.fr(s), *vkp)
    ear, ebs, ekrs = G.prove_points(pk, w, a, b, c, r, s, r1cs.nb_public)
    assert pkg.groth16_proof_bytes(ar, bs, krs) == G.proof_bytes(ear, ebs, ekrs)


@pytest.mark.gpu
def test_bls12_377_points_from_raw(pkg):
    from oracle import bls12_377 as B3
    from tests import util377 as U
    c = pkg.Context(0, curve=pkg.BLS12_377)
    try:
        pts = [B3.g1_mul(B3.G1_GEN, k) for k in (1, 2, 12345678901234567)] + [None]
        got = c.points_from_raw(b"".join(pk_raw.g1_raw(p, 48) for p in pts), 4, 1)
        assert np.array_equal(got, U.g1_arr(pts))
        q = [B3.g2_mul(B3.G2_GEN, k) for k in (1, 77)] + [None]
        got = c.points_from_raw(b"".join(pk_raw.g2_raw(p, 48) for p in q), 3, 2)
        assert np.array_equal(got, U.g2_arr(q))
        with pytest.raises(pkg.B200Error):      # 0b100 in the top three bits = a compressed form on the BLS curves
            c.points_from_raw(bytes([0x80]) + bytes(95), 1, 1)
    finally:
        c.close()


# ---- PlonK proving key: ProvingKey.WriteRawTo (plonk/bn254/marshal.go:101-180) ---------------------------------------
def _plonk_key_file(nb_mul, nb_commit, seed, lines_bytes=33280):
    from oracle import plonk_prove as PP
    from oracle import plonk_raw
    rng = random.Random(seed)
    spr = PP.commit_chain_scs(nb_mul, nb_commit) if nb_commit else PP.mul_chain_scs(nb_mul)
    pk = PP.setup(spr, rng.randrange(1, B.R))
    mono, lag = PP.srs_scalars(pk)
    srs = [B.g1_mul(B.G1_GEN, k) for k in mono]
    lagp = [B.g1_mul(B.G1_GEN, k) for k in lag]
    raw = plonk_raw.write_raw(pk, B, srs, lagp, g2_points=(B.G2_GEN, B.g2_mul(B.G2_GEN, pk.tau)), lines_bytes=lines_bytes)
    return PP, spr, pk, raw, rng


@pytest.mark.parametrize("nb_commit,lines_bytes", [(0, 33280), (2, 33280), (0, 1000)])
def test_plonk_parser_walks_the_writer_layout(pkg, nb_commit, lines_bytes):
    PP, spr, pk, raw, _ = _plonk_key_file(10, nb_commit, 5, lines_bytes)
    F = pkg.plonk_pk_format
    p = F.parse(raw, pkg.BN254)
    n = pk.d0.cardinality
    assert p.file_bytes == len(raw) and (p.size, p.size_big, p.nb_public) == (n, pk.d1.cardinality, pk.nb_public)
    assert (p.generator, p.generator_big, p.coset_shift) == (pk.d0.generator, pk.d1.generator, pk.d0.coset_gen)
    assert p.lines_bytes == lines_bytes and p.nb_qcp == nb_commit and p.commitment_indexes == [c["index"] for c in pk.commitments]
    off, ln = p.sections["ql"]
    assert ln == 32 * n and raw[off:off + 32] == (pk.ql[0] % B.R).to_bytes(32, "big")
    off, ln = p.sections["perm"]
    assert off + ln == len(raw) and int.from_bytes(raw[off + 8:off + 16], "big", signed=True) == pk.perm[1]
    off, ln = p.sections["kzg_lagrange_g1"]
    assert ln == 64 * n
    with pytest.raises(pkg.pk_format.PkFormatError):
        F.parse(raw[:-8], pkg.BN254)
    bad = bytearray(raw); bad[7] = 3
    with pytest.raises(pkg.pk_format.PkFormatError):
        F.parse(bytes(bad), pkg.BN254)


@pytest.mark.gpu
@pytest.mark.parametrize("nb_mul,nb_commit,ranks", [(200, 0, 1), (40, 2, 1), (100, 0, 4)])
def test_plonk_key_file_to_proof(ctx, pkg, cref, nb_mul, nb_commit, ranks):
    """A WriteRawTo file goes to the device as it is (scalars, permutation and points converted by kernels) and the key
    proves: proof bytes == oracle, on one context and sharded over a device group; a damaged scalar is refused."""
    from tests import plonk_util as PU
    PP, spr, pk, raw, rng = _plonk_key_file(nb_mul, nb_commit, 77 + nb_mul)
    if nb_commit:
        sol, bsb = PP.commit_chain_solution(spr, pk, rng.randrange(2, B.R), [(rng.randrange(B.R), rng.randrange(B.R)) for _ in range(nb_commit)])
    else:
        sol, bsb = PP.mul_chain_solution(nb_mul, 5), None
    blinding = tuple([rng.randrange(B.R) for _ in range(k)] for k in (2, 2, 2, 3))
    want, _ = PP.prove(spr, pk, sol, blinding, bsb)
    L, Rr, O = (util.fr_arr(v) for v in PP.evaluate_lro(spr, sol))
    prover = pkg.Group([0] * ranks) if ranks > 1 else ctx
    try:
        dpk, parsed = pkg.plonk_pk_format.upload(prover, raw)
        try:
            got, _ = prover.plonk_prove(dpk, L, Rr, O, util.fr_arr(sol[:1]), PU.blinding_array(blinding), PU.bsb_arrays(bsb))
        finally:
            dpk.free()
        assert pkg.plonk_proof_bytes(got) == PP.proof_bytes(want)
        if ranks == 1:
            bad = bytearray(raw)
            off, _ = parsed.sections["qm"]
            bad[off:off + 32] = b"\xff" * 32                    # a scalar >= r
            with pytest.raises(pkg.B200Error, match="not canonical"):
                pkg.plonk_pk_format.upload(prover, bytes(bad))
    finally:
        if ranks > 1:
            prover.close()
