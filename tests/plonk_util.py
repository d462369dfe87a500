"""Helpers of the PlonK prover tests: oracle key/solution -> the arrays the C ABI takes, proof conversions."""
import numpy as np

from oracle import bn254 as B
from oracle import cref
from oracle import plonk_prove as PP
from tests import util


def vk_points(pk: PP.ProvingKey):
    vk = pk.vk
    return [vk["S"][0], vk["S"][1], vk["S"][2], vk["Ql"], vk["Qr"], vk["Qm"], vk["Qo"], vk["Qk"]]


def srs_arrays(pk: PP.ProvingKey):
    """pk.Kzg.G1 (n + 3) and pk.KzgLagrange.G1 (n) as affine Montgomery arrays (oracle fixed-base batch mul)."""
    mono, lag = PP.srs_scalars(pk)
    gen = cref.g1_points_to_array([B.G1_GEN])[0]
    return cref.g1_batch_mul(gen, cref.ints_to_mont(mono, B.R)), cref.g1_batch_mul(gen, cref.ints_to_mont(lag, B.R))


def upload_pk(ctx, pk: PP.ProvingKey, shard_index=0, shard_count=1):
    kzg, kzg_lag = srs_arrays(pk)
    f = util.fr_arr
    return ctx.plonk_pk_upload(
        log_n=pk.d0.log_n, log_big=pk.d1.log_n, generator=cref.fr(pk.d0.generator), generator_big=cref.fr(pk.d1.generator),
        coset_gen=cref.fr(pk.d0.coset_gen), ql=f(pk.ql), qr=f(pk.qr), qm=f(pk.qm), qo=f(pk.qo), qk=f(pk.qk),
        s1=f(pk.s1), s2=f(pk.s2), s3=f(pk.s3), perm=np.array(pk.perm, dtype=np.int64), kzg_g1=kzg, kzg_lagrange_g1=kzg_lag,
        vk_commitments=util.g1_arr(vk_points(pk)), nb_public=pk.nb_public,
        qcp=[f(q) for q in pk.qcp], vk_qcp=util.g1_arr(pk.vk["Qcp"]) if pk.qcp else None,
        commitment_index=[c["index"] for c in pk.commitments], shard_index=shard_index, shard_count=shard_count)


def bsb_arrays(bsb):
    """oracle bsb dict -> the arrays Context.plonk_prove takes."""
    if not bsb or not bsb["pi2"]:
        return None
    return {"pi2": [util.fr_arr(c) for c in bsb["pi2"]], "commitments": util.g1_arr(bsb["commitments"]), "values": util.fr_arr(bsb["values"])}


def blinding_array(blinding):
    bl, br, bo, bz = blinding
    return util.fr_arr(list(bl) + list(br) + list(bo) + list(bz))


def proof_to_struct(pkg, proof: PP.Proof):
    """oracle Proof -> PlonkProof (ctypes)."""
    s = pkg.PlonkProof()

    def put(dst, arr):
        for i, v in enumerate(np.asarray(arr).reshape(-1)):
            dst[i] = int(v)
    for k in range(3):
        put(s.lro[k], util.g1_arr([proof.LRO[k]]))
        put(s.h[k], util.g1_arr([proof.H[k]]))
    put(s.z, util.g1_arr([proof.Z]))
    put(s.batched_h, util.g1_arr([proof.batched_H]))
    put(s.zshift_h, util.g1_arr([proof.zshift_H]))
    for k in range(7):
        put(s.claimed[k], util.fr_arr([proof.claimed[k]]))
    put(s.zshift_value, util.fr_arr([proof.zshift_value]))
    s.nb_bsb22 = len(proof.bsb22)
    for k in range(len(proof.bsb22)):
        put(s.bsb22[k], util.g1_arr([proof.bsb22[k]]))
        put(s.claimed_qcp[k], util.fr_arr([proof.claimed[7 + k]]))
    return s


def struct_to_proof(s) -> PP.Proof:
    d = s.as_dict()

    def pt(a):
        p = util.g1_aff_ints(a)
        return None if p is None else (p[0], p[1])
    nb = int(s.nb_bsb22)
    return PP.Proof([pt(d["lro"][k]) for k in range(3)], pt(d["z"]), [pt(d["h"][k]) for k in range(3)], pt(d["batched_h"]),
                    util.fr_ints(s.all_claimed()), pt(d["zshift_h"]), util.fr_ints(d["zshift_value"].reshape(1, 4))[0],
                    [pt(d["bsb22"][k]) for k in range(nb)])
