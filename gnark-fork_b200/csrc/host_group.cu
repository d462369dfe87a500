// Host-side group arithmetic of the C ABI: partial-sum combination, FromJacobian, the Groth16
// fix-ups of /root/reference/backend/groth16/bn254/prove.go:178-321 and Proof.WriteTo
// (/root/reference/backend/groth16/bn254/marshal.go:41-66).  Pure host code (the portable path of
// field.cuh); O(1) work per proof, which the reference also keeps on the CPU (icicle.go:306-391).
#include <future>
#include "host_group.cuh"

using namespace b200;

using namespace b200::hostgrp;

extern "C" {

int b200_g1_jac_add(const uint64_t* p, const uint64_t* q, uint64_t* out) try { return jac_add<Fp>(p, q, out); } B2_ABI_CATCH_INT
int b200_g2_jac_add(const uint64_t* p, const uint64_t* q, uint64_t* out) try { return jac_add<Fp2>(p, q, out); } B2_ABI_CATCH_INT
int b200_g1_jac_to_affine(const uint64_t* p, uint64_t* out) try { return jac_aff<Fp>(p, out); } B2_ABI_CATCH_INT
int b200_g2_jac_to_affine(const uint64_t* p, uint64_t* out) try { return jac_aff<Fp2>(p, out); } B2_ABI_CATCH_INT

int b200_groth16_finalize(const b200_groth16_out* msm, const uint64_t* r, const uint64_t* s,
                          const uint64_t* g1_alpha, const uint64_t* g1_beta, const uint64_t* g1_delta,
                          const uint64_t* g2_beta, const uint64_t* g2_delta, uint64_t* ar_affine,
                          uint64_t* bs_affine, uint64_t* krs_affine) try {
    if (!msm || !r || !s || !g1_alpha || !g1_beta || !g1_delta || !g2_beta || !g2_delta || !ar_affine ||
        !bs_affine || !krs_affine)
        return B200_ERR_ARG;
    Fr fr_r, fr_s;
    memcpy(fr_r.l, r, 32);
    memcpy(fr_s.l, s, 32);
    const Fr kr = (fr_r * fr_s).neg();  // _kr = -(r*s), prove.go:187-189
    const G1XYZZ delta = load_aff<Fp>(g1_delta);
    // The seven scalar multiplications are O(1) work but serial they are ~4 ms of host time inside the end-to-end
    // step; the four that depend on r, s and the key only start at once on host tasks, s*ar and r*bs1 follow in parallel.
    // deltas = BatchScalarMultiplicationG1(&pk.G1.Delta, {r, s, kr}), prove.go:192
    auto f_dr = std::async(std::launch::async, [&]() { return scalar_mul(delta, fr_r); });
    auto f_ds = std::async(std::launch::async, [&]() { return scalar_mul(delta, fr_s); });
    auto f_dkr = std::async(std::launch::async, [&]() { return scalar_mul(delta, kr); });
    auto f_d2s = std::async(std::launch::async, [&]() { return scalar_mul(load_aff<Fp2>(g2_delta), fr_s); });
    // ar = MSM_A + alpha + r*delta, prove.go:219-220
    G1XYZZ ar = load_jac<Fp>(msm->ar);
    ar.add(load_aff<Fp>(g1_alpha));
    ar.add(f_dr.get());
    auto f_sar = std::async(std::launch::async, [&]() { return scalar_mul(ar, fr_s); });
    // bs1 = MSM_B + beta + s*delta, prove.go:206-207
    G1XYZZ bs1 = load_jac<Fp>(msm->bs1);
    bs1.add(load_aff<Fp>(g1_beta));
    bs1.add(f_ds.get());
    auto f_rbs = std::async(std::launch::async, [&]() { return scalar_mul(bs1, fr_r); });
    // krs = MSM_K + kr*delta + MSM_Z + s*ar + r*bs1, prove.go:248-271
    G1XYZZ krs = load_jac<Fp>(msm->krs);
    krs.add(f_dkr.get());
    krs.add(load_jac<Fp>(msm->krs2));
    krs.add(f_sar.get());
    krs.add(f_rbs.get());
    // Bs = MSM_B2 + s*delta2 + beta2, prove.go:294-297
    G2XYZZ bs = load_jac<Fp2>(msm->bs2);
    bs.add(f_d2s.get());
    bs.add(load_aff<Fp2>(g2_beta));
    const G1Affine a_aff = ar.to_affine(), k_aff = krs.to_affine();
    const G2Affine b_aff = bs.to_affine();
    memcpy(ar_affine, &a_aff, sizeof a_aff);
    memcpy(bs_affine, &b_aff, sizeof b_aff);
    memcpy(krs_affine, &k_aff, sizeof k_aff);
    return B200_OK;
} B2_ABI_CATCH_INT

int b200_groth16_proof_bytes(const uint64_t* ar_affine, const uint64_t* bs_affine, const uint64_t* krs_affine,
                             uint8_t* out, size_t out_cap) try {
    if (!ar_affine || !bs_affine || !krs_affine || !out) return B200_ERR_ARG;
    const size_t FB = FP_BYTES, total = 5 * FB + 4;   // 164 bytes on BN254, 244 on BLS12-377
    if (out_cap < total) return B200_ERR_ARG;
    G1Affine ar, krs;
    G2Affine bs;
    memcpy(&ar, ar_affine, sizeof ar);
    memcpy(&bs, bs_affine, sizeof bs);
    memcpy(&krs, krs_affine, sizeof krs);
    g1_compress(ar, out);          // marshal.go:49
    g2_compress(bs, out + FB);     // marshal.go:52
    g1_compress(krs, out + 3 * FB);    // marshal.go:55
    memset(out + 4 * FB, 0, 4);       // u32 len(Commitments) = 0, marshal.go:58
    g1_compress(G1Affine::inf(), out + 4 * FB + 4);  // CommitmentPok, marshal.go:61
    return (int)total;
} B2_ABI_CATCH_INT

}  // extern "C"
