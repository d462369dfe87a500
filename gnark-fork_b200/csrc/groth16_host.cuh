// Resident Groth16 proving key and the host entry points of groth16.cu / host_group.cu.
#pragma once
#include <functional>
#include "ntt_host.cuh"
#include "msm_host.cuh"

// Device mirror of groth16_bn254.ProvingKey (setup.go:35-58) -- the analogue of the icicle
// deviceInfo (icicle/provingkey.go:10-28).  With shard_count > 1 it holds one point range.
struct b200_groth16_pk {
    unsigned log_n = 0;
    size_t n_wires = 0, nb_public = 0;
    b200::NttDomain* dom = nullptr;  // owned by the ctx cache
    b200::DevBuf A, B, K, Z, G2B;    // affine bases of this shard (or their folded-window tables)
    b200::MsmFold fA, fB, fK, fZ, fG2B;  // c == 0: plain bases
    b200::DevBuf mapA, mapB;         // wire index of each A / B base (prove.go:154-175 as a gather)
    size_t n_a = 0, n_b = 0, n_k = 0, n_z = 0;  // shard-local counts
    size_t k_lo = 0, z_lo = 0;       // first K / Z index of this shard
    size_t full_n_k = 0;             // len(pk.G1.K)
    size_t w_lo = 0, w_hi = 0;       // wire range the A / B maps of this shard reference (maps are relative to w_lo)
    ~b200_groth16_pk() { if (dom) --dom->refs; }
};

namespace b200 {
// Setup's source of the base arrays: the scalars behind them (setup.go:120-275) instead of points.  a, b, k: host arrays
// of the full (filtered) lengths; z_dev: device array, bit-reversed order, 2^log_n entries (the first n - 1 are used);
// out_*: optional host arrays that receive the generated points (full length; each shard fills its own range).
struct PkScalarSource {
    const Fr *a = nullptr, *b = nullptr, *k = nullptr, *z_dev = nullptr;
    G1Affine g1;
    G2Affine g2;
    uint64_t *out_a = nullptr, *out_b = nullptr, *out_k = nullptr, *out_z = nullptr, *out_g2b = nullptr;
};
int groth16_pk_upload(b200_ctx* ctx, const b200_groth16_pk_desc* d, b200_groth16_pk** out, const PkScalarSource* gen_src = nullptr);
// setup.cu: the scalar side of Setup (host + one device batch inversion), derived once, and the point side per shard
struct SetupScalars {
    std::vector<Fr> A, B, pkK, vkK;               // filtered A_i(tau), B_i(tau); K scalars of pk.G1.K and vk.G1.K
    std::vector<std::vector<Fr>> ckK;             // Pedersen bases' scalars per commitment
    std::vector<uint8_t> infA, infB;
    Fr tau, alpha, beta, gamma, delta, z_mult;    // z_mult = (tau^n - 1) / delta
};
int groth16_setup_scalars(b200_ctx* ctx, const b200_groth16_setup_desc* d, SetupScalars& S);
int groth16_setup_points(b200_ctx* ctx, const b200_groth16_setup_desc* d, const SetupScalars& S, int shard_index, int shard_count,
                         b200_groth16_pk** pk_out, b200_groth16_setup_out* out, bool vk_side);
int groth16_setup(b200_ctx* ctx, const b200_groth16_setup_desc* d, b200_groth16_pk** pk_out, b200_groth16_setup_out* out);
// computeH supplied by the caller and enqueued at the point of the schedule where the prove core runs its own
// (the group's distributed computeH, group.cu): copy_abc enqueues this rank's share of a, b, c on the copy stream `cp`
// (after W's copy); compute_h enqueues the transforms on ctx->stream; h_out will hold this shard's range of h.
struct ProveHHook {
    std::function<int(cudaStream_t)> copy_abc;
    std::function<int()> compute_h;
    const Fr* h_out = nullptr;
    bool inputs_on_device = false;     // computeH first (nothing to wait for), as for device-resident a, b, c
};
int groth16_prove_core(b200_ctx* ctx, const b200_groth16_pk* pk, const uint64_t* a, const uint64_t* b,
                       const uint64_t* c, size_t n_constraints, const uint64_t* w, const uint32_t* k_remove,
                       size_t n_remove, const uint64_t* h_local, b200_groth16_out* out, b200_timings* tm,
                       const ProveHHook* hook = nullptr);
}  // namespace b200
