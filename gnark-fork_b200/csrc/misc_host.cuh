#pragma once
#include "msm.cuh"
namespace b200 {
int vec_mul_fr(b200_ctx* ctx, const Fr* a, const Fr* b, Fr* out, size_t n);
int vec_mul_fp(b200_ctx* ctx, const Fp* a, const Fp* b, Fp* out, size_t n);
int imad_peak(b200_ctx* ctx, double* mac32_per_s, double* imad32_per_s, double* sm_mhz);
int batch_mul_g1(b200_ctx* ctx, const G1Affine& base, const Fr* d_scalars, size_t n, G1Affine* d_out);
int batch_mul_g2(b200_ctx* ctx, const G2Affine& base, const Fr* d_scalars, size_t n, G2Affine* d_out);
// kzg_lagrange.cu: kzg.ToLagrangeG1 (plonk setup.go:134); host affine points in and out
int kzg_to_lagrange_g1(b200_ctx* ctx, const uint64_t* srs_host, unsigned log_n, const Fr& gen, uint64_t* out_host);
// gnark-crypto RawEncoding point bytes -> affine Montgomery limbs, in place (both are sizeof(Affine) per point).
// *bad (device word, may be null) gets 1 if any point carries a compressed-form flag or a coordinate >= p.
// big-endian canonical fr.Element / big-endian int64 sections of a key file -> device layout, in place
int fr_from_be_inplace(b200_ctx* ctx, Fr* d, size_t n, uint32_t* d_bad);
int i64_from_be_inplace(b200_ctx* ctx, int64_t* d, size_t n);
int points_from_raw_inplace(b200_ctx* ctx, void* d_points, size_t n, int group, uint32_t* d_bad);
}  // namespace b200
