/* b200prover.h -- C ABI of the B200-native (sm_100a) proving backend for gnark's Groth16 hot path.
 *
 * This is the boundary a cgo shim binds.  It replaces, call for call, the
 * device-library surface the reference reaches through its `icicle` build tag
 * (/root/reference/backend/groth16/bn254/icicle/icicle.go) and the gnark-crypto calls of the CPU
 * prover (/root/reference/backend/groth16/bn254/prove.go).  Each entry point cites what it
 * replaces.
 *
 * Conventions
 *   - Every field element is 4 little-endian u64 limbs in MONTGOMERY form (R = 2^256), exactly as a
 *     gnark-crypto fr.Element / fp.Element sits in Go memory, so the glue passes
 *     unsafe.Pointer(&slice[0]) with no conversion.
 *   - G1Affine = X,Y (8 u64; (0,0) = infinity); G1Jac = X,Y,Z (12 u64; Z = 0 = infinity);
 *     G2Affine = X.A0,X.A1,Y.A0,Y.A1 (16 u64); G2Jac = 24 u64.
 *   - BLS12-377 (BASELINE.json config 5) is a SECOND library built from the same sources,
 *     libb200prover_bls12377.so (make CURVE=bls12_377; bind it with -DB200_CURVE_BLS12_377): same entry
 *     points, curve argument B200_BLS12_377, Fr still 4 u64 but Fp = 6 u64 (R = 2^384), so G1Affine = 12,
 *     G1Jac = 18, G2Affine = 24, G2Jac = 36 u64 -- B200_FP_WORDS below sizes the structs.  Each library
 *     rejects the other curve's id with B200_ERR_UNSUPPORTED.  Both libraries hold the Groth16 and the PlonK prover.
 *   - Jacobian results are SOME representative of the point (not gnark-crypto's particular
 *     one); equality with MultiExp is defined on the affine-normalised point (FromJacobian).
 *   - Return value: 0 on success, a negative b200_status otherwise; b200_last_error(ctx) holds a
 *     message.  No exception, abort or callback crosses the boundary.
 *   - The caller owns host buffers for the duration of the call only.  Handles (ctx, domain,
 *     bases, pk) own device memory and are freed explicitly.
 *   - A ctx is bound to ONE device and is not re-entrant (serialise calls per ctx); distinct
 *     ctx are independent.  Multi-GPU: a b200_group (one process, one call per proof over all devices; below), or
 *     one ctx per device/rank with shards selected by (shard_index, shard_count) and the partial sums combined by
 *     the caller (b200_g1_jac_add ...).
 *   - There is no CPU fallback: without a CUDA device b200_init fails with B200_ERR_CUDA.
 *   - Zero-length MSM returns the point at infinity; infinity bases (0,0) and zero scalars are
 *     legal inputs.
 */
#ifndef B200PROVER_H
#define B200PROVER_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Entry points are the only exported symbols of libb200prover.so. */
#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

/* BEGIN per-curve symbol names (generated: scripts/gen_symbol_renames.py) */
#if defined(B200_CURVE_BLS12_377) || defined(B2_CURVE_BLS12_377)
#define b200_bases_free b200_bls12377_bases_free
#define b200_bases_info b200_bls12377_bases_info
#define b200_bases_upload b200_bls12377_bases_upload
#define b200_bases_upload_raw b200_bls12377_bases_upload_raw
#define b200_batch_scalar_mul b200_bls12377_batch_scalar_mul
#define b200_build_curve b200_bls12377_build_curve
#define b200_build_fp_words b200_bls12377_build_fp_words
#define b200_d2h b200_bls12377_d2h
#define b200_destroy b200_bls12377_destroy
#define b200_dev_alloc b200_bls12377_dev_alloc
#define b200_dev_free b200_bls12377_dev_free
#define b200_device_count b200_bls12377_device_count
#define b200_domain_cache_trim b200_bls12377_domain_cache_trim
#define b200_fp_vec_mul b200_bls12377_fp_vec_mul
#define b200_fr_batch_inverse_dev b200_bls12377_fr_batch_inverse_dev
#define b200_fr_mul_sub_dev b200_bls12377_fr_mul_sub_dev
#define b200_fr_poly_div_linear_dev b200_bls12377_fr_poly_div_linear_dev
#define b200_fr_poly_eval_dev b200_bls12377_fr_poly_eval_dev
#define b200_fr_prefix_product_dev b200_bls12377_fr_prefix_product_dev
#define b200_fr_prefix_product_static b200_bls12377_fr_prefix_product_static
#define b200_fr_vec_mul b200_bls12377_fr_vec_mul
#define b200_g1_jac_add b200_bls12377_g1_jac_add
#define b200_g1_jac_to_affine b200_bls12377_g1_jac_to_affine
#define b200_g2_jac_add b200_bls12377_g2_jac_add
#define b200_g2_jac_to_affine b200_bls12377_g2_jac_to_affine
#define b200_groth16_compute_h b200_bls12377_groth16_compute_h
#define b200_groth16_finalize b200_bls12377_groth16_finalize
#define b200_groth16_msms_dev b200_bls12377_groth16_msms_dev
#define b200_groth16_pk_free b200_bls12377_groth16_pk_free
#define b200_groth16_pk_info b200_bls12377_groth16_pk_info
#define b200_groth16_pk_upload b200_bls12377_groth16_pk_upload
#define b200_groth16_pk_wire_range b200_bls12377_groth16_pk_wire_range
#define b200_groth16_proof_bytes b200_bls12377_groth16_proof_bytes
#define b200_groth16_prove_core b200_bls12377_groth16_prove_core
#define b200_groth16_prove_core_dev b200_bls12377_groth16_prove_core_dev
#define b200_groth16_setup b200_bls12377_groth16_setup
#define b200_group_ctx b200_bls12377_group_ctx
#define b200_group_destroy b200_bls12377_group_destroy
#define b200_group_groth16_pk_free b200_bls12377_group_groth16_pk_free
#define b200_group_groth16_pk_info b200_bls12377_group_groth16_pk_info
#define b200_group_groth16_pk_upload b200_bls12377_group_groth16_pk_upload
#define b200_group_groth16_prove_core b200_bls12377_group_groth16_prove_core
#define b200_group_groth16_prove_core_dev b200_bls12377_group_groth16_prove_core_dev
#define b200_group_groth16_setup b200_bls12377_group_groth16_setup
#define b200_group_init b200_bls12377_group_init
#define b200_group_last_error b200_bls12377_group_last_error
#define b200_group_plonk_pk_free b200_bls12377_group_plonk_pk_free
#define b200_group_plonk_pk_upload b200_bls12377_group_plonk_pk_upload
#define b200_group_plonk_prove b200_bls12377_group_plonk_prove
#define b200_group_set_option b200_bls12377_group_set_option
#define b200_group_size b200_bls12377_group_size
#define b200_h2d b200_bls12377_h2d
#define b200_host_alloc b200_bls12377_host_alloc
#define b200_host_free b200_bls12377_host_free
#define b200_imad_peak b200_bls12377_imad_peak
#define b200_init b200_bls12377_init
#define b200_kzg_commit_dev b200_bls12377_kzg_commit_dev
#define b200_kzg_to_lagrange_g1 b200_bls12377_kzg_to_lagrange_g1
#define b200_last_error b200_bls12377_last_error
#define b200_msm_dev b200_bls12377_msm_dev
#define b200_msm_g1 b200_bls12377_msm_g1
#define b200_msm_g2 b200_bls12377_msm_g2
#define b200_msm_resident b200_bls12377_msm_resident
#define b200_ntt b200_bls12377_ntt
#define b200_ntt_dev b200_bls12377_ntt_dev
#define b200_ntt_dist_phase b200_bls12377_ntt_dist_phase
#define b200_plonk_build_ratio_dev b200_bls12377_plonk_build_ratio_dev
#define b200_plonk_derive_challenges b200_bls12377_plonk_derive_challenges
#define b200_plonk_divide_by_xn_minus_one_dev b200_bls12377_plonk_divide_by_xn_minus_one_dev
#define b200_plonk_pk_free b200_bls12377_plonk_pk_free
#define b200_plonk_pk_upload b200_bls12377_plonk_pk_upload
#define b200_plonk_proof_bytes b200_bls12377_plonk_proof_bytes
#define b200_plonk_prove b200_bls12377_plonk_prove
#define b200_plonk_prove_bsb22 b200_bls12377_plonk_prove_bsb22
#define b200_plonk_quotient_rows_dev b200_bls12377_plonk_quotient_rows_dev
#define b200_plonk_session_begin b200_bls12377_plonk_session_begin
#define b200_plonk_session_finish b200_bls12377_plonk_session_finish
#define b200_plonk_session_numerator b200_bls12377_plonk_session_numerator
#define b200_plonk_session_step b200_bls12377_plonk_session_step
#define b200_points_from_raw b200_bls12377_points_from_raw
#define b200_profile_read b200_bls12377_profile_read
#define b200_profile_start b200_bls12377_profile_start
#define b200_set_option b200_bls12377_set_option
#define b200_sync b200_bls12377_sync
#define b200_use_stream b200_bls12377_use_stream
#define b200_version b200_bls12377_version
#endif
/* END per-curve symbol names */

typedef struct b200_ctx b200_ctx;
typedef struct b200_bases b200_bases;             /* resident MSM base set (G1 or G2)         */
typedef struct b200_groth16_pk b200_groth16_pk;   /* resident Groth16 proving key (or shard)  */

enum b200_status {
    B200_OK = 0,
    B200_ERR_CUDA = -1,        /* CUDA runtime error (message in b200_last_error)              */
    B200_ERR_ARG = -2,         /* invalid argument                                             */
    B200_ERR_NOMEM = -3,       /* device or pinned-host allocation failed                      */
    B200_ERR_UNSUPPORTED = -4, /* curve / size not supported by this build                     */
    B200_ERR_INTERNAL = -5     /* a C++ exception was stopped at the boundary (message kept)   */
};

enum b200_curve { B200_BN254 = 0, B200_BLS12_377 = 1 };

/* u64 words of one base-field element in the library this header is bound to */
#if defined(B200_CURVE_BLS12_377) || defined(B2_CURVE_BLS12_377)
#define B200_FP_WORDS 6
#else
#define B200_FP_WORDS 4
#endif
/* curve id and Fp words the loaded library was built for (a binder checks them once at start-up) */
int b200_build_curve(void);
int b200_build_fp_words(void);

#define B200_DIF 0 /* fft.DIF: natural order in, bit-reversed out */
#define B200_DIT 1 /* fft.DIT: bit-reversed in, natural out       */

/* ---- context ---------------------------------------------------------------------------------
 * icicle has no context (implicit default device, icicle.go:31-130); cgo calls hop OS threads, so
 * every entry point sets the device itself. */
int b200_device_count(void);
int b200_init(int device, b200_ctx **out);
void b200_destroy(b200_ctx *ctx);
const char *b200_last_error(const b200_ctx *ctx);
/* Version string of the library build (arch, git describe when available). */
const char *b200_version(void);

/* Tuning knobs (defaults are the measured best): "msm_c" (window width of the plain path, 0 =
 * automatic), "msm_precompute" (1 = resident base sets get folded-window tables 2^(c w) P_i, W x
 * the base memory, built once at upload; falls back to plain bases when the table does not fit),
 * "msm_fold_c" (window width of the folded path, 0 = automatic), "msm_overlap" (1 = the prove core
 * sorts the next MSM on a second stream while the current one accumulates; 0 = strictly serial), "msm_acc_smem" (bit 0 /
 * bit 1: G1 / G2 bucket accumulation keeps its running sum in shared memory; default G2 only on BN254, off on
 * BLS12-377), "msm_ba" (batched-affine pre-reduction rounds, 15 = chosen from the mean bucket load: the default), "aux_persist_min" (entries = windows x points from which an MSM's sort and round-0 denominator pass, when they run beside
 * another MSM's bucket stage, are launched as small persistent grids; -1 = the measured default, 5e7), "h_pos" (proofs from host buffers: computeH is enqueued after the
 * accumulation of MSM 0 / 1 / 2 = A / Bs1 / Krs, so that more of the proof covers the copy of a, b, c; -1 = automatic), "ntt_tile_log" (log2 of the
 * shared-memory tile, 6..12), "ntt_strided_stages" (max stages of a strided NTT pass, 1..12),
 * "ntt_threads" (threads per NTT CTA, power of two), "ntt_tma" (1 = NTT passes staged by TMA: one cp.async.bulk.tensor
 * per tile in, one out; 0 = per-thread vector loads). */
int b200_set_option(b200_ctx *ctx, const char *key, long value);
/* Run all subsequent work of the context on the caller's CUDA stream (a cudaStream_t of the
 * context's device; NULL: back to the context's own stream).  Drains the previous stream first. */
int b200_use_stream(b200_ctx *ctx, void *cuda_stream);

/* Pinned host memory for the solver outputs (prove.go:119-125 W, A, B, C) so that the per-proof
 * H2D copies run at PCIe speed; plain Go slices also work (pageable path). */
void *b200_host_alloc(size_t bytes);
void b200_host_free(void *p);

/* Raw device buffers for the *_dev entry points (iciclegnark.CopyToDevice / FreeDevicePointer,
 * icicle.go:44-65,231-278,416-418). */
int b200_dev_alloc(b200_ctx *ctx, size_t bytes, void **dptr);
int b200_dev_free(b200_ctx *ctx, void *dptr);
int b200_h2d(b200_ctx *ctx, void *dst_dev, const void *src_host, size_t bytes);
int b200_d2h(b200_ctx *ctx, void *dst_host, const void *src_dev, size_t bytes);
int b200_sync(b200_ctx *ctx);

/* ---- element-wise ----------------------------------------------------------------------------
 * out[i] = a[i] * b[i] in Fr / Fp (gnark-crypto fr.Element.Mul / fp.Element.Mul); host buffers. */
int b200_fr_vec_mul(b200_ctx *ctx, int curve, const uint64_t *a, const uint64_t *b, uint64_t *out, size_t n);
int b200_fp_vec_mul(b200_ctx *ctx, int curve, const uint64_t *a, const uint64_t *b, uint64_t *out, size_t n);
/* Roofline denominators measured on this device: register-resident mad.lo.cc/madc.hi.cc chains
 * (32x32+64 multiply-accumulates per second, chip-wide) and the same for dependent-free plain
 * 32-bit IMAD; sm_mhz = clock observed by the kernel (cycles / elapsed). */
int b200_imad_peak(b200_ctx *ctx, double *mac32_per_s, double *imad32_per_s, double *sm_mhz);

/* Per-kernel-class device times, accumulated between b200_profile_start and b200_profile_read
 * from CUDA event pairs recorded on the ctx stream around each launch (used by bench.py for the
 * roofline of the dominant kernel; adds two event records per launch while enabled). */
enum b200_prof_slot {
    B200_PROF_MSM_SORT = 0,    /* digits + histogram + scan + scatter + task lists               */
    B200_PROF_MSM_ACC_G1 = 1,  /* msm_accumulate_kernel<Fp>  (bucket accumulation, G1)           */
    B200_PROF_MSM_ACC_G2 = 2,  /* msm_accumulate_kernel<Fp2> (bucket accumulation, G2)           */
    B200_PROF_MSM_REDUCE = 3,  /* split-bucket combine + bucket reduction tree                   */
    B200_PROF_NTT_PASS = 4,    /* ntt_pass_kernel launches                                       */
    B200_PROF_POINTWISE = 5,   /* a*b - c                                                        */
    B200_PROF_COPY_IN = 6,     /* H2D / D2D staging of a, b, c, W                                */
    B200_PROF_SLOTS = 8
};
typedef struct b200_profile {
    double ms[B200_PROF_SLOTS];       /* summed device time per slot                             */
    uint64_t launches[B200_PROF_SLOTS]; /* spans per slot                                        */
    double units[B200_PROF_SLOTS];    /* work units: points (MSM slots), elements*stages/2 =
                                         butterflies (NTT), elements (pointwise), bytes (copy)   */
    uint64_t kernel_launches;         /* every kernel of this library launched since start       */
} b200_profile;
int b200_profile_start(b200_ctx *ctx);
int b200_profile_read(b200_ctx *ctx, b200_profile *out); /* synchronises, returns and clears    */

/* ---- NTT -------------------------------------------------------------------------------------
 * fft.Domain.FFT / FFTInverse (prove.go:369-375,393); replaces iciclegnark.NttOnDevice /
 * INttOnDevice (icicle.go:489,490,502) and GenerateTwiddleFactors (:68,:73).
 *   inverse    0 = Domain.FFT, 1 = Domain.FFTInverse (scaled by 1/n)
 *   decimation B200_DIF or B200_DIT (gnark-crypto semantics: no reordering pass is applied)
 *   on_coset   fft.OnCoset(): evaluate on / interpolate from coset_gen * <generator>
 *   generator  Domain.Generator (primitive 2^log_n-th root), coset_gen = Domain.FrMultiplicativeGen;
 *              both are INPUTS taken from pk.Domain so the device never recomputes a convention.
 * Twiddle and coset tables are built on the device once per (log_n, generator, coset_gen) and
 * cached in the ctx.
 * (see b200_domain_cache_trim below) */
int b200_ntt(b200_ctx *ctx, int curve, uint64_t *host_inout, unsigned log_n, int inverse, int decimation,
             int on_coset, const uint64_t *generator, const uint64_t *coset_gen);
/* The domain cache only grows by itself (a 2^24 domain with its computeH tables is ~2 GiB): a long-lived prover that has
 * freed the keys of a size calls b200_domain_cache_trim, which releases every cached domain no resident key refers
 * to (bytes_freed optional). */
int b200_domain_cache_trim(b200_ctx *ctx, size_t *bytes_freed);
/* Same as b200_ntt on a device-resident vector; kernel_ms (optional) receives the CUDA-event time. */
int b200_ntt_dev(b200_ctx *ctx, int curve, void *dev_inout, unsigned log_n, int inverse, int decimation,
                 int on_coset, const uint64_t *generator, const uint64_t *coset_gen, float *kernel_ms);

/* computeH (prove.go:353-396; icicle.go:453-513): a, b, c hold n_constraints elements (host),
 * are zero-padded to n = 2^log_n on the device; h_out receives n elements in BIT-REVERSED order
 * (prove.go:393 leaves them so; Setup bit-reverses pk.G1.Z to match, setup.go:265). */
int b200_groth16_compute_h(b200_ctx *ctx, int curve, const uint64_t *a, const uint64_t *b, const uint64_t *c,
                           size_t n_constraints, unsigned log_n, const uint64_t *generator,
                           const uint64_t *coset_gen, uint64_t *h_out);

/* ---- MSM -------------------------------------------------------------------------------------
 * G1Jac.MultiExp / G2Jac.MultiExp (prove.go:201,214,234,244,290); replaces
 * iciclegnark.MsmOnDevice / MsmG2OnDevice (icicle.go:302,315,332,355,382).  The MultiExpConfig
 * (NbTasks) of the reference has no meaning on the device and is not part of the ABI. */
int b200_msm_g1(b200_ctx *ctx, int curve, const uint64_t *scalars, const uint64_t *bases_affine, size_t n,
                uint64_t *out_jac);
int b200_msm_g2(b200_ctx *ctx, int curve, const uint64_t *scalars, const uint64_t *bases_affine, size_t n,
                uint64_t *out_jac);
/* Resident base sets (iciclegnark.CopyPointsToDevice / CopyG2PointsToDevice, icicle.go:88-126).
 * group = 1 (G1) or 2 (G2). */
int b200_bases_upload(b200_ctx *ctx, int curve, int group, const uint64_t *bases_affine, size_t n,
                      b200_bases **out);
/* The same from the ON-DISK form of a point slice: `raw` = n points as gnark-crypto's RawEncoding writes them
 * (curve.NewEncoder(w, curve.RawEncoding()), used by ProvingKey.WriteRawTo, groth16/bn254/marshal.go:246-304, and
 * decoded point by point on the CPU by ReadFrom / UnsafeReadFrom, marshal.go:319-374): X || Y, each coordinate
 * 8*B200_FP_WORDS big-endian canonical bytes (G2: X.A1 || X.A0 || Y.A1 || Y.A0); top bits of byte 0 = 00, or 01 for
 * the point at infinity (rest zero).  The bytes are copied as they are and converted on the device (byte order,
 * Montgomery form, G2 coordinate order); no curve or subgroup check (UnsafeReadFrom semantics).  A compressed-form
 * flag or a coordinate >= p fails with B200_ERR_ARG.  b200_points_from_raw returns the converted points in the
 * in-memory layout (n x G1Affine / G2Affine) instead of keeping them. */
int b200_bases_upload_raw(b200_ctx *ctx, int curve, int group, const uint8_t *raw, size_t n, b200_bases **out);
int b200_points_from_raw(b200_ctx *ctx, int curve, int group, const uint8_t *raw, size_t n, uint64_t *out_affine);
int b200_bases_free(b200_ctx *ctx, b200_bases *bases);
/* How a resident base set is laid out and which Pippenger plan its MSMs run (for roofline
 * accounting: a launch performs n * windows bucket additions). */
typedef struct b200_msm_set_info {
    uint64_t n;            /* points */
    uint64_t device_bytes; /* bases or folded-window table */
    int32_t c;             /* window width */
    int32_t windows;       /* signed-digit windows per scalar */
    int32_t folded;        /* 1: precomputed 2^(c w) P table, one shared bucket set */
    int32_t ba_rounds;     /* batched-affine pairwise rounds in front of the XYZZ accumulation (0: none): after them 1/2^rounds of
                            * the n * windows additions are left for msm_accumulate_kernel */
} b200_msm_set_info;
int b200_bases_info(b200_ctx *ctx, const b200_bases *bases, b200_msm_set_info *out);
/* scalars on the host (copied in) ... */
int b200_msm_resident(b200_ctx *ctx, const b200_bases *bases, const uint64_t *scalars, size_t n,
                      uint64_t *out_jac);
/* ... or already on the device; kernel_ms (optional) = CUDA-event time of the device work. */
int b200_msm_dev(b200_ctx *ctx, const b200_bases *bases, const void *scalars_dev, size_t n, uint64_t *out_jac,
                 float *kernel_ms);

/* Host-side group helpers for combining per-GPU partial sums and for tests
 * (G1Jac.AddAssign / FromJacobian semantics, prove.go:221,257-276,299). */
int b200_g1_jac_add(const uint64_t *p_jac, const uint64_t *q_jac, uint64_t *out_jac);
int b200_g2_jac_add(const uint64_t *p_jac, const uint64_t *q_jac, uint64_t *out_jac);
int b200_g1_jac_to_affine(const uint64_t *p_jac, uint64_t *out_affine);
int b200_g2_jac_to_affine(const uint64_t *p_jac, uint64_t *out_affine);

/* Fixed-base batch scalar multiplication on the device: out[i] = scalars[i] * base (affine),
 * curve.BatchScalarMultiplicationG1 / G2 (setup.go:251,308; prove.go:192).  group = 1 or 2.
 * Used to build large synthetic proving keys without Go. */
int b200_batch_scalar_mul(b200_ctx *ctx, int curve, int group, const uint64_t *base_affine,
                          const uint64_t *scalars, size_t n, uint64_t *out_affine);

/* kzg.ToLagrangeG1 (gnark-crypto kzg, as called by plonk Setup: /root/reference/backend/plonk/bn254/setup.go:134
 * `pk.KzgLagrange.G1, err = kzg.ToLagrangeG1(kzgSrs.Pk.G1[:int(vk.Size)])`): the SRS in Lagrange form,
 * out[i] = [L_i(tau)]G1, from the monomial SRS srs_g1_affine[j] = [tau^j]G1, j < 2^log_n, without knowing tau -- the
 * inverse DFT of the SRS as a vector of group elements (radix-2 DIF over the group with the inverse twiddles, bit
 * reversal, scaling by 1/n).  generator: the domain's Generator (fft.NewDomain(2^log_n)), Montgomery form.  Host
 * pointers, affine points in and out.   */
int b200_kzg_to_lagrange_g1(b200_ctx *ctx, int curve, const uint64_t *srs_g1_affine, unsigned log_n,
                            const uint64_t *generator, uint64_t *out_affine);

/* ---- Groth16 ---------------------------------------------------------------------------------
 * One-time residency of the proving key (setupDevicePointers, icicle.go:31-130).  All pointers are
 * host pointers into the Go ProvingKey (setup.go:35-58); the descriptor is only read during the
 * call.  With shard_count > 1 only the point range [i*len/shard_count, (i+1)*len/shard_count) of
 * every base array is uploaded and proved on this ctx (MSM sharding by point range). */
typedef struct b200_groth16_pk_desc {
    unsigned log_n;               /* pk.Domain.Cardinality = 2^log_n                           */
    const uint64_t *generator;    /* pk.Domain.Generator            (4 u64)                    */
    const uint64_t *coset_gen;    /* pk.Domain.FrMultiplicativeGen  (4 u64)                    */
    const uint64_t *g1_a; size_t n_a;    /* pk.G1.A (infinity points already stripped)        */
    const uint64_t *g1_b; size_t n_b;    /* pk.G1.B                                            */
    const uint64_t *g1_k; size_t n_k;    /* pk.G1.K (private wires)                            */
    const uint64_t *g1_z; size_t n_z;    /* pk.G1.Z (bit-reversed, n-1 entries)                */
    const uint64_t *g2_b; size_t n_g2b;  /* pk.G2.B                                            */
    const uint8_t *infinity_a;    /* pk.InfinityA, one byte per wire (Go []bool)               */
    const uint8_t *infinity_b;    /* pk.InfinityB                                              */
    size_t n_wires;
    size_t nb_public;             /* r1cs.GetNbPublicVariables() (K scalars = W[nb_public:])   */
    int shard_index, shard_count; /* 0,1 for a single GPU                                      */
    int presliced;                /* non-zero: every base pointer already addresses the first
                                     element of this shard's range (counts stay the full ones)  */
    int raw_points;               /* non-zero: the five base pointers address gnark-crypto RawEncoding
                                     bytes -- sections of a ProvingKey.WriteRawTo file
                                     (groth16/bn254/marshal.go:246-304) -- instead of in-memory
                                     limbs; same size per point, converted on the device
                                     (see b200_points_from_raw)                                 */
} b200_groth16_pk_desc;

int b200_groth16_pk_upload(b200_ctx *ctx, int curve, const b200_groth16_pk_desc *desc, b200_groth16_pk **out);
int b200_groth16_pk_free(b200_ctx *ctx, b200_groth16_pk *pk);
/* out[0..4] = G1.A, G1.B, G1.K, G1.Z, G2.B of this shard. */
int b200_groth16_pk_info(b200_ctx *ctx, const b200_groth16_pk *pk, b200_msm_set_info out[5]);
/* Wire range [lo, hi) of the solution vector W this shard reads (its A, B and K ranges): only
 * that slice of `w` is copied per proof. */
int b200_groth16_pk_wire_range(b200_ctx *ctx, const b200_groth16_pk *pk, size_t *lo, size_t *hi);

/* ---- Groth16 Setup on the device -----------------------------------------
 * groth16_bn254.Setup (/root/reference/backend/groth16/bn254/setup.go:85-337) with the toxic waste supplied by the
 * caller (sampleToxicWaste stays in Go): the Lagrange evaluations L_j(tau) for the whole domain (one batch inversion on
 * the device, where setupABC :352-433 walks a sequential recurrence), A_i(tau), B_i(tau), C_i(tau) accumulated over the
 * constraint matrices on host threads, the K scalars (beta A + alpha B + C) / gamma or / delta (:144-200), the Z scalars
 * (tau^n - 1) tau^i / delta bit-reversed (:203-214,265-267), the infinity filters (:217-243), and the two
 * BatchScalarMultiplication calls (:251,308) on the device fixed-base kernel.  The proving key is left RESIDENT (the
 * same handle b200_groth16_pk_upload returns, this shard's point ranges only); the verifying-key side and, when asked
 * for, the proving key's point arrays come back to the host so that Go can fill its own ProvingKey / VerifyingKey.
 * A matrix is CSR over the constraints; a term is constraint.Term {CID, VID} (constraint/term.go:31-33). */
typedef struct b200_r1cs_matrix {
    const uint64_t *row_ptr;   /* nb_constraints + 1 term offsets                                     */
    const uint32_t *terms;     /* 2 u32 per term: coefficient id (index into coeffs), wire id          */
} b200_r1cs_matrix;

typedef struct b200_groth16_setup_desc {
    unsigned log_n;                     /* fft.NewDomain(nbConstraints): Cardinality = 2^log_n            */
    const uint64_t *generator;          /* Domain.Generator (4 u64)                                       */
    const uint64_t *coset_gen;          /* Domain.FrMultiplicativeGen                                     */
    size_t nb_constraints, nb_wires, nb_public;   /* nb_public = r1cs.GetNbPublicVariables() (ONE included) */
    b200_r1cs_matrix l, r, o;           /* the L, R, O linear expressions of every R1C                    */
    const uint64_t *coeffs; size_t nb_coeffs;     /* r1cs.Coefficients (Montgomery Fr)                     */
    /* api.Commit circuits (setup.go:99-107,155-200): one byte per wire, NULL = none.  0: ordinary wire;
     * 1: a commitment wire (its K goes to vk.G1.K after the public ones, divided by gamma);
     * 2 + j: a private wire committed to by commitment j (its K goes to the j-th Pedersen basis, divided by gamma). */
    const uint8_t *wire_class;
    size_t nb_commitments;
    uint64_t tau[4], alpha[4], beta[4], gamma[4], delta[4];   /* toxic waste (Montgomery Fr)               */
    const uint64_t *g1_gen, *g2_gen;    /* curve.Generators(): affine G1 (2 Fp) and G2 (4 Fp)              */
    int shard_index, shard_count;       /* as in b200_groth16_pk_desc: 0, 1 = the whole key on this device  */
} b200_groth16_setup_desc;

typedef struct b200_groth16_setup_out {
    uint64_t g1_alpha[2 * B200_FP_WORDS], g1_beta[2 * B200_FP_WORDS], g1_delta[2 * B200_FP_WORDS];
    uint64_t g2_beta[4 * B200_FP_WORDS], g2_delta[4 * B200_FP_WORDS], g2_gamma[4 * B200_FP_WORDS];
    size_t n_a, n_b, n_k, n_z;          /* len(pk.G1.A), len(pk.G1.B) = len(pk.G2.B), len(pk.G1.K), len(pk.G1.Z) */
    size_t n_vk_k;                      /* len(vk.G1.K) = nb_public + nb_commitments                        */
    /* caller-allocated, each may be NULL: */
    uint8_t *infinity_a, *infinity_b;   /* pk.InfinityA / InfinityB, nb_wires bytes each                    */
    uint64_t *vk_k;                     /* vk.G1.K: (nb_public + nb_commitments) G1Affine                    */
    uint64_t *const *ck_k;              /* nb_commitments arrays: the Pedersen bases of commitment j (as many G1Affine as
                                           wires of class 2 + j), input of pedersen.Setup (setup.go:277-290)  */
    uint64_t *g1_a, *g1_b, *g1_k, *g1_z, *g2_b;   /* the proving key's point arrays, capacity nb_wires (a, b, k, g2_b) /
                                           2^log_n - 1 (z) points; with shards each call fills its own range   */
} b200_groth16_setup_out;

int b200_groth16_setup(b200_ctx *ctx, int curve, const b200_groth16_setup_desc *desc, b200_groth16_pk **pk_out,
                       b200_groth16_setup_out *out);

/* Per-stage device times in milliseconds (CUDA events).  The input copies run on a second stream:
 * W first, then a, b, c, overlapped with the four witness MSMs (order on the device: A, B1, K, B2,
 * computeH, Z).  h2d_ms spans the copy stream and is NOT a summand of total_ms. */
typedef struct b200_timings {
    float h2d_ms, compute_h_ms, msm_a_ms, msm_b1_ms, msm_k_ms, msm_z_ms, msm_b2_ms, total_ms;
} b200_timings;

/* ---- PlonK building blocks on device vectors ---------
 * Fr elements as everywhere: 4 x u64 Montgomery limbs.  All calls return after completion. */
/* out[i] = 1 / in[i], and 0 where in[i] = 0 (fr.BatchInvert; plonk/bn254/prove.go:1272).  No aliasing. */
int b200_fr_batch_inverse_dev(b200_ctx *ctx, int curve, const void *in_dev, void *out_dev, size_t n);
/* out[i] = in[0] * ... * in[i]: the accumulation of the copy-constraint ratio Z
 * (iop.BuildRatioCopyConstraint, prove.go:610-621).  in may alias out. */
int b200_fr_prefix_product_dev(b200_ctx *ctx, int curve, const void *in_dev, void *out_dev, size_t n);
/* The same prefix product for callers that cannot allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * The call performs no cudaMalloc / cudaFree / pinned allocation: its chunk-summary scratch is a file-scope static
 * __device__ array inside the library (csrc/noalloc.cu), 4 329 760 bytes that every GPU which loads the library
 * holds until the process exits.  Limit: n <= 2^22 elements per call; larger n fails with B200_ERR_ARG and a message
 * (use b200_fr_prefix_product_dev).  The call only ENQUEUES on the context's stream (b200_sync, or any synchronous
 * entry point, orders after it); the first call on a device also looks up the array's device address
 * (cudaGetSymbolAddress), so make that one outside a capture.  All calls on one device share the array: nothing in it
 * has to start at zero (every word read is written earlier in the same call), but two calls that overlap on the GPU
 * (two streams or two contexts of one device) race -- serialise them.  in may alias out. */
int b200_fr_prefix_product_static(b200_ctx *ctx, int curve, const void *in_dev, void *out_dev, size_t n);
/* value = sum coeffs[i] x^i (Polynomial.Evaluate, prove.go:723,1127,1300-1305). */
int b200_fr_poly_eval_dev(b200_ctx *ctx, int curve, const void *coeffs_dev, size_t n, const uint64_t *x, uint64_t *value);
/* quotient_dev[0..n-2]: p(X) = q(X) (X - z) + p(z) -- the synthetic division of kzg.Open /
 * kzg.BatchOpenSinglePoint (prove.go:646,823).  No aliasing. */
int b200_fr_poly_div_linear_dev(b200_ctx *ctx, int curve, const void *coeffs_dev, size_t n, const uint64_t *z,
                                void *quotient_dev);
/* divideByXMinusOne (prove.go:1223-1250): vec (2^log_big elements, basis LagrangeCoset, layout
 * BitReverse) is multiplied by 1 / (X^(2^log_small) - 1) on the big coset -- rho = 2^(log_big -
 * log_small) distinct values, evaluateXnMinusOneDomainBigCoset prove.go:1253-1276 -- and taken to
 * Canonical / Regular with the inverse coset transform of the big domain, in place. */
int b200_plonk_divide_by_xn_minus_one_dev(b200_ctx *ctx, int curve, void *vec_dev, unsigned log_big, unsigned log_small,
                                          const uint64_t *generator_big, const uint64_t *coset_gen);
/* kzg.Commit (prove.go:494,1203-1213): MSM of n coefficients against a resident SRS, affine result
 * (8 u64; (0,0) = infinity). */
int b200_kzg_commit_dev(b200_ctx *ctx, const b200_bases *srs, const void *coeffs_dev, size_t n, uint64_t *out_affine);

/* Z of the copy constraints (iop.BuildRatioCopyConstraint, prove.go:610-621): l, r, o in Lagrange /
 * Regular form (2^log_n elements each), perm_dev = the 3 * 2^log_n int64 permutation of the trace
 * (pk.trace.S), generator = Domain[0].Generator, coset_gen = Domain[0].FrMultiplicativeGen.
 * z_out_dev[0] = 1, z_out_dev[i+1] = z_out_dev[i] * num_i / den_i. */
int b200_plonk_build_ratio_dev(b200_ctx *ctx, int curve, const void *l_dev, const void *r_dev, const void *o_dev,
                               const void *perm_dev, const uint64_t *beta, const uint64_t *gamma, unsigned log_n,
                               const uint64_t *generator, const uint64_t *coset_gen, void *z_out_dev);
/* One coset of the quotient numerator (computeNumerator's iop.Evaluate(allConstraints, ...),
 * prove.go:837-1044).  x[k]: the polynomial with prove.go's id k (id_L = 0 ... id_LOne = 14; id_ZS = 4 is
 * not read: ZS is Z shifted by one row) in Lagrange / Regular form on the current coset; qci: the
 * BSB22 pairs [Qcp_0, Pi_0, ...]; bl/br/bo/bz: the blinding polynomials' coefficients as scaled for
 * this coset (prove.go:985-993); cs = Domain[1].FrMultiplicativeGen.  Row i goes to
 * cres_dev[bitrev_{log_big}(rho * i + coset_index)], rho = 2^(log_big - log_n). */
typedef struct b200_plonk_rows_desc {
    const void *x[15];
    const void *qci[16];
    unsigned nb_bsb;
    unsigned log_n, log_big, coset_index;
    const uint64_t *generator_small; /* Domain[0].Generator */
    uint64_t beta[4], gamma[4], alpha[4], cs[4];
    uint64_t bl[2][4], br[2][4], bo[2][4], bz[3][4];
} b200_plonk_rows_desc;
int b200_plonk_quotient_rows_dev(b200_ctx *ctx, int curve, const b200_plonk_rows_desc *desc, void *cres_dev);

/* ---- multi-GPU computeH (NTT with one exchange step per transform) -------------
 * One process per GPU holds n >> log_ranks elements of each vector in two device buffers (`data`,
 * `tmp`).  A transform is [local phase] -> all-to-all between the ranks (the caller's: NCCL, or
 * peer copies) -> [cross phase]; the phases below are the device work on either side.  Chunk q of
 * the all-to-all is the q-th contiguous (n >> 2 log_ranks)-element slice of the send buffer and
 * lands as slice `source rank` of the receive buffer.
 *   PACK        data (this rank's contiguous block of a natural-order vector) -> tmp; send tmp, receive into data
 *   DIF_LOCAL   data in place;                       send data, receive into tmp
 *   DIF_CROSS   tmp -> data (block of the bit-reversed result; `scale` applied)
 *   DIT_LOCAL   data -> tmp;                         send tmp, receive into data
 *   DIT_CROSS   data in place (cyclic slice of the natural-order result: element j of rank r is index j * ranks + r)
 * computeH = PACK x3, inverse DIF x3 (scales H_A, H_BC, H_A), forward DIT x3, b200_fr_mul_sub_dev,
 * inverse DIF (COSETINV_BR): rank r ends with h[r n / ranks, (r+1) n / ranks) in the bit-reversed
 * order G1.Z is stored in (prove.go:232-234,353-396), ready for b200_groth16_msms_dev.
 * The phase calls and b200_fr_mul_sub_dev only ENQUEUE on the context's stream: either b200_sync
 * before handing a buffer to the exchange, or -- the intended use -- b200_use_stream(ctx, s) with the
 * stream the exchange is ordered on (torch.cuda.current_stream().cuda_stream for torch.distributed),
 * so that phases and all-to-alls are stream-ordered with no host round trip. */
enum b200_dist_phase { B200_DIST_PACK = 0, B200_DIST_DIF_LOCAL = 1, B200_DIST_DIF_CROSS = 2, B200_DIST_DIT_LOCAL = 3, B200_DIST_DIT_CROSS = 4 };
enum b200_dist_scale { B200_SCALE_NONE = 0, B200_SCALE_H_A = 1, B200_SCALE_H_BC = 2, B200_SCALE_COSETINV_BR = 3, B200_SCALE_NINV = 4 };
int b200_ntt_dist_phase(b200_ctx *ctx, int curve, unsigned log_n, const uint64_t *generator, const uint64_t *coset_gen,
                        void *data_dev, void *tmp_dev, unsigned log_ranks, unsigned rank, int phase, int inverse, int scale);
/* a[i] = a[i] * b[i] - c[i] on device vectors (prove.go:377-390 without den, which H_A carries). */
int b200_fr_mul_sub_dev(b200_ctx *ctx, int curve, void *a_dev, const void *b_dev, const void *c_dev, size_t n);

/* The five MultiExp results BEFORE the alpha/beta/delta/r/s fix-ups, which stay in Go
 * (prove.go:206-207,219-221,248-271,294-299). */
typedef struct b200_groth16_out {
    uint64_t ar[3 * B200_FP_WORDS];    /* MSM(pk.G1.A, wireValuesA)       prove.go:214 */
    uint64_t bs1[3 * B200_FP_WORDS];   /* MSM(pk.G1.B, wireValuesB)       prove.go:201 */
    uint64_t krs[3 * B200_FP_WORDS];   /* MSM(pk.G1.K, W[nb_public:])     prove.go:244 */
    uint64_t krs2[3 * B200_FP_WORDS];  /* MSM(pk.G1.Z, h[:n-1])           prove.go:234 */
    uint64_t bs2[6 * B200_FP_WORDS];   /* MSM(pk.G2.B, wireValuesB)       prove.go:290 */
} b200_groth16_out;

/* computeH + wire filtering + the five MSMs on the device (prove.go:139-175,199-301;
 * icicle.go:196-393).  a, b, c: solver outputs, n_constraints elements each (host); w: full wire
 * vector, n_wires elements (host).  k_remove (optional, sorted ascending, indices relative to
 * W[nb_public:]) lists committed wires to drop from the K scalars (prove.go:238-242).
 * For a sharded pk the outputs are this shard's partial sums. */
int b200_groth16_prove_core(b200_ctx *ctx, const b200_groth16_pk *pk, const uint64_t *a, const uint64_t *b,
                            const uint64_t *c, size_t n_constraints, const uint64_t *w,
                            const uint32_t *k_remove, size_t n_remove, b200_groth16_out *out,
                            b200_timings *timings);

/* Same with a, b, c (n_constraints elements each) and w (n_wires) already resident on the device
 * (they are copied device-to-device into the workspace; the caller's buffers are not modified). */
int b200_groth16_prove_core_dev(b200_ctx *ctx, const b200_groth16_pk *pk, const void *a_dev, const void *b_dev,
                                const void *c_dev, size_t n_constraints, const void *w_dev,
                                b200_groth16_out *out, b200_timings *timings);
/* The five MSMs of this shard with h supplied by the distributed computeH: h_dev holds this
 * shard's range of h (len = the shard's G1.Z count). */
int b200_groth16_msms_dev(b200_ctx *ctx, const b200_groth16_pk *pk, const void *w_dev, const void *h_dev,
                          b200_groth16_out *out, b200_timings *timings);

/* ---- multi-GPU in ONE process: the device group ------------------------------------------------------------
 * What groth16.Prove calls to prove on every GPU of the box with one call (the reference's drop-in shape is one Go
 * process making one call per proof: backend/groth16/groth16.go:178-182 -> icicle/icicle.go:133; option plumbing
 * backend/backend.go:132-137 -- the b200 option carries the device list).  A group owns one b200_ctx per device and
 * runs them from its own host threads; no MPI / NCCL / torchrun around the caller.
 *   - the five MultiExp run on point-range shards (one per device); partial sums are added on the host inside the call;
 *   - computeH is distributed with one exchange per transform, fused into the kernels over peer-mapped memory
 *     (NVLink P2P stores / loads; csrc/group.cu) when the group has a power-of-two number >= 2 of devices with peer
 *     access and 2 log2(devices) <= log_n; otherwise every device runs the whole computeH (replicated);
 *   - `devices` may name one device several times (virtual ranks; what the tests do on a single GPU).
 * A group is not re-entrant (one call at a time), like a ctx.  Errors: negative b200_status, message in
 * b200_group_last_error. */
typedef struct b200_group b200_group;
typedef struct b200_group_groth16_pk b200_group_groth16_pk;
int b200_group_init(const int *devices, int ndev, b200_group **out);
void b200_group_destroy(b200_group *group);
const char *b200_group_last_error(const b200_group *group);
int b200_group_size(const b200_group *group);
/* The context of one rank (options, profiling, single-device calls); owned by the group. */
b200_ctx *b200_group_ctx(b200_group *group, int rank);
int b200_group_set_option(b200_group *group, const char *key, long value);
/* b200_groth16_pk_upload / b200_groth16_setup for the whole group: the descriptor's shard_index / shard_count /
 * presliced are ignored -- rank r keeps the point range r of every base array on its device. */
int b200_group_groth16_pk_upload(b200_group *group, int curve, const b200_groth16_pk_desc *desc, b200_group_groth16_pk **out);
int b200_group_groth16_setup(b200_group *group, int curve, const b200_groth16_setup_desc *desc, b200_group_groth16_pk **out,
                             b200_groth16_setup_out *setup_out);
int b200_group_groth16_pk_free(b200_group *group, b200_group_groth16_pk *pk);
/* out[0..4] of one rank's shard; *distributed_h = 1 when computeH runs distributed over the group. */
int b200_group_groth16_pk_info(b200_group *group, const b200_group_groth16_pk *pk, int rank, b200_msm_set_info out[5],
                               int *distributed_h);
/* b200_groth16_prove_core on the group: host a, b, c (n_constraints each), host W; `out` = the five MultiExp results
 * of the WHOLE proof (partial sums already combined); timings = per-stage maximum over the ranks. */
int b200_group_groth16_prove_core(b200_group *group, const b200_group_groth16_pk *pk, const uint64_t *a, const uint64_t *b,
                                  const uint64_t *c, size_t n_constraints, const uint64_t *w, const uint32_t *k_remove,
                                  size_t n_remove, b200_groth16_out *out, b200_timings *timings);
/* The same with inputs resident on the devices (distributed computeH only): a_dev[r], b_dev[r], c_dev[r] = rank r's
 * block [r n / R, (r+1) n / R) of the zero-padded vectors on device r, w_dev[r] = the full W on device r. */
int b200_group_groth16_prove_core_dev(b200_group *group, const b200_group_groth16_pk *pk, const void *const *a_dev,
                                      const void *const *b_dev, const void *const *c_dev, const void *const *w_dev,
                                      b200_groth16_out *out, b200_timings *timings);

/* Host mirror of groth16_bn254.Prove's tail (prove.go:178-321) for callers without Go: applies
 * the fix-ups to the five MSM results with the given blinding scalars r, s (Montgomery Fr) and
 * the key's alpha/beta/delta points, and returns the affine proof (Ar: 8 u64, Bs: 16 u64,
 * Krs: 8 u64).  vk-side points are host pointers: g1_alpha, g1_beta, g1_delta (8 u64 each),
 * g2_beta, g2_delta (16 u64 each). */
int b200_groth16_finalize(const b200_groth16_out *msm, const uint64_t *r, const uint64_t *s,
                          const uint64_t *g1_alpha, const uint64_t *g1_beta, const uint64_t *g1_delta,
                          const uint64_t *g2_beta, const uint64_t *g2_delta, uint64_t *ar_affine,
                          uint64_t *bs_affine, uint64_t *krs_affine);

/* Proof.WriteTo without commitments (marshal.go:41-66): compressed Ar | Bs | Krs | u32 0 | Pok=inf.
 * out must hold 5 * 8 * B200_FP_WORDS + 4 bytes (164 on BN254, 244 on BLS12-377); returns the number of bytes written
 * or a negative status. */
int b200_groth16_proof_bytes(const uint64_t *ar_affine, const uint64_t *bs_affine, const uint64_t *krs_affine,
                             uint8_t *out, size_t out_cap);

/* ---- PlonK prover (BN254; the BLS12-377 library runs the same code on its own field sizes:
 * /root/reference/backend/plonk/bls12-377/prove.go is the same template) ------------------------
 * plonk_bn254.Prove (/root/reference/backend/plonk/bn254/prove.go:116-180) behind one call, with or
 * without BSB22 commitments (api.Commit circuits, at most B200_PLONK_MAX_BSB22 of them).  The solver stays
 * with the caller (prove.go:365-392 gives L, R, O); everything after it
 * runs here: nine KZG commitments as G1 MSMs against two resident SRS (the tenth, the linearised polynomial's, is formed from
 * existing commitments as the verifier does, verify.go:210-250), the copy-constraint ratio,
 * the quotient over the rho cosets of the big domain, the openings, and the Fiat-Shamir transcript
 * (SHA-256, host) that orders them. */
typedef struct b200_plonk_pk b200_plonk_pk;
#define B200_PLONK_MAX_BSB22 8

/* What the glue reads out of plonk_bn254.ProvingKey (setup.go:79-107).  All pointers are host memory. */
typedef struct b200_plonk_pk_desc {
    unsigned log_n;                 /* pk.Domain[0].Cardinality = 2^log_n                                  */
    unsigned log_big;               /* pk.Domain[1].Cardinality = 2^log_big (4n, or 8n when n+pub < 6)       */
    const uint64_t *generator;      /* pk.Domain[0].Generator                                               */
    const uint64_t *generator_big;  /* pk.Domain[1].Generator                                               */
    const uint64_t *coset_gen;      /* FrMultiplicativeGen (= pk.Vk.CosetShift)                             */
    const uint64_t *ql, *qr, *qm, *qo, *qk;   /* pk.trace.Q*: canonical / regular, n each; qk incomplete     */
    const uint64_t *s1, *s2, *s3;             /* pk.trace.S1..S3: canonical / regular                         */
    const int64_t *perm;            /* pk.trace.S, 3n entries                                                */
    const uint64_t *kzg_g1;         /* pk.Kzg.G1: n + 3 affine points                                       */
    const uint64_t *kzg_lagrange_g1;/* pk.KzgLagrange.G1: n affine points                                   */
    const uint64_t *vk_commitments; /* pk.Vk.S[0], S[1], S[2], Ql, Qr, Qm, Qo, Qk: 8 affine points          */
    size_t nb_public;               /* pk.Vk.NbPublicVariables                                              */
    size_t nb_bsb22;                /* len(pk.Vk.Qcp) = len(spr.CommitmentInfo), <= B200_PLONK_MAX_BSB22      */
    const uint64_t *const *qcp;     /* pk.trace.Qcp[i]: canonical / regular, n each (NULL when nb_bsb22 = 0)  */
    const uint64_t *vk_qcp;         /* pk.Vk.Qcp: nb_bsb22 affine points                                     */
    const size_t *commitment_index; /* CommitmentInfo[i].CommitmentIndex (constraint index; the row is + nb_public) */
    int raw_file;                   /* non-zero: the arrays are sections of a ProvingKey.WriteRawTo file
                                       (plonk/bn254/marshal.go:101-180), converted on the device: ql..s3 and qcp[i] = n
                                       32-byte big-endian canonical scalars, perm = 3n big-endian int64, kzg_g1 /
                                       kzg_lagrange_g1 / vk_commitments / vk_qcp = raw-encoded points (see
                                       b200_points_from_raw); generators and coset_gen stay Montgomery limbs          */
    int shard_index, shard_count;   /* multi-GPU: this rank's share (0, 0 or 0, 1 = the whole key).  The arrays above are
                                       always the full ones; the rank keeps the SRS point range [N k / R, N (k+1) / R) on its
                                       device and evaluates the quotient on the cosets k = shard_index (mod shard_count).      */
} b200_plonk_pk_desc;

/* One-time residency: trace polynomials, their evaluations on the rho cosets of the big domain
 * (the precomputation prove.go:995-999 leaves out for memory: 8 polynomials x rho x n elements),
 * the permutation, and both SRS with folded-window MSM tables. */
int b200_plonk_pk_upload(b200_ctx *ctx, int curve, const b200_plonk_pk_desc *desc, b200_plonk_pk **out);
int b200_plonk_pk_free(b200_ctx *ctx, b200_plonk_pk *pk);

/* plonk_bn254.Proof (prove.go:96-114) with points affine (2 * B200_FP_WORDS u64: 8 on BN254, 12 on BLS12-377) and
 * scalars Montgomery (4 u64). */
typedef struct b200_plonk_proof {
    uint64_t lro[3][2 * B200_FP_WORDS];
    uint64_t z[2 * B200_FP_WORDS];
    uint64_t h[3][2 * B200_FP_WORDS];
    uint64_t batched_h[2 * B200_FP_WORDS];      /* BatchedProof.H                                           */
    uint64_t claimed[7][4];         /* BatchedProof.ClaimedValues: folded H, linearised, l, r, o, s1, s2     */
    uint64_t zshift_h[2 * B200_FP_WORDS];       /* ZShiftedOpening.H                                        */
    uint64_t zshift_value[4];       /* ZShiftedOpening.ClaimedValue                                         */
    uint64_t challenges[5][4];      /* gamma, beta, alpha, zeta, KZG folding gamma (for tests; not serialised) */
    uint64_t nb_bsb22;              /* len(Bsb22Commitments)                                                */
    uint64_t claimed_qcp[B200_PLONK_MAX_BSB22][4];   /* BatchedProof.ClaimedValues[7:]: Qcp_i(zeta)          */
    uint64_t bsb22[B200_PLONK_MAX_BSB22][2 * B200_FP_WORDS];         /* Bsb22Commitments                     */
} b200_plonk_proof;

/* What the solver's bsb22Hint (prove.go:303-352) produced for an api.Commit circuit: for each commitment the
 * committed-values column in Lagrange / regular form (s.cCommitments[i] before ToCanonical; n elements), its
 * KZG commitment against pk.KzgLagrange (the hint can compute it with b200_msm_resident / b200_kzg_commit_dev)
 * and the hashed value injected in Qk (s.commitmentVal[i]). */
typedef struct b200_plonk_bsb22 {
    const uint64_t *const *pi2;     /* nb_bsb22 host vectors                                                */
    const uint64_t *commitments;    /* nb_bsb22 affine points                                               */
    const uint64_t *values;         /* nb_bsb22 Fr                                                          */
} b200_plonk_bsb22;

typedef struct b200_plonk_timings {
    float h2d_ms, commit_lro_ms, ratio_ms, commit_z_ms, quotient_ms, commit_h_ms, open_z_ms, linearize_ms,
          batch_open_ms, total_ms;
} b200_plonk_timings;

/* l, r, o: the solution in Lagrange / regular form, n elements each (SparseR1CSSolution,
 * constraint/bn254/system.go:221-265); public_witness: nb_public elements; blinding: the 9 coefficients
 * of getRandomPolynomial (prove.go:295-301) in the order bl0 bl1 br0 br1 bo0 bo1 bz0 bz1 bz2, or NULL
 * to draw them from the OS entropy source as the reference does. */
int b200_plonk_prove(b200_ctx *ctx, const b200_plonk_pk *pk, const uint64_t *l, const uint64_t *r, const uint64_t *o,
                     const uint64_t *public_witness, const uint64_t *blinding, b200_plonk_proof *out,
                     b200_plonk_timings *timings);
/* Same for a key with BSB22 commitments (bsb may be NULL when the key has none). */
int b200_plonk_prove_bsb22(b200_ctx *ctx, const b200_plonk_pk *pk, const uint64_t *l, const uint64_t *r, const uint64_t *o,
                           const uint64_t *public_witness, const uint64_t *blinding, const b200_plonk_bsb22 *bsb,
                           b200_plonk_proof *out, b200_plonk_timings *timings);

/* ---- the same proof as a session of phases: multi-GPU (PlonK P1 range-sharded MSMs, P4 coset-
 * parallel quotient).  One process per GPU, each with its shard of the key; every rank runs the same phases on the
 * same inputs and blinding; a phase returns this rank's PARTIAL commitments (G1Jac, 3 * B200_FP_WORDS u64 each) and the next phase
 * takes the TOTALS (the caller adds the ranks' partials: b200_g1_jac_add after an all-gather of the 96-byte (144 on BLS12-377) points --
 * no data-path collective for the MSMs).  After phase 2 the caller all-gathers the numerator's coset blocks in
 * place (b200_plonk_session_numerator: coset k lives in block bitrev(k) of 2^log_blocks blocks of block_elems Fr).
 *   begin   -> 3 partials [L] [R] [O]
 *   step 1  (3 totals) -> 1 partial [Z]
 *   step 2  (1 total)  -> nothing; this rank's coset blocks of the numerator are written
 *   step 3  (nothing)  -> 3 partials [H1] [H2] [H3]     (needs every block of the numerator)
 *   step 4  (3 totals) -> 2 partials [Z opening quotient] [linearised polynomial]
 *   step 5  (2 totals) -> 1 partial [batch opening quotient]
 *   finish  (1 total)  -> the proof (identical on every rank); frees the session.
 * b200_plonk_prove_bsb22 is exactly this sequence with totals = partials.  A sharded proof requires `blinding`. */
/* Re-entrancy: the per-proof workspace lives in the proving key and in the ctx, so a key (and a ctx) proves ONE proof
 * at a time.  b200_plonk_session_begin (and therefore b200_plonk_prove*) returns B200_ERR_ARG while another session on
 * the same key is open; a Go glue serialises Prove calls per key with a mutex or uploads the key once
 * per concurrent prover. */
typedef struct b200_plonk_session b200_plonk_session;
int b200_plonk_session_begin(b200_ctx *ctx, const b200_plonk_pk *pk, const uint64_t *l, const uint64_t *r, const uint64_t *o,
                             const uint64_t *public_witness, const uint64_t *blinding, const b200_plonk_bsb22 *bsb,
                             b200_plonk_session **out, uint64_t *partials);
int b200_plonk_session_step(b200_ctx *ctx, b200_plonk_session *session, int phase, const uint64_t *totals, uint64_t *partials);
int b200_plonk_session_numerator(b200_ctx *ctx, b200_plonk_session *session, void **cres_dev, size_t *block_elems,
                                 unsigned *log_blocks);
/* totals = proof = NULL abandons a session at any point. */
int b200_plonk_session_finish(b200_ctx *ctx, b200_plonk_session *session, const uint64_t *totals, b200_plonk_proof *out,
                              b200_plonk_timings *timings);

/* The same proof over a device group (one process, one call): the key is sharded over the ranks as above, every rank
 * runs the session phases from its own host thread, partial commitments are added on the host inside the call and the
 * numerator's coset blocks move between the devices by peer copies.  blinding = NULL draws it once for all ranks. */
typedef struct b200_group_plonk_pk b200_group_plonk_pk;
int b200_group_plonk_pk_upload(b200_group *group, int curve, const b200_plonk_pk_desc *desc, b200_group_plonk_pk **out);
int b200_group_plonk_pk_free(b200_group *group, b200_group_plonk_pk *pk);
int b200_group_plonk_prove(b200_group *group, const b200_group_plonk_pk *pk, const uint64_t *l, const uint64_t *r, const uint64_t *o,
                           const uint64_t *public_witness, const uint64_t *blinding, const b200_plonk_bsb22 *bsb,
                           b200_plonk_proof *out, b200_plonk_timings *timings);

/* The four Fiat-Shamir challenges gamma, beta, alpha, zeta (16 u64, Montgomery) of a proof's commitments:
 * bindPublicData + deriveRandomness (plonk/bn254/verify.go:296-360) over SHA-256, gnark-crypto's
 * fiat-shamir.Transcript layout (restated in-tree by plonk/bn254/solidity.go:407-538).  Host only. */
int b200_plonk_derive_challenges(const uint64_t *vk_commitments, const uint64_t *vk_qcp, const uint64_t *public_witness,
                                 size_t nb_public, const b200_plonk_proof *proof, uint64_t *out);

/* Proof.WriteTo (plonk/bn254/marshal.go:41-66): 8 compressed G1 | u32 7+k | (7+k) Fr | G1 | Fr | u32 k | k G1
 * = 552 + 64 k bytes for k BSB22 commitments on BN254 (9 x 48 + 8 + 32 (8 + k) + 48 k on BLS12-377).  Returns the number of bytes written or a negative status. */
int b200_plonk_proof_bytes(const b200_plonk_proof *proof, uint8_t *out, size_t out_cap);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* B200PROVER_H */
