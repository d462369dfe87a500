/* Pure-C caller of include/b200prover.h (compiled with gcc -std=c11): what a cgo shim does, without Go.
 *   b200_init -> b200_groth16_setup (refCircuit Mul chain) -> b200_groth16_prove_core -> b200_groth16_finalize
 *   -> b200_groth16_proof_bytes, proof printed as hex on stdout.
 * Input file (argv[1], little-endian): u32 log_n | generator | coset_gen | tau alpha beta gamma delta | r | s   (Fr, 4 u64
 * each) | G1 generator (2 Fp) | G2 generator (4 Fp) | W (n + 2 Fr) | a | b | c (n Fr each).
 * Built twice by tests/test_abi_c.py: for BN254, and with -DB200_CURVE_BLS12_377 -DSMOKE_MAIN=smoke_bls12377 for the
 * second library; both objects are linked into ONE binary with both libraries (the per-curve symbol names at work). */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "b200prover.h"

#ifndef SMOKE_MAIN
#define SMOKE_MAIN smoke_bn254
#endif
#if defined(B200_CURVE_BLS12_377)
#define SMOKE_CURVE B200_BLS12_377
#else
#define SMOKE_CURVE B200_BN254
#endif

static int rd(FILE *f, void *dst, size_t bytes) { return fread(dst, 1, bytes, f) == bytes ? 0 : -1; }

int SMOKE_MAIN(const char *path) {
    FILE *f = fopen(path, "rb");
    if (!f) { fprintf(stderr, "cannot open %s\n", path); return 2; }
    uint32_t log_n;
    uint64_t gen[4], coset[4], r[4], s[4], g1[2 * B200_FP_WORDS], g2[4 * B200_FP_WORDS];
    b200_groth16_setup_desc d;
    memset(&d, 0, sizeof d);
    int bad = rd(f, &log_n, 4) | rd(f, gen, 32) | rd(f, coset, 32) | rd(f, d.tau, 32) | rd(f, d.alpha, 32) | rd(f, d.beta, 32) |
              rd(f, d.gamma, 32) | rd(f, d.delta, 32) | rd(f, r, 32) | rd(f, s, 32) | rd(f, g1, sizeof g1) | rd(f, g2, sizeof g2);
    const size_t n = (size_t)1 << log_n, nw = n + 2;
    uint64_t *w = malloc(nw * 32), *a = malloc(n * 32), *b = malloc(n * 32), *c = malloc(n * 32);
    bad |= rd(f, w, nw * 32) | rd(f, a, n * 32) | rd(f, b, n * 32) | rd(f, c, n * 32);
    fclose(f);
    if (bad) { fprintf(stderr, "short input file\n"); return 2; }

    /* refCircuit as CSR terms {coefficient id, wire id}: constraint j < n-1: X_j * X_j = X_{j+1}; last: ONE * X_N = Y */
    uint64_t *row_ptr = malloc((n + 1) * 8);
    uint32_t *tl = malloc(n * 8), *tr = malloc(n * 8), *to = malloc(n * 8);
    for (size_t j = 0; j <= n; ++j) row_ptr[j] = j;
    for (size_t j = 0; j + 1 < n; ++j) {
        tl[2 * j] = tr[2 * j] = to[2 * j] = 0;
        tl[2 * j + 1] = tr[2 * j + 1] = (uint32_t)(2 + j);
        to[2 * j + 1] = (uint32_t)(3 + j);
    }
    tl[2 * (n - 1)] = tr[2 * (n - 1)] = to[2 * (n - 1)] = 0;
    tl[2 * (n - 1) + 1] = 0; tr[2 * (n - 1) + 1] = (uint32_t)(n + 1); to[2 * (n - 1) + 1] = 1;

    b200_ctx *ctx = NULL;
    int st = b200_init(0, &ctx);
    if (st != B200_OK) { fprintf(stderr, "b200_init: %d %s\n", st, b200_last_error(NULL)); return 3; }
    if (b200_build_curve() != SMOKE_CURVE || b200_build_fp_words() != B200_FP_WORDS) { fprintf(stderr, "wrong library bound\n"); return 3; }
    d.log_n = log_n; d.generator = gen; d.coset_gen = coset;
    d.nb_constraints = n; d.nb_wires = nw; d.nb_public = 2;
    d.l.row_ptr = d.r.row_ptr = d.o.row_ptr = row_ptr;
    d.l.terms = tl; d.r.terms = tr; d.o.terms = to;
    d.coeffs = w; d.nb_coeffs = 1;                   /* W[0] is the ONE wire: the field's 1 in Montgomery form */
    d.g1_gen = g1; d.g2_gen = g2;
    d.shard_index = 0; d.shard_count = 1;
    b200_groth16_setup_out so;
    memset(&so, 0, sizeof so);
    b200_groth16_pk *pk = NULL;
    st = b200_groth16_setup(ctx, SMOKE_CURVE, &d, &pk, &so);
    if (st != B200_OK) { fprintf(stderr, "setup: %d %s\n", st, b200_last_error(ctx)); return 4; }
    b200_groth16_out out;
    b200_timings tm;
    st = b200_groth16_prove_core(ctx, pk, a, b, c, n, w, NULL, 0, &out, &tm);
    if (st != B200_OK) { fprintf(stderr, "prove_core: %d %s\n", st, b200_last_error(ctx)); return 5; }
    uint64_t ar[2 * B200_FP_WORDS], bs[4 * B200_FP_WORDS], krs[2 * B200_FP_WORDS];
    st = b200_groth16_finalize(&out, r, s, so.g1_alpha, so.g1_beta, so.g1_delta, so.g2_beta, so.g2_delta, ar, bs, krs);
    if (st != B200_OK) { fprintf(stderr, "finalize: %d\n", st); return 6; }
    uint8_t bytes[5 * 8 * B200_FP_WORDS + 4];
    int nb = b200_groth16_proof_bytes(ar, bs, krs, bytes, sizeof bytes);
    if (nb != (int)sizeof bytes) { fprintf(stderr, "proof_bytes: %d\n", nb); return 7; }
    for (int i = 0; i < nb; ++i) printf("%02x", bytes[i]);
    printf("\n");
    /* error path through the C ABI: an unsupported curve id is a status and a message, nothing else */
    if (b200_fr_vec_mul(ctx, 7, NULL, NULL, NULL, 0) != B200_ERR_UNSUPPORTED || !b200_last_error(ctx)[0]) return 8;
    b200_groth16_pk_free(ctx, pk);
    b200_destroy(ctx);
    free(w); free(a); free(b); free(c); free(row_ptr); free(tl); free(tr); free(to);
    return 0;
}
