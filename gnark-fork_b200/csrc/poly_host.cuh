// Host entry points of poly.cu (Fr vector / polynomial primitives of the PlonK prover).
#pragma once
#include "ntt_host.cuh"

namespace b200 {
// out[i] = 1 / in[i], 0 for in[i] = 0 (fr.BatchInvert).  in and out must not alias.
int fr_batch_inverse(b200_ctx* ctx, const Fr* in, Fr* out, size_t n);
// out[i] = in[0] * ... * in[i]   (in may alias out)
int fr_prefix_product(b200_ctx* ctx, const Fr* in, Fr* out, size_t n);
// sum coeffs[i] x^i
int fr_poly_eval(b200_ctx* ctx, const Fr* coeffs, size_t n, const Fr& x, Fr* out_host);
// q (n - 1 coefficients) with p(X) = q(X) (X - z) + p(z)
int fr_poly_div_linear(b200_ctx* ctx, const Fr* a, size_t n, const Fr& z, Fr* q);
// v[i] *= factors[(i >> shift) mod count], count in {1, 2, 4, 8}
int fr_scale_cyclic(b200_ctx* ctx, Fr* v, size_t n, const Fr* factors, uint32_t count, uint32_t shift);
// Z of the copy constraints (Lagrange / Regular): iop.BuildRatioCopyConstraint over l, r, o with
// the permutation `perm` (3 n int64 on the device), domain d (generator w, shift g = d->coset)
int plonk_build_ratio(b200_ctx* ctx, NttDomain* d, const Fr* l, const Fr* r, const Fr* o, const int64_t* perm, const Fr& beta,
                      const Fr& gamma, Fr* z_out, Fr* ws_den = nullptr, Fr* ws_inv = nullptr);   // ws_*: optional n-element scratch
struct PlonkRowArgs {
    const Fr* x[15];          // id_L .. id_LOne (prove.go:60-77); x[id_ZS] is not read
    const Fr* qci[16];        // [Qcp_0, Pi_0, Qcp_1, Pi_1, ...]
    uint32_t nb_bsb;
    Fr beta, gamma, alpha, cs, css;
    Fr bl[2], br[2], bo[2], bz[3];   // blinding polynomials, coefficients as scaled for this coset (prove.go:985-993)
    const Fr* tw;             // w^j, j < n/2, of the small domain
    uint32_t n, log_big, rho, coset_index;
    Fr* cres;
};
enum { PID_L = 0, PID_R, PID_O, PID_Z, PID_ZS, PID_QL, PID_QR, PID_QM, PID_QO, PID_QK, PID_S1, PID_S2, PID_S3, PID_ID, PID_LONE };

int plonk_quotient_rows(b200_ctx* ctx, PlonkRowArgs& args);

// divideByXMinusOne on the big domain d (plonk.cu)
int plonk_divide_xn(b200_ctx* ctx, NttDomain* d, unsigned log_small, Fr* vec);
// PlonK prover glue (poly.cu)
int fr_bitrev_copy(b200_ctx* ctx, const Fr* in, Fr* out, unsigned log_n);
int fr_scaled_omega(b200_ctx* ctx, NttDomain* d, const Fr& c, Fr* out);
int fr_blind(b200_ctx* ctx, Fr* p, size_t n, const Fr* b, uint32_t nb);
int plonk_fold_h(b200_ctx* ctx, const Fr* h1, const Fr* h2, const Fr* h3, const Fr& z, size_t m, Fr* out);
struct PlonkLinArgs {
    const Fr *bz, *s3, *ql, *qr, *qm, *qo, *qk;   // blinded Z (len coefficients), trace polynomials (n each), canonical
    const Fr* pi2[8];                              // BSB22 commitment polynomials, canonical
    Fr qcp_zeta[8];
    uint32_t nb_bsb;
    Fr c1, c2, alpha, rl, l, r, o, lag;            // prove.go:1293-1334
    uint32_t n, len;
    Fr* out;
};
int plonk_linearize(b200_ctx* ctx, const PlonkLinArgs& a);
struct PlonkFoldArgs {
    const Fr* p[16];
    uint32_t len[16];
    Fr g[16];            // g[k] = gamma^k
    uint32_t count, width;
    Fr* out;
};
int plonk_fold_polys(b200_ctx* ctx, const PlonkFoldArgs& a);
}  // namespace b200
