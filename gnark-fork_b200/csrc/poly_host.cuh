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
}  // namespace b200
