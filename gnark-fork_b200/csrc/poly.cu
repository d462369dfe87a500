// Fr vector / polynomial primitives of the PlonK prover:
//   batch inversion        fr.BatchInvert (prove.go:1272, iop ratio construction)     zero stays zero
//   prefix product         the grand-product accumulation of iop.BuildRatioCopyConstraint (prove.go:610-621)
//   evaluation at a point  Polynomial.Evaluate / evaluateBlinded (prove.go:723,1118-1145,1300-1305)
//   quotient by (X - z)    the synthetic division inside kzg.Open / BatchOpenSinglePoint (prove.go:646,823)
//   cyclic scaling         divideByXMinusOne's product with the rho values of 1/(X^n - 1) (prove.go:1223-1250)
// All are HBM-bound integer kernels: chunks per thread, a short recursion over chunk summaries.
#include "poly_host.cuh"

namespace b200 {

__device__ __forceinline__ Fr pl_ld(const Fr* p) {
    const uint4* q = reinterpret_cast<const uint4*>(p);
    const uint4 a = q[0], b = q[1];
    Fr f;
    f.l[0] = a.x; f.l[1] = a.y; f.l[2] = a.z; f.l[3] = a.w;
    f.l[4] = b.x; f.l[5] = b.y; f.l[6] = b.z; f.l[7] = b.w;
    return f;
}
__device__ __forceinline__ void pl_st(Fr* p, const Fr& f) {
    uint4* q = reinterpret_cast<uint4*>(p);
    q[0] = make_uint4(f.l[0], f.l[1], f.l[2], f.l[3]);
    q[1] = make_uint4(f.l[4], f.l[5], f.l[6], f.l[7]);
}

constexpr int PL_CHUNK = 32;      // elements per thread
constexpr int PL_THREADS = 128;

// ---- batch inversion: per-thread Montgomery trick over PL_CHUNK elements ----------------------
__global__ void __launch_bounds__(PL_THREADS) batch_inverse_kernel(const Fr* __restrict__ in, Fr* __restrict__ out, size_t n) {
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t lo = t * PL_CHUNK;
    if (lo >= n) return;
    const int cnt = (int)(n - lo < (size_t)PL_CHUNK ? n - lo : (size_t)PL_CHUNK);
    // forward: out[i] = product of the non-zero elements before i
    Fr acc = Fr::one();
    for (int i = 0; i < cnt; ++i) {
        const Fr x = pl_ld(in + lo + i);
        pl_st(out + lo + i, acc);
        if (!x.is_zero()) acc = acc * x;
    }
    Fr inv = acc.inv();
    for (int i = cnt - 1; i >= 0; --i) {
        const Fr x = pl_ld(in + lo + i);
        if (x.is_zero()) { pl_st(out + lo + i, Fr::zero()); continue; }
        const Fr pre = pl_ld(out + lo + i);
        pl_st(out + lo + i, inv * pre);
        inv = inv * x;
    }
}

// ---- prefix product: chunk products, recursion over them, then the chunks again -----------------
// pass 1: prod[t] = product of chunk t
__global__ void __launch_bounds__(PL_THREADS) chunk_product_kernel(const Fr* __restrict__ in, size_t n, Fr* __restrict__ prod) {
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t lo = t * PL_CHUNK;
    if (lo >= n) return;
    const int cnt = (int)(n - lo < (size_t)PL_CHUNK ? n - lo : (size_t)PL_CHUNK);
    Fr acc = pl_ld(in + lo);
    for (int i = 1; i < cnt; ++i) acc = acc * pl_ld(in + lo + i);
    pl_st(prod + t, acc);
}
// pass 3: out[i] = carry[t] * in[lo] * ... * in[i]   (carry[t] = exclusive prefix of the chunk products; null for t = 0 only)
__global__ void __launch_bounds__(PL_THREADS) chunk_scan_kernel(const Fr* __restrict__ in, size_t n, const Fr* __restrict__ incl_prev, Fr* __restrict__ out) {
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t lo = t * PL_CHUNK;
    if (lo >= n) return;
    const int cnt = (int)(n - lo < (size_t)PL_CHUNK ? n - lo : (size_t)PL_CHUNK);
    Fr acc = (t == 0 || incl_prev == nullptr) ? Fr::one() : pl_ld(incl_prev + t - 1);
    for (int i = 0; i < cnt; ++i) {
        acc = acc * pl_ld(in + lo + i);
        pl_st(out + lo + i, acc);
    }
}

// inclusive prefix product of n elements, in -> out (may alias); scratch: ceil(n / PL_CHUNK) * 2 elements at least
static int prefix_product_rec(b200_ctx* ctx, const Fr* in, Fr* out, size_t n, Fr* scratch) {
    const size_t chunks = (n + PL_CHUNK - 1) / PL_CHUNK;
    const unsigned blocks = (unsigned)((chunks + PL_THREADS - 1) / PL_THREADS);
    if (chunks == 1) {
        chunk_scan_kernel<<<1, PL_THREADS, 0, ctx->stream>>>(in, n, nullptr, out);
        B2_CUDA(ctx, cudaGetLastError());
        ctx->n_launches += 1;
        return B200_OK;
    }
    Fr* prod = scratch;                 // chunks elements, scanned in place
    chunk_product_kernel<<<blocks, PL_THREADS, 0, ctx->stream>>>(in, n, prod);
    B2_CUDA(ctx, cudaGetLastError());
    B2_TRY(prefix_product_rec(ctx, prod, prod, chunks, scratch + chunks));
    chunk_scan_kernel<<<blocks, PL_THREADS, 0, ctx->stream>>>(in, n, prod, out);
    B2_CUDA(ctx, cudaGetLastError());
    ctx->n_launches += 2;
    return B200_OK;
}

size_t poly_scratch_elems(size_t n) {
    size_t tot = 0, m = n;
    while (m > 1) { m = (m + PL_CHUNK - 1) / PL_CHUNK; tot += m; }
    return tot + 4;
}

int fr_batch_inverse(b200_ctx* ctx, const Fr* in, Fr* out, size_t n) {
    if (n == 0) return B200_OK;
    const size_t threads = (n + PL_CHUNK - 1) / PL_CHUNK;
    batch_inverse_kernel<<<(unsigned)((threads + PL_THREADS - 1) / PL_THREADS), PL_THREADS, 0, ctx->stream>>>(in, out, n);
    B2_CUDA(ctx, cudaGetLastError());
    ctx->n_launches += 1;
    return B200_OK;
}

int fr_prefix_product(b200_ctx* ctx, const Fr* in, Fr* out, size_t n) {
    if (n == 0) return B200_OK;
    void* sp;
    B2_TRY(ctx_poly_scratch(ctx, poly_scratch_elems(n) * sizeof(Fr), &sp));
    return prefix_product_rec(ctx, in, out, n, (Fr*)sp);
}

// ---- evaluation at a point: Horner per chunk, then the chunk values are a polynomial in x^PL_CHUNK -----
__global__ void __launch_bounds__(PL_THREADS) chunk_horner_kernel(const Fr* __restrict__ coeffs, size_t n, const Fr x, Fr* __restrict__ part) {
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t lo = t * PL_CHUNK;
    if (lo >= n) return;
    const int cnt = (int)(n - lo < (size_t)PL_CHUNK ? n - lo : (size_t)PL_CHUNK);
    Fr acc = pl_ld(coeffs + lo + cnt - 1);
    for (int i = cnt - 2; i >= 0; --i) acc = acc * x + pl_ld(coeffs + lo + i);
    pl_st(part + t, acc);
}

// Coalesced evaluation: thread t of T takes the coefficients t, t + T, t + 2T, ... (a warp reads 1 KiB
// contiguous per step) and returns part[t] = sum_j c[t + jT] y^j with y = x^T; then p(x) = sum_t part[t] x^t is
// the same problem on T coefficients.  The last <= 1024 values are finished on the host.
__global__ void __launch_bounds__(128) strided_horner_kernel(const Fr* __restrict__ c, size_t n, uint32_t T, const Fr y, Fr* __restrict__ part) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T || t >= n) return;
    size_t j = (n - 1 - t) / T;                  // highest j with t + j T < n
    Fr acc = pl_ld(c + t + j * T);
    while (j > 0) {
        --j;
        acc = acc * y + pl_ld(c + t + j * T);
    }
    pl_st(part + t, acc);
}

int fr_poly_eval(b200_ctx* ctx, const Fr* coeffs, size_t n, const Fr& x, Fr* out_host) {
    if (n == 0) { *out_host = Fr::zero(); return B200_OK; }
    constexpr size_t HOST_TAIL = 1024, T_MAX = (size_t)1 << 17;
    const Fr* cur = coeffs;
    size_t m = n;
    if (m > HOST_TAIL) {
        void* sp;
        B2_TRY(ctx_poly_scratch(ctx, (T_MAX + T_MAX / 16 + 4 * HOST_TAIL) * sizeof(Fr), &sp));   // levels shrink by >= 32x down to HOST_TAIL
        Fr* nxt = (Fr*)sp;
        while (m > HOST_TAIL) {
            size_t T = HOST_TAIL;
            while (T * 2 <= T_MAX && T * 64 <= m) T *= 2;      // the largest T <= T_MAX with >= 32 coefficients per thread
            strided_horner_kernel<<<(unsigned)((T + 127) / 128), 128, 0, ctx->stream>>>(cur, m, (uint32_t)T, x.pow_u64(T), nxt);
            B2_CUDA(ctx, cudaGetLastError());
            ctx->n_launches += 1;
            cur = nxt;
            nxt += T;
            m = T < m ? T : m;
        }
    }
    std::vector<Fr> tail(m);
    B2_CUDA(ctx, cudaMemcpyAsync(tail.data(), cur, m * sizeof(Fr), cudaMemcpyDeviceToHost, ctx->stream));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    Fr acc = tail[m - 1];
    for (size_t i = m - 1; i-- > 0;) acc = acc * x + tail[i];
    *out_host = acc;
    return B200_OK;
}

// ---- quotient by (X - z): q[n-2] = a[n-1], q[i-1] = a[i] + z q[i]; remainder a[0] + z q[0] = p(z) ------
// A chunk's outgoing carry is  carry_in * z^cnt + local,  local = Horner of the chunk at z: the carries
// of all chunks are themselves a quotient recursion over the chunk Horner values with z^PL_CHUNK.
// pass A (chunk_horner_kernel): local[t].   pass B (recursion): carry_in[t] = value entering chunk t from above.
// pass C: replay each chunk from its carry and write the quotient coefficients.
__global__ void __launch_bounds__(PL_THREADS) chunk_divide_kernel(const Fr* __restrict__ a, size_t n, const Fr z, const Fr* __restrict__ carry_in,
                                                                  Fr* __restrict__ q) {
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t lo = t * PL_CHUNK;
    if (lo >= n) return;
    const int cnt = (int)(n - lo < (size_t)PL_CHUNK ? n - lo : (size_t)PL_CHUNK);
    // acc = the value  a[i] + z * (acc of i+1)  running downwards; q[i-1] = acc at i
    Fr acc = carry_in ? pl_ld(carry_in + t) : Fr::zero();
    for (int i = cnt - 1; i >= 0; --i) {
        acc = acc * z + pl_ld(a + lo + i);
        if (lo + i >= 1) pl_st(q + lo + i - 1, acc);
    }
}
// carries of a level: carry_in[t] for chunk t = Horner state after all chunks above t.
// Given local[] (length m) this is again "divide local by (X - zc)" with the roles shifted by one:
//   carry_in[m-1] = 0, carry_in[t] = carry_in[t+1] * zc + local[t+1]            (zc = z^PL_CHUNK)
__global__ void __launch_bounds__(PL_THREADS) chunk_carry_kernel(const Fr* __restrict__ local, size_t m, const Fr zc, const Fr* __restrict__ carry_up,
                                                                 Fr* __restrict__ carry) {
    // thread t handles positions [lo, lo + cnt) of `local`; carry_up[t] enters from above the chunk
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t lo = t * PL_CHUNK;
    if (lo >= m) return;
    const int cnt = (int)(m - lo < (size_t)PL_CHUNK ? m - lo : (size_t)PL_CHUNK);
    Fr acc = carry_up ? pl_ld(carry_up + t) : Fr::zero();
    for (int i = cnt - 1; i >= 0; --i) {
        pl_st(carry + lo + i, acc);            // value entering element lo + i from above
        acc = acc * zc + pl_ld(local + lo + i);
    }
}

// carry[t] for every chunk of a level with `m` chunk values local[0..m), chunk factor zc
static int divide_carries(b200_ctx* ctx, const Fr* local, size_t m, const Fr& zc, Fr* carry, Fr* scratch) {
    const size_t chunks = (m + PL_CHUNK - 1) / PL_CHUNK;
    const unsigned blocks = (unsigned)((chunks + PL_THREADS - 1) / PL_THREADS);
    if (chunks == 1) {
        chunk_carry_kernel<<<1, PL_THREADS, 0, ctx->stream>>>(local, m, zc, nullptr, carry);
        B2_CUDA(ctx, cudaGetLastError());
        ctx->n_launches += 1;
        return B200_OK;
    }
    Fr* loc2 = scratch;                    // chunks values: Horner of each chunk of `local` at zc
    Fr* car2 = scratch + chunks;           // chunks carries
    chunk_horner_kernel<<<blocks, PL_THREADS, 0, ctx->stream>>>(local, m, zc, loc2);
    B2_CUDA(ctx, cudaGetLastError());
    B2_TRY(divide_carries(ctx, loc2, chunks, zc.pow_u64(PL_CHUNK), car2, scratch + 2 * chunks));
    chunk_carry_kernel<<<blocks, PL_THREADS, 0, ctx->stream>>>(local, m, zc, car2, carry);
    B2_CUDA(ctx, cudaGetLastError());
    ctx->n_launches += 2;
    return B200_OK;
}

int fr_poly_div_linear(b200_ctx* ctx, const Fr* a, size_t n, const Fr& z, Fr* q) {
    if (n <= 1) return B200_OK;   // constant: empty quotient
    const size_t chunks = (n + PL_CHUNK - 1) / PL_CHUNK;
    const unsigned blocks = (unsigned)((chunks + PL_THREADS - 1) / PL_THREADS);
    if (chunks == 1) {
        chunk_divide_kernel<<<1, PL_THREADS, 0, ctx->stream>>>(a, n, z, nullptr, q);
        B2_CUDA(ctx, cudaGetLastError());
        ctx->n_launches += 1;
        return B200_OK;
    }
    void* sp;
    B2_TRY(ctx_poly_scratch(ctx, (3 * poly_scratch_elems(n) + 8) * sizeof(Fr), &sp));
    Fr* local = (Fr*)sp;
    Fr* carry = local + chunks;
    chunk_horner_kernel<<<blocks, PL_THREADS, 0, ctx->stream>>>(a, n, z, local);
    B2_CUDA(ctx, cudaGetLastError());
    B2_TRY(divide_carries(ctx, local, chunks, z.pow_u64(PL_CHUNK), carry, carry + chunks));
    chunk_divide_kernel<<<blocks, PL_THREADS, 0, ctx->stream>>>(a, n, z, carry, q);
    B2_CUDA(ctx, cudaGetLastError());
    ctx->n_launches += 2;
    return B200_OK;
}

// ---- cyclic scaling: v[i] *= f[(i >> shift) mod count]  (count <= 8 factors) ----------------------
struct CycFactors { Fr f[8]; };
__global__ void scale_cyclic_kernel(Fr* __restrict__ v, size_t n, const CycFactors fac, uint32_t shift, uint32_t mask) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    pl_st(v + i, pl_ld(v + i) * fac.f[(uint32_t)(i >> shift) & mask]);
}

int fr_scale_cyclic(b200_ctx* ctx, Fr* v, size_t n, const Fr* factors, uint32_t count, uint32_t shift) {
    if (count == 0 || count > 8 || (count & (count - 1))) return ctx->fail(B200_ERR_ARG, "factor count must be 1, 2, 4 or 8");
    CycFactors fac;
    for (uint32_t i = 0; i < 8; ++i) fac.f[i] = factors[i % count];
    const int bs = 256;
    scale_cyclic_kernel<<<(unsigned)((n + bs - 1) / bs), bs, 0, ctx->stream>>>(v, n, fac, shift, count - 1);
    B2_CUDA(ctx, cudaGetLastError());
    ctx->n_launches += 1;
    return B200_OK;
}

// ---------------------------------------------------------------------------------------
// PlonK row kernels.  w^i of the small domain comes from the NTT twiddle table (w^j, j < n/2):
// w^i = -w^(i - n/2) in the upper half.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ Fr omega_pow(const Fr* __restrict__ tw, uint32_t i, uint32_t half) {
    if (half == 0) return Fr::one();                      // n = 1
    return i < half ? pl_ld(tw + i) : pl_ld(tw + (i - half)).neg();
}

// copy-constraint ratio (iop.BuildRatioCopyConstraint, called at prove.go:610-621), entries in
// Lagrange / Regular form:  num[0] = den[0] = 1 and for i < n - 1
//   num[i+1] = prod_j (beta * id[j n + i] + gamma + w_j[i]),  den[i+1] = prod_j (beta * id[perm[j n + i]] + gamma + w_j[i])
// with id = [w^i | g w^i | g^2 w^i].  Z is then prefix(num) / prefix(den).
__global__ void plonk_ratio_terms_kernel(const Fr* __restrict__ l, const Fr* __restrict__ r, const Fr* __restrict__ o,
                                         const int64_t* __restrict__ perm, const Fr* __restrict__ tw, uint32_t n, const Fr beta,
                                         const Fr gamma, const Fr g, const Fr g2, Fr* __restrict__ num, Fr* __restrict__ den) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (i == n - 1) { pl_st(num, Fr::one()); pl_st(den, Fr::one()); return; }   // slot 0
    const uint32_t half = n >> 1;
    const Fr* w[3] = {l, r, o};
    const Fr shift[3] = {Fr::one(), g, g2};
    Fr a = Fr::one(), b = Fr::one();
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        const Fr v = pl_ld(w[j] + i) + gamma;
        Fr id = omega_pow(tw, i, half);
        if (j) id = id * shift[j];
        a = a * (beta * id + v);
        const uint64_t k = (uint64_t)perm[(size_t)j * n + i];
        const uint32_t kc = (uint32_t)(k / n), ki = (uint32_t)(k % n);
        Fr idp = omega_pow(tw, ki, half);
        if (kc) idp = idp * shift[kc];
        b = b * (beta * idp + v);
    }
    pl_st(num + i + 1, a);
    pl_st(den + i + 1, b);
}

__global__ void vec_mul_inplace_kernel(Fr* __restrict__ a, const Fr* __restrict__ b, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    pl_st(a + i, pl_ld(a + i) * pl_ld(b + i));
}

int plonk_build_ratio(b200_ctx* ctx, NttDomain* d, const Fr* l, const Fr* r, const Fr* o, const int64_t* perm, const Fr& beta,
                      const Fr& gamma, Fr* z_out, Fr* ws_den, Fr* ws_inv) {
    const size_t n = d->n();
    DevBuf den_b, inv_b;
    if (!ws_den) { B2_CUDA(ctx, den_b.alloc(n * sizeof(Fr))); ws_den = den_b.as<Fr>(); }
    if (!ws_inv) { B2_CUDA(ctx, inv_b.alloc(n * sizeof(Fr))); ws_inv = inv_b.as<Fr>(); }
    const int bs = 128;
    plonk_ratio_terms_kernel<<<(unsigned)((n + bs - 1) / bs), bs, 0, ctx->stream>>>(l, r, o, perm, d->tw_fwd.as<Fr>(), (uint32_t)n, beta, gamma,
                                                                                  d->coset, d->coset * d->coset, z_out, ws_den);
    B2_CUDA(ctx, cudaGetLastError());
    ctx->n_launches += 1;
    B2_TRY(fr_prefix_product(ctx, z_out, z_out, n));
    B2_TRY(fr_prefix_product(ctx, ws_den, ws_den, n));
    B2_TRY(fr_batch_inverse(ctx, ws_den, ws_inv, n));
    vec_mul_inplace_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(z_out, ws_inv, n);
    B2_CUDA(ctx, cudaGetLastError());
    ctx->n_launches += 1;
    if (den_b.p || inv_b.p) B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));   // the local buffers die here
    return B200_OK;
}

// Quotient numerator rows (computeNumerator / allConstraints, prove.go:837-1044): for row i of the
// current coset, with every input in Lagrange / Regular form on that coset,
//   blind L, R, O, Z with the (pre-scaled) blinding polynomials at w^i, ZS = Z shifted by one row,
//   gate + alpha * ordering + alpha^2 * (Z - 1) * LOne,
// scattered to cres[bitrev_{log_big}(rho * i + coset_index)] (prove.go:1029-1035).
__global__ void __launch_bounds__(128) plonk_quotient_rows_kernel(const PlonkRowArgs a) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    const uint32_t half = a.n >> 1;
    const Fr wi = omega_pow(a.tw, i, half);
    const uint32_t i1 = i + 1 == a.n ? 0u : i + 1;
    const Fr wi1 = omega_pow(a.tw, i1, half);
    const Fr L = pl_ld(a.x[PID_L] + i) + (a.bl[1] * wi + a.bl[0]);
    const Fr R = pl_ld(a.x[PID_R] + i) + (a.br[1] * wi + a.br[0]);
    const Fr O = pl_ld(a.x[PID_O] + i) + (a.bo[1] * wi + a.bo[0]);
    const Fr Z = pl_ld(a.x[PID_Z] + i) + ((a.bz[2] * wi + a.bz[1]) * wi + a.bz[0]);
    const Fr ZS = pl_ld(a.x[PID_Z] + i1) + ((a.bz[2] * wi1 + a.bz[1]) * wi1 + a.bz[0]);
    // gate (prove.go:850-867)
    Fr gate = pl_ld(a.x[PID_QL] + i) * L + pl_ld(a.x[PID_QR] + i) * R + pl_ld(a.x[PID_QM] + i) * L * R +
              pl_ld(a.x[PID_QO] + i) * O + pl_ld(a.x[PID_QK] + i);
    for (uint32_t k = 0; k < a.nb_bsb; ++k) gate = gate + pl_ld(a.qci[2 * k] + i) * pl_ld(a.qci[2 * k + 1] + i);
    // ordering (prove.go:873-891)
    const Fr id = pl_ld(a.x[PID_ID] + i);
    const Fr r = (a.gamma + L + id) * (id * a.cs + R + a.gamma) * (id * a.css + O + a.gamma) * Z;
    const Fr l = (pl_ld(a.x[PID_S1] + i) * a.beta + L + a.gamma) * (pl_ld(a.x[PID_S2] + i) * a.beta + R + a.gamma) *
                 (pl_ld(a.x[PID_S3] + i) * a.beta + O + a.gamma) * ZS;
    // local ratio (prove.go:893-900) and the alpha combination (prove.go:952)
    Fr c = (Z - Fr::one()) * pl_ld(a.x[PID_LONE] + i);
    c = (c * a.alpha + (l - r)) * a.alpha + gate;
    const uint32_t pos = __brev(a.rho * i + a.coset_index) >> (32 - a.log_big);
    pl_st(a.cres + pos, c);
}

int plonk_quotient_rows(b200_ctx* ctx, PlonkRowArgs& args) {
    const int bs = 128;
    plonk_quotient_rows_kernel<<<(args.n + bs - 1) / bs, bs, 0, ctx->stream>>>(args);
    B2_CUDA(ctx, cudaGetLastError());
    ctx->n_launches += 1;
    return B200_OK;
}

// ---------------------------------------------------------------------------------------
// PlonK prover glue kernels (HBM-bound, one element per thread)
// ---------------------------------------------------------------------------------------
// out[bitrev(i)] = in[i]  (fft.BitReverse / Polynomial.ToRegular as a copy)
__global__ void bitrev_copy_kernel(const Fr* __restrict__ in, Fr* __restrict__ out, uint32_t log_n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >> log_n) return;
    const uint32_t j = log_n ? (__brev(i) >> (32 - log_n)) : 0u;
    pl_st(out + j, pl_ld(in + i));
}
int fr_bitrev_copy(b200_ctx* ctx, const Fr* in, Fr* out, unsigned log_n) {
    const size_t n = (size_t)1 << log_n;
    bitrev_copy_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(in, out, log_n);
    B2_CUDA(ctx, cudaGetLastError());
    ctx->n_launches += 1;
    return B200_OK;
}

// out[i] = c * w^i: the polynomial beta*X (x[id_ID], prove.go:567-570) on the coset whose shift is folded in c
__global__ void scaled_omega_kernel(Fr* __restrict__ out, const Fr* __restrict__ tw, uint32_t n, const Fr c) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    pl_st(out + i, c * omega_pow(tw, i, n >> 1));
}
int fr_scaled_omega(b200_ctx* ctx, NttDomain* d, const Fr& c, Fr* out) {
    const uint32_t n = (uint32_t)d->n();
    scaled_omega_kernel<<<(n + 255) / 256, 256, 0, ctx->stream>>>(out, d->tw_fwd.as<Fr>(), n, c);
    B2_CUDA(ctx, cudaGetLastError());
    ctx->n_launches += 1;
    return B200_OK;
}

// getBlindedCoefficients (prove.go:1148-1156): p <- p + b (X^n - 1); p has room for n + nb coefficients
struct BlindCoefs { Fr b[4]; };
__global__ void blind_kernel(Fr* __restrict__ p, uint32_t n, const BlindCoefs bc, uint32_t nb) {
    const uint32_t j = threadIdx.x;
    if (j >= nb) return;
    pl_st(p + n + j, bc.b[j]);
    pl_st(p + j, pl_ld(p + j) - bc.b[j]);      // nb <= n is checked by the caller
}
int fr_blind(b200_ctx* ctx, Fr* p, size_t n, const Fr* b, uint32_t nb) {
    if (nb > 4 || nb > n) return ctx->fail(B200_ERR_ARG, "blinding order");
    BlindCoefs bc;
    for (uint32_t j = 0; j < 4; ++j) bc.b[j] = j < nb ? b[j] : Fr::zero();
    blind_kernel<<<1, 32, 0, ctx->stream>>>(p, (uint32_t)n, bc, nb);
    B2_CUDA(ctx, cudaGetLastError());
    ctx->n_launches += 1;
    return B200_OK;
}

// foldH (prove.go:670-705): out[i] = (h3[i] * z + h2[i]) * z + h1[i]
__global__ void fold_h_kernel(const Fr* __restrict__ h1, const Fr* __restrict__ h2, const Fr* __restrict__ h3, const Fr z, size_t m,
                              Fr* __restrict__ out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    pl_st(out + i, (pl_ld(h3 + i) * z + pl_ld(h2 + i)) * z + pl_ld(h1 + i));
}
int plonk_fold_h(b200_ctx* ctx, const Fr* h1, const Fr* h2, const Fr* h3, const Fr& z, size_t m, Fr* out) {
    fold_h_kernel<<<(unsigned)((m + 255) / 256), 256, 0, ctx->stream>>>(h1, h2, h3, z, m, out);
    B2_CUDA(ctx, cudaGetLastError());
    ctx->n_launches += 1;
    return B200_OK;
}

// computeLinearizedPolynomial's loop (prove.go:1336-1385), same order of operations
__global__ void __launch_bounds__(128) linearize_kernel(const PlonkLinArgs a) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.len) return;
    const Fr bz = pl_ld(a.bz + i);
    Fr t = bz * a.c2;
    if (i < a.n) t = t + pl_ld(a.s3 + i) * a.c1;
    t = t * a.alpha;
    if (i < a.n) {
        Fr t0 = pl_ld(a.ql + i) * a.l + pl_ld(a.qm + i) * a.rl;
        t = t + t0;
        t = t + pl_ld(a.qr + i) * a.r;
        t0 = pl_ld(a.qo + i) * a.o + pl_ld(a.qk + i);
        t = t + t0;
        for (uint32_t j = 0; j < a.nb_bsb; ++j) t = t + pl_ld(a.pi2[j] + i) * a.qcp_zeta[j];
    }
    pl_st(a.out + i, t + bz * a.lag);
}
int plonk_linearize(b200_ctx* ctx, const PlonkLinArgs& a) {
    linearize_kernel<<<(a.len + 127) / 128, 128, 0, ctx->stream>>>(a);
    B2_CUDA(ctx, cudaGetLastError());
    ctx->n_launches += 1;
    return B200_OK;
}

// the fold of kzg.BatchOpenSinglePoint (call site prove.go:823): out[i] = sum_k g^k p_k[i], i < len_k
__global__ void __launch_bounds__(128) fold_polys_kernel(const PlonkFoldArgs a) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.width) return;
    Fr acc = Fr::zero();
    for (uint32_t k = 0; k < a.count; ++k) {
        if (i >= a.len[k]) continue;
        const Fr v = pl_ld(a.p[k] + i);
        acc = acc + (k ? v * a.g[k] : v);
    }
    pl_st(a.out + i, acc);
}
int plonk_fold_polys(b200_ctx* ctx, const PlonkFoldArgs& a) {
    if (a.count == 0 || a.count > 16) return ctx->fail(B200_ERR_ARG, "fold: 1..16 polynomials");
    fold_polys_kernel<<<(a.width + 127) / 128, 128, 0, ctx->stream>>>(a);
    B2_CUDA(ctx, cudaGetLastError());
    ctx->n_launches += 1;
    return B200_OK;
}

}  // namespace b200
