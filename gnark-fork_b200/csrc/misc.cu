// Element-wise field kernels, the IMAD roofline microbenchmark and the fixed-base batch scalar
// multiplication (curve.BatchScalarMultiplicationG1/G2, /root/reference/backend/groth16/bn254/
// setup.go:251,308; prove.go:192) used to build synthetic proving keys on the device.
#include "misc_host.cuh"

namespace b200 {

template <class F>
__global__ void vec_mul_kernel(const F* __restrict__ a, const F* __restrict__ b, F* __restrict__ out, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    st_vec(out + i, ld_vec(a + i) * ld_vec(b + i));
}

int vec_mul_fr(b200_ctx* ctx, const Fr* a, const Fr* b, Fr* out, size_t n) {
    if (!n) return B200_OK;
    vec_mul_kernel<Fr><<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(a, b, out, n);
    B2_CUDA(ctx, cudaGetLastError());
    return B200_OK;
}
int vec_mul_fp(b200_ctx* ctx, const Fp* a, const Fp* b, Fp* out, size_t n) {
    if (!n) return B200_OK;
    vec_mul_kernel<Fp><<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(a, b, out, n);
    B2_CUDA(ctx, cudaGetLastError());
    return B200_OK;
}

// ---------------------------------------------------------------------------------------
// IMAD-pipe peak.  (a) the instruction stream the field multiplication is made of, timed as
// register-resident Montgomery products (SASS: IMAD.WIDE.U32.X chains);
// (b) independent plain 32-bit mad.lo chains (SASS: IMAD).
// ---------------------------------------------------------------------------------------
constexpr int PEAK_ITERS = 4096;

// (a) four independent chains of Montgomery products per thread, register resident: every
// product is 128 IMAD.WIDE.U32(.X) + 8 IMAD (= 136 MAC32, the algorithmic count 2n^2 + n).
constexpr int PEAK_MUL_ITERS = 512;
__global__ void __launch_bounds__(256) peak_wide_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, long long* cycles) {
    Fr v[4], w[4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            v[k].l[j] = in[(threadIdx.x + 8 * k + j) & 127] >> 3;   // < 2^29: reduced
            w[k].l[j] = in[(threadIdx.x + 40 * k + 5 * j) & 127] >> 3;
        }
    unsigned long long g0, g1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g0));
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < PEAK_MUL_ITERS; ++it) {
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = v[k] * w[k];
    }
    const long long t1 = clock64();
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g1));
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int j = 0; j < 8; ++j) s ^= v[k].l[j];
    if (s == 0x12345678u) out[0] = s;  // keep the chains live
    if (blockIdx.x == 0 && threadIdx.x == 0) { cycles[0] = t1 - t0; cycles[1] = (long long)(g1 - g0); }
}

__global__ void __launch_bounds__(256) peak_imad_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out) {
    uint32_t a[16];
    const uint32_t b = in[threadIdx.x & 31] | 1u, c = in[32 + (threadIdx.x & 31)];
#pragma unroll
    for (int j = 0; j < 16; ++j) a[j] = c + j;
#pragma unroll 1
    for (int it = 0; it < PEAK_ITERS; ++it) {
#pragma unroll
        for (int j = 0; j < 16; ++j) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[j]) : "r"(b), "r"(c));
    }
    uint32_t s = 0;
#pragma unroll
    for (int j = 0; j < 16; ++j) s ^= a[j];
    if (s == 0x12345678u) out[0] = s;
}

int imad_peak(b200_ctx* ctx, double* mac32_per_s, double* imad32_per_s, double* sm_mhz) {
    DevBuf in, out, cyc;
    B2_CUDA(ctx, in.alloc(128 * 4));
    B2_CUDA(ctx, out.alloc(16));
    B2_CUDA(ctx, cyc.alloc(16));
    uint32_t h[128];
    for (int i = 0; i < 128; ++i) h[i] = 0x9e3779b9u * (i + 1) + 12345u;
    B2_CUDA(ctx, cudaMemcpyAsync(in.p, h, sizeof h, cudaMemcpyHostToDevice, ctx->stream));
    const int blocks = ctx->sm_count * 4, threads = 256;
    float best_w = 1e30f, best_i = 1e30f;
    long long cycles[2] = {0, 1};
    for (int rep = 0; rep < 5; ++rep) {
        B2_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
        peak_wide_kernel<<<blocks, threads, 0, ctx->stream>>>(in.as<uint32_t>(), out.as<uint32_t>(), cyc.as<long long>());
        B2_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
        B2_CUDA(ctx, cudaEventSynchronize(ctx->ev1));
        float ms;
        B2_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
        if (rep && ms < best_w) {
            best_w = ms;
            B2_CUDA(ctx, cudaMemcpy(cycles, cyc.p, 16, cudaMemcpyDeviceToHost));
        }
        B2_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
        peak_imad_kernel<<<blocks, threads, 0, ctx->stream>>>(in.as<uint32_t>(), out.as<uint32_t>());
        B2_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
        B2_CUDA(ctx, cudaEventSynchronize(ctx->ev1));
        B2_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
        if (rep && ms < best_i) best_i = ms;
    }
    const double lanes = (double)blocks * threads;
    if (mac32_per_s) *mac32_per_s = lanes * PEAK_MUL_ITERS * 4.0 * 136.0 / (best_w * 1e-3);
    if (imad32_per_s) *imad32_per_s = lanes * PEAK_ITERS * 16.0 / (best_i * 1e-3);
    // SM clock seen by the kernel: clock64 delta over the globaltimer delta (ns) of one thread
    if (sm_mhz) *sm_mhz = (double)cycles[0] / (double)(cycles[1] ? cycles[1] : 1) * 1e3;
    return B200_OK;
}

// ---------------------------------------------------------------------------------------
// fixed-base batch scalar multiplication: 32 windows of 8 bits, table T[w][d-1] = d * 2^(8w) * G
// (affine, built on the host), one thread per scalar, per-thread Fermat inversion to normalise.
// ---------------------------------------------------------------------------------------
template <class F>
__global__ void __launch_bounds__(128) batch_mul_kernel(const Affine<F>* __restrict__ table, const Fr* __restrict__ scalars, size_t n,
                                 Affine<F>* __restrict__ out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Fr s = ld_vec(scalars + i).from_mont();
    XYZZ<F> acc = XYZZ<F>::inf();
    for (int w = 0; w < 32; ++w) {
        const uint32_t d = (s.l[w >> 2] >> ((w & 3) * 8)) & 0xffu;
        if (d) acc.add_mixed(ldg_vec(table + (size_t)w * 255 + d - 1));
    }
    st_vec(out + i, acc.to_affine());
}

template <class F>
static int batch_mul_impl(b200_ctx* ctx, const Affine<F>& base, const Fr* d_scalars, size_t n, Affine<F>* d_out) {
    if (!n) return B200_OK;
    // host table: projective accumulation, then one batched inversion
    const int TW = 32, TD = 255;
    std::vector<XYZZ<F>> proj((size_t)TW * TD);
    XYZZ<F> b = XYZZ<F>::from_affine(base);
    for (int w = 0; w < TW; ++w) {
        XYZZ<F> acc = b;
        for (int d = 1; d <= TD; ++d) {
            proj[(size_t)w * TD + d - 1] = acc;
            acc.add(b);
        }
        b = acc;  // 256 * previous
    }
    // batch inversion of ZZZ (Montgomery trick); infinity entries (base = inf) stay (0,0)
    std::vector<Affine<F>> table(proj.size());
    std::vector<F> prefix(proj.size());
    F run = F::one();
    for (size_t k = 0; k < proj.size(); ++k) {
        prefix[k] = run;
        if (!proj[k].is_inf()) run = run * proj[k].ZZZ;
    }
    F inv = run.inv();
    for (size_t k = proj.size(); k-- > 0;) {
        if (proj[k].is_inf()) { table[k] = Affine<F>::inf(); continue; }
        const F zi = inv * prefix[k];          // 1/ZZZ_k
        inv = inv * proj[k].ZZZ;
        const F zzi = (zi * proj[k].ZZ).sqr();  // 1/ZZ_k
        table[k] = {proj[k].X * zzi, proj[k].Y * zi};
    }
    DevBuf dt;
    B2_CUDA(ctx, dt.alloc(table.size() * sizeof(Affine<F>)));
    B2_CUDA(ctx, cudaMemcpyAsync(dt.p, table.data(), table.size() * sizeof(Affine<F>), cudaMemcpyHostToDevice, ctx->stream));
    batch_mul_kernel<F><<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(dt.as<Affine<F>>(), d_scalars, n, d_out);
    B2_CUDA(ctx, cudaGetLastError());
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B200_OK;
}

int batch_mul_g1(b200_ctx* ctx, const G1Affine& base, const Fr* d_scalars, size_t n, G1Affine* d_out) {
    return batch_mul_impl<Fp>(ctx, base, d_scalars, n, d_out);
}
int batch_mul_g2(b200_ctx* ctx, const G2Affine& base, const Fr* d_scalars, size_t n, G2Affine* d_out) {
    return batch_mul_impl<Fp2>(ctx, base, d_scalars, n, d_out);
}


// ---------------------------------------------------------------------------------------------------------
// On-disk point format -> device layout (the data format in front of the path).
// ProvingKey.WriteRawTo (/root/reference/backend/groth16/bn254/marshal.go:246-304) stores every point slice with
// gnark-crypto's RawEncoding: X || Y (G2: X.A1 || X.A0 || Y.A1 || Y.A0), each coordinate FP_BYTES big-endian
// CANONICAL bytes; the top bits of byte 0 are flags: 00 = uncompressed, 01 = infinity (rest zero); anything else is
// a compressed form and is refused.  One thread per point: the point's bytes are read as words, byte-swapped into
// little-endian limbs, brought into Montgomery form (x * R^2 / R) and written back over the same bytes -- the raw
// point and the Affine struct have the same size, so a file section DMA'd to the device becomes the base array.
// HBM-bound: sizeof(Affine) read + written per point, 300 (G1) / 600 (G2) MAC32 on BN254.
// ---------------------------------------------------------------------------------------------------------
#if defined(B2_CURVE_BLS12_377)
constexpr uint32_t RAW_FLAG_MASK = 0xE0000000u, RAW_FLAG_INF = 0x40000000u;   // three flag bits on the BLS curves
#else
constexpr uint32_t RAW_FLAG_MASK = 0xC0000000u, RAW_FLAG_INF = 0x40000000u;   // two on BN254
#endif

template <int COORDS>
__global__ void __launch_bounds__(128) points_from_raw_kernel(uint32_t* __restrict__ pts, size_t n, uint32_t* __restrict__ bad) {
    constexpr int NL = Fp::N;
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t* p = pts + i * (size_t)(COORDS * NL);
    uint32_t w[COORDS * NL];
#pragma unroll
    for (int k = 0; k < COORDS * NL / 4; ++k) {
        const uint4 v = reinterpret_cast<const uint4*>(p)[k];
        w[4 * k] = v.x; w[4 * k + 1] = v.y; w[4 * k + 2] = v.z; w[4 * k + 3] = v.w;
    }
    // word 0 holds byte 0 in its LOW byte (little-endian load of big-endian data): swap first
    const uint32_t first = __byte_perm(w[0], 0, 0x0123);
    const uint32_t flags = first & RAW_FLAG_MASK;
    Fp c[COORDS];
    bool is_bad = flags != 0 && flags != RAW_FLAG_INF;
    const bool inf = flags == RAW_FLAG_INF;
#pragma unroll
    for (int k = 0; k < COORDS; ++k) {
        Fp x;
#pragma unroll
        for (int j = 0; j < NL; ++j) x.l[j] = __byte_perm(w[k * NL + (NL - 1 - j)], 0, 0x0123);
        if (k == 0) x.l[NL - 1] &= ~RAW_FLAG_MASK;
        // canonical input required: x < p
        bool lt = false, decided = false;
#pragma unroll
        for (int j = NL - 1; j >= 0; --j) {
            const uint32_t pj = FpParams::p(j);
            if (!decided && x.l[j] != pj) { lt = x.l[j] < pj; decided = true; }
        }
        is_bad |= !lt;
        if (inf && !x.is_zero()) is_bad = true;
        c[k] = x.to_mont();
    }
    if (inf || is_bad) {
#pragma unroll
        for (int k = 0; k < COORDS; ++k) c[k] = Fp::zero();
    }
    if (is_bad && bad) atomicOr(bad, 1u);
    // G1: X, Y.  G2 on disk: X.A1, X.A0, Y.A1, Y.A0 -> memory: X.A0, X.A1, Y.A0, Y.A1
#pragma unroll
    for (int k = 0; k < COORDS; ++k) {
        const int dst = COORDS == 4 ? (k ^ 1) : k;
#pragma unroll
        for (int q = 0; q < NL / 4; ++q)
            reinterpret_cast<uint4*>(p + dst * NL)[q] = make_uint4(c[k].l[4 * q], c[k].l[4 * q + 1], c[k].l[4 * q + 2], c[k].l[4 * q + 3]);
    }
}

// []fr.Element as curve.Encoder writes it (plonk/bn254/marshal.go:158-176: the trace polynomials of a proving key):
// 32-byte big-endian CANONICAL integers -> Montgomery limbs, in place.  *bad is set when an element is >= r.
__global__ void __launch_bounds__(256) fr_from_be_kernel(uint32_t* __restrict__ v, size_t n, uint32_t* __restrict__ bad) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint4* p = reinterpret_cast<uint4*>(v + i * 8);
    const uint4 a = p[0], b = p[1];
    const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    Fr x;
#pragma unroll
    for (int j = 0; j < 8; ++j) x.l[j] = __byte_perm(w[7 - j], 0, 0x0123);
    bool lt = false, decided = false;
#pragma unroll
    for (int j = 7; j >= 0; --j) {
        const uint32_t rj = FrParams::p(j);
        if (!decided && x.l[j] != rj) { lt = x.l[j] < rj; decided = true; }
    }
    if (!lt) { if (bad) atomicOr(bad, 1u); x = Fr::zero(); }
    x = x.to_mont();
    p[0] = make_uint4(x.l[0], x.l[1], x.l[2], x.l[3]);
    p[1] = make_uint4(x.l[4], x.l[5], x.l[6], x.l[7]);
}
// []int64 as encoding/binary writes it (big-endian), in place (pk.trace.S, marshal.go:175)
__global__ void __launch_bounds__(256) i64_from_be_kernel(uint2* __restrict__ v, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint2 w = v[i];
    v[i] = make_uint2(__byte_perm(w.y, 0, 0x0123), __byte_perm(w.x, 0, 0x0123));
}
int fr_from_be_inplace(b200_ctx* ctx, Fr* d, size_t n, uint32_t* d_bad) {
    if (n == 0) return B200_OK;
    fr_from_be_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(reinterpret_cast<uint32_t*>(d), n, d_bad);
    B2_CUDA(ctx, cudaGetLastError());
    ctx->n_launches++;
    return B200_OK;
}
int i64_from_be_inplace(b200_ctx* ctx, int64_t* d, size_t n) {
    if (n == 0) return B200_OK;
    i64_from_be_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(reinterpret_cast<uint2*>(d), n);
    B2_CUDA(ctx, cudaGetLastError());
    ctx->n_launches++;
    return B200_OK;
}

int points_from_raw_inplace(b200_ctx* ctx, void* d_points, size_t n, int group, uint32_t* d_bad) {
    if (n == 0) return B200_OK;
    const unsigned blocks = (unsigned)((n + 127) / 128);
    if (group == 1) points_from_raw_kernel<2><<<blocks, 128, 0, ctx->stream>>>((uint32_t*)d_points, n, d_bad);
    else points_from_raw_kernel<4><<<blocks, 128, 0, ctx->stream>>>((uint32_t*)d_points, n, d_bad);
    B2_CUDA(ctx, cudaGetLastError());
    ctx->n_launches++;
    return B200_OK;
}

}  // namespace b200
