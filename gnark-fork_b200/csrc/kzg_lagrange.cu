// kzg.ToLagrangeG1 on the device: the SRS in Lagrange form, [L_i(tau)]G1 for i < n, from the monomial
// SRS [tau^j]G1 WITHOUT tau -- what plonk Setup calls at /root/reference/backend/plonk/bn254/setup.go:134
// (pk.KzgLagrange.G1, err = kzg.ToLagrangeG1(kzgSrs.Pk.G1[:int(vk.Size)])).  gnark-crypto (module absent, algorithm
// restated): copy the points, radix-2 DIF over the GROUP with the domain's inverse twiddles (butterfly: a' = a + b,
// b' = [w^-k](a - b)), bit-reverse, scale every point by 1/n.  L(tau) = F^-1 (tau^j), so the result is the inverse DFT of
// the SRS as a vector of group elements.
//
// Device shape: one launch per stage, one thread per butterfly, points affine in HBM between stages.  A butterfly is two
// XYZZ additions, one 254-bit double-and-add (skipped where the twiddle is 1: the last stage, and lane 0 of every block)
// and two Fermat normalisations; the final kernel permutes and scales.  n log n / 2 scalar multiplications are the cost
// the reference pays too (there: ScalarMultiplication per butterfly on CPU threads); 2^20 points take seconds here.
#include "misc_host.cuh"
#include "ntt_host.cuh"

namespace b200 {
namespace {

// [s]P for an affine P and a canonical (non-Montgomery) scalar, MSB first; 0 -> infinity
__device__ __forceinline__ G1XYZZ scalar_mul_affine(const G1Affine& p, const Fr& s) {
    G1XYZZ acc = G1XYZZ::inf();
    if (p.is_inf()) return acc;
    int top = Fr::N * 32 - 1;
    while (top >= 0 && !((s.l[top >> 5] >> (top & 31)) & 1u)) --top;
    for (int i = top; i >= 0; --i) {
        acc = acc.dbl();
        if ((s.l[i >> 5] >> (i & 31)) & 1u) acc.add_mixed(p);
    }
    return acc;
}

// stage with blocks of m points (half = m / 2): thread t -> block t / half, lane j = t % half;
// twiddle = tw_inv[j * (n / m)] = w^-(j n / m), Montgomery form
__global__ void __launch_bounds__(128) lagrange_stage_kernel(G1Affine* __restrict__ x, const Fr* __restrict__ tw_inv, uint32_t n, uint32_t m) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t half = m >> 1;
    if (t >= (n >> 1)) return;
    const uint32_t j = t % half, k = (t / half) * m + j;
    const G1Affine u = ld_vec(x + k), v = ld_vec(x + k + half);
    G1XYZZ sum = G1XYZZ::from_affine(u);
    sum.add_mixed(v);
    G1XYZZ diff = G1XYZZ::from_affine(u);
    diff.add_mixed(v.neg());
    G1Affine d = diff.to_affine();
    if (j != 0) {
        const Fr w = ld_vec(tw_inv + (size_t)j * (n / m)).from_mont();
        d = scalar_mul_affine(d, w).to_affine();
    }
    st_vec(x + k, sum.to_affine());
    st_vec(x + k + half, d);
}

// out[i] = [1/n] in[bitrev(i)]
__global__ void __launch_bounds__(128) lagrange_finish_kernel(const G1Affine* __restrict__ in, G1Affine* __restrict__ out, uint32_t n, uint32_t log_n, Fr ninv_canonical) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t r = log_n ? (__brev(i) >> (32 - log_n)) : 0u;
    st_vec(out + i, scalar_mul_affine(ld_vec(in + r), ninv_canonical).to_affine());
}

}  // namespace

int kzg_to_lagrange_g1(b200_ctx* ctx, const uint64_t* srs_host, unsigned log_n, const Fr& gen, uint64_t* out_host) {
    if (log_n > 26) return ctx->fail(B200_ERR_UNSUPPORTED, "ToLagrangeG1 above 2^26 points");
    const size_t n = (size_t)1 << log_n;
    {   // the generator must have order exactly n (fft.NewDomain's Generator for this size)
        Fr g = gen;
        for (unsigned s = 0; s + 1 < log_n; ++s) g = g * g;
        if (log_n == 0 ? gen != Fr::one() : (g == Fr::one() || g * g != Fr::one()))
            return ctx->fail(B200_ERR_ARG, "generator is not a primitive n-th root of unity");
    }
    DevBuf a, b, tw;
    B2_CUDA(ctx, a.alloc(n * sizeof(G1Affine)));
    B2_CUDA(ctx, b.alloc(n * sizeof(G1Affine)));
    B2_CUDA(ctx, tw.alloc((n / 2 + 1) * sizeof(Fr)));
    cudaStream_t st = ctx->stream;
    B2_CUDA(ctx, cudaMemcpyAsync(a.p, srs_host, n * sizeof(G1Affine), cudaMemcpyHostToDevice, st));
    if (n > 1) B2_TRY(gen_powers(ctx, tw.as<Fr>(), gen.inv(), Fr::one(), n / 2, 0));
    for (size_t m = n; m >= 2; m >>= 1) {
        lagrange_stage_kernel<<<(unsigned)((n / 2 + 127) / 128), 128, 0, st>>>(a.as<G1Affine>(), tw.as<Fr>(), (uint32_t)n, (uint32_t)m);
        B2_CUDA(ctx, cudaGetLastError());
        ctx->n_launches += 1;
    }
    Fr nf = Fr::zero();
    nf.l[0] = (uint32_t)n;                      // n < 2^27 as a canonical integer
    if (Fr::N > 1) nf.l[1] = (uint32_t)((uint64_t)n >> 32);
    const Fr ninv = nf.to_mont().inv().from_mont();
    lagrange_finish_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(a.as<G1Affine>(), b.as<G1Affine>(), (uint32_t)n, log_n, ninv);
    B2_CUDA(ctx, cudaGetLastError());
    ctx->n_launches += 1;
    B2_CUDA(ctx, cudaMemcpyAsync(out_host, b.p, n * sizeof(G1Affine), cudaMemcpyDeviceToHost, st));
    B2_CUDA(ctx, cudaStreamSynchronize(st));
    return B200_OK;
}

}  // namespace b200
