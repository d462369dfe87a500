// Host-side NTT domain object (device tables) shared by ntt.cu, groth16.cu and capi.cu.
#pragma once
#include "common.cuh"
#include "ntt.cuh"

namespace b200 {

// Device mirror of fft.Domain (gnark-crypto; fields the reference touches are listed in
// Cardinality, Generator, FrMultiplicativeGen, CardinalityInv, Twiddles, CosetTable).
struct NttDomain {
    enum Table {
        COSET_NAT = 0,     // g^i                     pre-scale, FFT DIF OnCoset
        COSET_BR,          // g^bitrev(i)             pre-scale, FFT DIT OnCoset      (prove.go:373-375)
        COSETINV_NAT,      // g^-i / n                post-scale, FFTInverse DIT OnCoset
        COSETINV_BR,       // g^-bitrev(i) / n        post-scale, FFTInverse DIF OnCoset (prove.go:393)
        H_BC,              // g^bitrev(i) / n         FFTInverse DIF fused with the following coset FFT DIT (vector b)
        H_A,               // g^bitrev(i) / n * den   same for vectors a and c, den = 1/(g^n - 1) folded in (prove.go:377-390)
        NUM_TABLES
    };
    unsigned log_n = 0;
    int refs = 0;           // resident keys that point at this domain (b200_domain_cache_trim frees the unreferenced ones)
    Fr gen, coset, gen_inv, coset_inv, ninv, den;
    DevBuf tw_fwd, tw_inv;  // w^j, w^-j for j < n/2
    DevBuf tab[NUM_TABLES];
    size_t n() const { return (size_t)1 << log_n; }
};

// ntt.cu
// out[i] = mult * base^e(i), e(i) = i, or the bit reversal of i over bitrev_log_n bits when that is not 0
int gen_powers(b200_ctx* ctx, Fr* out, const Fr& base, const Fr& mult, size_t count, int bitrev_log_n);
int ntt_domain_get(b200_ctx* ctx, unsigned log_n, const Fr& gen, const Fr& coset, NttDomain** out);
int ntt_domain_table(b200_ctx* ctx, NttDomain* d, NttDomain::Table t, const Fr** out);
void ntt_domain_free_all(b200_ctx* ctx);
// frees every cached domain no resident key refers to; returns the bytes released
size_t ntt_domain_trim(b200_ctx* ctx);
// one transform on a device-resident vector.  pre / post: optional per-element scale tables
// (indexed by memory position) applied when loading the first pass / storing the last pass;
// post_scalar: optional single factor applied with the last store.
int ntt_run(b200_ctx* ctx, NttDomain* d, Fr* data, int inverse, int dit, const Fr* pre, const Fr* post,
            const Fr* post_scalar);
// one phase of a transform distributed over 2^lg ranks (ntt.cuh: NttDistPhaseKind)
// peer_src (cross phases only, optional): the 2^lg ranks' send buffers (peer-mapped device pointers); the phase then
// LOADS its input from them -- chunk `rank` of each -- instead of reading a received buffer, and stores to `data`.
int ntt_dist_phase(b200_ctx* ctx, NttDomain* d, Fr* data, Fr* tmp, uint32_t lg, uint32_t rank, int kind, int inverse,
                   const Fr* post, const Fr* post_scalar, const Fr* const* peer_src = nullptr);
// PACK fused with its all-to-all: this rank's natural-order block is scattered into the ranks' buffers peer_dst[q]
// (slot (rank << m) | (i >> lg) for element i, q = i mod 2^lg).
int ntt_dist_pack_push(b200_ctx* ctx, NttDomain* d, const Fr* block, Fr* const* peer_dst, uint32_t lg, uint32_t rank);
// gnark-crypto semantics (Domain.FFT / FFTInverse with optional OnCoset)
int ntt_gnark(b200_ctx* ctx, NttDomain* d, Fr* data, int inverse, int dit, int on_coset);
// a[i] = a[i]*b[i] - c[i]
int ntt_mul_sub(b200_ctx* ctx, Fr* a, const Fr* b, const Fr* c, size_t n);
// computeH on device vectors a, b, c of n elements (zero padded); result in a, bit-reversed.
int ntt_compute_h(b200_ctx* ctx, NttDomain* d, Fr* a, Fr* b, Fr* c);

}  // namespace b200
