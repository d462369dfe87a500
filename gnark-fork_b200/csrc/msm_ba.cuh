// Batched-affine pre-reduction of the MSM buckets (an optional stage in front of msm_accumulate_kernel).
//
// A mixed XYZZ addition costs 8 products + one a*b - c*d (1288 MAC32).  An AFFINE addition costs one inversion
// and 3 products; with the inversions of a whole launch shared (Montgomery's trick: prefix products, one inversion
// per 1024 additions, back-substitution) that is 6 products (816 MAC32) + ~0.6 for the shared inversion.
// The bucket contents are reduced pairwise for R rounds -- after round r every bucket holds ceil(count / 2^r)
// affine points -- and msm_accumulate_kernel finishes from 1/2^R of the additions.
//
// Layout trick that makes the rounds bucket-agnostic: the counting sort places every bucket's entries in a region
// padded to a multiple of 2^R slots (unused slots keep the 0xffffffff sentinel = the point at infinity), so
// out[i] = in[2i] + in[2i+1] never mixes two buckets and no per-round offsets exist: offset_r = offset_0 >> r.
//
// One round = three launches over n_out additions, BA_K per thread:
//   ba_prep_kernel     denominators d_i, pref[i] = product of the d before i inside the thread's chunk, cprod[t]
//   ba_inverse_kernel  cinv[t] = 1 / cprod[t]  (chunks of BA_INV_CHUNK with one Fermat inversion each)
//   ba_apply_kernel    walks the chunk backwards: 1/d_i = running * pref[i], then lambda, x3, y3
// Cases: P or Q at infinity (copy), P = Q (doubling: d = 2y, lambda = 3x^2/d; a = 0), P = -Q (infinity), else
// d = x2 - x1.  d is never zero, so the shared product is invertible.
#pragma once
#include "msm.cuh"

namespace b200 {

constexpr int BA_K = 16;            // additions per thread
constexpr int BA_INV_CHUNK = 64;    // chunk products per Fermat inversion

#if defined(__CUDACC__)

template <class F>
__device__ __forceinline__ F ba_denominator(const Affine<F>& P, const Affine<F>& Q) {
    if (P.is_inf() || Q.is_inf()) return F::one();
    if (P.x != Q.x) return Q.x - P.x;
    if (P.y == Q.y && !P.y.is_zero()) return P.y.dbl();
    return F::one();
}

// operand `idx` of a round: the first round gathers through the sorted entries (sign bit = negate,
// 0xffffffff = empty slot), later rounds read the previous round's points
template <class F, bool FIRST>
__device__ __forceinline__ Affine<F> ba_load(const Affine<F>* __restrict__ pts, const uint32_t* __restrict__ sorted, uint32_t idx) {
    if (!FIRST) return ld_vec(pts + idx);
    const uint32_t e = sorted[idx];
    if (e == 0xffffffffu) return Affine<F>::inf();
    Affine<F> p = ldg_vec(pts + (e & 0x7fffffffu));
    if (e >> 31) p.y = p.y.neg();
    return p;
}

template <class F, bool FIRST>
__global__ void __launch_bounds__(128) ba_prep_kernel(const Affine<F>* __restrict__ pts, const uint32_t* __restrict__ sorted, uint32_t n_out,
                                                      F* __restrict__ pref, F* __restrict__ cprod) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t lo = t * BA_K;
    if (lo >= n_out) return;
    F acc = F::one();
    for (int j = 0; j < BA_K; ++j) {
        const uint32_t i = lo + j;
        const Affine<F> P = ba_load<F, FIRST>(pts, sorted, 2 * i), Q = ba_load<F, FIRST>(pts, sorted, 2 * i + 1);
        st_vec(pref + i, acc);
        acc = acc * ba_denominator(P, Q);
    }
    st_vec(cprod + t, acc);
}

template <class F>
__global__ void __launch_bounds__(64) ba_inverse_kernel(const F* __restrict__ in, F* __restrict__ out, uint32_t m) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t lo = t * BA_INV_CHUNK;
    if (lo >= m) return;
    const int cnt = (int)(m - lo < (uint32_t)BA_INV_CHUNK ? m - lo : (uint32_t)BA_INV_CHUNK);
    F acc = F::one();
    for (int i = 0; i < cnt; ++i) {
        st_vec(out + lo + i, acc);
        acc = acc * ld_vec(in + lo + i);
    }
    F inv = acc.inv();
    for (int i = cnt - 1; i >= 0; --i) {
        const F pre = ld_vec(out + lo + i);
        st_vec(out + lo + i, inv * pre);
        inv = inv * ld_vec(in + lo + i);
    }
}

template <class F, bool FIRST>
__global__ void __launch_bounds__(128) ba_apply_kernel(const Affine<F>* __restrict__ pts, const uint32_t* __restrict__ sorted, uint32_t n_out,
                                                       const F* __restrict__ pref, const F* __restrict__ cinv, Affine<F>* __restrict__ out) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t lo = t * BA_K;
    if (lo >= n_out) return;
    F inv = ld_vec(cinv + t);
    for (int j = BA_K - 1; j >= 0; --j) {
        const uint32_t i = lo + j;
        const Affine<F> P = ba_load<F, FIRST>(pts, sorted, 2 * i), Q = ba_load<F, FIRST>(pts, sorted, 2 * i + 1);
        Affine<F> R;
        if (P.is_inf()) R = Q;                       // d = 1: the running inverse is unchanged
        else if (Q.is_inf()) R = P;
        else {
            const bool same_x = P.x == Q.x;
            if (same_x && !(P.y == Q.y && !P.y.is_zero())) R = Affine<F>::inf();     // P = -Q (or a 2-torsion point): d = 1
            else {
                const F d = same_x ? P.y.dbl() : Q.x - P.x;
                const F dinv = inv * ld_vec(pref + i);
                inv = inv * d;
                F num;
                if (same_x) { const F xx = P.x.sqr(); num = xx.dbl() + xx; }      // 3 x^2 (a = 0)
                else num = Q.y - P.y;
                const F lam = num * dinv;
                R.x = lam.sqr() - P.x - Q.x;
                R.y = lam * (P.x - R.x) - P.y;
            }
        }
        st_vec(out + i, R);
    }
}

#endif  // __CUDACC__

}  // namespace b200
