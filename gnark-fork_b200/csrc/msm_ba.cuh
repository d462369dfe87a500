// Batched-affine pre-reduction of the MSM buckets (an optional stage in front of msm_accumulate_kernel).
//
// A mixed XYZZ addition costs 8 products + one a*b - c*d (1288 MAC32 in Fp, 3568 in Fp2).  An AFFINE addition costs one
// inversion and 3 products; with the inversions of a whole launch shared (Montgomery's trick: prefix products, one
// inversion per BA_K x chunk additions, back-substitution) that is 6 products (816 MAC32 in Fp, 1952 in Fp2: 0.63 / 0.55
// of the XYZZ addition) + ~0.4 for the shared inversion.
// The bucket contents are reduced pairwise for R rounds -- after round r every bucket holds ceil(count / 2^r)
// affine points -- and msm_accumulate_kernel finishes from 1/2^R of the additions.
//
// Layout trick that makes the rounds bucket-agnostic: the counting sort places every bucket's entries in a region
// padded to a multiple of 2^R slots (unused slots keep the 0xffffffff sentinel = the point at infinity), so
// out[i] = in[2i] + in[2i+1] never mixes two buckets and no per-round offsets exist: offset_r = offset_0 >> r.
//
// One round = three launches over n_out additions, BA_K per thread.  Thread t of T = n_out / BA_K owns the additions
// t, t + T, t + 2T, ...: neighbouring lanes work on neighbouring pairs, so every load and store of a warp is one
// contiguous run (the first version gave each thread a contiguous chunk: 2 KB between lanes, L1-bound).
//   ba_prep_kernel     denominators d_i from the x coordinates alone, pref[i] = product of the thread's earlier d, cprod[t]
//   ba_inverse_kernel  cinv[t] = 1 / cprod[t]  (strided chunks with one Fermat inversion each)
//   ba_apply_kernel    walks the thread's additions backwards: 1/d_i = running * pref[i], then lambda, x3, y3
// Between rounds the points are stored as two planes (all x, then all y): the denominator pass streams the x plane only.
// The last round writes (x, y) pairs, which msm_accumulate_kernel reads as its bases.
// Cases: P or Q at infinity (copy), P = Q (doubling: d = 2y, lambda = 3x^2/d; a = 0), P = -Q (infinity), else
// d = x2 - x1.  d is never zero, so the shared product is invertible.
#pragma once
#include "msm.cuh"

namespace b200 {

constexpr int BA_K = 32;             // additions per thread
constexpr int BA_INV_CHUNK = 128;    // at most this many thread products per Fermat inversion

#if defined(__CUDACC__)

constexpr uint32_t BA_EMPTY = 0xffffffffu;
#ifndef BA_INLINE_FP2
#define BA_INLINE_FP2 0   // measured twice (3 blocks; 4 blocks with the low-liveness loads): inlined 15.6-15.8 ms per 5.6e7 additions against 12.8-13.6 ms with the calls (fma pipe 33 % against 42 %: the loop body leaves the instruction cache)
#endif
#ifndef BA_APPLY_BLOCKS_FP2
#define BA_APPLY_BLOCKS_FP2 4
#endif
#ifndef BA_LOW_LIVENESS_FP2
#define BA_LOW_LIVENESS_FP2 1
#endif
#define BA_LOW_LIVENESS(F) (sizeof(F) > 48 && BA_LOW_LIVENESS_FP2)
// resident blocks asked of ptxas for ba_apply_kernel.  BN254 Fp2 at 2^24 points (sum of the stage's kernels): 2 blocks (222 registers) 86.9 ms,
// 3 blocks 76.1, 3 blocks + low-liveness loads 74.9 (no stack), 4 blocks + low-liveness (128 registers, 48 B stack) 72.6, 5 blocks 78.5;
// 4 blocks WITHOUT the low-liveness loads spill to DRAM: 84.5.  BLS12-377 Fp2 (96-byte elements): 3 blocks.
#define BA_APPLY_BLOCKS(F) (sizeof(F) == 64 ? BA_APPLY_BLOCKS_FP2 : sizeof(F) > 64 ? 3 : 1)

// Products of the round kernels.  Fp2: inlined (Karatsuba with lazy reduction) instead of the out-of-line calls the XYZZ
// kernels use -- an affine addition has 5 products and a square, its loop body stays inside the instruction cache, and
// the argument shuffling of a call (IMAD.MOV on the pipe the kernel is bound by) goes away.
template <class F> __device__ __forceinline__ F ba_mul(const F& a, const F& b) { return a * b; }
template <class F> __device__ __forceinline__ F ba_sqr(const F& a) { return a.sqr(); }
#if BA_INLINE_FP2
template <> __device__ __forceinline__ Fp2 ba_mul<Fp2>(const Fp2& a, const Fp2& b) { return fp2_mul_lazy(a, b); }
template <> __device__ __forceinline__ Fp2 ba_sqr<Fp2>(const Fp2& a) { return a.sqr_inline(); }
#endif

// One operand pair of a round.  FIRST: gathered from the (folded) base table through the sorted entries (sign bit =
// negate, 0xffffffff = empty slot); later rounds: elements 2i, 2i+1 of the previous round's planes.
template <class F, bool FIRST>
struct BaPair {
    const Affine<F>* table;
    const uint32_t* sorted;
    const F* xs;
    const F* ys;
    uint32_t e0, e1;
    // the pair's entries: FIRST: two sorted entries (one 8-byte load), else the positions 2i, 2i+1
    static __device__ __forceinline__ uint2 entries(const uint32_t* sorted_, uint32_t i) {
        if (FIRST) return reinterpret_cast<const uint2*>(sorted_)[i];
        return make_uint2(2u * i, 2u * i + 1u);
    }
    __device__ __forceinline__ BaPair(const Affine<F>* table_, const uint32_t* sorted_, const F* xs_, const F* ys_, uint2 e)
        : table(table_), sorted(sorted_), xs(xs_), ys(ys_), e0(e.x), e1(e.y) {}
    __device__ __forceinline__ BaPair(const Affine<F>* table_, const uint32_t* sorted_, const F* xs_, const F* ys_, uint32_t i)
        : BaPair(table_, sorted_, xs_, ys_, entries(sorted_, i)) {}
    // start the DRAM -> L2 fetch of the pair's operands (no destination registers): issued one iteration ahead, it turns
    // the dependent gather of the next addition into an L2 hit.  with_y = false: the denominator pass reads x only.
    static __device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
    static __device__ __forceinline__ void prefetch(const Affine<F>* table_, const F* xs_, const F* ys_, uint2 e, bool with_y) {
        if (FIRST) {
            // nothing: prefetch.global.L2 pulls whole 128-byte lines, and the random gathers of the first round sit at the
            // DRAM's random-access rate (measured: 46 G 64-byte accesses/s with ld.global.nc.L2::64B, 29 G lines/s with the
            // prefetch in front -- the denominator pass went from 5.0 to 8.8 ms)
        } else {
            prefetch_l2(xs_ + e.x);
            if (sizeof(F) >= 64) prefetch_l2(xs_ + e.y);
            if (with_y) { prefetch_l2(ys_ + e.x); if (sizeof(F) >= 64) prefetch_l2(ys_ + e.y); }
        }
    }
    __device__ __forceinline__ bool empty() const { return FIRST && (e0 == BA_EMPTY || e1 == BA_EMPTY); }
    __device__ __forceinline__ F x(int k) const {
        const uint32_t e = k ? e1 : e0;
        if (FIRST) return ldg_vec(&table[e & 0x7fffffffu].x);
        return ld_vec(xs + e);
    }
    __device__ __forceinline__ F y(int k) const {
        const uint32_t e = k ? e1 : e0;
        if (FIRST) {
            if (e == BA_EMPTY) return F::zero();
            const F v = ldg_vec(&table[e & 0x7fffffffu].y);
            return (e >> 31) ? v.neg() : v;
        }
        return ld_vec(ys + e);
    }
    // The same coordinate read AGAIN where the formula uses it (low-liveness path of ba_apply_kernel): volatile, so that the
    // compiler cannot merge it with the earlier load and keep 16 words live across the out-of-line products.  It hits
    // L1 / L2 -- the line was fetched a few hundred instructions ago.
    static __device__ __forceinline__ F reload(const F* p) {
        F r;
        const uint4* s = reinterpret_cast<const uint4*>(p);
        uint4* d = reinterpret_cast<uint4*>(&r);
#pragma unroll
        for (int i = 0; i < (int)(sizeof(F) / 16); ++i)
            asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(d[i].x), "=r"(d[i].y), "=r"(d[i].z), "=r"(d[i].w) : "l"(s + i) : "memory");
        return r;
    }
    __device__ __forceinline__ F x_again(int k) const {
        const uint32_t e = k ? e1 : e0;
        return FIRST ? reload(&table[e & 0x7fffffffu].x) : reload(xs + e);
    }
    __device__ __forceinline__ F y_again(int k) const {
        const uint32_t e = k ? e1 : e0;
        if (FIRST) {
            const F v = reload(&table[e & 0x7fffffffu].y);
            return (e >> 31) ? v.neg() : v;
        }
        return reload(ys + e);
    }
    __device__ __forceinline__ Affine<F> point(int k) const {
        const uint32_t e = k ? e1 : e0;
        if (FIRST) {
            if (e == BA_EMPTY) return Affine<F>::inf();
            Affine<F> p = ldg_vec(table + (e & 0x7fffffffu));
            if (e >> 31) p.y = p.y.neg();
            return p;
        }
        return {ld_vec(xs + e), ld_vec(ys + e)};
    }
};

// Kind of an addition: 0 = no division (an operand at infinity, or P = -Q), 1 = chord (d = x2 - x1), 2 = tangent (d = 2y).
// Decided from x alone whenever the two x differ and neither is zero (the common case); y is read only otherwise.
template <class F, bool FIRST>
__device__ __forceinline__ int ba_denominator(const BaPair<F, FIRST>& pr, F* d) {
    if (pr.empty()) return 0;
    const F x0 = pr.x(0), x1 = pr.x(1);
    if (x0.is_zero() || x1.is_zero()) {          // (0, 0) is infinity; x = 0 with y != 0 can be a curve point
        const F y0 = pr.y(0), y1 = pr.y(1);
        if ((x0.is_zero() && y0.is_zero()) || (x1.is_zero() && y1.is_zero())) return 0;
        if (x0 != x1) { *d = x1 - x0; return 1; }
        if (y0 == y1 && !y0.is_zero()) { *d = y0.dbl(); return 2; }
        return 0;
    }
    if (x0 != x1) { *d = x1 - x0; return 1; }
    const F y0 = pr.y(0), y1 = pr.y(1);
    if (y0 == y1 && !y0.is_zero()) { *d = y0.dbl(); return 2; }
    return 0;
}

template <class F, bool FIRST>
__global__ void __launch_bounds__(128) ba_prep_kernel(const Affine<F>* __restrict__ table, const uint32_t* __restrict__ sorted,
                                                      const F* __restrict__ xs, const F* __restrict__ ys, uint32_t T,
                                                      F* __restrict__ pref, F* __restrict__ cprod) {
    // grid-stride over the chains: the round-0 pass is also launched as a small persistent grid (msm_enqueue_prep0) that
    // shares the SMs with the bucket stage of the previous MSM
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < T; t += gridDim.x * blockDim.x) {
    F acc = F::one();
    bool have = false;
    typedef BaPair<F, FIRST> Pair;
    uint2 e_cur = Pair::entries(sorted, t);
    uint2 e_nxt = Pair::entries(sorted, T + t);
    Pair::prefetch(table, xs, ys, e_cur, false);
#pragma unroll 1
    for (int j = 0; j < BA_K; ++j) {
        const uint32_t i = (uint32_t)j * T + t;
        uint2 e_far = e_nxt;
        if (j + 2 < BA_K) e_far = Pair::entries(sorted, (uint32_t)(j + 2) * T + t);
        if (j + 1 < BA_K) Pair::prefetch(table, xs, ys, e_nxt, false);
        if (j) st_vec(pref + i, acc);
        F d;
        const Pair pr(table, sorted, xs, ys, e_cur);
        if (ba_denominator<F, FIRST>(pr, &d)) {
            acc = have ? ba_mul(acc, d) : d;
            have = true;
        }
        e_cur = e_nxt;
        e_nxt = e_far;
    }
    st_vec(cprod + t, acc);
    }
}

// thread u inverts the elements u, u + U, u + 2U, ... (at most `chunk` of them) with one Fermat inversion
template <class F>
__global__ void __launch_bounds__(64) ba_inverse_kernel(const F* __restrict__ in, F* __restrict__ out, uint32_t m, uint32_t U, uint32_t chunk) {
    const uint32_t u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= U) return;
    F acc = F::one();
    uint32_t cnt = 0;
    for (uint32_t e = u; e < m && cnt < chunk; e += U, ++cnt) {
        st_vec(out + e, acc);
        acc = acc * ld_vec(in + e);
    }
    F inv = acc.inv();
    for (uint32_t k = cnt; k-- > 0;) {
        const uint32_t e = u + k * U;
        const F pre = ld_vec(out + e);
        st_vec(out + e, inv * pre);
        inv = inv * ld_vec(in + e);
    }
}

// LAST: the round's output is read by msm_accumulate_kernel: (x, y) pairs; otherwise two planes
template <class F, bool FIRST, bool LAST>
__global__ void __launch_bounds__(128, BA_APPLY_BLOCKS(F)) ba_apply_kernel(const Affine<F>* __restrict__ table, const uint32_t* __restrict__ sorted,
                                                       const F* __restrict__ xs, const F* __restrict__ ys, uint32_t T,
                                                       const F* __restrict__ pref, const F* __restrict__ cinv,
                                                       F* __restrict__ out_x, F* __restrict__ out_y, Affine<F>* __restrict__ out_xy) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T) return;
    F inv = ld_vec(cinv + t);
    typedef BaPair<F, FIRST> Pair;
    uint2 e_cur = Pair::entries(sorted, (uint32_t)(BA_K - 1) * T + t);
    uint2 e_nxt = Pair::entries(sorted, (uint32_t)(BA_K - 2) * T + t);
    Pair::prefetch(table, xs, ys, e_cur, true);
#pragma unroll 1
    for (int j = BA_K - 1; j >= 0; --j) {
        const uint32_t i = (uint32_t)j * T + t;
        uint2 e_far = e_nxt;
        if (j >= 2) e_far = Pair::entries(sorted, (uint32_t)(j - 2) * T + t);
        if (j >= 1) {
            Pair::prefetch(table, xs, ys, e_nxt, true);
            if (j >= 2) Pair::prefetch_l2(pref + (i - T));
        }
        const Pair pr(table, sorted, xs, ys, e_cur);
        e_cur = e_nxt;
        e_nxt = e_far;
        Affine<F> R;
        if (BA_LOW_LIVENESS(F)) {
            // Fp2: a coordinate is 16 words; holding P and Q through five out-of-line products is what pushed the kernel to
            // 222 registers (168 + stack at 3 resident blocks).  Here every coordinate is loaded where it is consumed.
            F d;
            const int kind = ba_denominator<F, FIRST>(pr, &d);        // the same decision ba_prep_kernel took
            if (kind == 0) {
                const Affine<F> P = pr.point(0), Q = pr.point(1);
                R = P.is_inf() ? Q : (Q.is_inf() ? P : Affine<F>::inf());
            } else {
                F dinv = inv;
                if (j) { dinv = ba_mul(inv, ld_vec(pref + i)); inv = ba_mul(inv, d); }
                F num;
                if (kind == 2) { const F xx = ba_sqr(pr.x_again(0)); num = xx.dbl() + xx; }
                else num = pr.y_again(1) - pr.y_again(0);
                const F lam = ba_mul(num, dinv);
                R.x = ba_sqr(lam) - pr.x_again(0) - pr.x_again(1);
                R.y = ba_mul(lam, pr.x_again(0) - R.x) - pr.y_again(0);
            }
            if (LAST) st_vec(out_xy + i, R);
            else { st_vec(out_x + i, R.x); st_vec(out_y + i, R.y); }
            continue;
        }
        const Affine<F> P = pr.point(0), Q = pr.point(1);
        const bool p_inf = P.is_inf(), q_inf = Q.is_inf();
        if (p_inf) R = Q;                            // no division: the running inverse is unchanged
        else if (q_inf) R = P;
        else {
            const bool same_x = P.x == Q.x;
            if (same_x && !(P.y == Q.y && !P.y.is_zero())) R = Affine<F>::inf();     // P = -Q (or a 2-torsion point)
            else {
                const F d = same_x ? P.y.dbl() : Q.x - P.x;
                F dinv = inv;
                if (j) { dinv = ba_mul(inv, ld_vec(pref + i)); inv = ba_mul(inv, d); }
                F num;
                if (same_x) { const F xx = ba_sqr(P.x); num = xx.dbl() + xx; }      // 3 x^2 (a = 0)
                else num = Q.y - P.y;
                const F lam = ba_mul(num, dinv);
                R.x = ba_sqr(lam) - P.x - Q.x;
                R.y = ba_mul(lam, P.x - R.x) - P.y;
            }
        }
        if (LAST) st_vec(out_xy + i, R);
        else { st_vec(out_x + i, R.x); st_vec(out_y + i, R.y); }
    }
}

#endif  // __CUDACC__

}  // namespace b200
