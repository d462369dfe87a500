// BN254 G1 / G2 group arithmetic for the MSM kernels and the host-side fix-ups.
//
// Device-side mirror of gnark-crypto's G1Affine / G1Jac / G2Affine / G2Jac as they are used by
// /root/reference/backend/groth16/bn254/prove.go:192-301 (MultiExp inputs are affine bases with
// (0,0) = infinity; results are Jacobian X,Y,Z with Z = 0 = infinity).  Buckets are kept in
// extended Jacobian "XYZZ" coordinates (x = X/ZZ, y = Y/ZZZ, ZZ^3 = ZZZ^2), whose mixed addition
// costs 8M + 2S; formulas are the EFD ones (madd-2008-s, add-2008-s, dbl-2008-s-1, mdbl-2008-s-1)
// for y^2 = x^3 + b with a = 0, with the P = +-Q and infinity cases handled explicitly because
// gnark's DummySetup (setup.go:532-546) makes every base the same point.
#pragma once
#include "field.cuh"

namespace b200 {

struct Fp2;
#if defined(__CUDACC__)
// Out-of-line Fp2 product / square: the three (two) Fp products inside are independent, so ptxas
// interleaves six (four) carry chains -- enough ILP to keep the IMAD.WIDE pipe busy at the two
// warps per scheduler that a 228-register G2 accumulator allows -- while the hot loop stays
// inside the instruction cache.
template <int TAG = 0> __device__ __noinline__ Fp2 fp2_mul_call(Fp2 a, Fp2 b);
template <int TAG = 0> __device__ __noinline__ Fp2 fp2_sqr_call(Fp2 a);
#endif

// Fp2 = Fp[u]/(u^2 - beta), beta = -1 on BN254 and -5 on BLS12-377 (reference: std/algebra/native/fields_bls12377/e12.go:79)
// (gnark-crypto E2{A0, A1}; memory order A0 then A1)
#if defined(B2_CURVE_BLS12_377)
B2_HD Fp fp2_mul_by_minus_beta(const Fp& x) { const Fp x4 = x.dbl().dbl(); return x4 + x; }   // 5 x
#else
B2_HD Fp fp2_mul_by_minus_beta(const Fp& x) { return x; }
#endif
struct Fp2 {
    Fp a0, a1;
    B2_HD static Fp2 zero() { return {Fp::zero(), Fp::zero()}; }
    B2_HD static Fp2 one() { return {Fp::one(), Fp::zero()}; }
    B2_HD bool is_zero() const { return a0.is_zero() && a1.is_zero(); }
    B2_HD bool operator==(const Fp2& o) const { return a0 == o.a0 && a1 == o.a1; }
    B2_HD bool operator!=(const Fp2& o) const { return !(*this == o); }
    B2_HD Fp2 operator+(const Fp2& o) const { return {a0 + o.a0, a1 + o.a1}; }
    B2_HD Fp2 operator-(const Fp2& o) const { return {a0 - o.a0, a1 - o.a1}; }
    B2_HD Fp2 neg() const { return {a0.neg(), a1.neg()}; }
    B2_HD Fp2 dbl() const { return {a0.dbl(), a1.dbl()}; }
    B2_HD Fp2 mul_inline(const Fp2& o) const {  // Karatsuba, 3 Fp mul
        const Fp t0 = a0 * o.a0, t1 = a1 * o.a1;
        const Fp t2 = (a0 + a1) * (o.a0 + o.a1);
        return {t0 - fp2_mul_by_minus_beta(t1), t2 - t0 - t1};
    }
#if defined(B2_CURVE_BLS12_377)
    B2_HD Fp2 sqr_inline() const {  // (a0 + a1 u)^2 = a0^2 - 5 a1^2 + 2 a0 a1 u: (a0 + a1)(a0 - 5 a1) + 4 a0 a1, 2 Fp mul
        const Fp t = a0 * a1;
        const Fp c = (a0 + a1) * (a0 - fp2_mul_by_minus_beta(a1));
        return {c + t.dbl().dbl(), t.dbl()};
    }
#else
    B2_HD Fp2 sqr_inline() const {  // complex squaring, 2 Fp mul
        const Fp t = a0 * a1;
        return {(a0 + a1) * (a0 - a1), t.dbl()};
    }
#endif
    B2_HD Fp2 operator*(const Fp2& o) const {
#if defined(__CUDA_ARCH__)
        return fp2_mul_call<0>(*this, o);
#else
        return mul_inline(o);
#endif
    }
    B2_HD Fp2 sqr() const {
#if defined(__CUDA_ARCH__)
        return fp2_sqr_call<0>(*this);
#else
        return sqr_inline();
#endif
    }
    B2_HD Fp2 inv() const {
        const Fp d = (a0.sqr() + fp2_mul_by_minus_beta(a1.sqr())).inv();     // norm = a0^2 - beta a1^2
        return {a0 * d, a1.neg() * d};
    }
};

#if defined(__CUDACC__)
// templates only to get vague linkage (one copy per translation unit that uses them)
// Karatsuba with lazy reduction: three 512-bit products, two Montgomery reductions
// (3 x 64 + 2 x 72 = 336 MAC32 instead of 3 x 136 = 408).  With a0, a1, b0, b1 < p:
//   c1 = (a0 + a1)(b0 + b1) - a0 b0 - a1 b1 = a0 b1 + a1 b0  in [0, 2 p^2)
//   c0 = a0 b0 - a1 b1 + p^2                                   in (0, 2 p^2)
// both below p * 2^256 * 0.38, the bound mont_reduce_dev needs for a canonical result.
#if defined(B2_CURVE_BLS12_377)
// u^2 = -5:  c0 = a0 b0 - 5 a1 b1 + 5 p^2 in (0, 6 p^2),  c1 = (a0 + a1)(b0 + b1) - a0 b0 - a1 b1 in [0, 2 p^2);
// 6 p^2 < p * 2^384 (p has 7 spare bits).  3 x 144 + 2 x 156 = 744 MAC32 instead of 3 x 300.
__device__ __forceinline__ Fp2 fp2_mul_lazy(const Fp2& a, const Fp2& b) {
    uint32_t t0[24], t1[24], t2[24], sa[12], sb[12];
    mul_wide_dev12(t0, a.a0.l, b.a0.l);
    mul_wide_dev12(t1, a.a1.l, b.a1.l);
    add_chain12(sa, a.a0.l, a.a1.l);
    add_chain12(sb, b.a0.l, b.a1.l);
    mul_wide_dev12(t2, sa, sb);
    sub24_dev(t2, t2, t0);
    sub24_dev(t2, t2, t1);
    uint32_t x[24];
    add24_dev(x, t1, t1);
    add24_dev(x, x, x);
    add24_dev(x, x, t1);          // 5 a1 b1 < 5 p^2
    uint32_t p2[24];
#pragma unroll
    for (int i = 0; i < 24; ++i) p2[i] = FpParams::p_squared_x5(i);
    add24_dev(t0, t0, p2);
    sub24_dev(t0, t0, x);
    Fp2 r;
    mont_reduce_dev12<FpParams>(r.a0.l, t0);
    mont_reduce_dev12<FpParams>(r.a1.l, t2);
    return r;
}
#else
__device__ __forceinline__ Fp2 fp2_mul_lazy(const Fp2& a, const Fp2& b) {
    uint32_t t0[16], t1[16], t2[16], sa[8], sb[8];
    mul_wide_dev(t0, a.a0.l, b.a0.l);
    mul_wide_dev(t1, a.a1.l, b.a1.l);
    add8_raw_dev(sa, a.a0.l, a.a1.l);
    add8_raw_dev(sb, b.a0.l, b.a1.l);
    mul_wide_dev(t2, sa, sb);
    sub16_dev(t2, t2, t0);
    sub16_dev(t2, t2, t1);
    uint32_t p2[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) p2[i] = FpParams::p_squared(i);
    add16_dev(t0, t0, p2);
    sub16_dev(t0, t0, t1);
    Fp2 r;
    mont_reduce_dev<FpParams>(r.a0.l, t0);
    mont_reduce_dev<FpParams>(r.a1.l, t2);
    return r;
}
#endif
template <int TAG> __device__ __noinline__ Fp2 fp2_mul_call(Fp2 a, Fp2 b) { return fp2_mul_lazy(a, b); }

template <int TAG> __device__ __noinline__ Fp2 fp2_sqr_call(Fp2 a) { return a.sqr_inline(); }
#endif

// In Fp2 the six-wide-product form (two reductions for a*b - c*d) measured 1 % SLOWER than two
// lazy products (register pressure in the 255-register G2 accumulator): kept as two products.
// Re-measured under msm_accumulate_smem_kernel as one out-of-line call with four Fp2 arguments: 25.3 vs 23.8 ms per
// 2^22 points (the 64 argument registers spill around the call) -- still two products.
B2_HD Fp2 mul_sub_mul(const Fp2& a, const Fp2& b, const Fp2& c, const Fp2& d) { return a * b - c * d; }

template <class F>
struct Affine {
    F x, y;
    B2_HD bool is_inf() const { return x.is_zero() && y.is_zero(); }
    B2_HD static Affine inf() { return {F::zero(), F::zero()}; }
    B2_HD Affine neg() const { return {x, y.neg()}; }
};

template <class F>
struct Jac {
    F X, Y, Z;
    B2_HD static Jac inf() { return {F::one(), F::one(), F::zero()}; }
    B2_HD bool is_inf() const { return Z.is_zero(); }
};

template <class F>
struct XYZZ {
    F X, Y, ZZ, ZZZ;
    B2_HD static XYZZ inf() { return {F::zero(), F::zero(), F::zero(), F::zero()}; }
    B2_HD bool is_inf() const { return ZZ.is_zero(); }
    B2_HD static XYZZ from_affine(const Affine<F>& p) {
        if (p.is_inf()) return inf();
        return {p.x, p.y, F::one(), F::one()};
    }
    B2_HD XYZZ neg() const { return {X, Y.neg(), ZZ, ZZZ}; }

    // doubling of an affine point (mdbl-2008-s-1, a = 0)
    B2_HD static XYZZ dbl_affine(const Affine<F>& p) {
        if (p.is_inf() || p.y.is_zero()) return inf();
        const F U = p.y.dbl(), V = U.sqr(), W = U * V, S = p.x * V;
        const F x2 = p.x.sqr(), M = x2.dbl() + x2;
        XYZZ r;
        r.X = M.sqr() - S.dbl();
        r.Y = mul_sub_mul(M, S - r.X, W, p.y);
        r.ZZ = V;
        r.ZZZ = W;
        return r;
    }
    // dbl-2008-s-1, a = 0
    B2_HD XYZZ dbl() const {
        if (is_inf()) return *this;
        const F U = Y.dbl(), V = U.sqr(), W = U * V, S = X * V;
        const F x2 = X.sqr(), M = x2.dbl() + x2;
        XYZZ r;
        r.X = M.sqr() - S.dbl();
        r.Y = mul_sub_mul(M, S - r.X, W, Y);
        r.ZZ = V * ZZ;
        r.ZZZ = W * ZZZ;
        return r;
    }
    // this += q (affine), madd-2008-s
    B2_HD void add_mixed(const Affine<F>& q) {
        if (q.is_inf()) return;
        if (is_inf()) { X = q.x; Y = q.y; ZZ = F::one(); ZZZ = F::one(); return; }
        const F U2 = q.x * ZZ, S2 = q.y * ZZZ;
        const F P = U2 - X, R = S2 - Y;
        if (P.is_zero()) {
            if (R.is_zero()) *this = dbl_affine(q);
            else *this = inf();
            return;
        }
        const F PP = P.sqr(), PPP = P * PP, Q = X * PP;
        const F X3 = R.sqr() - PPP - Q.dbl();
        Y = mul_sub_mul(R, Q - X3, Y, PPP);
        X = X3;
        ZZ = ZZ * PP;
        ZZZ = ZZZ * PPP;
    }
    // this += q, add-2008-s
    B2_HD void add(const XYZZ& q) {
        if (q.is_inf()) return;
        if (is_inf()) { *this = q; return; }
        const F U1 = X * q.ZZ, U2 = q.X * ZZ, S1 = Y * q.ZZZ, S2 = q.Y * ZZZ;
        const F P = U2 - U1, R = S2 - S1;
        if (P.is_zero()) {
            if (R.is_zero()) *this = dbl();
            else *this = inf();
            return;
        }
        const F PP = P.sqr(), PPP = P * PP, Q = U1 * PP;
        const F X3 = R.sqr() - PPP - Q.dbl();
        Y = mul_sub_mul(R, Q - X3, S1, PPP);
        X = X3;
        ZZ = ZZ * q.ZZ * PP;
        ZZZ = ZZZ * q.ZZZ * PPP;
    }
    // (X/ZZ, Y/ZZZ) == (X*ZZ / ZZ^2, Y*ZZZ / ZZ^3): a Jacobian representative with Z = ZZ
    B2_HD Jac<F> to_jac() const {
        if (is_inf()) return Jac<F>::inf();
        return {X * ZZ, Y * ZZZ, ZZ};
    }
    B2_HD Affine<F> to_affine() const {
        if (is_inf()) return Affine<F>::inf();
        const F zi = ZZZ.inv();        // 1/ZZZ
        const F zzi = (zi * ZZ).sqr(); // (ZZ/ZZZ)^2 = 1/ZZ  (ZZ^3 = ZZZ^2)
        return {X * zzi, Y * zi};
    }
};

// Jacobian helpers used on the host for the O(1) fix-ups (prove.go:206-207,219-221,248-271,294-299)
template <class F>
B2_HD XYZZ<F> xyzz_from_jac(const Jac<F>& j) {
    if (j.is_inf()) return XYZZ<F>::inf();
    const F zz = j.Z.sqr();
    return {j.X, j.Y, zz, zz * j.Z};
}
template <class F>
B2_HD Affine<F> jac_to_affine(const Jac<F>& j) {
    if (j.is_inf()) return Affine<F>::inf();
    const F zi = j.Z.inv(), zi2 = zi.sqr();
    return {j.X * zi2, j.Y * zi2 * zi};
}

typedef Affine<Fp> G1Affine;
typedef Affine<Fp2> G2Affine;
typedef Jac<Fp> G1Jac;
typedef Jac<Fp2> G2Jac;
typedef XYZZ<Fp> G1XYZZ;
typedef XYZZ<Fp2> G2XYZZ;

}  // namespace b200
