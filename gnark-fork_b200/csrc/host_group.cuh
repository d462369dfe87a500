// Host-side group helpers shared by host_group.cu (Groth16 fix-ups) and plonk.cu (PlonK host stages):
// scalar multiplication, Jacobian/affine loads, big-endian field bytes, point compression
// (gnark-crypto's encoding as used by Proof.WriteTo, groth16/bn254/marshal.go:41-66, plonk/bn254/marshal.go:41-66).
#pragma once
#include <cstring>
#include "common.cuh"

namespace b200 {
namespace hostgrp {


template <class F>
XYZZ<F> scalar_mul(const XYZZ<F>& p, const Fr& k_mont) {
    const Fr k = k_mont.from_mont();
    XYZZ<F> acc = XYZZ<F>::inf();
    for (int i = 255; i >= 0; --i) {
        acc = acc.dbl();
        if ((k.l[i >> 5] >> (i & 31)) & 1) acc.add(p);
    }
    return acc;
}

template <class F>
int jac_add(const uint64_t* p, const uint64_t* q, uint64_t* out) {
    if (!p || !q || !out) return B200_ERR_ARG;
    Jac<F> a, b;
    memcpy(&a, p, sizeof a);
    memcpy(&b, q, sizeof b);
    XYZZ<F> x = xyzz_from_jac(a);
    x.add(xyzz_from_jac(b));
    const Jac<F> r = x.to_jac();
    memcpy(out, &r, sizeof r);
    return B200_OK;
}

template <class F>
int jac_aff(const uint64_t* p, uint64_t* out) {
    if (!p || !out) return B200_ERR_ARG;
    Jac<F> a;
    memcpy(&a, p, sizeof a);
    const Affine<F> r = jac_to_affine(a);
    memcpy(out, &r, sizeof r);
    return B200_OK;
}

template <class F>
XYZZ<F> load_jac(const uint64_t* p) { Jac<F> a; memcpy(&a, p, sizeof a); return xyzz_from_jac(a); }
template <class F>
XYZZ<F> load_aff(const uint64_t* p) { Affine<F> a; memcpy(&a, p, sizeof a); return XYZZ<F>::from_affine(a); }

// canonical big-endian bytes of an Fp element (FP_BYTES = 32 on BN254, 48 on BLS12-377)
constexpr int FP_LIMBS = Fp::N;
constexpr int FP_BYTES = 4 * FP_LIMBS;
inline void fp_be(const Fp& x_mont, uint8_t* out) {
    const Fp x = x_mont.from_mont();
    for (int i = 0; i < FP_LIMBS; ++i) {
        const uint32_t v = x.l[FP_LIMBS - 1 - i];
        out[4 * i] = (uint8_t)(v >> 24); out[4 * i + 1] = (uint8_t)(v >> 16);
        out[4 * i + 2] = (uint8_t)(v >> 8); out[4 * i + 3] = (uint8_t)v;
    }
}
// y > (p-1)/2  (gnark-crypto fp.Element.LexicographicallyLargest)
inline bool lex_largest(const Fp& y_mont) {
    const Fp y = y_mont.from_mont();
    // (p-1)/2 limbs
    uint32_t half[FP_LIMBS];
    uint32_t carry = 0;
    for (int i = FP_LIMBS - 1; i >= 0; --i) {
        const uint32_t v = FpParams::p(i) - (i == 0 ? 1u : 0u);  // p is odd: p-1 only clears bit 0
        half[i] = (v >> 1) | (carry << 31);
        carry = v & 1u;
    }
    for (int i = FP_LIMBS - 1; i >= 0; --i) {
        if (y.l[i] != half[i]) return y.l[i] > half[i];
    }
    return false;
}
// flags in the top bits of byte 0: BN254 has two spare bits (0b10 / 0b11 / 0b01), the BLS curves three
// (0b100 / 0b101 / 0b110), gnark-crypto ecc/<curve>/marshal.go
#if defined(B2_CURVE_BLS12_377)
const uint8_t M_SMALLEST = 0x80, M_LARGEST = 0xA0, M_INFINITY = 0xC0;
#else
const uint8_t M_SMALLEST = 0x80, M_LARGEST = 0xC0, M_INFINITY = 0x40;
#endif

inline void g1_compress(const G1Affine& p, uint8_t* out) {
    if (p.is_inf()) { memset(out, 0, FP_BYTES); out[0] = M_INFINITY; return; }
    fp_be(p.x, out);
    out[0] |= lex_largest(p.y) ? M_LARGEST : M_SMALLEST;
}
inline void g2_compress(const G2Affine& p, uint8_t* out) {
    if (p.is_inf()) { memset(out, 0, 2 * FP_BYTES); out[0] = M_INFINITY; return; }
    fp_be(p.x.a1, out);
    fp_be(p.x.a0, out + FP_BYTES);
    const bool largest = p.y.a1.is_zero() ? lex_largest(p.y.a0) : lex_largest(p.y.a1);
    out[0] |= largest ? M_LARGEST : M_SMALLEST;
}


}  // namespace hostgrp
}  // namespace b200
