// Montgomery prime-field arithmetic on 8 x 32-bit limbs for sm_100a (BN254 Fp and Fr).
//
// Replaces, for the device path, gnark-crypto's fp.Element / fr.Element arithmetic that
// /root/reference/backend/groth16/bn254/prove.go reaches through MultiExp and fft.Domain
// (prove.go:201-290,369-393).  Storage is bit-compatible with the Go side: an element is
// 4 little-endian u64 limbs in Montgomery form (R = 2^256), viewed here as 8 u32 limbs.
//
// Device multiplication: operand-scanning Montgomery (CIOS) with the running sum split into
// an "even" and an "odd" accumulator so that every row is one uninterrupted
// mad.lo.cc / madc.hi.cc carry chain (ptxas fuses each lo/hi pair into one IMAD.WIDE.U32.X).
// Invariant: T = E + O*2^32 < 2p < 2^255 between rows (p has two spare top bits), so E needs
// 9 limbs inside a row and O never overflows 8.  After the reduction row the low limb of E is
// zero, T/2^32 = (E>>32) + O, and the two accumulators swap parity: O becomes the new even
// accumulator (plus the stray limb e1), E>>64 the new odd one.
//
// Host builds of the same templates (used by the host-side fix-ups of prove.go:192-301 and by
// CPU unit tests) take the portable u64 path.
#pragma once
#include <cstdint>
#include <cstring>

#if defined(__CUDACC__)
#define B2_HD __host__ __device__ __forceinline__
#define B2_D __device__ __forceinline__
#else
#define B2_HD inline
#define B2_D inline
#endif

namespace b200 {

// ---------------------------------------------------------------------------------------
// Field parameter packs.  Constants are derived from the moduli pinned in the reference at
// backend/groth16/bn254/solidity.go:41-42 (P, R);  INV = -p^-1 mod 2^32, ONE = 2^256 mod p,
// R2 = 2^512 mod p.
// ---------------------------------------------------------------------------------------
#if defined(B2_CURVE_BLS12_377)
#if defined(__CUDACC__)
static __constant__ uint32_t b2_opaque_inv = 0xffffffffu;   // see FpParams::inv_dev
#endif
// BLS12-377 (BASELINE.json config 5; p and r pinned in the reference at std/math/emulated/emparams/emparams.go:123,137):
// Fp has 12 x 32-bit limbs (377 bits, R = 2^384), Fr 8 limbs (253 bits, R = 2^256).  Built as a second library from
// the same sources with -DB2_CURVE_BLS12_377.
// 12-limb copies of the hand-scheduled even/odd chains below are generated into field_chains12.cuh
// (scripts/gen_field_chains.py).
struct FpParams {
    static constexpr int N = 12;
    B2_HD static constexpr uint32_t p(int i) {
        constexpr uint32_t v[12] = {0x00000001u, 0x8508c000u, 0x30000000u, 0x170b5d44u, 0xba094800u, 0x1ef3622fu,
                                    0x00f5138fu, 0x1a22d9f3u, 0x6ca1493bu, 0xc63b05c0u, 0x17c510eau, 0x01ae3a46u};
        return v[i];
    }
    B2_HD static constexpr uint32_t one(int i) {
        constexpr uint32_t v[12] = {0xffffff68u, 0x02cdffffu, 0x7fffffb1u, 0x51409f83u, 0x8a7d3ff2u, 0x9f7db3a9u,
                                    0x6e7c6305u, 0x7b4e97b7u, 0x803c84e8u, 0x4cf495bfu, 0xe2fdf49au, 0x008d6661u};
        return v[i];
    }
    B2_HD static constexpr uint32_t r2(int i) {
        constexpr uint32_t v[12] = {0x9400cd22u, 0xb786686cu, 0xb00431b1u, 0x0329fcaau, 0x62d6b46du, 0x22a5f111u,
                                    0x827dc3acu, 0xbfdf7d03u, 0x41790bf9u, 0x837e92f0u, 0x1e914b88u, 0x006dfccbu};
        return v[i];
    }
    // p^2 and 5 p^2 (24 limbs): offsets that keep a*b - c*d and the Fp2 product a0 b0 - 5 a1 b1 non-negative before
    // the single reduction; 6 p^2 < p * 2^384, the bound mont_reduce_dev12 needs for a canonical result
    B2_HD static constexpr uint32_t p_squared(int i) {
        constexpr uint32_t v[24] = {0x00000001u, 0x0a118000u, 0xf0000001u, 0x7338d254u, 0x2e1bd800u, 0x4ada268fu, 0x35f1c09au, 0x6bcbfbd2u,
                                    0x58638c9du, 0x318324b9u, 0x8bb70ae0u, 0x460aaaaau, 0x502a4d6cu, 0xc014e712u, 0xb90660cdu, 0x09d018afu,
                                    0x3dda4d5cu, 0x1f5e7141u, 0xa4aee93fu, 0x4bb8b87du, 0xb361263cu, 0x2256913bu, 0xd0bbaffbu, 0x0002d307u};
        return v[i];
    }
    B2_HD static constexpr uint32_t p_squared_x5(int i) {
        constexpr uint32_t v[24] = {0x00000005u, 0x32578000u, 0xb0000005u, 0x401c1ba8u, 0xe68b3802u, 0x7642c0cbu, 0x0db8c303u, 0x1afbeb1bu,
                                    0xb9f1bf13u, 0xf78fb79eu, 0xba933660u, 0x5e355554u, 0x90d3831du, 0xc068835bu, 0x9d1fe404u, 0x31107b6eu,
                                    0x354382ccu, 0x9cd83646u, 0x376a8e3bu, 0x7a9b9a74u, 0x80e5bf2du, 0xabb0d62au, 0x13aa6fe7u, 0x000e1f27u};
        return v[i];
    }
    static constexpr uint32_t INV = 0xffffffffu;
    // Device code reads INV through constant memory: with the literal 0xffffffff ptxas rewrites m = t0 * INV as -t0 and
    // the m * p rows no longer match its mad.lo.cc / madc.hi.cc -> IMAD.WIDE pattern -- every pair became IMAD.X +
    // IMAD.HI.X (424 multiplier instructions per 12-limb product instead of 303; cuobjdump, profiles/r02_sass_histograms.txt)
    B2_D static uint32_t inv_dev() {
#if defined(__CUDA_ARCH__)
        return b2_opaque_inv;
#else
        return INV;
#endif
    }
};

struct FrParams {
    static constexpr int N = 8;
    B2_HD static constexpr uint32_t p(int i) {
        constexpr uint32_t v[8] = {0x00000001u, 0x0a118000u, 0xd0000001u, 0x59aa76feu,
                                   0x5c37b001u, 0x60b44d1eu, 0x9a2ca556u, 0x12ab655eu};
        return v[i];
    }
    B2_HD static constexpr uint32_t one(int i) {
        constexpr uint32_t v[8] = {0xfffffff3u, 0x7d1c7fffu, 0x6ffffff2u, 0x7257f50fu,
                                   0x512c0feeu, 0x16d81575u, 0x2bbb9a9du, 0x0d4bda32u};
        return v[i];
    }
    B2_HD static constexpr uint32_t r2(int i) {
        constexpr uint32_t v[8] = {0xb861857bu, 0x25d577bau, 0x8860591fu, 0xcc2c27b5u,
                                   0xe5dc8593u, 0xa7cc008fu, 0xeff1c939u, 0x011fdae7u};
        return v[i];
    }
    static constexpr uint32_t INV = 0xffffffffu;
    // Device code reads INV through constant memory: with the literal 0xffffffff ptxas rewrites m = t0 * INV as -t0 and
    // the m * p rows no longer match its mad.lo.cc / madc.hi.cc -> IMAD.WIDE pattern -- every pair became IMAD.X +
    // IMAD.HI.X (424 multiplier instructions per 12-limb product instead of 303; cuobjdump, profiles/r02_sass_histograms.txt)
    B2_D static uint32_t inv_dev() {
#if defined(__CUDA_ARCH__)
        return b2_opaque_inv;
#else
        return INV;
#endif
    }
};

#else
struct FpParams {
    static constexpr int N = 8;
    B2_HD static constexpr uint32_t p(int i) {
        constexpr uint32_t v[8] = {0xd87cfd47u, 0x3c208c16u, 0x6871ca8du, 0x97816a91u,
                                   0x8181585du, 0xb85045b6u, 0xe131a029u, 0x30644e72u};
        return v[i];
    }
    B2_HD static constexpr uint32_t one(int i) {
        constexpr uint32_t v[8] = {0xc58f0d9du, 0xd35d438du, 0xf5c70b3du, 0x0a78eb28u,
                                   0x7879462cu, 0x666ea36fu, 0x9a07df2fu, 0x0e0a77c1u};
        return v[i];
    }
    B2_HD static constexpr uint32_t r2(int i) {
        constexpr uint32_t v[8] = {0x538afa89u, 0xf32cfc5bu, 0xd44501fbu, 0xb5e71911u,
                                   0x0a417ff6u, 0x47ab1effu, 0xcab8351fu, 0x06d89f71u};
        return v[i];
    }
    static constexpr uint32_t INV = 0xe4866389u;
    B2_D static uint32_t inv_dev() { return INV; }
    // p^2 (16 words): added to a difference of wide products to keep it non-negative (Fp2 lazy reduction)
    B2_HD static constexpr uint32_t p_squared(int i) {
        constexpr uint32_t v[16] = {0x275d69b1u, 0x3b5458a2u, 0x09eac101u, 0xa602072du, 0x6d96cadcu, 0x4a50189cu, 0x7a1242c8u, 0x04689e95u, 0x34c6b38du, 0x26edfa5cu, 0x16375606u, 0xb00b8551u, 0x0348d21cu, 0x599a6f7cu, 0x763cbf9cu, 0x0925c4b8u};
        return v[i];
    }
};

struct FrParams {
    static constexpr int N = 8;
    B2_HD static constexpr uint32_t p(int i) {
        constexpr uint32_t v[8] = {0xf0000001u, 0x43e1f593u, 0x79b97091u, 0x2833e848u,
                                   0x8181585du, 0xb85045b6u, 0xe131a029u, 0x30644e72u};
        return v[i];
    }
    B2_HD static constexpr uint32_t one(int i) {
        constexpr uint32_t v[8] = {0x4ffffffbu, 0xac96341cu, 0x9f60cd29u, 0x36fc7695u,
                                   0x7879462eu, 0x666ea36fu, 0x9a07df2fu, 0x0e0a77c1u};
        return v[i];
    }
    B2_HD static constexpr uint32_t r2(int i) {
        constexpr uint32_t v[8] = {0xae216da7u, 0x1bb8e645u, 0xe35c59e3u, 0x53fe3ab1u,
                                   0x53bb8085u, 0x8c49833du, 0x7f4e44a5u, 0x0216d0b1u};
        return v[i];
    }
    static constexpr uint32_t INV = 0xefffffffu;
    B2_D static uint32_t inv_dev() { return INV; }
};

#endif

// ---------------------------------------------------------------------------------------
// Device carry-chain helpers (8 limbs).  Every helper is one self-contained asm block: the
// carry flag never has to survive between two asm statements.
// ---------------------------------------------------------------------------------------
#if defined(__CUDACC__)

// acc[0..7] += {x0,x1,x2,x3} * y laid out as 64-bit products at limb pairs (0,1),(2,3),(4,5),(6,7).
// No carry leaves the top (caller guarantees the bound).
B2_D void mad_chain8(uint32_t* acc, uint32_t x0, uint32_t x1, uint32_t x2, uint32_t x3, uint32_t y) {
    asm("mad.lo.cc.u32 %0, %8, %12, %0;\n\t"
        "madc.hi.cc.u32 %1, %8, %12, %1;\n\t"
        "madc.lo.cc.u32 %2, %9, %12, %2;\n\t"
        "madc.hi.cc.u32 %3, %9, %12, %3;\n\t"
        "madc.lo.cc.u32 %4, %10, %12, %4;\n\t"
        "madc.hi.cc.u32 %5, %10, %12, %5;\n\t"
        "madc.lo.cc.u32 %6, %11, %12, %6;\n\t"
        "madc.hi.u32 %7, %11, %12, %7;\n\t"
        : "+r"(acc[0]), "+r"(acc[1]), "+r"(acc[2]), "+r"(acc[3]), "+r"(acc[4]), "+r"(acc[5]),
          "+r"(acc[6]), "+r"(acc[7])
        : "r"(x0), "r"(x1), "r"(x2), "r"(x3), "r"(y));
}

// Same, and the carry out of limb 7 is written to (NEW=true) or added into (NEW=false) acc[8].
template <bool NEW>
B2_D void mad_chain8_top(uint32_t* acc, uint32_t x0, uint32_t x1, uint32_t x2, uint32_t x3, uint32_t y) {
    if (NEW) {
        asm("mad.lo.cc.u32 %0, %9, %13, %0;\n\t"
            "madc.hi.cc.u32 %1, %9, %13, %1;\n\t"
            "madc.lo.cc.u32 %2, %10, %13, %2;\n\t"
            "madc.hi.cc.u32 %3, %10, %13, %3;\n\t"
            "madc.lo.cc.u32 %4, %11, %13, %4;\n\t"
            "madc.hi.cc.u32 %5, %11, %13, %5;\n\t"
            "madc.lo.cc.u32 %6, %12, %13, %6;\n\t"
            "madc.hi.cc.u32 %7, %12, %13, %7;\n\t"
            "addc.u32 %8, 0, 0;\n\t"
            : "+r"(acc[0]), "+r"(acc[1]), "+r"(acc[2]), "+r"(acc[3]), "+r"(acc[4]), "+r"(acc[5]),
              "+r"(acc[6]), "+r"(acc[7]), "=r"(acc[8])
            : "r"(x0), "r"(x1), "r"(x2), "r"(x3), "r"(y));
    } else {
        asm("mad.lo.cc.u32 %0, %9, %13, %0;\n\t"
            "madc.hi.cc.u32 %1, %9, %13, %1;\n\t"
            "madc.lo.cc.u32 %2, %10, %13, %2;\n\t"
            "madc.hi.cc.u32 %3, %10, %13, %3;\n\t"
            "madc.lo.cc.u32 %4, %11, %13, %4;\n\t"
            "madc.hi.cc.u32 %5, %11, %13, %5;\n\t"
            "madc.lo.cc.u32 %6, %12, %13, %6;\n\t"
            "madc.hi.cc.u32 %7, %12, %13, %7;\n\t"
            "addc.u32 %8, %8, 0;\n\t"
            : "+r"(acc[0]), "+r"(acc[1]), "+r"(acc[2]), "+r"(acc[3]), "+r"(acc[4]), "+r"(acc[5]),
              "+r"(acc[6]), "+r"(acc[7]), "+r"(acc[8])
            : "r"(x0), "r"(x1), "r"(x2), "r"(x3), "r"(y));
    }
}

// Parity swap + next odd row in one chain:
//   ne0      = o0 + e[1]                       (stray limb folded into the new even accumulator)
//   no[0..7] = (e[2..8], 0) + {x1,x3,x5,x7}*y  (+ the carry of the fold, same weight 2^32)
B2_D void fold_mad_shift8(uint32_t& ne0, uint32_t* no, const uint32_t* e, uint32_t o0,
                          uint32_t x1, uint32_t x3, uint32_t x5, uint32_t x7, uint32_t y) {
    asm("add.cc.u32 %0, %9, %10;\n\t"
        "madc.lo.cc.u32 %1, %18, %22, %11;\n\t"
        "madc.hi.cc.u32 %2, %18, %22, %12;\n\t"
        "madc.lo.cc.u32 %3, %19, %22, %13;\n\t"
        "madc.hi.cc.u32 %4, %19, %22, %14;\n\t"
        "madc.lo.cc.u32 %5, %20, %22, %15;\n\t"
        "madc.hi.cc.u32 %6, %20, %22, %16;\n\t"
        "madc.lo.cc.u32 %7, %21, %22, %17;\n\t"
        "madc.hi.u32 %8, %21, %22, 0;\n\t"
        : "=&r"(ne0), "=&r"(no[0]), "=&r"(no[1]), "=&r"(no[2]), "=&r"(no[3]), "=&r"(no[4]),
          "=&r"(no[5]), "=&r"(no[6]), "=&r"(no[7])
        : "r"(o0), "r"(e[1]), "r"(e[2]), "r"(e[3]), "r"(e[4]), "r"(e[5]), "r"(e[6]), "r"(e[7]),
          "r"(e[8]), "r"(x1), "r"(x3), "r"(x5), "r"(x7), "r"(y));
}

template <class PRM>
B2_D void mont_mul_dev(uint32_t* r, const uint32_t* a, const uint32_t* b) {
    uint32_t e[9], o[8];
    // row 0: plain products
    {
        const uint32_t y = b[0];
#pragma unroll
        for (int j = 0; j < 8; j += 2) {
            asm("mul.lo.u32 %0, %2, %3;\n\tmul.hi.u32 %1, %2, %3;" : "=r"(e[j]), "=r"(e[j + 1]) : "r"(a[j]), "r"(y));
            asm("mul.lo.u32 %0, %2, %3;\n\tmul.hi.u32 %1, %2, %3;" : "=r"(o[j]), "=r"(o[j + 1]) : "r"(a[j + 1]), "r"(y));
        }
        const uint32_t m = e[0] * PRM::inv_dev();
        mad_chain8(o, PRM::p(1), PRM::p(3), PRM::p(5), PRM::p(7), m);
        mad_chain8_top<true>(e, PRM::p(0), PRM::p(2), PRM::p(4), PRM::p(6), m);
    }
#pragma unroll
    for (int i = 1; i < 8; ++i) {
        const uint32_t y = b[i];
        uint32_t ne[9], no[8];
        fold_mad_shift8(ne[0], no, e, o[0], a[1], a[3], a[5], a[7], y);
#pragma unroll
        for (int k = 1; k < 8; ++k) ne[k] = o[k];
        mad_chain8_top<true>(ne, a[0], a[2], a[4], a[6], y);
        const uint32_t m = ne[0] * PRM::inv_dev();
        mad_chain8(no, PRM::p(1), PRM::p(3), PRM::p(5), PRM::p(7), m);
        mad_chain8_top<false>(ne, PRM::p(0), PRM::p(2), PRM::p(4), PRM::p(6), m);
#pragma unroll
        for (int k = 0; k < 9; ++k) e[k] = ne[k];
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = no[k];
    }
    // T = (E >> 32) + O  (< 2p), then one conditional subtraction
    uint32_t t[8];
    asm("add.cc.u32 %0, %8, %16;\n\t"
        "addc.cc.u32 %1, %9, %17;\n\t"
        "addc.cc.u32 %2, %10, %18;\n\t"
        "addc.cc.u32 %3, %11, %19;\n\t"
        "addc.cc.u32 %4, %12, %20;\n\t"
        "addc.cc.u32 %5, %13, %21;\n\t"
        "addc.cc.u32 %6, %14, %22;\n\t"
        "addc.u32 %7, %15, %23;\n\t"
        : "=&r"(t[0]), "=&r"(t[1]), "=&r"(t[2]), "=&r"(t[3]), "=&r"(t[4]), "=&r"(t[5]), "=&r"(t[6]), "=&r"(t[7])
        : "r"(o[0]), "r"(o[1]), "r"(o[2]), "r"(o[3]), "r"(o[4]), "r"(o[5]), "r"(o[6]), "r"(o[7]),
          "r"(e[1]), "r"(e[2]), "r"(e[3]), "r"(e[4]), "r"(e[5]), "r"(e[6]), "r"(e[7]), "r"(e[8]));
    uint32_t s[8], borrow;
    asm("sub.cc.u32 %0, %9, %17;\n\t"
        "subc.cc.u32 %1, %10, %18;\n\t"
        "subc.cc.u32 %2, %11, %19;\n\t"
        "subc.cc.u32 %3, %12, %20;\n\t"
        "subc.cc.u32 %4, %13, %21;\n\t"
        "subc.cc.u32 %5, %14, %22;\n\t"
        "subc.cc.u32 %6, %15, %23;\n\t"
        "subc.cc.u32 %7, %16, %24;\n\t"
        "subc.u32 %8, 0, 0;\n\t"
        : "=&r"(s[0]), "=&r"(s[1]), "=&r"(s[2]), "=&r"(s[3]), "=&r"(s[4]), "=&r"(s[5]), "=&r"(s[6]), "=&r"(s[7]), "=&r"(borrow)
        : "r"(t[0]), "r"(t[1]), "r"(t[2]), "r"(t[3]), "r"(t[4]), "r"(t[5]), "r"(t[6]), "r"(t[7]),
          "r"(PRM::p(0)), "r"(PRM::p(1)), "r"(PRM::p(2)), "r"(PRM::p(3)), "r"(PRM::p(4)), "r"(PRM::p(5)), "r"(PRM::p(6)), "r"(PRM::p(7)));
#pragma unroll
    for (int k = 0; k < 8; ++k) r[k] = borrow ? t[k] : s[k];
}

// ---------------------------------------------------------------------------------------
// Product / reduction split (lazy reduction in Fp2: three wide products, two reductions).
// ---------------------------------------------------------------------------------------

// t[0..15] = a * b for a, b < 2^255 (same even/odd row chains as mont_mul_dev, without the m*p rows:
// the low word of the window is an output word instead of being cancelled).  64 MAC32.
B2_D void mul_wide_dev(uint32_t* t, const uint32_t* a, const uint32_t* b) {
    uint32_t e[9], o[8];
    {
        const uint32_t y = b[0];
#pragma unroll
        for (int j = 0; j < 8; j += 2) {
            asm("mul.lo.u32 %0, %2, %3;\n\tmul.hi.u32 %1, %2, %3;" : "=r"(e[j]), "=r"(e[j + 1]) : "r"(a[j]), "r"(y));
            asm("mul.lo.u32 %0, %2, %3;\n\tmul.hi.u32 %1, %2, %3;" : "=r"(o[j]), "=r"(o[j + 1]) : "r"(a[j + 1]), "r"(y));
        }
        e[8] = 0;
        t[0] = e[0];
    }
#pragma unroll
    for (int i = 1; i < 8; ++i) {
        const uint32_t y = b[i];
        uint32_t ne[9], no[8];
        fold_mad_shift8(ne[0], no, e, o[0], a[1], a[3], a[5], a[7], y);
#pragma unroll
        for (int k = 1; k < 8; ++k) ne[k] = o[k];
        mad_chain8_top<true>(ne, a[0], a[2], a[4], a[6], y);
        t[i] = ne[0];
#pragma unroll
        for (int k = 0; k < 9; ++k) e[k] = ne[k];
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = no[k];
    }
    asm("add.cc.u32 %0, %8, %16;\n\t"
        "addc.cc.u32 %1, %9, %17;\n\t"
        "addc.cc.u32 %2, %10, %18;\n\t"
        "addc.cc.u32 %3, %11, %19;\n\t"
        "addc.cc.u32 %4, %12, %20;\n\t"
        "addc.cc.u32 %5, %13, %21;\n\t"
        "addc.cc.u32 %6, %14, %22;\n\t"
        "addc.u32 %7, %15, %23;\n\t"
        : "=&r"(t[8]), "=&r"(t[9]), "=&r"(t[10]), "=&r"(t[11]), "=&r"(t[12]), "=&r"(t[13]), "=&r"(t[14]), "=&r"(t[15])
        : "r"(o[0]), "r"(o[1]), "r"(o[2]), "r"(o[3]), "r"(o[4]), "r"(o[5]), "r"(o[6]), "r"(o[7]),
          "r"(e[1]), "r"(e[2]), "r"(e[3]), "r"(e[4]), "r"(e[5]), "r"(e[6]), "r"(e[7]), "r"(e[8]));
}

// 16-word r = a + b and r = a - b (no carry / borrow leaves the top: caller's bound).  Two asm
// blocks of 8 words each; the carry crosses them in a register and is re-created with one add.
B2_D void add16_dev(uint32_t* r, const uint32_t* a, const uint32_t* b) {
    uint32_t c;
    asm("add.cc.u32 %0, %9, %17;\n\taddc.cc.u32 %1, %10, %18;\n\taddc.cc.u32 %2, %11, %19;\n\taddc.cc.u32 %3, %12, %20;\n\t"
        "addc.cc.u32 %4, %13, %21;\n\taddc.cc.u32 %5, %14, %22;\n\taddc.cc.u32 %6, %15, %23;\n\taddc.cc.u32 %7, %16, %24;\n\t"
        "addc.u32 %8, 0, 0;\n\t"
        : "=&r"(r[0]), "=&r"(r[1]), "=&r"(r[2]), "=&r"(r[3]), "=&r"(r[4]), "=&r"(r[5]), "=&r"(r[6]), "=&r"(r[7]), "=&r"(c)
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(a[7]),
          "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3]), "r"(b[4]), "r"(b[5]), "r"(b[6]), "r"(b[7]));
    uint32_t junk;   // the carry-recreating add needs a destination
    (void)junk;
    asm("add.cc.u32 %8, %25, 0xffffffff;\n\t"
        "addc.cc.u32 %0, %9, %17;\n\taddc.cc.u32 %1, %10, %18;\n\taddc.cc.u32 %2, %11, %19;\n\taddc.cc.u32 %3, %12, %20;\n\t"
        "addc.cc.u32 %4, %13, %21;\n\taddc.cc.u32 %5, %14, %22;\n\taddc.cc.u32 %6, %15, %23;\n\taddc.u32 %7, %16, %24;\n\t"
        : "=&r"(r[8]), "=&r"(r[9]), "=&r"(r[10]), "=&r"(r[11]), "=&r"(r[12]), "=&r"(r[13]), "=&r"(r[14]), "=&r"(r[15]), "=&r"(junk)
        : "r"(a[8]), "r"(a[9]), "r"(a[10]), "r"(a[11]), "r"(a[12]), "r"(a[13]), "r"(a[14]), "r"(a[15]),
          "r"(b[8]), "r"(b[9]), "r"(b[10]), "r"(b[11]), "r"(b[12]), "r"(b[13]), "r"(b[14]), "r"(b[15]), "r"(c));
}
B2_D void sub16_dev(uint32_t* r, const uint32_t* a, const uint32_t* b) {
    uint32_t c;
    asm("sub.cc.u32 %0, %9, %17;\n\tsubc.cc.u32 %1, %10, %18;\n\tsubc.cc.u32 %2, %11, %19;\n\tsubc.cc.u32 %3, %12, %20;\n\t"
        "subc.cc.u32 %4, %13, %21;\n\tsubc.cc.u32 %5, %14, %22;\n\tsubc.cc.u32 %6, %15, %23;\n\tsubc.cc.u32 %7, %16, %24;\n\t"
        "subc.u32 %8, 0, 0;\n\t"
        : "=&r"(r[0]), "=&r"(r[1]), "=&r"(r[2]), "=&r"(r[3]), "=&r"(r[4]), "=&r"(r[5]), "=&r"(r[6]), "=&r"(r[7]), "=&r"(c)
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(a[7]),
          "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3]), "r"(b[4]), "r"(b[5]), "r"(b[6]), "r"(b[7]));
    // c = 0 or 0xffffffff: 0 - (c & 1) borrows exactly when the low half borrowed
    const uint32_t bflag = c & 1u;
    uint32_t junk;   // the carry-recreating add needs a destination
    (void)junk;
    asm("sub.cc.u32 %8, 0, %25;\n\t"
        "subc.cc.u32 %0, %9, %17;\n\tsubc.cc.u32 %1, %10, %18;\n\tsubc.cc.u32 %2, %11, %19;\n\tsubc.cc.u32 %3, %12, %20;\n\t"
        "subc.cc.u32 %4, %13, %21;\n\tsubc.cc.u32 %5, %14, %22;\n\tsubc.cc.u32 %6, %15, %23;\n\tsubc.u32 %7, %16, %24;\n\t"
        : "=&r"(r[8]), "=&r"(r[9]), "=&r"(r[10]), "=&r"(r[11]), "=&r"(r[12]), "=&r"(r[13]), "=&r"(r[14]), "=&r"(r[15]), "=&r"(junk)
        : "r"(a[8]), "r"(a[9]), "r"(a[10]), "r"(a[11]), "r"(a[12]), "r"(a[13]), "r"(a[14]), "r"(a[15]),
          "r"(b[8]), "r"(b[9]), "r"(b[10]), "r"(b[11]), "r"(b[12]), "r"(b[13]), "r"(b[14]), "r"(b[15]), "r"(bflag));
}
// 8-word r = a + b without reduction (a, b < p: the sum is below 2^255)
B2_D void add8_raw_dev(uint32_t* r, const uint32_t* a, const uint32_t* b) {
    asm("add.cc.u32 %0, %8, %16;\n\taddc.cc.u32 %1, %9, %17;\n\taddc.cc.u32 %2, %10, %18;\n\taddc.cc.u32 %3, %11, %19;\n\t"
        "addc.cc.u32 %4, %12, %20;\n\taddc.cc.u32 %5, %13, %21;\n\taddc.cc.u32 %6, %14, %22;\n\taddc.u32 %7, %15, %23;\n\t"
        : "=&r"(r[0]), "=&r"(r[1]), "=&r"(r[2]), "=&r"(r[3]), "=&r"(r[4]), "=&r"(r[5]), "=&r"(r[6]), "=&r"(r[7])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(a[7]),
          "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3]), "r"(b[4]), "r"(b[5]), "r"(b[6]), "r"(b[7]));
}

// parity swap of the reduction window + the next Montgomery factor + its odd row, one carry chain:
//   ne0 = o0 + e[1];  m = ne0 * INV;  no[0..7] = (e[2..8], 0) + {p1,p3,p5,p7} * m + carry
template <class PRM>
B2_D void fold_reduce_shift8(uint32_t& ne0, uint32_t& m, uint32_t* no, const uint32_t* e, uint32_t o0) {
    asm("add.cc.u32 %0, %10, %11;\n\t"
        "mul.lo.u32 %9, %0, %23;\n\t"
        "madc.lo.cc.u32 %1, %19, %9, %12;\n\t"
        "madc.hi.cc.u32 %2, %19, %9, %13;\n\t"
        "madc.lo.cc.u32 %3, %20, %9, %14;\n\t"
        "madc.hi.cc.u32 %4, %20, %9, %15;\n\t"
        "madc.lo.cc.u32 %5, %21, %9, %16;\n\t"
        "madc.hi.cc.u32 %6, %21, %9, %17;\n\t"
        "madc.lo.cc.u32 %7, %22, %9, %18;\n\t"
        "madc.hi.u32 %8, %22, %9, 0;\n\t"
        : "=&r"(ne0), "=&r"(no[0]), "=&r"(no[1]), "=&r"(no[2]), "=&r"(no[3]), "=&r"(no[4]),
          "=&r"(no[5]), "=&r"(no[6]), "=&r"(no[7]), "=&r"(m)
        : "r"(o0), "r"(e[1]), "r"(e[2]), "r"(e[3]), "r"(e[4]), "r"(e[5]), "r"(e[6]), "r"(e[7]),
          "r"(e[8]), "r"(PRM::p(1)), "r"(PRM::p(3)), "r"(PRM::p(5)), "r"(PRM::p(7)), "r"(PRM::inv_dev()));
}

// r = t / 2^256 mod p for t = t_lo + 2^256 t_hi < p * 2^256  (r < p)
template <class PRM>
B2_D void mont_reduce_dev(uint32_t* r, const uint32_t* t) {
    uint32_t e[9], o[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) e[k] = t[k];
    {
        const uint32_t m = e[0] * PRM::inv_dev();
#pragma unroll
        for (int j = 0; j < 8; j += 2)
            asm("mul.lo.u32 %0, %2, %3;\n\tmul.hi.u32 %1, %2, %3;" : "=r"(o[j]), "=r"(o[j + 1]) : "r"(PRM::p(j + 1)), "r"(m));
        mad_chain8_top<true>(e, PRM::p(0), PRM::p(2), PRM::p(4), PRM::p(6), m);
    }
#pragma unroll
    for (int i = 1; i < 8; ++i) {
        uint32_t ne[9], no[8], m;
        fold_reduce_shift8<PRM>(ne[0], m, no, e, o[0]);
#pragma unroll
        for (int k = 1; k < 8; ++k) ne[k] = o[k];
        mad_chain8_top<true>(ne, PRM::p(0), PRM::p(2), PRM::p(4), PRM::p(6), m);
#pragma unroll
        for (int k = 0; k < 9; ++k) e[k] = ne[k];
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = no[k];
    }
    // u = (E >> 32) + O <= p;  v = u + t_hi < 2p + 1;  one conditional subtraction
    uint32_t u[8];
    asm("add.cc.u32 %0, %8, %16;\n\t"
        "addc.cc.u32 %1, %9, %17;\n\t"
        "addc.cc.u32 %2, %10, %18;\n\t"
        "addc.cc.u32 %3, %11, %19;\n\t"
        "addc.cc.u32 %4, %12, %20;\n\t"
        "addc.cc.u32 %5, %13, %21;\n\t"
        "addc.cc.u32 %6, %14, %22;\n\t"
        "addc.u32 %7, %15, %23;\n\t"
        : "=&r"(u[0]), "=&r"(u[1]), "=&r"(u[2]), "=&r"(u[3]), "=&r"(u[4]), "=&r"(u[5]), "=&r"(u[6]), "=&r"(u[7])
        : "r"(o[0]), "r"(o[1]), "r"(o[2]), "r"(o[3]), "r"(o[4]), "r"(o[5]), "r"(o[6]), "r"(o[7]),
          "r"(e[1]), "r"(e[2]), "r"(e[3]), "r"(e[4]), "r"(e[5]), "r"(e[6]), "r"(e[7]), "r"(e[8]));
    uint32_t v[8];
    asm("add.cc.u32 %0, %8, %16;\n\t"
        "addc.cc.u32 %1, %9, %17;\n\t"
        "addc.cc.u32 %2, %10, %18;\n\t"
        "addc.cc.u32 %3, %11, %19;\n\t"
        "addc.cc.u32 %4, %12, %20;\n\t"
        "addc.cc.u32 %5, %13, %21;\n\t"
        "addc.cc.u32 %6, %14, %22;\n\t"
        "addc.u32 %7, %15, %23;\n\t"
        : "=&r"(v[0]), "=&r"(v[1]), "=&r"(v[2]), "=&r"(v[3]), "=&r"(v[4]), "=&r"(v[5]), "=&r"(v[6]), "=&r"(v[7])
        : "r"(u[0]), "r"(u[1]), "r"(u[2]), "r"(u[3]), "r"(u[4]), "r"(u[5]), "r"(u[6]), "r"(u[7]),
          "r"(t[8]), "r"(t[9]), "r"(t[10]), "r"(t[11]), "r"(t[12]), "r"(t[13]), "r"(t[14]), "r"(t[15]));
    uint32_t s[8], borrow;
    asm("sub.cc.u32 %0, %9, %17;\n\t"
        "subc.cc.u32 %1, %10, %18;\n\t"
        "subc.cc.u32 %2, %11, %19;\n\t"
        "subc.cc.u32 %3, %12, %20;\n\t"
        "subc.cc.u32 %4, %13, %21;\n\t"
        "subc.cc.u32 %5, %14, %22;\n\t"
        "subc.cc.u32 %6, %15, %23;\n\t"
        "subc.cc.u32 %7, %16, %24;\n\t"
        "subc.u32 %8, 0, 0;\n\t"
        : "=&r"(s[0]), "=&r"(s[1]), "=&r"(s[2]), "=&r"(s[3]), "=&r"(s[4]), "=&r"(s[5]), "=&r"(s[6]), "=&r"(s[7]), "=&r"(borrow)
        : "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
          "r"(PRM::p(0)), "r"(PRM::p(1)), "r"(PRM::p(2)), "r"(PRM::p(3)), "r"(PRM::p(4)), "r"(PRM::p(5)), "r"(PRM::p(6)), "r"(PRM::p(7)));
#pragma unroll
    for (int k = 0; k < 8; ++k) r[k] = borrow ? v[k] : s[k];
}

template <class PRM>
B2_D void mod_add_dev(uint32_t* r, const uint32_t* a, const uint32_t* b) {
    uint32_t t[8];
    asm("add.cc.u32 %0, %8, %16;\n\t"
        "addc.cc.u32 %1, %9, %17;\n\t"
        "addc.cc.u32 %2, %10, %18;\n\t"
        "addc.cc.u32 %3, %11, %19;\n\t"
        "addc.cc.u32 %4, %12, %20;\n\t"
        "addc.cc.u32 %5, %13, %21;\n\t"
        "addc.cc.u32 %6, %14, %22;\n\t"
        "addc.u32 %7, %15, %23;\n\t"
        : "=&r"(t[0]), "=&r"(t[1]), "=&r"(t[2]), "=&r"(t[3]), "=&r"(t[4]), "=&r"(t[5]), "=&r"(t[6]), "=&r"(t[7])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(a[7]),
          "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3]), "r"(b[4]), "r"(b[5]), "r"(b[6]), "r"(b[7]));
    // a + b < 2p < 2^255: no carry out of limb 7
    uint32_t s[8], borrow;
    asm("sub.cc.u32 %0, %9, %17;\n\t"
        "subc.cc.u32 %1, %10, %18;\n\t"
        "subc.cc.u32 %2, %11, %19;\n\t"
        "subc.cc.u32 %3, %12, %20;\n\t"
        "subc.cc.u32 %4, %13, %21;\n\t"
        "subc.cc.u32 %5, %14, %22;\n\t"
        "subc.cc.u32 %6, %15, %23;\n\t"
        "subc.cc.u32 %7, %16, %24;\n\t"
        "subc.u32 %8, 0, 0;\n\t"
        : "=&r"(s[0]), "=&r"(s[1]), "=&r"(s[2]), "=&r"(s[3]), "=&r"(s[4]), "=&r"(s[5]), "=&r"(s[6]), "=&r"(s[7]), "=&r"(borrow)
        : "r"(t[0]), "r"(t[1]), "r"(t[2]), "r"(t[3]), "r"(t[4]), "r"(t[5]), "r"(t[6]), "r"(t[7]),
          "r"(PRM::p(0)), "r"(PRM::p(1)), "r"(PRM::p(2)), "r"(PRM::p(3)), "r"(PRM::p(4)), "r"(PRM::p(5)), "r"(PRM::p(6)), "r"(PRM::p(7)));
#pragma unroll
    for (int k = 0; k < 8; ++k) r[k] = borrow ? t[k] : s[k];
}

template <class PRM>
B2_D void mod_sub_dev(uint32_t* r, const uint32_t* a, const uint32_t* b) {
    uint32_t t[8], borrow;
    asm("sub.cc.u32 %0, %9, %17;\n\t"
        "subc.cc.u32 %1, %10, %18;\n\t"
        "subc.cc.u32 %2, %11, %19;\n\t"
        "subc.cc.u32 %3, %12, %20;\n\t"
        "subc.cc.u32 %4, %13, %21;\n\t"
        "subc.cc.u32 %5, %14, %22;\n\t"
        "subc.cc.u32 %6, %15, %23;\n\t"
        "subc.cc.u32 %7, %16, %24;\n\t"
        "subc.u32 %8, 0, 0;\n\t"
        : "=&r"(t[0]), "=&r"(t[1]), "=&r"(t[2]), "=&r"(t[3]), "=&r"(t[4]), "=&r"(t[5]), "=&r"(t[6]), "=&r"(t[7]), "=&r"(borrow)
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(a[4]), "r"(a[5]), "r"(a[6]), "r"(a[7]),
          "r"(b[0]), "r"(b[1]), "r"(b[2]), "r"(b[3]), "r"(b[4]), "r"(b[5]), "r"(b[6]), "r"(b[7]));
    // borrow is 0 or 0xffffffff: add (p & borrow)
    asm("add.cc.u32 %0, %0, %8;\n\t"
        "addc.cc.u32 %1, %1, %9;\n\t"
        "addc.cc.u32 %2, %2, %10;\n\t"
        "addc.cc.u32 %3, %3, %11;\n\t"
        "addc.cc.u32 %4, %4, %12;\n\t"
        "addc.cc.u32 %5, %5, %13;\n\t"
        "addc.cc.u32 %6, %6, %14;\n\t"
        "addc.u32 %7, %7, %15;\n\t"
        : "+r"(t[0]), "+r"(t[1]), "+r"(t[2]), "+r"(t[3]), "+r"(t[4]), "+r"(t[5]), "+r"(t[6]), "+r"(t[7])
        : "r"(PRM::p(0) & borrow), "r"(PRM::p(1) & borrow), "r"(PRM::p(2) & borrow), "r"(PRM::p(3) & borrow),
          "r"(PRM::p(4) & borrow), "r"(PRM::p(5) & borrow), "r"(PRM::p(6) & borrow), "r"(PRM::p(7) & borrow));
#pragma unroll
    for (int k = 0; k < 8; ++k) r[k] = t[k];
}
#if defined(B2_CURVE_BLS12_377)
#include "field_chains12.cuh"   // 12-limb copies of the chains above (BLS12-377 Fp), scripts/gen_field_chains.py
#endif
#endif  // __CUDACC__

// ---------------------------------------------------------------------------------------
// Portable path (host): same results, u64 arithmetic.
// ---------------------------------------------------------------------------------------
template <class PRM>
B2_HD void mont_mul_host(uint32_t* r, const uint32_t* a, const uint32_t* b) {
    constexpr int N = PRM::N;
    uint32_t t[N + 2];
    for (int i = 0; i < N + 2; ++i) t[i] = 0;
    for (int i = 0; i < N; ++i) {
        uint64_t c = 0;
        for (int j = 0; j < N; ++j) {
            c += (uint64_t)a[j] * b[i] + t[j];
            t[j] = (uint32_t)c; c >>= 32;
        }
        c += t[N]; t[N] = (uint32_t)c; t[N + 1] = (uint32_t)(c >> 32);
        const uint32_t m = t[0] * PRM::INV;
        c = ((uint64_t)m * PRM::p(0) + t[0]) >> 32;
        for (int j = 1; j < N; ++j) {
            c += (uint64_t)m * PRM::p(j) + t[j];
            t[j - 1] = (uint32_t)c; c >>= 32;
        }
        c += t[N]; t[N - 1] = (uint32_t)c; t[N] = t[N + 1] + (uint32_t)(c >> 32);
    }
    uint32_t s[N]; uint64_t br = 0;
    for (int j = 0; j < N; ++j) {
        uint64_t d = (uint64_t)t[j] - PRM::p(j) - br;
        s[j] = (uint32_t)d; br = (d >> 32) & 1;
    }
    const bool ge = t[N] || !br;
    for (int j = 0; j < N; ++j) r[j] = ge ? s[j] : t[j];
}

#if !defined(__CUDA_ARCH__)
// host callers go through one out-of-line copy per field: the group formulas inline dozens of products, and
// gcc -O3 on a fully inlined 12-limb XYZZ addition takes minutes
template <class PRM>
#if defined(__GNUC__)
__attribute__((noinline))
#endif
void mont_mul_host_outlined(uint32_t* r, const uint32_t* a, const uint32_t* b) { mont_mul_host<PRM>(r, a, b); }
#endif

template <class PRM>
B2_HD void mod_add_host(uint32_t* r, const uint32_t* a, const uint32_t* b) {
    constexpr int N = PRM::N;
    uint32_t t[N], s[N]; uint64_t c = 0, br = 0;
    for (int j = 0; j < N; ++j) { c += (uint64_t)a[j] + b[j]; t[j] = (uint32_t)c; c >>= 32; }
    for (int j = 0; j < N; ++j) { uint64_t d = (uint64_t)t[j] - PRM::p(j) - br; s[j] = (uint32_t)d; br = (d >> 32) & 1; }
    const bool ge = c || !br;
    for (int j = 0; j < N; ++j) r[j] = ge ? s[j] : t[j];
}

template <class PRM>
B2_HD void mod_sub_host(uint32_t* r, const uint32_t* a, const uint32_t* b) {
    constexpr int N = PRM::N;
    uint32_t t[N]; uint64_t br = 0;
    for (int j = 0; j < N; ++j) { uint64_t d = (uint64_t)a[j] - b[j] - br; t[j] = (uint32_t)d; br = (d >> 32) & 1; }
    if (br) { uint64_t c = 0; for (int j = 0; j < N; ++j) { c += (uint64_t)t[j] + PRM::p(j); t[j] = (uint32_t)c; c >>= 32; } }
    for (int j = 0; j < N; ++j) r[j] = t[j];
}

// ---------------------------------------------------------------------------------------
// Field element
// ---------------------------------------------------------------------------------------
template <class PRM>
struct alignas(16) Fe {
    static constexpr int N = PRM::N;
    typedef PRM Params;
    uint32_t l[N];

    B2_HD static Fe zero() { Fe r; for (int i = 0; i < N; ++i) r.l[i] = 0; return r; }
    B2_HD static Fe one() { Fe r; for (int i = 0; i < N; ++i) r.l[i] = PRM::one(i); return r; }
    B2_HD static Fe r2() { Fe r; for (int i = 0; i < N; ++i) r.l[i] = PRM::r2(i); return r; }
    B2_HD static Fe modulus() { Fe r; for (int i = 0; i < N; ++i) r.l[i] = PRM::p(i); return r; }
    B2_HD bool is_zero() const { uint32_t x = 0; for (int i = 0; i < N; ++i) x |= l[i]; return x == 0; }
    B2_HD bool operator==(const Fe& o) const { uint32_t x = 0; for (int i = 0; i < N; ++i) x |= l[i] ^ o.l[i]; return x == 0; }
    B2_HD bool operator!=(const Fe& o) const { return !(*this == o); }

    B2_HD Fe operator*(const Fe& o) const {
        Fe r;
#if defined(__CUDA_ARCH__)
        if constexpr (N == 8) mont_mul_dev<PRM>(r.l, l, o.l);
#if defined(B2_CURVE_BLS12_377)
        else if constexpr (N == 12) mont_mul_dev12<PRM>(r.l, l, o.l);
#endif
        else mont_mul_host<PRM>(r.l, l, o.l);
#else
        mont_mul_host_outlined<PRM>(r.l, l, o.l);
#endif
        return r;
    }
    B2_HD Fe sqr() const { return *this * *this; }
    B2_HD Fe operator+(const Fe& o) const {
        Fe r;
#if defined(__CUDA_ARCH__)
        if constexpr (N == 8) mod_add_dev<PRM>(r.l, l, o.l);
#if defined(B2_CURVE_BLS12_377)
        else if constexpr (N == 12) mod_add_dev12<PRM>(r.l, l, o.l);
#endif
        else mod_add_host<PRM>(r.l, l, o.l);
#else
        mod_add_host<PRM>(r.l, l, o.l);
#endif
        return r;
    }
    B2_HD Fe operator-(const Fe& o) const {
        Fe r;
#if defined(__CUDA_ARCH__)
        if constexpr (N == 8) mod_sub_dev<PRM>(r.l, l, o.l);
#if defined(B2_CURVE_BLS12_377)
        else if constexpr (N == 12) mod_sub_dev12<PRM>(r.l, l, o.l);
#endif
        else mod_sub_host<PRM>(r.l, l, o.l);
#else
        mod_sub_host<PRM>(r.l, l, o.l);
#endif
        return r;
    }
    B2_HD Fe dbl() const { return *this + *this; }
    B2_HD Fe neg() const { return zero() - *this; }
    // Montgomery form <-> canonical integer
    B2_HD Fe to_mont() const { return *this * r2(); }
    B2_HD Fe from_mont() const { Fe o = zero(); o.l[0] = 1; return *this * o; }
    B2_HD static Fe from_u32(uint32_t x) { Fe r = zero(); r.l[0] = x; return r.to_mont(); }

    // x^e for a little-endian u32 exponent (square-and-multiply, MSB first)
    B2_HD Fe pow(const uint32_t* e, int nlimbs) const {
        Fe acc = one();
        for (int i = nlimbs * 32 - 1; i >= 0; --i) {
            acc = acc.sqr();
            if ((e[i >> 5] >> (i & 31)) & 1) acc = acc * *this;
        }
        return acc;
    }
    B2_HD Fe pow_u64(uint64_t e) const {
        uint32_t w[2] = {(uint32_t)e, (uint32_t)(e >> 32)};
        return pow(w, 2);
    }
    // Fermat inverse x^(p-2); 0 -> 0 (same convention as gnark-crypto's Inverse)
    B2_HD Fe inv() const {
        uint32_t e[N];
        uint32_t borrow = 2;   // p - 2 with the borrow carried up (both BLS12-377 moduli end in ...00000001)
        for (int i = 0; i < N; ++i) {
            const uint32_t v = PRM::p(i);
            e[i] = v - borrow;
            borrow = v < borrow ? 1u : 0u;
        }
        return pow(e, N);
    }
};

typedef Fe<FpParams> Fp;
typedef Fe<FrParams> Fr;

// a*b - c*d with one Montgomery reduction on the device (two wide products, their difference kept
// non-negative by adding p^2): 200 MAC32 instead of 272.  Canonical result either way.
B2_HD Fp mul_sub_mul(const Fp& a, const Fp& b, const Fp& c, const Fp& d) {
#if defined(__CUDA_ARCH__) && defined(B2_CURVE_BLS12_377)
    // 2 x 144 + 156 MAC32 instead of 2 x 300
    uint32_t t0[24], t1[24], p2[24];
    mul_wide_dev12(t0, a.l, b.l);
    mul_wide_dev12(t1, c.l, d.l);
#pragma unroll
    for (int i = 0; i < 24; ++i) p2[i] = FpParams::p_squared(i);
    add24_dev(t0, t0, p2);
    sub24_dev(t0, t0, t1);
    Fp r;
    mont_reduce_dev12<FpParams>(r.l, t0);
    return r;
#elif defined(__CUDA_ARCH__)
    uint32_t t0[16], t1[16], p2[16];
    mul_wide_dev(t0, a.l, b.l);
    mul_wide_dev(t1, c.l, d.l);
#pragma unroll
    for (int i = 0; i < 16; ++i) p2[i] = FpParams::p_squared(i);
    add16_dev(t0, t0, p2);
    sub16_dev(t0, t0, t1);
    Fp r;
    mont_reduce_dev<FpParams>(r.l, t0);
    return r;
#else
    return a * b - c * d;
#endif
}

// Out-of-line product (operands and result travel in registers).  Used by the Fp2 tower: a fully
// inlined G2 mixed addition is 28 Fp products = ~90 KB of SASS, far beyond the instruction cache
// (ncu: "no instruction" was the top stall of msm_accumulate_kernel<Fp2>), and 250 registers.
#if defined(__CUDACC__)
template <class PRM>
__device__ __noinline__ Fe<PRM> mont_mul_call(Fe<PRM> a, Fe<PRM> b) {
    return a * b;
}
#endif
template <class PRM>
B2_HD Fe<PRM> mul_outlined(const Fe<PRM>& a, const Fe<PRM>& b) {
#if defined(__CUDA_ARCH__)
    return mont_mul_call<PRM>(a, b);
#else
    return a * b;
#endif
}

}  // namespace b200
