// Montgomery multiplication with the 256 x 256 -> 512 bit product on the FP64 pipe.
//
// sm_100 issues IMAD.WIDE at half the plain IMAD rate, so the all-integer CIOS product (136 MAC32,
// field.cuh) keeps one pipe busy for ~544 SMSP cycles per warp-multiplication while the FP64 pipe
// (64 DFMA/clk/SM on B200) and half of the ALU pipe idle.  Here the operand product a*b runs on the
// FP64 pipe (5 x 5 limbs of 52 bits, exact hi/lo halves from two round-toward-zero FMAs per limb
// product), and only the Montgomery reduction (72 MAC32) stays on the IMAD pipe, so the three pipes
// share one multiplication.
//
// Exactness of the limb product (x, y integers < 2^52 held in doubles, P = x*y < 2^104):
//   hi = fma_rz(x, y, 2^104)              = 2^104 + floor(P / 2^52) * 2^52    (ulp of [2^104, 2^105) is 2^52)
//   lo = fma_rz(x, y, (2^104 + 2^52) - hi) = 2^52 + (P mod 2^52)                (the addend is exact; so is the sum)
// The raw bit patterns are then  bits(2^104) + H  and  bits(2^52) + L : integers that add up column
// by column in 64-bit registers, the exponent patterns cancelled by the column's start value.
#pragma once
#include "field.cuh"

namespace b200 {
#if defined(__CUDACC__)

struct Dpf5 { double l[5]; };   // 256-bit integer as five 52-bit limbs (the last holds 48 bits)

B2_D Dpf5 dpf_split(const uint32_t* a) {
    // limb k = bits [52k, 52k+52); as a double: (2^52 + limb) - 2^52 with the limb OR-ed into the mantissa
    const uint32_t lo0 = a[0], hi0 = a[1] & 0xFFFFFu;
    const uint32_t lo1 = __funnelshift_r(a[1], a[2], 20), hi1 = __funnelshift_r(a[2], a[3], 20) & 0xFFFFFu;
    const uint32_t lo2 = __funnelshift_r(a[3], a[4], 8), hi2 = (a[4] >> 8) & 0xFFFFFu;
    const uint32_t lo3 = __funnelshift_r(a[4], a[5], 28), hi3 = __funnelshift_r(a[5], a[6], 28) & 0xFFFFFu;
    const uint32_t lo4 = __funnelshift_r(a[6], a[7], 16), hi4 = a[7] >> 16;
    const double two52 = 4503599627370496.0;
    Dpf5 r;
    r.l[0] = __hiloint2double((int)(hi0 | 0x43300000u), (int)lo0) - two52;
    r.l[1] = __hiloint2double((int)(hi1 | 0x43300000u), (int)lo1) - two52;
    r.l[2] = __hiloint2double((int)(hi2 | 0x43300000u), (int)lo2) - two52;
    r.l[3] = __hiloint2double((int)(hi3 | 0x43300000u), (int)lo3) - two52;
    r.l[4] = __hiloint2double((int)(hi4 | 0x43300000u), (int)lo4) - two52;
    return r;
}

// t[0..15] = a * b
B2_D void dpf_mul_wide(uint32_t* t, const Dpf5& a, const Dpf5& b) {
    constexpr unsigned long long B52 = 0x4330000000000000ull, B104 = 0x4670000000000000ull;
    const double c1 = __longlong_as_double((long long)B104);
    const double c2 = __longlong_as_double((long long)(B104 + 1ull));   // 2^104 + 2^52
    unsigned long long col[10];
#pragma unroll
    for (int k = 0; k < 10; ++k) {
        const unsigned long long n_lo = k <= 4 ? k + 1 : (k <= 8 ? 9 - k : 0);
        const unsigned long long n_hi = k == 0 ? 0 : (k - 1 <= 4 ? k : 10 - k);
        col[k] = 0ull - (n_lo * B52 + n_hi * B104);
    }
#pragma unroll
    for (int i = 0; i < 5; ++i) {
#pragma unroll
        for (int j = 0; j < 5; ++j) {
            const double hi = __fma_rz(a.l[i], b.l[j], c1);
            const double lo = __fma_rz(a.l[i], b.l[j], c2 - hi);
            col[i + j] += (unsigned long long)__double_as_longlong(lo);
            col[i + j + 1] += (unsigned long long)__double_as_longlong(hi);
        }
    }
    // carry-resolve into 52-bit limbs, then repack into 64-bit words
    constexpr unsigned long long M52 = (1ull << 52) - 1ull;
    unsigned long long limb[10], carry = 0;
#pragma unroll
    for (int k = 0; k < 10; ++k) {
        const unsigned long long v = col[k] + carry;
        limb[k] = v & M52;
        carry = v >> 52;
    }
    unsigned long long w[8];
    w[0] = limb[0] | (limb[1] << 52);
    w[1] = (limb[1] >> 12) | (limb[2] << 40);
    w[2] = (limb[2] >> 24) | (limb[3] << 28);
    w[3] = (limb[3] >> 36) | (limb[4] << 16);
    w[4] = (limb[4] >> 48) | (limb[5] << 4) | (limb[6] << 56);
    w[5] = (limb[6] >> 8) | (limb[7] << 44);
    w[6] = (limb[7] >> 20) | (limb[8] << 32);
    w[7] = (limb[8] >> 32) | (limb[9] << 20);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        t[2 * k] = (uint32_t)w[k];
        t[2 * k + 1] = (uint32_t)(w[k] >> 32);
    }
}

template <class PRM>
B2_D void mont_mul_dpf_dev(uint32_t* r, const uint32_t* a, const uint32_t* b) {
    uint32_t t[16];
    dpf_mul_wide(t, dpf_split(a), dpf_split(b));
    mont_reduce_dev<PRM>(r, t);
}

#endif  // __CUDACC__
}  // namespace b200
