// Fr NTT on the device: fft.Domain.FFT / FFTInverse semantics of gnark-crypto as used by
// computeH (/root/reference/backend/groth16/bn254/prove.go:353-396).
//
// The transform is the in-place radix-2 DIF (natural -> bit-reversed) or DIT (bit-reversed ->
// natural) network, exactly the one whose structure is visible in-tree at
// backend/groth16/bn254/mpcsetup/lagrange.go:143-179; it is executed as a few HBM passes, each of
// which stages a tile of 2^TL elements in shared memory and runs S consecutive stages on it.
// No separate bit-reversal pass exists: tiles are chosen so that every pass is in place.
//
// The index helpers below are __host__ __device__ so the CPU unit tests can replay a pass
// sequentially with the very same mapping the kernel uses (tests/cpu_model).
#pragma once
#include "field.cuh"

namespace b200 {

// Geometry of one pass.  A tile owns the elements whose global index has
//   bits [0, LB)       free  (contiguous run of 2^LB elements -> coalesced 2^LB * 32 B segments)
//   bits [LB, L)       fixed by the tile id (low part)
//   bits [L, L+S)      free  (the S index bits the butterflies of this pass couple)
//   bits [L+S, log_n)  fixed by the tile id (high part)
// with LB = TL - S <= L.  Local index = (coupled bits << LB) | low bits.
struct NttPassGeom {
    uint32_t log_n, s0, S, TL, LB, L;
};

B2_HD uint32_t ntt_gidx(const NttPassGeom& g, uint32_t tile, uint32_t loc) {
    const uint32_t lo = loc & ((1u << g.LB) - 1u), hi = loc >> g.LB;
    const uint32_t midbits = g.L - g.LB;
    const uint32_t mid = tile & ((1u << midbits) - 1u), top = tile >> midbits;
    return lo | (mid << g.LB) | (hi << g.L) | (top << (g.L + g.S));
}

// local bit coupled by the k-th stage of the pass
B2_HD uint32_t ntt_stage_bit(const NttPassGeom& g, int dit, uint32_t k) {
    return dit ? g.LB + k : g.TL - 1u - k;
}

// butterfly number -> local index of its lower element (insert a 0 at position `bit`)
B2_HD uint32_t ntt_bf_lower(uint32_t bf, uint32_t bit) {
    const uint32_t m = (1u << bit) - 1u;
    return ((bf & ~m) << 1) | (bf & m);
}

// exponent of the twiddle (power of the domain generator) of the butterfly whose lower element
// has global index g0, at global stage s.
//   DIF stage s: pairs (i, i + h), h = n >> (s+1), twiddle w^((i mod h) << s)
//   DIT stage s: pairs (i, i + h), h = 1 << s,      twiddle w^((i mod h) << (log_n - s - 1))
B2_HD uint32_t ntt_twiddle_exp(uint32_t log_n, int dit, uint32_t s, uint32_t g0) {
    if (dit) {
        const uint32_t h = 1u << s;
        return (g0 & (h - 1u)) << (log_n - s - 1u);
    }
    const uint32_t h = (1u << log_n) >> (s + 1u);
    return (g0 & (h - 1u)) << s;
}

// Pass plan.  tl_max = log2 of the tile (<= 12: 2^12 * 32 B = 128 KiB of shared memory),
// s_strided_max = most stages a strided (non-contiguous) pass may take.
inline int ntt_plan(uint32_t log_n, int dit, uint32_t tl_max, uint32_t s_strided_max, NttPassGeom* out /*[8]*/) {
    int np = 0;
    if (log_n <= tl_max) {
        out[np++] = NttPassGeom{log_n, 0, log_n, log_n, 0, 0};
        return np;
    }
    const uint32_t rem = log_n - tl_max;
    const uint32_t k = (rem + s_strided_max - 1) / s_strided_max;
    uint32_t sizes[8];
    for (uint32_t i = 0; i < k; ++i) sizes[i] = rem / k + (i < rem % k ? 1 : 0);
    if (!dit) {
        uint32_t s0 = 0;
        for (uint32_t i = 0; i < k; ++i) {
            const uint32_t S = sizes[i];
            out[np++] = NttPassGeom{log_n, s0, S, tl_max, tl_max - S, log_n - s0 - S};
            s0 += S;
        }
        out[np++] = NttPassGeom{log_n, s0, tl_max, tl_max, 0, 0};
    } else {
        out[np++] = NttPassGeom{log_n, 0, tl_max, tl_max, 0, 0};
        uint32_t s0 = tl_max;
        for (uint32_t i = 0; i < k; ++i) {
            const uint32_t S = sizes[i];
            out[np++] = NttPassGeom{log_n, s0, S, tl_max, tl_max - S, s0};
            s0 += S;
        }
    }
    return np;
}

}  // namespace b200
