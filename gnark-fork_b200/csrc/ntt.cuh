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

// ---------------------------------------------------------------------------------------
// Distributed transform over R = 2^lg ranks (one exchange step per transform).
//
// Every rank holds n/R elements.  Two layouts of the n positions of the in-place network:
//   A (cyclic): rank = position mod R,          local index = position >> lg
//   B (block):  rank = position >> (log_n - lg), local index = position mod n/R
// A DIF transform (natural -> bit-reversed) starts in A and ends in B, a DIT transform
// (bit-reversed -> natural) starts in B and ends in A:
//   DIF: [local]  the log_n - lg stages on position bits >= lg          (in place, layout A)
//        [all-to-all]  chunk q = local indices with top lg bits q goes to rank q
//        [cross]  the lg stages on position bits < lg, over the received (r, j) buffer; the store
//                 rotates the index to (j << lg) | r: layout B in natural local order.
//   DIT: [local]  the log_n - lg stages on position bits < log_n - lg   (layout B); the last
//                 store rotates the index to ((i mod R) << m) | (i >> lg) so that the chunks of the
//        [all-to-all]  are contiguous: chunk q = local indices congruent to q mod R
//        [cross]  the lg stages on the top position bits over the received (r, j) buffer, whose
//                 index (r << m) | j is already the local index of layout A.          m = log_n - 2 lg
// Twiddles and scale tables are looked up by GLOBAL position, so the full-size tables serve all
// ranks.  The maps below turn a local buffer index into that position / into the store index.
// ---------------------------------------------------------------------------------------
struct NttIndexMap {   // i -> ((i >> m) << sh_hi) | ((i mod 2^m) << sh_lo) | add
    uint32_t m, sh_hi, sh_lo, add;
};
B2_HD NttIndexMap ntt_map_identity() { return NttIndexMap{31u, 0u, 0u, 0u}; }
B2_HD uint32_t ntt_map(const NttIndexMap& p, uint32_t i) {
    return ((i >> p.m) << p.sh_hi) | ((i & ((1u << p.m) - 1u)) << p.sh_lo) | p.add;
}

constexpr int NTT_MAX_RANKS = 16;   // ranks of a distributed transform whose exchange is fused over peer memory (ntt.cu: NttPeerPull)

enum NttDistPhaseKind { NTT_DIST_PACK = 0, NTT_DIST_DIF_LOCAL, NTT_DIST_DIF_CROSS, NTT_DIST_DIT_LOCAL, NTT_DIST_DIT_CROSS };

struct NttDistPhase {
    NttIndexMap gmap;    // local buffer index -> global position (twiddles, scale tables)
    NttIndexMap smap;    // local buffer index -> index the LAST pass of the phase stores to
    uint32_t s_off;      // global stage number = s_off + stage number of the local plan
    int dit;             // network of the phase's passes
    int cross;           // 1: a single pass over the top lg bits of the buffer
    int out_of_place;    // 1: the last pass stores to the other buffer
};

inline NttDistPhase ntt_dist_phase_desc(uint32_t log_n, uint32_t lg, uint32_t rank, int kind) {
    const uint32_t m = log_n - 2u * lg, log_loc = log_n - lg;
    NttDistPhase d;
    d.gmap = ntt_map_identity();
    d.smap = ntt_map_identity();
    d.s_off = 0; d.dit = 0; d.cross = 0; d.out_of_place = 0;
    switch (kind) {
        case NTT_DIST_PACK:         // layout B (natural block) -> chunks for layout A
            d.smap = NttIndexMap{lg, 0u, m, 0u};
            d.out_of_place = 1;
            break;
        case NTT_DIST_DIF_LOCAL:    // layout A
            d.gmap = NttIndexMap{31u, 0u, lg, rank};
            break;
        case NTT_DIST_DIF_CROSS:    // buffer (r, j): position (((rank << m) | j) << lg) | r
            d.gmap = NttIndexMap{m, 0u, lg, rank << (m + lg)};
            d.smap = NttIndexMap{m, 0u, lg, 0u};
            d.s_off = log_loc; d.cross = 1; d.out_of_place = 1;
            break;
        case NTT_DIST_DIT_LOCAL:    // layout B
            d.gmap = NttIndexMap{31u, 0u, 0u, rank << log_loc};
            d.smap = NttIndexMap{lg, 0u, m, 0u};
            d.dit = 1; d.out_of_place = 1;
            break;
        default:                    // NTT_DIST_DIT_CROSS: buffer (r, j): position (r << log_loc) | (j << lg) | rank
            d.gmap = NttIndexMap{m, log_loc, lg, rank};
            d.s_off = lg; d.dit = 1; d.cross = 1;
            break;
    }
    return d;
}

// the single pass of a cross phase: lg stages over the top lg bits of a buffer of 2^log_loc elements
inline NttPassGeom ntt_dist_cross_geom(uint32_t log_n, uint32_t lg, int dit, uint32_t tl_max) {
    const uint32_t log_loc = log_n - lg, m = log_n - 2u * lg;
    const uint32_t TL = log_loc < tl_max ? log_loc : tl_max;
    return NttPassGeom{log_loc, dit ? m : 0u, lg, TL, TL - lg, m};
}

}  // namespace b200
