// Signed-digit Pippenger multi-scalar multiplication on the device (G1 and G2 share the code).
//
// Replaces gnark-crypto's G1Jac.MultiExp / G2Jac.MultiExp as called from
// /root/reference/backend/groth16/bn254/prove.go:201,214,234,244,290 (and iciclegnark.MsmOnDevice /
// MsmG2OnDevice, /root/reference/backend/groth16/bn254/icicle/icicle.go:302-382).
//
// Pipeline (all on one stream):
//   1. digits   : scalar out of Montgomery form, c-bit signed-digit recoding, per-window
//                 bucket histogram                                  (msm_digits_kernel)
//   2. scan     : exclusive prefix sum of the W * 2^(c-1) bucket counts (msm_scan_*)
//   3. scatter  : counting sort of (bucket -> point index | sign) per window
//   4. accumulate: one thread per bucket, XYZZ += affine (8M + 2S), doubling / cancellation
//                 handled (DummySetup keys make every base equal, setup.go:532-546)
//   5. reduce   : sum_k k * B_k per window by a tree of running sums
//   6. the W window sums go back to the host, which does the c-doubling Horner combine.
#pragma once
#include "common.cuh"

namespace b200 {

// Window count for c-bit signed digits over a 254-bit scalar: W*c >= 255 absorbs the last carry.
B2_HD int msm_num_windows(int c) { return (255 + c - 1) / c; }

// Signed-digit recoding of a canonical (non-Montgomery) scalar: digit d_w in [-2^(c-1)+1, 2^(c-1)],
// sum_w d_w 2^(c w) = s.
B2_HD void msm_signed_digits(const uint32_t* s, int c, int W, int32_t* out, size_t stride) {
    int32_t carry = 0;
    const uint32_t mask = (1u << c) - 1u;
    for (int w = 0; w < W; ++w) {
        const int off = w * c, limb = off >> 5, sh = off & 31;
        const uint64_t lo = limb < 8 ? s[limb] : 0u, hi = limb + 1 < 8 ? s[limb + 1] : 0u;
        const uint32_t raw = (uint32_t)(((hi << 32) | lo) >> sh) & mask;
        int32_t d = (int32_t)raw + carry;
        carry = 0;
        if (d > (1 << (c - 1))) { d -= (1 << c); carry = 1; }
        out[(size_t)w * stride] = d;
    }
}

struct MsmPlan {
    int c, W;
    uint32_t nb;      // buckets per window = 2^(c-1)
    uint32_t L0;      // leaves per level-0 reduce node
    uint32_t L;       // fan-in of the upper reduce levels
    int fold;         // 1: all windows share ONE bucket set (bases precomputed as 2^(c w) P)
    int WB;           // bucket sets: W, or 1 when folded
};

// Folded windows (resident base sets only): with T[w][i] = 2^(c w) P_i stored next to the bases,
// digit d_w of scalar i adds sign(d_w) T[w][i] to bucket |d_w| of ONE shared bucket set, so the
// bucket reduction is paid once instead of W times and c can grow: 2^24 points run 13 windows of
// 20 bits (13 n additions) instead of 15 windows of 17 bits.  Costs W x the base memory -- the
// reason this design wants a 180 GB part.  The widths are the measured optimum of
// scripts/fold_sweep.py on a B200 (16, 17 and 20 leave a full or nearly full top window; 18, 19,
// 21 pile the top digits into a few buckets).  Returns 0 when the sorted-entry index (w n + i) does
// not fit 31 bits.
inline int msm_fold_window(size_t n) {
    int log_n = 0;
    while (((size_t)1 << log_n) < n) ++log_n;
    int c;
    if (log_n >= 19) c = 20;
    else if (log_n >= 17) c = 17;
    else if (log_n >= 15) c = log_n;
    else c = 13;
    if ((double)n * msm_num_windows(c) >= 2147483648.0) return 0;
    return c;
}

inline MsmPlan msm_plan(size_t n, int c_override) {
    int c;
    if (c_override > 0) c = c_override;
    else if (n <= ((size_t)1 << 10)) c = 8;
    else if (n <= ((size_t)1 << 14)) c = 10;
    else if (n <= ((size_t)1 << 17)) c = 12;
    else if (n <= ((size_t)1 << 19)) c = 14;
    else if (n <= ((size_t)1 << 21)) c = 15;
    else c = 17;   // 15 x 17 = 255: full top window; measured best at 2^22..2^24 (scripts/tune_sweep.py)
    MsmPlan p;
    p.c = c;
    p.W = msm_num_windows(c);
    p.nb = 1u << (c - 1);
    p.L0 = 16;
    p.L = 16;
    p.fold = 0;
    p.WB = p.W;
    return p;
}
inline MsmPlan msm_plan_folded(int c) {
    MsmPlan p;
    p.c = c;
    p.W = msm_num_windows(c);
    p.nb = 1u << (c - 1);
    p.L0 = 16;
    p.L = 16;
    p.fold = 1;
    p.WB = 1;
    return p;
}

#if defined(__CUDACC__)

template <class T>
__device__ __forceinline__ T ld_vec(const T* p) {
    static_assert(sizeof(T) % 16 == 0, "16-byte multiples only");
    T r;
    const uint4* s = reinterpret_cast<const uint4*>(p);
    uint4* d = reinterpret_cast<uint4*>(&r);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(T) / 16); ++i) d[i] = s[i];
    return r;
}
// Read-only gather of one record.  A miss of a plain ld.global(.nc) makes L2 fetch the whole 128-byte line from DRAM; with the
// L2::64B qualifier it fetches the 64-byte half that was asked for (scripts/gather_probe.cu, ncu dram__bytes_read: 2^26 random
// 64-byte records = 4.31 GB against 8.58 GB; cudaLimitMaxL2FetchGranularity changes nothing).  Records that are not whole
// lines (a 64-byte G1 point, a 32-byte coordinate, 96-byte BLS12-377 points) take the hint; whole-line records do not.
template <class T>
__device__ __forceinline__ T ldg_vec(const T* p) {
    static_assert(sizeof(T) % 16 == 0, "16-byte multiples only");
    T r;
    const uint4* s = reinterpret_cast<const uint4*>(p);
    uint4* d = reinterpret_cast<uint4*>(&r);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(T) / 16); ++i) {
        if (sizeof(T) % 128 == 0) d[i] = __ldg(s + i);
        else asm("ld.global.nc.L2::64B.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(d[i].x), "=r"(d[i].y), "=r"(d[i].z), "=r"(d[i].w) : "l"(s + i));
    }
    return r;
}
template <class T>
__device__ __forceinline__ void st_vec(T* p, const T& v) {
    static_assert(sizeof(T) % 16 == 0, "16-byte multiples only");
    uint4* d = reinterpret_cast<uint4*>(p);
    const uint4* s = reinterpret_cast<const uint4*>(&v);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(T) / 16); ++i) d[i] = s[i];
}

// 1. digits + histogram.  idx_map (optional): scalar i is scalars[idx_map[i]] (wire filtering by
// pk.InfinityA/B, prove.go:154-175, done as a gather instead of three host-side copies).
// Windows whose digits are expected to pile up on a few buckets: window 0 (real witnesses are
// full of 0/1/small values -- every higher digit of those is 0 and skipped) and a top window that
// holds only a few of the 254 scalar bits.  There, same-bucket lanes of a warp are merged with
// match.any before touching the counter, otherwise millions of atomics serialise on one L2
// address (measured: ~0.8 ns each); match.any itself costs a few hundred cycles per warp, so the
// other windows use plain atomics.
__device__ __forceinline__ bool msm_window_is_skewed(int w, int W, int c) {
    return w == 0 || (w == W - 1 && 254 - c * (W - 1) < c - 2);
}
static __global__ void msm_digits_kernel(const Fr* __restrict__ scalars, const uint32_t* __restrict__ idx_map, uint32_t n,
                                  int c, int W, int fold, int32_t* __restrict__ digits, uint32_t* __restrict__ counts, uint32_t nb) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = i < n;
    const int lane = threadIdx.x & 31;
    Fr s = Fr::zero();
    if (valid) s = ld_vec(scalars + (idx_map ? idx_map[i] : i)).from_mont();
    int32_t carry = 0;
    const uint32_t mask = (1u << c) - 1u;
    for (int w = 0; w < W; ++w) {
        const int off = w * c, limb = off >> 5, sh = off & 31;
        const uint64_t lo = limb < 8 ? s.l[limb] : 0u, hi = limb + 1 < 8 ? s.l[limb + 1] : 0u;
        const uint32_t raw = (uint32_t)(((hi << 32) | lo) >> sh) & mask;
        int32_t d = (int32_t)raw + carry;
        carry = 0;
        if (d > (1 << (c - 1))) { d -= (1 << c); carry = 1; }
        if (valid) digits[(size_t)w * n + i] = d;
        const uint32_t key = (valid && d != 0) ? (uint32_t)((d < 0 ? -d : d) - 1) : 0xffffffffu;
        if (msm_window_is_skewed(w, W, c)) {   // warp-uniform
            const uint32_t peers = __match_any_sync(0xffffffffu, key);
            if (key != 0xffffffffu && lane == __ffs(peers) - 1) atomicAdd(&counts[(fold ? 0 : (size_t)w * nb) + key], (uint32_t)__popc(peers));
        } else if (key != 0xffffffffu) {
            atomicAdd(&counts[(fold ? 0 : (size_t)w * nb) + key], 1u);
        }
    }
}

// 2. exclusive scan of `m` counters, three launches (block sums, scan of block sums, rescan).
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 16;
constexpr int SCAN_CHUNK = SCAN_THREADS * SCAN_ITEMS;

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* total) {
    __shared__ uint32_t warp_sums[32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
    }
    if (lane == 31) warp_sums[wid] = x;
    __syncthreads();
    if (wid == 0) {
        const int nw = blockDim.x >> 5;
        uint32_t s = lane < nw ? warp_sums[lane] : 0u;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, s, o);
            if (lane >= o) s += y;
        }
        warp_sums[lane] = s;  // inclusive
    }
    __syncthreads();
    const uint32_t base = wid ? warp_sums[wid - 1] : 0u;
    if (total) *total = warp_sums[(blockDim.x >> 5) - 1];
    const uint32_t r = base + x - v;
    __syncthreads();
    return r;
}

static __global__ void msm_scan_sums_kernel(const uint32_t* __restrict__ in, uint32_t m, uint32_t* __restrict__ block_sums) {
    const uint32_t base = blockIdx.x * SCAN_CHUNK + threadIdx.x * SCAN_ITEMS;
    uint32_t s = 0;
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; ++j) if (base + j < m) s += in[base + j];
    uint32_t total;
    block_exclusive_scan(s, &total);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}
// single block: in-place exclusive scan of nblocks block sums
static __global__ void msm_scan_top_kernel(uint32_t* block_sums, uint32_t nblocks) {
    __shared__ uint32_t carry_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (uint32_t base = 0; base < nblocks; base += blockDim.x) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = i < nblocks ? block_sums[i] : 0u;
        uint32_t total;
        const uint32_t ex = block_exclusive_scan(v, &total);
        const uint32_t carry = carry_s;
        if (i < nblocks) block_sums[i] = carry + ex;
        __syncthreads();
        if (threadIdx.x == 0) carry_s = carry + total;
        __syncthreads();
    }
}
static __global__ void msm_scan_final_kernel(const uint32_t* __restrict__ in, uint32_t m, const uint32_t* __restrict__ block_sums,
                                      uint32_t* out, uint32_t* out2) {
    const uint32_t base = blockIdx.x * SCAN_CHUNK + threadIdx.x * SCAN_ITEMS;
    uint32_t v[SCAN_ITEMS], s = 0;
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; ++j) { v[j] = base + j < m ? in[base + j] : 0u; s += v[j]; }
    uint32_t ex = block_exclusive_scan(s, nullptr) + block_sums[blockIdx.x];
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; ++j) {
        if (base + j < m) { out[base + j] = ex; if (out2) out2[base + j] = ex; }
        ex += v[j];
    }
}

// batched-affine mode (msm_ba.cuh): padded[b] = count rounded up to 2^ba slots, reduced[b] = padded >> ba points
static __global__ void msm_pad_counts_kernel(const uint32_t* __restrict__ counts, uint32_t m, int ba, uint32_t* __restrict__ padded,
                                             uint32_t* __restrict__ reduced) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= m) return;
    const uint32_t q = (1u << ba) - 1u;
    const uint32_t p = (counts[b] + q) & ~q;
    padded[b] = p;
    reduced[b] = p >> ba;
}
static __global__ void msm_shift_kernel(const uint32_t* __restrict__ in, uint32_t m, int sh, uint32_t* __restrict__ out) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < m) out[b] = in[b] >> sh;
}

// 3. scatter (counting sort): cursor starts as a copy of the offsets.
// Persistent form: a small grid (a few CTAs per SM) strides over the points, and every thread keeps G cursor atomics in
// flight (the G digit loads, then the G atomics, then the G stores are issued back to back).  The kernel sits at the L2's
// request rate for scattered 4-byte accesses, which needs requests in flight, not resident threads: G = 8 in 512 threads
// per SM holds as many as one atomic in each of 2048 -- and leaves the rest of the SM to the bucket stage of the previous
// MSM, which runs beside it on the main stream and is bound by the IMAD pipe.
template <int G>
static __global__ void __launch_bounds__(256) msm_scatter_kernel(const int32_t* __restrict__ digits, uint32_t n, int c, int W, int fold, uint32_t stride, uint32_t nb,
                                   uint32_t* __restrict__ cursor, uint32_t* __restrict__ sorted) {
    const int lane = threadIdx.x & 31;
    for (uint32_t base = blockIdx.x * blockDim.x; base < n; base += gridDim.x * blockDim.x) {   // block-uniform trip count
        const uint32_t i = base + threadIdx.x;
        const bool valid = i < n;
        for (int w0 = 0; w0 < W; w0 += G) {
            int32_t d[G];
            uint32_t pos[G];
#pragma unroll
            for (int k = 0; k < G; ++k) d[k] = (valid && w0 + k < W) ? digits[(size_t)(w0 + k) * n + i] : 0;
#pragma unroll
            for (int k = 0; k < G; ++k) {
                const int w = w0 + k;
                const uint32_t key = d[k] != 0 ? (uint32_t)((d[k] < 0 ? -d[k] : d[k]) - 1) : 0xffffffffu;
                pos[k] = 0xffffffffu;
                if (w < W && msm_window_is_skewed(w, W, c)) {   // warp-uniform
                    const uint32_t peers = __match_any_sync(0xffffffffu, key);
                    const int leader = __ffs(peers) - 1;
                    uint32_t first = 0;
                    if (key != 0xffffffffu && lane == leader) first = atomicAdd(&cursor[(fold ? 0 : (size_t)w * nb) + key], (uint32_t)__popc(peers));
                    first = __shfl_sync(0xffffffffu, first, leader);
                    if (key != 0xffffffffu) pos[k] = first + __popc(peers & ((1u << lane) - 1u));
                } else if (key != 0xffffffffu) {
                    pos[k] = atomicAdd(&cursor[(fold ? 0 : (size_t)w * nb) + key], 1u);
                }
            }
#pragma unroll
            for (int k = 0; k < G; ++k)
                if (pos[k] != 0xffffffffu) sorted[pos[k]] = (fold ? (uint32_t)(w0 + k) * stride + i : i) | (d[k] < 0 ? 0x80000000u : 0u);
        }
    }
}

// 4. bucket accumulation.  Buckets are cut into tasks of at most `chunk` points so that one
// oversized bucket (the top window only sees 254 - c*(W-1) scalar bits; real witnesses are full of
// 0/1 values) cannot serialise the launch: ntasks[b] = max(1, ceil(count[b] / chunk)), an
// exclusive scan gives task_off[b], msm_task_fill_kernel writes the owning bucket of every task.
// A bucket is cut into at most MSM_MAX_TASKS_PER_BUCKET equal pieces (so the warp-cooperative merge
// stays short); oversized pieces are scheduled first by the size ordering below.
constexpr uint32_t MSM_MAX_TASKS_PER_BUCKET = 4096;
// piece j of a bucket with cnt points cut into nt pieces: [j*cb, min(cnt, (j+1)*cb)), cb = ceil(cnt/nt)
__device__ __forceinline__ void msm_task_range(uint32_t cnt, uint32_t nt, uint32_t j, uint32_t* first, uint32_t* size) {
    const uint32_t cb = (cnt + nt - 1u) / nt;
    const uint32_t j0 = j * cb;
    *first = j0;
    *size = j0 >= cnt ? 0u : (cnt - j0 < cb ? cnt - j0 : cb);
}
// buckets cut into more than MSM_SMALL_SPLIT pieces are listed in big_list: up to MSM_WARP_SPLIT
// pieces from the front (big_count[0] entries, merged by one warp each), more from the back
// (big_count[1] entries, merged by one CTA each).
constexpr uint32_t MSM_SMALL_SPLIT = 8;
constexpr uint32_t MSM_WARP_SPLIT = 256;
static __global__ void msm_task_count_kernel(const uint32_t* __restrict__ counts, uint32_t n_buckets, uint32_t chunk,
                                             uint32_t* __restrict__ ntasks, uint32_t* __restrict__ big_count,
                                             uint32_t* __restrict__ big_list) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_buckets) return;
    const uint32_t c = counts[b];
    uint32_t nt = c <= chunk ? 1u : (c + chunk - 1u) / chunk;
    if (nt > MSM_MAX_TASKS_PER_BUCKET) nt = MSM_MAX_TASKS_PER_BUCKET;
    ntasks[b] = nt;
    if (nt > MSM_WARP_SPLIT) big_list[n_buckets - 1u - atomicAdd(big_count + 1, 1u)] = b;
    else if (nt > MSM_SMALL_SPLIT) big_list[atomicAdd(big_count, 1u)] = b;
}
static __global__ void msm_task_fill_kernel(const uint32_t* __restrict__ ntasks, const uint32_t* __restrict__ task_off,
                                            uint32_t n_buckets, uint32_t* __restrict__ task_bucket) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_buckets) return;
    const uint32_t nt = ntasks[b], o = task_off[b];
    for (uint32_t j = 0; j < nt; ++j) task_bucket[o + j] = b;
}

// Task order: tasks sorted by decreasing size (counting sort over size bins) so that the 32 lanes
// of a warp run the same trip count -- with a mean load of 32 points the Poisson spread alone
// costs a quarter of the lanes otherwise -- and the longest tasks start first.
constexpr int ORDER_BINS = 2048;
constexpr int ORDER_THREADS = 256;
__device__ __forceinline__ uint32_t msm_task_size(uint32_t t, const uint32_t* task_bucket, const uint32_t* task_off,
                                                  const uint32_t* ntasks, const uint32_t* counts, uint32_t chunk, uint32_t* b_out) {
    const uint32_t b = task_bucket[t];
    *b_out = b;
    if (b == 0xffffffffu) return 0;
    const uint32_t nt = ntasks[b], cnt_b = counts[b];
    if (nt == 1) return cnt_b;
    uint32_t first, size;
    msm_task_range(cnt_b, nt, t - task_off[b], &first, &size);
    return size;
}
// bin 0 holds the largest sizes; unused task slots go to the last bin
__device__ __forceinline__ uint32_t msm_size_bin(uint32_t size, bool used, uint32_t chunk, uint32_t shift) {
    if (!used) return ORDER_BINS - 1;
    const uint32_t q = (size > chunk ? chunk : size) >> shift;   // 0 .. chunk >> shift  (<= ORDER_BINS - 2)
    return (chunk >> shift) - q;
}
static __global__ void msm_order_hist_kernel(const uint32_t* __restrict__ task_bucket, const uint32_t* __restrict__ task_off,
                                             const uint32_t* __restrict__ ntasks, const uint32_t* __restrict__ counts,
                                             uint32_t max_tasks, uint32_t chunk, uint32_t shift, uint32_t* __restrict__ hist) {
    __shared__ uint32_t lh[ORDER_BINS];
    for (int k = threadIdx.x; k < ORDER_BINS; k += blockDim.x) lh[k] = 0;
    __syncthreads();
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < max_tasks) {
        uint32_t b;
        const uint32_t sz = msm_task_size(t, task_bucket, task_off, ntasks, counts, chunk, &b);
        atomicAdd(&lh[msm_size_bin(sz, b != 0xffffffffu, chunk, shift)], 1u);
    }
    __syncthreads();
    for (int k = threadIdx.x; k < ORDER_BINS; k += blockDim.x) if (lh[k]) atomicAdd(&hist[k], lh[k]);
}
// single block: exclusive scan of the ORDER_BINS bins, in place
static __global__ void msm_order_scan_kernel(uint32_t* hist) {
    __shared__ uint32_t carry_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < ORDER_BINS; base += blockDim.x) {
        const int k = base + threadIdx.x;
        const uint32_t v = hist[k];
        uint32_t total;
        const uint32_t ex = block_exclusive_scan(v, &total);
        const uint32_t carry = carry_s;
        hist[k] = carry + ex;
        __syncthreads();
        if (threadIdx.x == 0) carry_s = carry + total;
        __syncthreads();
    }
}
static __global__ void msm_order_fill_kernel(const uint32_t* __restrict__ task_bucket, const uint32_t* __restrict__ task_off,
                                             const uint32_t* __restrict__ ntasks, const uint32_t* __restrict__ counts,
                                             uint32_t max_tasks, uint32_t chunk, uint32_t shift, uint32_t* __restrict__ cursor,
                                             uint32_t* __restrict__ order) {
    __shared__ uint32_t lh[ORDER_BINS];
    __shared__ uint32_t lbase[ORDER_BINS];
    for (int k = threadIdx.x; k < ORDER_BINS; k += blockDim.x) lh[k] = 0;
    __syncthreads();
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t bin = 0, rank = 0;
    if (t < max_tasks) {
        uint32_t b;
        const uint32_t sz = msm_task_size(t, task_bucket, task_off, ntasks, counts, chunk, &b);
        bin = msm_size_bin(sz, b != 0xffffffffu, chunk, shift);
        rank = atomicAdd(&lh[bin], 1u);
    }
    __syncthreads();
    for (int k = threadIdx.x; k < ORDER_BINS; k += blockDim.x) if (lh[k]) lbase[k] = atomicAdd(&cursor[k], lh[k]);
    __syncthreads();
    if (t < max_tasks) order[lbase[bin] + rank] = t;
}

// one thread per task (taken in size order); a single-task bucket is written straight to
// buckets[], the pieces of a split bucket go to partials[task] and are merged by msm_combine_kernel.
// resident blocks asked of ptxas: 12-limb Fp (48 B) is capped at 168 registers so that 3 blocks = 12 warps stay resident
// (188 registers / 2 blocks left the IMAD chains' fixed latency exposed: ncu 'wait' stall 3.1 warps per issue, r01_bls377_*)
template <class F>
__global__ void __launch_bounds__(128, (sizeof(F) == 48 ? 3 : sizeof(F) > 32 ? 2 : 1)) msm_accumulate_kernel(const Affine<F>* __restrict__ bases, const uint32_t* __restrict__ sorted,
                                      const uint32_t* __restrict__ offsets, const uint32_t* __restrict__ counts,
                                      const uint32_t* __restrict__ task_bucket, const uint32_t* __restrict__ task_off,
                                      const uint32_t* __restrict__ ntasks, const uint32_t* __restrict__ order, uint32_t max_tasks,
                                      uint32_t chunk, XYZZ<F>* __restrict__ buckets, XYZZ<F>* __restrict__ partials, uint32_t direct) {
    // direct != 0: `bases` are the points left by the batched-affine rounds, already in bucket order (no entry list)
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= max_tasks) return;
    const uint32_t t = order[slot];
    const uint32_t b = task_bucket[t];
    if (b == 0xffffffffu) return;
    const uint32_t nt = ntasks[b];
    uint32_t first, cnt;
    msm_task_range(counts[b], nt, t - task_off[b], &first, &cnt);
    const uint32_t start = offsets[b] + first;
    XYZZ<F> acc = XYZZ<F>::inf();
    for (uint32_t j = 0; j < cnt; ++j) {
        const uint32_t e = direct ? start + j : sorted[start + j];
        Affine<F> p = ldg_vec(bases + (e & 0x7fffffffu));
        if (e >> 31) p.y = p.y.neg();
        acc.add_mixed(p);
    }
    st_vec(nt == 1 ? buckets + b : partials + t, acc);
}

// The same with the running bucket sum kept in SHARED memory (word-interleaved across the block: word k of thread t
// at smem[k * 128 + t], conflict-free).  For Fp2 the XYZZ accumulator alone is 64 registers (96 on BLS12-377) that stay
// live through the whole addition; here each coordinate is loaded where the formula uses it and stored as soon as its
// new value exists, so the live set is P, R, PP, PPP, Q and one operand.  Option msm_acc_smem: bit 0 = G1, bit 1 = G2.
template <class F> struct FieldWords { static constexpr int N = F::N; B2_D static uint32_t& w(F& x, int k) { return x.l[k]; } };
template <> struct FieldWords<Fp2> {
    static constexpr int N = 2 * Fp::N;
    B2_D static uint32_t& w(Fp2& x, int k) { return k < Fp::N ? x.a0.l[k] : x.a1.l[k - Fp::N]; }
};
template <class F>
struct SmemPoint {
    static constexpr int FW = FieldWords<F>::N;
    uint32_t* base;     // &smem[threadIdx.x]
    B2_D F ld(int field) const {
        F r;
#pragma unroll
        for (int k = 0; k < FW; ++k) FieldWords<F>::w(r, k) = base[(field * FW + k) * 128];
        return r;
    }
    B2_D void st(int field, F v) const {
#pragma unroll
        for (int k = 0; k < FW; ++k) base[(field * FW + k) * 128] = FieldWords<F>::w(v, k);
    }
};

// Resident blocks asked of ptxas.  Measured on BN254 G2 at 2^22 (scripts/g2_smem_ab.py; register kernel 26.4 ms):
// 2 blocks (248 registers, no spill) 25.9 ms, 3 (168 + 264 B stack) 24.5, 4 (128 + 392 B) 23.75, 5 (96 + 568 B) 24.3,
// 6 (80 + 656 B) 25.0 -- the Fp2 products are out-of-line calls, so the spills sit at call boundaries and 16 warps hide
// them.  On BLS12-377 G2 (96-byte field) nothing is gained (2 blocks 67.9 vs 68.1 ms, 3 blocks 70.2): the option is off.
template <class F> constexpr int msm_smem_blocks() { return sizeof(F) >= 96 ? 2 : 4; }
template <class F>
__global__ void __launch_bounds__(128, msm_smem_blocks<F>()) msm_accumulate_smem_kernel(const Affine<F>* __restrict__ bases, const uint32_t* __restrict__ sorted,
                                      const uint32_t* __restrict__ offsets, const uint32_t* __restrict__ counts,
                                      const uint32_t* __restrict__ task_bucket, const uint32_t* __restrict__ task_off,
                                      const uint32_t* __restrict__ ntasks, const uint32_t* __restrict__ order, uint32_t max_tasks,
                                      uint32_t chunk, XYZZ<F>* __restrict__ buckets, XYZZ<F>* __restrict__ partials, uint32_t direct) {
    __shared__ uint32_t sm[4 * FieldWords<F>::N * 128];
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= max_tasks) return;
    const uint32_t t = order[slot];
    const uint32_t b = task_bucket[t];
    if (b == 0xffffffffu) return;
    const uint32_t nt = ntasks[b];
    uint32_t first, cnt;
    msm_task_range(counts[b], nt, t - task_off[b], &first, &cnt);
    const uint32_t start = offsets[b] + first;
    const SmemPoint<F> acc{sm + threadIdx.x};
    enum { X = 0, Y = 1, ZZ = 2, ZZZ = 3 };
    bool inf = true;
    for (uint32_t j = 0; j < cnt; ++j) {
        const uint32_t e = direct ? start + j : sorted[start + j];
        Affine<F> q = ldg_vec(bases + (e & 0x7fffffffu));
        if (e >> 31) q.y = q.y.neg();
        if (q.is_inf()) continue;
        if (inf) { acc.st(X, q.x); acc.st(Y, q.y); acc.st(ZZ, F::one()); acc.st(ZZZ, F::one()); inf = false; continue; }
        const F P = q.x * acc.ld(ZZ) - acc.ld(X);
        const F R = q.y * acc.ld(ZZZ) - acc.ld(Y);
        if (P.is_zero()) {
            if (R.is_zero()) {
                const XYZZ<F> d = XYZZ<F>::dbl_affine(q);
                inf = d.is_inf();
                acc.st(X, d.X); acc.st(Y, d.Y); acc.st(ZZ, d.ZZ); acc.st(ZZZ, d.ZZZ);
            } else inf = true;
            continue;
        }
        const F PP = P.sqr(), PPP = P * PP, Q = acc.ld(X) * PP;
        const F X3 = R.sqr() - PPP - Q.dbl();
        acc.st(X, X3);
        acc.st(Y, mul_sub_mul(R, Q - X3, acc.ld(Y), PPP));
        acc.st(ZZ, acc.ld(ZZ) * PP);
        acc.st(ZZZ, acc.ld(ZZZ) * PPP);
    }
    XYZZ<F> r = XYZZ<F>::inf();
    if (!inf) { r.X = acc.ld(X); r.Y = acc.ld(Y); r.ZZ = acc.ld(ZZ); r.ZZZ = acc.ld(ZZZ); }
    st_vec(nt == 1 ? buckets + b : partials + t, r);
}

template <class F>
__device__ __forceinline__ XYZZ<F> shfl_down_point(const XYZZ<F>& p, int delta) {
    XYZZ<F> r;
    const uint32_t* s = reinterpret_cast<const uint32_t*>(&p);
    uint32_t* d = reinterpret_cast<uint32_t*>(&r);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(XYZZ<F>) / 4); ++i) d[i] = __shfl_down_sync(0xffffffffu, s[i], delta);
    return r;
}

// buckets cut into 2..MSM_SMALL_SPLIT pieces: the owning thread adds them itself
template <class F>
__global__ void __launch_bounds__(128) msm_combine_kernel(const uint32_t* __restrict__ task_off, const uint32_t* __restrict__ ntasks,
                                   uint32_t n_buckets, const XYZZ<F>* __restrict__ partials, XYZZ<F>* __restrict__ buckets) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_buckets) return;
    const uint32_t nt = ntasks[b];
    if (nt <= 1u || nt > MSM_SMALL_SPLIT) return;
    const uint32_t o = task_off[b];
    XYZZ<F> acc = ld_vec(partials + o);
    for (uint32_t j = 1; j < nt; ++j) acc.add(ld_vec(partials + o + j));
    st_vec(buckets + b, acc);
}

// buckets cut into MSM_SMALL_SPLIT+1..MSM_WARP_SPLIT pieces: one warp per bucket
template <class F>
__global__ void __launch_bounds__(128) msm_combine_warp_kernel(const uint32_t* __restrict__ big_count, const uint32_t* __restrict__ big_list,
                                        const uint32_t* __restrict__ task_off, const uint32_t* __restrict__ ntasks,
                                        const XYZZ<F>* __restrict__ partials, XYZZ<F>* __restrict__ buckets) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t n_mid = big_count[0];
    for (uint32_t k = warp; k < n_mid; k += n_warps) {
        const uint32_t b = big_list[k], nt = ntasks[b], o = task_off[b];
        XYZZ<F> acc = XYZZ<F>::inf();
        for (uint32_t j = lane; j < nt; j += 32u) acc.add(ld_vec(partials + o + j));
#pragma unroll 1
        for (int d = 16; d >= 1; d >>= 1) {
            const XYZZ<F> other = shfl_down_point(acc, d);
            if ((int)lane < d) acc.add(other);
        }
        if (lane == 0) st_vec(buckets + b, acc);
    }
}

// heavily split buckets (skewed digits): one CTA per bucket -- strided partial sums per thread, a
// shuffle tree per warp, the warp results through shared memory.
constexpr int COMBINE_BIG_THREADS = 256;
template <class F>
__global__ void __launch_bounds__(COMBINE_BIG_THREADS) msm_combine_big_kernel(const uint32_t* __restrict__ big_count, const uint32_t* __restrict__ big_list,
                                       const uint32_t* __restrict__ task_off, const uint32_t* __restrict__ ntasks,
                                       const XYZZ<F>* __restrict__ partials, XYZZ<F>* __restrict__ buckets) {
    __shared__ XYZZ<F> warp_sum[COMBINE_BIG_THREADS / 32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const uint32_t n_big = big_count[1];
    for (uint32_t k = blockIdx.x; k < n_big; k += gridDim.x) {
        const uint32_t b = big_list[-(int64_t)k], nt = ntasks[b], o = task_off[b];   // big_list = last slot, filled downwards
        XYZZ<F> acc = XYZZ<F>::inf();
        for (uint32_t j = threadIdx.x; j < nt; j += blockDim.x) acc.add(ld_vec(partials + o + j));
#pragma unroll 1
        for (int d = 16; d >= 1; d >>= 1) {
            const XYZZ<F> other = shfl_down_point(acc, d);
            if (lane < d) acc.add(other);
        }
        if (lane == 0) warp_sum[wid] = acc;
        __syncthreads();
        if (threadIdx.x == 0) {
#pragma unroll 1
            for (int w2 = 1; w2 < COMBINE_BIG_THREADS / 32; ++w2) acc.add(warp_sum[w2]);
            st_vec(buckets + b, acc);
        }
        __syncthreads();
    }
}

// 5. reduction tree.  A node over a range of `len` consecutive buckets carries
//      S  = sum of the buckets,   Wt = sum of (position in range + first_weight) * bucket.
// Level 0 builds nodes over L0 buckets with the running-sum trick (first_weight = 1: bucket k-1
// of a window weighs k).  Upper levels merge L children of length `len`:
//      S' = sum S_c,   Wt' = sum Wt_c + len * sum c * S_c.
template <class F>
__global__ void __launch_bounds__(64) msm_reduce_leaf_kernel(const XYZZ<F>* __restrict__ buckets, uint32_t n_nodes, uint32_t L0,
                                       XYZZ<F>* __restrict__ outS, XYZZ<F>* __restrict__ outW) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_nodes) return;
    const XYZZ<F>* b = buckets + (size_t)t * L0;
    XYZZ<F> run = XYZZ<F>::inf(), acc = XYZZ<F>::inf();
    for (int j = (int)L0 - 1; j >= 0; --j) {
        run.add(ld_vec(b + j));
        acc.add(run);
    }
    st_vec(outS + t, run);
    st_vec(outW + t, acc);
}

// Upper part of the reduction, latency-optimised: with T leaf nodes (S_t, Wt_t) per bucket set,
//   sum_k k B_k = sum_t Wt_t + L0 * sum_t t * S_t = V_0 + L0 * sum_b 2^b V_{1+b},
//   V_{1+b} = sum over the nodes t whose index has bit b set of S_t.
// The Q = 1 + log2 T plain sums V_q are binary trees (log2 T launches of independent additions,
// depth one addition each) instead of a chain of ~50 dependent additions per level; the last
// Q values per set go to the host, which does the Horner step.
template <class F>
__global__ void __launch_bounds__(128) msm_tree_first_kernel(const XYZZ<F>* __restrict__ S, const XYZZ<F>* __restrict__ Wt, uint32_t T,
                                                            uint32_t Q, uint32_t total, XYZZ<F>* __restrict__ out) {
    const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const uint32_t half = T >> 1;
    const uint32_t tp = idx % half, q = (idx / half) % Q, set = idx / half / Q;
    const uint32_t t0 = 2u * tp, a = set * T + t0;
    XYZZ<F> r;
    if (q == 0) {
        r = ld_vec(Wt + a);
        r.add(ld_vec(Wt + a + 1));
    } else {
        const uint32_t bit = q - 1u;
        r = XYZZ<F>::inf();
        if ((t0 >> bit) & 1u) r = ld_vec(S + a);
        if (((t0 + 1u) >> bit) & 1u) r.add(ld_vec(S + a + 1));
    }
    st_vec(out + idx, r);
}
// out[g * (m/2) + t] = in[g * m + 2t] + in[g * m + 2t + 1]
template <class F>
__global__ void __launch_bounds__(128) msm_tree_pass_kernel(const XYZZ<F>* __restrict__ in, uint32_t m, uint32_t total,
                                                           XYZZ<F>* __restrict__ out) {
    const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const uint32_t half = m >> 1;
    const uint32_t g = idx / half, t = idx % half;
    XYZZ<F> r = ld_vec(in + (size_t)g * m + 2u * t);
    r.add(ld_vec(in + (size_t)g * m + 2u * t + 1u));
    st_vec(out + idx, r);
}

// Folded-window tables.  Pass 1: thread i walks 2^c P_i, 2^(2c) P_i, ... by repeated doubling and
// parks the XYZZ results; pass 2 normalises the W-1 results of a point with ONE inversion
// (Montgomery's trick over their ZZZ) and writes T[w][i] affine.  One-time cost at upload.
template <class F>
__global__ void __launch_bounds__(128) msm_fold_double_kernel(const Affine<F>* __restrict__ bases, uint32_t n0, uint32_t cnt, int c, int W,
                                                             XYZZ<F>* __restrict__ tmp) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= cnt) return;
    XYZZ<F> p = XYZZ<F>::from_affine(ld_vec(bases + n0 + j));
    for (int w = 1; w < W; ++w) {
        for (int k = 0; k < c; ++k) p = p.dbl();
        st_vec(tmp + (size_t)(w - 1) * cnt + j, p);
    }
}
template <class F>
__global__ void __launch_bounds__(128) msm_fold_normalize_kernel(const XYZZ<F>* __restrict__ tmp, uint32_t n0, uint32_t cnt, int W,
                                                                uint32_t stride, Affine<F>* __restrict__ table) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= cnt) return;
    F pre[31];
    F run = F::one();
    for (int w = 1; w < W; ++w) {
        pre[w - 1] = run;
        const F zzz = ld_vec(&tmp[(size_t)(w - 1) * cnt + j].ZZZ);
        if (!zzz.is_zero()) run = run * zzz;
    }
    F inv = run.inv();
    for (int w = W - 1; w >= 1; --w) {
        const XYZZ<F> p = ld_vec(tmp + (size_t)(w - 1) * cnt + j);
        Affine<F> a = Affine<F>::inf();
        if (!p.ZZZ.is_zero()) {
            const F zi = inv * pre[w - 1];       // 1 / ZZZ_w
            inv = inv * p.ZZZ;
            const F zzi = (zi * p.ZZ).sqr();     // 1 / ZZ_w
            a.x = p.X * zzi;
            a.y = p.Y * zi;
        }
        st_vec(table + (size_t)w * stride + n0 + j, a);
    }
}

#endif  // __CUDACC__

}  // namespace b200
