// Host driver of the MSM pipeline (template over the coordinate field: Fp for G1, Fp2 for G2).
#pragma once
#include <algorithm>
#include <cstdlib>
#include <vector>
#include "msm.cuh"
#include "msm_ba.cuh"

namespace b200 {

// msm_g1.cu / msm_g2.cu
// d_bases: n affine points on the device; d_scalars: Montgomery Fr on the device, read through
// d_idx_map when it is not null (scalar i = d_scalars[d_idx_map[i]]).  The workspace is carved
// from `ws` when ws != nullptr (the caller reserved msm_workspace_bytes), else from the ctx.
// Folded-window table of a resident base set (c == 0: none).  d_bases then points at the table
// T[w][i] = 2^(c w) P_i, w-major with `stride` points per window (T[0] = the bases themselves).
struct MsmFold {
    int c = 0;
    size_t stride = 0;
};
size_t msm_workspace_bytes(size_t n, int group, int c_override, int fold_c);
int msm_g1_run(b200_ctx* ctx, const G1Affine* d_bases, const Fr* d_scalars, const uint32_t* d_idx_map, size_t n,
               G1Jac* out_host, void* ws, const MsmFold* fold);
int msm_g2_run(b200_ctx* ctx, const G2Affine* d_bases, const Fr* d_scalars, const uint32_t* d_idx_map, size_t n,
               G2Jac* out_host, void* ws, const MsmFold* fold);
// The same pipeline in three steps, for callers that overlap the sort of one MSM with the bucket
// accumulation of another (groth16.cu): sizes of the two workspace groups and of the host staging,
// then job_new / job_sort(stream) / job_acc(stream) / [sync] / job_finish / job_delete.
struct MsmSizes { size_t sort_bytes, acc_bytes, result_bytes, ba_bytes; };   // ba_bytes: batched-affine round buffers (one region serves every MSM of a stream: only one accumulates at a time)
MsmSizes msm_sizes(size_t n, int group, int c_override, int fold_c);
int msm_ba_rounds(size_t n, int c_override, int fold_c);   // batched-affine rounds an MSM of n points runs under this mode (the same for G1 and G2)
void* msm_g1_job_new(b200_ctx* ctx, const G1Affine* d_bases, const Fr* d_scalars, const uint32_t* d_idx_map, size_t n,
                     const MsmFold* fold, void* sort_ws, void* acc_ws, void* host_staging, void* ba_ws = nullptr);
int msm_g1_job_sort(b200_ctx* ctx, void* job, cudaStream_t st);
// st_post / ev_split (optional): the bucket reduction (leaf sums, trees, result copy) is enqueued on st_post after
// ev_split, recorded on st behind the accumulation and the split-bucket merges -- it then overlaps whatever st runs next
int msm_g1_job_acc(b200_ctx* ctx, void* job, cudaStream_t st, cudaStream_t st_post = nullptr, cudaEvent_t ev_split = nullptr);
void msm_g1_job_finish(void* job, G1Jac* out_host);
void msm_g1_job_delete(void* job);
void* msm_g2_job_new(b200_ctx* ctx, const G2Affine* d_bases, const Fr* d_scalars, const uint32_t* d_idx_map, size_t n,
                     const MsmFold* fold, void* sort_ws, void* acc_ws, void* host_staging, void* ba_ws = nullptr);
int msm_g2_job_sort(b200_ctx* ctx, void* job, cudaStream_t st);
// the part of job_sort that depends on the job's own bases: for a job that reuses another job's sorted entries (groth16.cu: Bs2 after Bs1)
int msm_g2_job_prep0(b200_ctx* ctx, void* job, cudaStream_t st);
int msm_g2_job_acc(b200_ctx* ctx, void* job, cudaStream_t st, cudaStream_t st_post = nullptr, cudaEvent_t ev_split = nullptr);
void msm_g2_job_finish(void* job, G2Jac* out_host);
void msm_g2_job_delete(void* job);
// builds the table for n resident points (device) into `table` ((W x n) affine points); returns
// B200_ERR_NOMEM without side effects when the table does not fit.
int msm_fold_build_g1(b200_ctx* ctx, const G1Affine* d_bases, size_t n, int c, size_t reserve_bytes, DevBuf& table);
int msm_fold_build_g2(b200_ctx* ctx, const G2Affine* d_bases, size_t n, int c, size_t reserve_bytes, DevBuf& table);

#if defined(__CUDACC__)

// Workspace of one MSM in two groups, so that the sort of the next MSM can run (on a second stream)
// while the buckets of the current one are accumulated:
//   sort group: digits, sorted entries, histogram / offsets / cursors, task lists, size order
//   acc group:  buckets, partial sums of split buckets, reduction buffers
// Offsets are relative to the group's base.
template <class F>
struct MsmLayout {
    MsmPlan plan;
    size_t n, M;  // M = WB * nb buckets
    size_t off_digits, off_sorted, off_counts, off_offsets, off_cursor, off_bsums;
    size_t off_ntasks, off_task_off, off_task_bucket, off_order, off_order_hist, off_big, sort_total;
    size_t off_buckets, off_partials, off_S[2], off_W[2], acc_total;
    size_t ba_total;   // batched-affine round buffers (offsets off_ba_* are relative to their own base)
    size_t total;   // sort_total + acc_total + ba_total
    uint32_t scan_blocks, chunk, max_tasks, order_shift, T, logT, Q;
    // batched-affine pre-reduction (msm_ba.cuh): ba rounds, entry slots padded per bucket to 2^ba
    int ba;
    size_t pad_entries;                       // upper bound of the padded slot count (multiple of 2^ba * BA_K)
    size_t off_rcounts, off_roffsets;         // per-bucket counts / offsets after the rounds
    size_t off_ba_pts[2], off_ba_pref, off_ba_cprod, off_ba_cinv;   // in the batched-affine group
    size_t off_ba_pref0, off_ba_cprod0;       // in the SORT group: the first round's denominator pass runs with the sort (msm_enqueue_prep0)
    // c_override carries the MSM mode: bits 0-7 the window override, bits 8-11 the batched-affine rounds
    MsmLayout(size_t n_, int c_override, int fold_c) {
        n = n_;
        ba = (c_override >> 8) & 0xf;
        const bool ba_force = (c_override & 0x1000) != 0;
        c_override &= 0xff;
        plan = fold_c ? msm_plan_folded(fold_c) : msm_plan(n, c_override);
        M = (size_t)plan.WB * plan.nb;
        if (ba == 0xf) {
            // automatic: rounds until ~24 points are left per bucket (measured flat optimum at 2^24 points, c = 20: 3-4 rounds
            // of a mean load of 416), none below 2^22 entries -- the rounds' fixed costs (one latency-bound inversion
            // kernel each) outweigh them
            ba = 0;
            const size_t mean = ((size_t)plan.W * n) / M;
            while (ba < 6 && (mean >> (ba + 1)) >= 24) ++ba;
            if ((size_t)plan.W * n < ((size_t)1 << 24)) ba = 0;
        }
        if (!ba_force && ((size_t)plan.W * n < ((size_t)1 << 22) || ((size_t)plan.W * n) / M < ((size_t)2 << ba))) ba = 0;   // too little work per round / per bucket
        pad_entries = (size_t)plan.W * n;
        if (ba) {
            const size_t q = ((size_t)1 << ba) * BA_K;
            pad_entries += M * (((size_t)1 << ba) - 1);
            pad_entries = (pad_entries + q - 1) / q * q;
        }
        // task splitting: pieces of at most twice the mean bucket load, but small enough that the
        // launch has ~400 K tasks (148 SMs x 20 warps x 32 lanes x 4 waves) even when few buckets exist
        const size_t entries = ba ? (pad_entries >> ba) : (size_t)plan.W * n, mean = entries / M;
        size_t ch = entries / 400000;
        const size_t hi = 2 * mean < 64 ? 64 : 2 * mean;
        if (ch > hi) ch = hi;
        if (ch < 32) ch = 32;
        chunk = (uint32_t)ch;
        max_tasks = (uint32_t)(M + entries / chunk + 1);
        scan_blocks = (uint32_t)((M + SCAN_CHUNK - 1) / SCAN_CHUNK);
        order_shift = 0;
        while ((chunk >> order_shift) > (uint32_t)ORDER_BINS - 2) ++order_shift;
        Arena a;
        off_digits = a.off; a.take((size_t)plan.W * n * 4);
        off_sorted = a.off; a.take(pad_entries * 4);
        off_counts = a.off; a.take(M * 4);
        off_offsets = a.off; a.take(M * 4);
        off_cursor = a.off; a.take(M * 4);
        off_bsums = a.off; a.take((size_t)scan_blocks * 4);
        off_ntasks = a.off; a.take(M * 4);
        off_task_off = a.off; a.take(M * 4);
        off_task_bucket = a.off; a.take((size_t)max_tasks * 4);
        off_big = a.off; a.take((M + 2) * 4);   // [0], [1] = counts, then the list (front: warp-merged, back: CTA-merged)
        off_order = a.off; a.take((size_t)max_tasks * 4);
        off_order_hist = a.off; a.take((size_t)ORDER_BINS * 4 * 2);
        off_rcounts = a.off; a.take(ba ? M * 4 : 0);
        off_roffsets = a.off; a.take(ba ? M * 4 : 0);
        off_ba_pref0 = a.off; a.take(ba ? (pad_entries / 2) * sizeof(F) : 0);
        off_ba_cprod0 = a.off; a.take(ba ? (pad_entries / 2 / BA_K + 1) * sizeof(F) : 0);
        sort_total = a.off;
        Arena g;
        off_buckets = g.off; g.take(M * sizeof(XYZZ<F>));
        off_partials = g.off; g.take((size_t)max_tasks * sizeof(XYZZ<F>));
        const size_t nodes0 = M / plan.L0;
        T = plan.nb / plan.L0;
        logT = 0;
        while ((1u << logT) < T) ++logT;
        Q = 1 + logT;
        off_S[0] = g.off; g.take(nodes0 * sizeof(XYZZ<F>));
        off_W[0] = g.off; g.take(nodes0 * sizeof(XYZZ<F>));
        // tree ping-pong buffers: WB * Q arrays of T/2 and T/4 points
        off_S[1] = g.off; g.take((size_t)plan.WB * Q * (T / 2 + 1) * sizeof(XYZZ<F>));
        off_W[1] = g.off; g.take((size_t)plan.WB * Q * (T / 4 + 1) * sizeof(XYZZ<F>));
        acc_total = g.off;
        Arena q;
        off_ba_pts[0] = q.off; q.take(ba ? (pad_entries / 2) * sizeof(Affine<F>) : 0);
        off_ba_pts[1] = q.off; q.take(ba > 1 ? (pad_entries / 4) * sizeof(Affine<F>) : 0);
        off_ba_pref = q.off; q.take(ba > 1 ? (pad_entries / 4) * sizeof(F) : 0);
        off_ba_cprod = q.off; q.take(ba > 1 ? (pad_entries / 4 / BA_K + 1) * sizeof(F) : 0);
        off_ba_cinv = q.off; q.take(ba ? (pad_entries / 2 / BA_K + 1) * sizeof(F) : 0);
        ba_total = ba ? q.off : 0;
        total = sort_total + acc_total + ba_total;
    }
    size_t result_points() const { return (size_t)plan.WB * Q; }   // XYZZ values the host finishes from
};

// One MSM in flight: enqueue the sort, enqueue the accumulation (possibly on another stream, after
// an event), and finish on the host once the result values have arrived.
template <class F>
struct MsmJob {
    MsmLayout<F> lay;
    const Affine<F>* d_bases;
    const Fr* d_scalars;
    const uint32_t* d_idx_map;
    size_t n;
    uint32_t stride;
    char* sort_base;
    char* acc_base;
    char* ba_base;     // batched-affine round buffers (lay.ba_total bytes)
    XYZZ<F>* h_result;   // host staging of lay.result_points() values (page-locked for a truly async copy)
    MsmJob(b200_ctx* ctx, const Affine<F>* bases, const Fr* scalars, const uint32_t* idx_map, size_t n_, const MsmFold* fold,
           void* sort_ws, void* acc_ws, XYZZ<F>* host_staging, void* ba_ws = nullptr)
        : lay(n_ ? n_ : 1, ctx->msm_mode(), fold ? fold->c : 0), d_bases(bases), d_scalars(scalars), d_idx_map(idx_map), n(n_),
          stride(fold && fold->c ? (uint32_t)fold->stride : 0u), sort_base((char*)sort_ws), acc_base((char*)acc_ws),
          ba_base((char*)ba_ws), h_result(host_staging) {}
};

// First batched-affine round, denominator pass: needs the sorted entries and the base table only, and is bound by the
// DRAM's random-access rate with the multiplier idle -- so it is enqueued with the sort (on the stream that runs beside the
// previous MSM's arithmetic) and leaves pref0 / cprod0 in the sort group.
// Small persistent grids for the sort's scatter and the round-0 denominator pass: only where the caller runs them beside
// another MSM's bucket stage (groth16.cu) and the MSM is large enough that the slower aux chain still ends before that
// stage does (measured at 2^24 points x 13 windows: 233.8 -> 226.3 ms per proof, and at 2^22: Groth16 64.2 -> 63.3 ms, PlonK
// 130.6 -> 128.6 ms, BLS12-377 140.8 -> 139.5 ms; not swept below that, so the 2^21-point MSMs of an 8-GPU group keep full grids).
template <class F>
inline bool msm_aux_persistent(const b200_ctx* ctx, const MsmLayout<F>& lay) {
    static const long long min_env = [] { const char* e = getenv("B200_AUX_MIN_ENTRIES"); return e && *e ? atoll(e) : -1LL; }();
    const size_t min_entries = ctx->aux_persist_min >= 0 ? (size_t)ctx->aux_persist_min : min_env >= 0 ? (size_t)min_env : ((size_t)3 << 24);   // default: 2^22 points x 13 windows and up
    return ctx->aux_persist && lay.ba && (size_t)lay.plan.W * lay.n >= min_entries;
}

template <class F>
int msm_enqueue_prep0(b200_ctx* ctx, const MsmJob<F>& job, cudaStream_t st) {
    const MsmLayout<F>& lay = job.lay;
    if (job.n == 0 || !lay.ba) return B200_OK;
    char* base = job.sort_base;
    const uint32_t n_out = (uint32_t)(lay.pad_entries >> 1);
    const uint32_t threads = n_out / BA_K;
    uint32_t blocks = (threads + 127) / 128;
    // Beside the bucket stage of the previous MSM (ctx->aux_persist) the pass runs as 2 CTAs per SM striding over the chains:
    // it is bound by the DRAM's random-access rate, which 256 threads per SM with two gathers in flight each sustain, and
    // every CTA it does not occupy stays with the IMAD-bound stage.  Measured on the 2^24 prove core together with the
    // persistent scatter: 233.8 -> 226.3 ms (profiles/r02_persistent_aux_ab.txt).  B200_PREP0_BPS overrides (0 = full grid).
    static const int prep0_env = [] { const char* e = getenv("B200_PREP0_BPS"); return e && *e ? atoi(e) : -1; }();
    static const int prep0_env_g2 = [] { const char* e = getenv("B200_PREP0_BPS_G2"); return e && *e ? atoi(e) : -1; }();
    int prep0_bps = prep0_env >= 0 ? prep0_env : (msm_aux_persistent(ctx, lay) ? 2 : 0);
    if (sizeof(F) > sizeof(Fp) && prep0_env_g2 >= 0) prep0_bps = prep0_env_g2;
    if (prep0_bps > 0) blocks = std::min(blocks, (uint32_t)(ctx->sm_count * prep0_bps));
    ba_prep_kernel<F, true><<<blocks, 128, 0, st>>>(job.d_bases, (const uint32_t*)(base + lay.off_sorted), nullptr, nullptr, threads,
                                                    (F*)(base + lay.off_ba_pref0), (F*)(base + lay.off_ba_cprod0));
    B2_CUDA(ctx, cudaGetLastError());
    ctx->n_launches += 1;
    return B200_OK;
}

template <class F>
int msm_enqueue_sort(b200_ctx* ctx, const MsmJob<F>& job, cudaStream_t st) {
    if (job.n == 0) return B200_OK;
    const MsmLayout<F>& lay = job.lay;
    const MsmPlan& pl = lay.plan;
    const size_t n = job.n;
    char* base = job.sort_base;
    int32_t* digits = (int32_t*)(base + lay.off_digits);
    uint32_t* sorted = (uint32_t*)(base + lay.off_sorted);
    uint32_t* counts = (uint32_t*)(base + lay.off_counts);
    uint32_t* offsets = (uint32_t*)(base + lay.off_offsets);
    uint32_t* cursor = (uint32_t*)(base + lay.off_cursor);
    uint32_t* bsums = (uint32_t*)(base + lay.off_bsums);
    uint32_t* ntasks = (uint32_t*)(base + lay.off_ntasks);
    uint32_t* task_off = (uint32_t*)(base + lay.off_task_off);
    uint32_t* task_bucket = (uint32_t*)(base + lay.off_task_bucket);
    uint32_t* big = (uint32_t*)(base + lay.off_big);
    uint32_t* order = (uint32_t*)(base + lay.off_order);
    uint32_t* order_hist = (uint32_t*)(base + lay.off_order_hist);
    const uint32_t M = (uint32_t)lay.M;
    const int span = ctx->prof_begin(B200_PROF_MSM_SORT, (double)n, st);
    B2_CUDA(ctx, cudaMemsetAsync(counts, 0, lay.M * 4, st));
    {
        const int bs = 256;
        msm_digits_kernel<<<(unsigned)((n + bs - 1) / bs), bs, 0, st>>>(job.d_scalars, job.d_idx_map, (uint32_t)n, pl.c, pl.W, pl.fold, digits, counts, pl.nb);
        B2_CUDA(ctx, cudaGetLastError());
    }
    // batched-affine mode: every bucket's slot region is padded to a multiple of 2^ba; `ntasks` holds the
    // padded counts until the task lists are built, `rcounts` the counts after the rounds
    uint32_t* rcounts = lay.ba ? (uint32_t*)(base + lay.off_rcounts) : counts;
    uint32_t* roffsets = lay.ba ? (uint32_t*)(base + lay.off_roffsets) : offsets;
    const uint32_t* scan_in = counts;
    if (lay.ba) {
        msm_pad_counts_kernel<<<(M + 255) / 256, 256, 0, st>>>(counts, M, lay.ba, ntasks, rcounts);
        B2_CUDA(ctx, cudaGetLastError());
        B2_CUDA(ctx, cudaMemsetAsync(sorted, 0xff, lay.pad_entries * 4, st));
        scan_in = ntasks;
    }
    msm_scan_sums_kernel<<<lay.scan_blocks, SCAN_THREADS, 0, st>>>(scan_in, M, bsums);
    B2_CUDA(ctx, cudaGetLastError());
    msm_scan_top_kernel<<<1, 1024, 0, st>>>(bsums, lay.scan_blocks);
    B2_CUDA(ctx, cudaGetLastError());
    msm_scan_final_kernel<<<lay.scan_blocks, SCAN_THREADS, 0, st>>>(scan_in, M, bsums, offsets, cursor);
    B2_CUDA(ctx, cudaGetLastError());
    if (lay.ba) {
        msm_shift_kernel<<<(M + 255) / 256, 256, 0, st>>>(offsets, M, lay.ba, roffsets);
        B2_CUDA(ctx, cudaGetLastError());
    }
    {
        const int bs = 256;
        // B200_SCATTER_BPS = CTAs per SM of the persistent form (0: one thread per point, one atomic in flight each)
        static const int scatter_env = [] { const char* e = getenv("B200_SCATTER_BPS"); return e && *e ? atoi(e) : -1; }();
        const int scatter_bps = scatter_env >= 0 ? scatter_env : (msm_aux_persistent(ctx, lay) ? 1 : 0);
        const unsigned full = (unsigned)((n + bs - 1) / bs);
        static const int scatter_bs = [] { const char* e = getenv("B200_SCATTER_BS"); return e && *e ? atoi(e) : 256; }();
        static const int scatter_g = [] { const char* e = getenv("B200_SCATTER_G"); return e && *e ? atoi(e) : 8; }();
        if (scatter_bps > 0) {
            const unsigned grid = std::min((unsigned)((n + scatter_bs - 1) / scatter_bs), (unsigned)(ctx->sm_count * scatter_bps));
            if (scatter_g >= 16) msm_scatter_kernel<16><<<grid, scatter_bs, 0, st>>>(digits, (uint32_t)n, pl.c, pl.W, pl.fold, job.stride, pl.nb, cursor, sorted);
            else msm_scatter_kernel<8><<<grid, scatter_bs, 0, st>>>(digits, (uint32_t)n, pl.c, pl.W, pl.fold, job.stride, pl.nb, cursor, sorted);
        } else {
            msm_scatter_kernel<1><<<full, bs, 0, st>>>(digits, (uint32_t)n, pl.c, pl.W, pl.fold, job.stride, pl.nb, cursor, sorted);
        }
        B2_CUDA(ctx, cudaGetLastError());
    }
    {
        const int bs = 256;
        B2_CUDA(ctx, cudaMemsetAsync(big, 0, 8, st));
        msm_task_count_kernel<<<(M + bs - 1) / bs, bs, 0, st>>>(rcounts, M, lay.chunk, ntasks, big, big + 2);
        B2_CUDA(ctx, cudaGetLastError());
        msm_scan_sums_kernel<<<lay.scan_blocks, SCAN_THREADS, 0, st>>>(ntasks, M, bsums);
        B2_CUDA(ctx, cudaGetLastError());
        msm_scan_top_kernel<<<1, 1024, 0, st>>>(bsums, lay.scan_blocks);
        B2_CUDA(ctx, cudaGetLastError());
        msm_scan_final_kernel<<<lay.scan_blocks, SCAN_THREADS, 0, st>>>(ntasks, M, bsums, task_off, nullptr);
        B2_CUDA(ctx, cudaGetLastError());
        B2_CUDA(ctx, cudaMemsetAsync(task_bucket, 0xff, (size_t)lay.max_tasks * 4, st));
        msm_task_fill_kernel<<<(M + bs - 1) / bs, bs, 0, st>>>(ntasks, task_off, M, task_bucket);
        B2_CUDA(ctx, cudaGetLastError());
        // size-ordered task list
        const uint32_t ob = (lay.max_tasks + ORDER_THREADS - 1) / ORDER_THREADS;
        B2_CUDA(ctx, cudaMemsetAsync(order_hist, 0, (size_t)ORDER_BINS * 4, st));
        msm_order_hist_kernel<<<ob, ORDER_THREADS, 0, st>>>(task_bucket, task_off, ntasks, rcounts, lay.max_tasks, lay.chunk, lay.order_shift, order_hist);
        B2_CUDA(ctx, cudaGetLastError());
        msm_order_scan_kernel<<<1, 1024, 0, st>>>(order_hist);
        B2_CUDA(ctx, cudaGetLastError());
        msm_order_fill_kernel<<<ob, ORDER_THREADS, 0, st>>>(task_bucket, task_off, ntasks, rcounts, lay.max_tasks, lay.chunk, lay.order_shift, order_hist, order);
        B2_CUDA(ctx, cudaGetLastError());
    }
    B2_TRY(msm_enqueue_prep0(ctx, job, st));
    ctx->prof_end(span, st);
    ctx->n_launches += 13;  // digits, 2 x 3 scan, scatter, task count/fill, 3 x order
    return B200_OK;
}

template <class F>
int msm_enqueue_acc(b200_ctx* ctx, const MsmJob<F>& job, cudaStream_t st, cudaStream_t st_post = nullptr, cudaEvent_t ev_split = nullptr) {
    if (job.n == 0) return B200_OK;
    const MsmLayout<F>& lay = job.lay;
    const MsmPlan& pl = lay.plan;
    const size_t n = job.n;
    char* sb = job.sort_base;
    char* ab = job.acc_base;
    uint32_t* sorted = (uint32_t*)(sb + lay.off_sorted);
    uint32_t* counts = (uint32_t*)(sb + lay.off_counts);
    uint32_t* offsets = (uint32_t*)(sb + lay.off_offsets);
    uint32_t* ntasks = (uint32_t*)(sb + lay.off_ntasks);
    uint32_t* task_off = (uint32_t*)(sb + lay.off_task_off);
    uint32_t* task_bucket = (uint32_t*)(sb + lay.off_task_bucket);
    uint32_t* big = (uint32_t*)(sb + lay.off_big);
    uint32_t* order = (uint32_t*)(sb + lay.off_order);
    XYZZ<F>* buckets = (XYZZ<F>*)(ab + lay.off_buckets);
    XYZZ<F>* partials = (XYZZ<F>*)(ab + lay.off_partials);
    XYZZ<F>* S[2] = {(XYZZ<F>*)(ab + lay.off_S[0]), (XYZZ<F>*)(ab + lay.off_S[1])};
    XYZZ<F>* Wt[2] = {(XYZZ<F>*)(ab + lay.off_W[0]), (XYZZ<F>*)(ab + lay.off_W[1])};
    const uint32_t M = (uint32_t)lay.M;
    constexpr int ACC_SLOT = sizeof(F) == sizeof(Fp) ? B200_PROF_MSM_ACC_G1 : B200_PROF_MSM_ACC_G2;
    int span;
    {
        const int bs = 128;
        span = ctx->prof_begin(ACC_SLOT, (double)n, st);
        const Affine<F>* acc_bases = job.d_bases;
        const uint32_t *acc_offsets = offsets, *acc_counts = counts;
        if (lay.ba) {
            // pairwise affine rounds: round r turns pad_entries >> r slots into half as many points
            char* bb = job.ba_base;
            if (!bb) return ctx->fail(B200_ERR_INTERNAL, "batched-affine rounds without their workspace");
            Affine<F>* pts[2] = {(Affine<F>*)(bb + lay.off_ba_pts[0]), (Affine<F>*)(bb + lay.off_ba_pts[1])};
            F *pref = nullptr, *cprod = nullptr;
            F* cinv = (F*)(bb + lay.off_ba_cinv);
            // round r reads the base table through the sorted entries (r = 0) or the planes of round r - 1, and writes
            // planes (x then y) -- the last round (x, y) pairs for msm_accumulate_kernel
            const F *in_x = nullptr, *in_y = nullptr;
            const Affine<F>* in = job.d_bases;
            for (int r = 0; r < lay.ba; ++r) {
                const uint32_t n_out = (uint32_t)(lay.pad_entries >> (r + 1));
                const uint32_t threads = n_out / BA_K, blocks = (threads + 127) / 128;
                const bool last = r == lay.ba - 1;
                Affine<F>* out = pts[r & 1];
                F* out_x = (F*)out;
                F* out_y = out_x + n_out;
                if (r == 0) { pref = (F*)(sb + lay.off_ba_pref0); cprod = (F*)(sb + lay.off_ba_cprod0); }   // msm_enqueue_prep0 ran with the sort
                else {
                    pref = (F*)(bb + lay.off_ba_pref); cprod = (F*)(bb + lay.off_ba_cprod);
                    ba_prep_kernel<F, false><<<blocks, 128, 0, st>>>(nullptr, nullptr, in_x, in_y, threads, pref, cprod);
                    B2_CUDA(ctx, cudaGetLastError());
                }
                // thread products per Fermat inversion: as many as keep ~32 K inversion threads busy
                uint32_t chunk = threads / 32768u;
                if (chunk < 8u) chunk = 8u;
                if (chunk > (uint32_t)BA_INV_CHUNK) chunk = (uint32_t)BA_INV_CHUNK;
                const uint32_t inv_threads = (threads + chunk - 1) / chunk;
                ba_inverse_kernel<F><<<(inv_threads + 63) / 64, 64, 0, st>>>(cprod, cinv, threads, inv_threads, chunk);
                B2_CUDA(ctx, cudaGetLastError());
                if (r == 0) {
                    if (last) ba_apply_kernel<F, true, true><<<blocks, 128, 0, st>>>(job.d_bases, sorted, nullptr, nullptr, threads, pref, cinv, nullptr, nullptr, out);
                    else ba_apply_kernel<F, true, false><<<blocks, 128, 0, st>>>(job.d_bases, sorted, nullptr, nullptr, threads, pref, cinv, out_x, out_y, nullptr);
                } else {
                    if (last) ba_apply_kernel<F, false, true><<<blocks, 128, 0, st>>>(nullptr, nullptr, in_x, in_y, threads, pref, cinv, nullptr, nullptr, out);
                    else ba_apply_kernel<F, false, false><<<blocks, 128, 0, st>>>(nullptr, nullptr, in_x, in_y, threads, pref, cinv, out_x, out_y, nullptr);
                }
                B2_CUDA(ctx, cudaGetLastError());
                ctx->n_launches += 3;
                in = out;
                in_x = out_x;
                in_y = out_y;
            }
            acc_bases = in;
            acc_offsets = (const uint32_t*)(sb + lay.off_roffsets);
            acc_counts = (const uint32_t*)(sb + lay.off_rcounts);
        }
        if (((ctx->msm_acc_smem >> (sizeof(F) > sizeof(Fp) ? 1 : 0)) & 1) && bs == 128)
            msm_accumulate_smem_kernel<F><<<(lay.max_tasks + bs - 1) / bs, bs, 0, st>>>(acc_bases, sorted, acc_offsets, acc_counts, task_bucket, task_off,
                                                                                     ntasks, order, lay.max_tasks, lay.chunk, buckets, partials, lay.ba ? 1u : 0u);
        else
        msm_accumulate_kernel<F><<<(lay.max_tasks + bs - 1) / bs, bs, 0, st>>>(acc_bases, sorted, acc_offsets, acc_counts, task_bucket, task_off,
                                                                            ntasks, order, lay.max_tasks, lay.chunk, buckets, partials, lay.ba ? 1u : 0u);
        ctx->prof_end(span, st);
        B2_CUDA(ctx, cudaGetLastError());
        span = ctx->prof_begin(B200_PROF_MSM_REDUCE, (double)n, st);
        msm_combine_kernel<F><<<(M + bs - 1) / bs, bs, 0, st>>>(task_off, ntasks, M, partials, buckets);
        B2_CUDA(ctx, cudaGetLastError());
        msm_combine_warp_kernel<F><<<ctx->sm_count * 8, 128, 0, st>>>(big, big + 2, task_off, ntasks, partials, buckets);
        B2_CUDA(ctx, cudaGetLastError());
        msm_combine_big_kernel<F><<<ctx->sm_count * 2, COMBINE_BIG_THREADS, 0, st>>>(big, big + 2 + (M - 1), task_off, ntasks, partials, buckets);
        B2_CUDA(ctx, cudaGetLastError());
    }
    // reduction: leaf running sums, then binary trees for the Q plain sums of every bucket set.  It reads the buckets
    // only (the sort group is free from here): with a post stream it runs beside the next MSM's accumulation.
    if (st_post && ev_split) {
        ctx->prof_end(span, st);
        B2_CUDA(ctx, cudaEventRecord(ev_split, st));
        B2_CUDA(ctx, cudaStreamWaitEvent(st_post, ev_split, 0));
        st = st_post;
        span = ctx->prof_begin(B200_PROF_MSM_REDUCE, 0.0, st);
    }
    const uint32_t T = lay.T, Q = lay.Q;
    {
        const uint32_t nodes = (uint32_t)pl.WB * T;
        msm_reduce_leaf_kernel<F><<<(nodes + 63) / 64, 64, 0, st>>>(buckets, nodes, pl.L0, S[0], Wt[0]);
        B2_CUDA(ctx, cudaGetLastError());
    }
    ctx->n_launches += 5;  // accumulate, 3 x combine, leaf
    const size_t out_bytes = lay.result_points() * sizeof(XYZZ<F>);
    if (T >= 2) {
        XYZZ<F>* buf[2] = {S[1], Wt[1]};
        const uint32_t groups = (uint32_t)pl.WB * Q;
        uint32_t m = T >> 1;  // length of every array after the first pass
        {
            const uint32_t total = groups * m;
            msm_tree_first_kernel<F><<<(total + 127) / 128, 128, 0, st>>>(S[0], Wt[0], T, Q, total, buf[0]);
            B2_CUDA(ctx, cudaGetLastError());
            ctx->n_launches += 1;
        }
        int cur = 0;
        while (m > 1) {
            const uint32_t total = groups * (m >> 1);
            msm_tree_pass_kernel<F><<<(total + 127) / 128, 128, 0, st>>>(buf[cur], m, total, buf[cur ^ 1]);
            B2_CUDA(ctx, cudaGetLastError());
            ctx->n_launches += 1;
            m >>= 1;
            cur ^= 1;
        }
        ctx->prof_end(span, st);
        B2_CUDA(ctx, cudaMemcpyAsync(job.h_result, buf[cur], out_bytes, cudaMemcpyDeviceToHost, st));
    } else {
        ctx->prof_end(span, st);
        B2_CUDA(ctx, cudaMemcpyAsync(job.h_result, Wt[0], out_bytes, cudaMemcpyDeviceToHost, st));
    }
    return B200_OK;
}

// host, after the stream that ran msm_enqueue_acc has been synchronised: per bucket set
// V_0 + L0 * sum_b 2^b V_{1+b}; then the c-doubling Horner over the sets
template <class F>
void msm_finish(const MsmJob<F>& job, Jac<F>* out_host) {
    if (job.n == 0) { *out_host = Jac<F>::inf(); return; }
    const MsmPlan& pl = job.lay.plan;
    const uint32_t Q = job.lay.Q;
    uint32_t logL0 = 0;
    while ((1u << logL0) < pl.L0) ++logL0;
    XYZZ<F> acc = XYZZ<F>::inf();
    for (int w = pl.WB - 1; w >= 0; --w) {
        for (int k = 0; k < pl.c; ++k) acc = acc.dbl();
        const XYZZ<F>* v = job.h_result + (size_t)w * Q;
        XYZZ<F> hi = XYZZ<F>::inf();
        for (int bit = (int)Q - 2; bit >= 0; --bit) {
            hi = hi.dbl();
            hi.add(v[1 + bit]);
        }
        for (uint32_t k = 0; k < logL0; ++k) hi = hi.dbl();
        hi.add(v[0]);
        acc.add(hi);
    }
    *out_host = acc.to_jac();
}

template <class F>
int msm_run_impl(b200_ctx* ctx, const Affine<F>* d_bases, const Fr* d_scalars, const uint32_t* d_idx_map, size_t n,
                 Jac<F>* out_host, void* ws, const MsmFold* fold) {
    if (n == 0) { *out_host = Jac<F>::inf(); return B200_OK; }
    if (n >= ((size_t)1 << 31)) return ctx->fail(B200_ERR_UNSUPPORTED, "MSM above 2^31 points");
    const int fold_c = fold ? fold->c : 0;
    MsmLayout<F> lay(n, ctx->msm_mode(), fold_c);
    char* base;
    if (ws) base = (char*)ws;
    else { B2_TRY(ctx_reserve(ctx, lay.total)); base = (char*)ctx->scratch; }
    std::vector<XYZZ<F>> fin(lay.result_points());
    MsmJob<F> job(ctx, d_bases, d_scalars, d_idx_map, n, fold, base, base + lay.sort_total, fin.data(), base + lay.sort_total + lay.acc_total);
    B2_TRY(msm_enqueue_sort(ctx, job, ctx->stream));
    B2_TRY(msm_enqueue_acc(ctx, job, ctx->stream));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    msm_finish(job, out_host);
    return B200_OK;
}

template <class F>
int msm_fold_build_impl(b200_ctx* ctx, const Affine<F>* d_bases, size_t n, int c, size_t reserve_bytes, DevBuf& table) {
    const int W = msm_num_windows(c);
    if (n == 0 || W > 32 || (double)n * W >= 2147483648.0) return ctx->fail(B200_ERR_ARG, "folded table: bad size");
    const size_t chunk = n < ((size_t)1 << 20) ? n : ((size_t)1 << 20);
    {   // the table must leave `reserve_bytes` of device memory to the caller's later allocations
        size_t free_b = 0, total_b = 0;
        if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) { cudaGetLastError(); return B200_ERR_NOMEM; }
        const size_t need = (size_t)W * n * sizeof(Affine<F>) + (size_t)(W - 1) * chunk * sizeof(XYZZ<F>) + reserve_bytes;
        if (need > free_b) return B200_ERR_NOMEM;
    }
    DevBuf tmp;
    if (table.alloc((size_t)W * n * sizeof(Affine<F>)) != cudaSuccess || tmp.alloc((size_t)(W - 1) * chunk * sizeof(XYZZ<F>)) != cudaSuccess) {
        cudaGetLastError();
        table.release();
        return B200_ERR_NOMEM;
    }
    cudaStream_t st = ctx->stream;
    B2_CUDA(ctx, cudaMemcpyAsync(table.p, d_bases, n * sizeof(Affine<F>), cudaMemcpyDeviceToDevice, st));
    for (size_t n0 = 0; n0 < n; n0 += chunk) {
        const uint32_t cnt = (uint32_t)(n - n0 < chunk ? n - n0 : chunk);
        msm_fold_double_kernel<F><<<(cnt + 127) / 128, 128, 0, st>>>(d_bases, (uint32_t)n0, cnt, c, W, tmp.as<XYZZ<F>>());
        B2_CUDA(ctx, cudaGetLastError());
        msm_fold_normalize_kernel<F><<<(cnt + 127) / 128, 128, 0, st>>>(tmp.as<XYZZ<F>>(), (uint32_t)n0, cnt, W, (uint32_t)n, table.as<Affine<F>>());
        B2_CUDA(ctx, cudaGetLastError());
    }
    B2_CUDA(ctx, cudaStreamSynchronize(st));
    return B200_OK;
}

#endif  // __CUDACC__
}  // namespace b200

// A resident base set (b200_bases_upload): the bases, or their folded-window table.
struct b200_bases {
    int group = 1;
    size_t n = 0;
    b200::DevBuf pts;
    b200::MsmFold fold;     // c == 0: plain
};
