// Host driver of the MSM pipeline (template over the coordinate field: Fp for G1, Fp2 for G2).
#pragma once
#include "msm.cuh"

namespace b200 {

// msm_g1.cu / msm_g2.cu
// d_bases: n affine points on the device; d_scalars: Montgomery Fr on the device, read through
// d_idx_map when it is not null (scalar i = d_scalars[d_idx_map[i]]).  The workspace is carved
// from `ws` when ws != nullptr (the caller reserved msm_workspace_bytes), else from the ctx.
size_t msm_workspace_bytes(size_t n, int group, int c_override);
int msm_g1_run(b200_ctx* ctx, const G1Affine* d_bases, const Fr* d_scalars, const uint32_t* d_idx_map, size_t n,
               G1Jac* out_host, void* ws);
int msm_g2_run(b200_ctx* ctx, const G2Affine* d_bases, const Fr* d_scalars, const uint32_t* d_idx_map, size_t n,
               G2Jac* out_host, void* ws);

#if defined(__CUDACC__)

template <class F>
struct MsmLayout {
    MsmPlan plan;
    size_t n, M;  // M = W * nb buckets
    size_t off_digits, off_sorted, off_counts, off_offsets, off_cursor, off_bsums, off_buckets, off_S[2], off_W[2], total;
    size_t off_ntasks, off_task_off, off_task_bucket, off_partials, off_order, off_order_hist;
    uint32_t scan_blocks, chunk, max_tasks, order_shift;
    MsmLayout(size_t n_, int c_override) {
        n = n_;
        plan = msm_plan(n, c_override);
        M = (size_t)plan.W * plan.nb;
        Arena a;
        off_digits = a.off; a.take((size_t)plan.W * n * 4);
        off_sorted = a.off; a.take((size_t)plan.W * n * 4);
        off_counts = a.off; a.take(M * 4);
        off_offsets = a.off; a.take(M * 4);
        off_cursor = a.off; a.take(M * 4);
        scan_blocks = (uint32_t)((M + SCAN_CHUNK - 1) / SCAN_CHUNK);
        off_bsums = a.off; a.take((size_t)scan_blocks * 4);
        off_buckets = a.off; a.take(M * sizeof(XYZZ<F>));
        // task splitting: chunk = twice the mean bucket load (at least 64 points)
        const size_t mean = n / plan.nb;
        chunk = (uint32_t)(2 * mean < 64 ? 64 : 2 * mean);
        max_tasks = (uint32_t)(M + ((size_t)plan.W * n) / chunk + 1);
        off_ntasks = a.off; a.take(M * 4);
        off_task_off = a.off; a.take(M * 4);
        off_task_bucket = a.off; a.take((size_t)max_tasks * 4);
        off_partials = a.off; a.take((size_t)max_tasks * sizeof(XYZZ<F>));
        off_order = a.off; a.take((size_t)max_tasks * 4);
        off_order_hist = a.off; a.take((size_t)ORDER_BINS * 4 * 2);
        order_shift = 0;
        while ((chunk >> order_shift) > (uint32_t)ORDER_BINS - 2) ++order_shift;
        const size_t nodes0 = M / plan.L0;
        for (int i = 0; i < 2; ++i) {
            off_S[i] = a.off; a.take((i == 0 ? nodes0 : nodes0 / 2 + 1) * sizeof(XYZZ<F>));
            off_W[i] = a.off; a.take((i == 0 ? nodes0 : nodes0 / 2 + 1) * sizeof(XYZZ<F>));
        }
        total = a.off;
    }
};

template <class F>
int msm_run_impl(b200_ctx* ctx, const Affine<F>* d_bases, const Fr* d_scalars, const uint32_t* d_idx_map, size_t n,
                 Jac<F>* out_host, void* ws) {
    if (n == 0) { *out_host = Jac<F>::inf(); return B200_OK; }
    if (n >= ((size_t)1 << 31)) return ctx->fail(B200_ERR_UNSUPPORTED, "MSM above 2^31 points");
    MsmLayout<F> lay(n, ctx->msm_c_override);
    const MsmPlan& pl = lay.plan;
    char* base;
    if (ws) base = (char*)ws;
    else { B2_TRY(ctx_reserve(ctx, lay.total)); base = (char*)ctx->scratch; }
    int32_t* digits = (int32_t*)(base + lay.off_digits);
    uint32_t* sorted = (uint32_t*)(base + lay.off_sorted);
    uint32_t* counts = (uint32_t*)(base + lay.off_counts);
    uint32_t* offsets = (uint32_t*)(base + lay.off_offsets);
    uint32_t* cursor = (uint32_t*)(base + lay.off_cursor);
    uint32_t* bsums = (uint32_t*)(base + lay.off_bsums);
    XYZZ<F>* buckets = (XYZZ<F>*)(base + lay.off_buckets);
    uint32_t* ntasks = (uint32_t*)(base + lay.off_ntasks);
    uint32_t* task_off = (uint32_t*)(base + lay.off_task_off);
    uint32_t* task_bucket = (uint32_t*)(base + lay.off_task_bucket);
    XYZZ<F>* partials = (XYZZ<F>*)(base + lay.off_partials);
    uint32_t* order = (uint32_t*)(base + lay.off_order);
    uint32_t* order_hist = (uint32_t*)(base + lay.off_order_hist);
    XYZZ<F>* S[2] = {(XYZZ<F>*)(base + lay.off_S[0]), (XYZZ<F>*)(base + lay.off_S[1])};
    XYZZ<F>* Wt[2] = {(XYZZ<F>*)(base + lay.off_W[0]), (XYZZ<F>*)(base + lay.off_W[1])};
    cudaStream_t st = ctx->stream;
    const uint32_t M = (uint32_t)lay.M;

    constexpr int ACC_SLOT = sizeof(F) == sizeof(Fp) ? B200_PROF_MSM_ACC_G1 : B200_PROF_MSM_ACC_G2;
    int span = ctx->prof_begin(B200_PROF_MSM_SORT, (double)n);
    B2_CUDA(ctx, cudaMemsetAsync(counts, 0, lay.M * 4, st));
    {
        const int bs = 256;
        msm_digits_kernel<<<(unsigned)((n + bs - 1) / bs), bs, 0, st>>>(d_scalars, d_idx_map, (uint32_t)n, pl.c, pl.W, digits, counts, pl.nb);
        B2_CUDA(ctx, cudaGetLastError());
    }
    msm_scan_sums_kernel<<<lay.scan_blocks, SCAN_THREADS, 0, st>>>(counts, M, bsums);
    B2_CUDA(ctx, cudaGetLastError());
    msm_scan_top_kernel<<<1, 1024, 0, st>>>(bsums, lay.scan_blocks);
    B2_CUDA(ctx, cudaGetLastError());
    msm_scan_final_kernel<<<lay.scan_blocks, SCAN_THREADS, 0, st>>>(counts, M, bsums, offsets, cursor);
    B2_CUDA(ctx, cudaGetLastError());
    {
        const int bs = 256;
        msm_scatter_kernel<<<(unsigned)((n + bs - 1) / bs), bs, 0, st>>>(digits, (uint32_t)n, pl.c, pl.W, pl.nb, cursor, sorted);
        B2_CUDA(ctx, cudaGetLastError());
    }
    {
        const int bs = 256;
        msm_task_count_kernel<<<(M + bs - 1) / bs, bs, 0, st>>>(counts, M, lay.chunk, ntasks);
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
        msm_order_hist_kernel<<<ob, ORDER_THREADS, 0, st>>>(task_bucket, task_off, ntasks, counts, lay.max_tasks, lay.chunk, lay.order_shift, order_hist);
        B2_CUDA(ctx, cudaGetLastError());
        msm_order_scan_kernel<<<1, 1024, 0, st>>>(order_hist);
        B2_CUDA(ctx, cudaGetLastError());
        msm_order_fill_kernel<<<ob, ORDER_THREADS, 0, st>>>(task_bucket, task_off, ntasks, counts, lay.max_tasks, lay.chunk, lay.order_shift, order_hist, order);
        B2_CUDA(ctx, cudaGetLastError());
    }
    ctx->prof_end(span);
    {
        const int bs = 128;
        span = ctx->prof_begin(ACC_SLOT, (double)n);
        msm_accumulate_kernel<F><<<(lay.max_tasks + bs - 1) / bs, bs, 0, st>>>(d_bases, sorted, offsets, counts, task_bucket, task_off,
                                                                            ntasks, order, lay.max_tasks, lay.chunk, buckets, partials);
        ctx->prof_end(span);
        B2_CUDA(ctx, cudaGetLastError());
        span = ctx->prof_begin(B200_PROF_MSM_REDUCE, (double)n);
        msm_combine_kernel<F><<<(M + bs - 1) / bs, bs, 0, st>>>(task_off, ntasks, M, partials, buckets);
        B2_CUDA(ctx, cudaGetLastError());
    }
    // reduction tree
    uint32_t T = pl.nb / pl.L0;  // nodes per window
    uint32_t log_len = 0;
    while ((1u << log_len) < pl.L0) ++log_len;
    {
        const uint32_t nodes = (uint32_t)pl.W * T;
        msm_reduce_leaf_kernel<F><<<(nodes + 63) / 64, 64, 0, st>>>(buckets, nodes, pl.L0, S[0], Wt[0]);
        B2_CUDA(ctx, cudaGetLastError());
    }
    int cur = 0;
    ctx->n_launches += 16;  // digits, 2 x 3 scan, scatter, task count/fill, 3 x order, accumulate, combine, leaf
    while (T > 1) {
        ctx->n_launches += 1;
        const uint32_t L = T < pl.L ? T : pl.L;
        uint32_t logL = 0;
        while ((1u << logL) < L) ++logL;
        const uint32_t nodes = (uint32_t)pl.W * (T / L);
        msm_reduce_node_kernel<F><<<(nodes + 63) / 64, 64, 0, st>>>(S[cur], Wt[cur], nodes, L, log_len, S[cur ^ 1], Wt[cur ^ 1]);
        B2_CUDA(ctx, cudaGetLastError());
        T /= L;
        log_len += logL;
        cur ^= 1;
    }
    ctx->prof_end(span);
    // window sums -> host, Horner combine (c doublings per window)
    XYZZ<F> win[32];
    B2_CUDA(ctx, cudaMemcpyAsync(win, Wt[cur], (size_t)pl.W * sizeof(XYZZ<F>), cudaMemcpyDeviceToHost, st));
    B2_CUDA(ctx, cudaStreamSynchronize(st));
    XYZZ<F> acc = XYZZ<F>::inf();
    for (int w = pl.W - 1; w >= 0; --w) {
        for (int k = 0; k < pl.c; ++k) acc = acc.dbl();
        acc.add(win[w]);
    }
    *out_host = acc.to_jac();
    return B200_OK;
}

#endif  // __CUDACC__
}  // namespace b200
