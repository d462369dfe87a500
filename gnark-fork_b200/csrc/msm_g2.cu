// G2 instantiation of the MSM pipeline (G2Jac.MultiExp, prove.go:290).
#include "msm_host.cuh"
namespace b200 {
int msm_g2_run(b200_ctx* ctx, const G2Affine* d_bases, const Fr* d_scalars, const uint32_t* d_idx_map, size_t n,
               G2Jac* out_host, void* ws, const MsmFold* fold) {
    return msm_run_impl<Fp2>(ctx, d_bases, d_scalars, d_idx_map, n, out_host, ws, fold);
}
int msm_fold_build_g2(b200_ctx* ctx, const G2Affine* d_bases, size_t n, int c, size_t reserve_bytes, DevBuf& table) {
    return msm_fold_build_impl<Fp2>(ctx, d_bases, n, c, reserve_bytes, table);
}
void* msm_g2_job_new(b200_ctx* ctx, const G2Affine* d_bases, const Fr* d_scalars, const uint32_t* d_idx_map, size_t n,
                     const MsmFold* fold, void* sort_ws, void* acc_ws, void* host_staging, void* ba_ws) {
    return new MsmJob<Fp2>(ctx, d_bases, d_scalars, d_idx_map, n, fold, sort_ws, acc_ws, (XYZZ<Fp2>*)host_staging, ba_ws);
}
int msm_g2_job_sort(b200_ctx* ctx, void* job, cudaStream_t st) { return msm_enqueue_sort(ctx, *(MsmJob<Fp2>*)job, st); }
int msm_g2_job_prep0(b200_ctx* ctx, void* job, cudaStream_t st) { return msm_enqueue_prep0(ctx, *(MsmJob<Fp2>*)job, st); }
int msm_g2_job_acc(b200_ctx* ctx, void* job, cudaStream_t st, cudaStream_t st_post, cudaEvent_t ev_split) { return msm_enqueue_acc(ctx, *(MsmJob<Fp2>*)job, st, st_post, ev_split); }
void msm_g2_job_finish(void* job, G2Jac* out_host) { msm_finish(*(MsmJob<Fp2>*)job, out_host); }
void msm_g2_job_delete(void* job) { delete (MsmJob<Fp2>*)job; }
}  // namespace b200
