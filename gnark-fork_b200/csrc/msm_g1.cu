// G1 instantiation of the MSM pipeline (G1Jac.MultiExp, prove.go:201,214,234,244).
#include "msm_host.cuh"
namespace b200 {
size_t msm_workspace_bytes(size_t n, int group, int c_override, int fold_c) {
    if (n == 0) return 0;
    return group == 2 ? MsmLayout<Fp2>(n, c_override, fold_c).total : MsmLayout<Fp>(n, c_override, fold_c).total;
}
int msm_ba_rounds(size_t n, int c_override, int fold_c) { return n ? MsmLayout<Fp>(n, c_override, fold_c).ba : 0; }
MsmSizes msm_sizes(size_t n, int group, int c_override, int fold_c) {
    MsmSizes z{0, 0, 0, 0};
    if (n == 0) return z;
    if (group == 2) { MsmLayout<Fp2> l(n, c_override, fold_c); z = MsmSizes{l.sort_total, l.acc_total, l.result_points() * sizeof(XYZZ<Fp2>), l.ba_total}; }
    else { MsmLayout<Fp> l(n, c_override, fold_c); z = MsmSizes{l.sort_total, l.acc_total, l.result_points() * sizeof(XYZZ<Fp>), l.ba_total}; }
    return z;
}
int msm_g1_run(b200_ctx* ctx, const G1Affine* d_bases, const Fr* d_scalars, const uint32_t* d_idx_map, size_t n,
               G1Jac* out_host, void* ws, const MsmFold* fold) {
    return msm_run_impl<Fp>(ctx, d_bases, d_scalars, d_idx_map, n, out_host, ws, fold);
}
int msm_fold_build_g1(b200_ctx* ctx, const G1Affine* d_bases, size_t n, int c, size_t reserve_bytes, DevBuf& table) {
    return msm_fold_build_impl<Fp>(ctx, d_bases, n, c, reserve_bytes, table);
}
void* msm_g1_job_new(b200_ctx* ctx, const G1Affine* d_bases, const Fr* d_scalars, const uint32_t* d_idx_map, size_t n,
                     const MsmFold* fold, void* sort_ws, void* acc_ws, void* host_staging, void* ba_ws) {
    return new MsmJob<Fp>(ctx, d_bases, d_scalars, d_idx_map, n, fold, sort_ws, acc_ws, (XYZZ<Fp>*)host_staging, ba_ws);
}
int msm_g1_job_sort(b200_ctx* ctx, void* job, cudaStream_t st) { return msm_enqueue_sort(ctx, *(MsmJob<Fp>*)job, st); }
int msm_g1_job_acc(b200_ctx* ctx, void* job, cudaStream_t st, cudaStream_t st_post, cudaEvent_t ev_split) { return msm_enqueue_acc(ctx, *(MsmJob<Fp>*)job, st, st_post, ev_split); }
void msm_g1_job_finish(void* job, G1Jac* out_host) { msm_finish(*(MsmJob<Fp>*)job, out_host); }
void msm_g1_job_delete(void* job) { delete (MsmJob<Fp>*)job; }
}  // namespace b200
