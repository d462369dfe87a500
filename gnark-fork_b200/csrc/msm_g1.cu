// G1 instantiation of the MSM pipeline (G1Jac.MultiExp, prove.go:201,214,234,244).
#include "msm_host.cuh"
namespace b200 {
size_t msm_workspace_bytes(size_t n, int group, int c_override, int fold_c) {
    if (n == 0) return 0;
    return group == 2 ? MsmLayout<Fp2>(n, c_override, fold_c).total : MsmLayout<Fp>(n, c_override, fold_c).total;
}
int msm_g1_run(b200_ctx* ctx, const G1Affine* d_bases, const Fr* d_scalars, const uint32_t* d_idx_map, size_t n,
               G1Jac* out_host, void* ws, const MsmFold* fold) {
    return msm_run_impl<Fp>(ctx, d_bases, d_scalars, d_idx_map, n, out_host, ws, fold);
}
int msm_fold_build_g1(b200_ctx* ctx, const G1Affine* d_bases, size_t n, int c, size_t reserve_bytes, DevBuf& table) {
    return msm_fold_build_impl<Fp>(ctx, d_bases, n, c, reserve_bytes, table);
}
}  // namespace b200
