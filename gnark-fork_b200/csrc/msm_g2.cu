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
}  // namespace b200
