// extern "C" surface of libb200prover.so (declared in include/b200prover.h).  Thin: argument
// checks, host<->device staging, and calls into ntt.cu / msm_g*.cu / groth16.cu / misc.cu.
// No CPU fallback exists anywhere below this file: without a CUDA device b200_init fails.
#include <cstdlib>
#include "groth16_host.cuh"
#include "poly_host.cuh"
#include "misc_host.cuh"

using namespace b200;

namespace {
thread_local std::string g_init_error;

int set_device(b200_ctx* ctx) {
    cudaError_t e = cudaSetDevice(ctx->device);
    if (e != cudaSuccess) return ctx->fail(B200_ERR_CUDA, "cudaSetDevice", cudaGetErrorString(e));
    return B200_OK;
}
#if defined(B2_CURVE_BLS12_377)
constexpr int BUILD_CURVE = B200_BLS12_377;
#else
constexpr int BUILD_CURVE = B200_BN254;
#endif
int check_curve(b200_ctx* ctx, int curve) {
    if (curve != BUILD_CURVE)
        return ctx->fail(B200_ERR_UNSUPPORTED, BUILD_CURVE == B200_BN254 ? "this library is the BN254 build (BLS12-377: libb200prover_bls12377.so)"
                                                                         : "this library is the BLS12-377 build (BN254: libb200prover.so)");
    return B200_OK;
}
#define B2_ENTER(ctx, curve)                   \
    if (!(ctx)) return B200_ERR_ARG;           \
    B2_TRY(set_device(ctx));                   \
    B2_TRY(check_curve(ctx, curve))
}  // namespace

extern "C" {

const char* b200_version(void) try { return BUILD_CURVE == B200_BN254 ? "b200prover 0.1 (sm_100a, BN254)" : "b200prover 0.1 (sm_100a, BLS12-377)"; } B2_ABI_CATCH_NULL
int b200_build_curve(void) try { return BUILD_CURVE; } B2_ABI_CATCH_INT
int b200_build_fp_words(void) try { return (int)(sizeof(Fp) / 8); } B2_ABI_CATCH_INT

int b200_device_count(void) try {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
} B2_ABI_CATCH_INT

int b200_init(int device, b200_ctx** out) try {
    if (!out) return B200_ERR_ARG;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0 || device < 0 || device >= n) {
        g_init_error = e != cudaSuccess ? cudaGetErrorString(e) : "no such CUDA device";
        cudaGetLastError();
        return B200_ERR_CUDA;
    }
    if (cudaSetDevice(device) != cudaSuccess) return B200_ERR_CUDA;
    b200_ctx* ctx = new b200_ctx();
    ctx->device = device;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { delete ctx; return B200_ERR_CUDA; }
    ctx->sm_count = prop.multiProcessorCount;
    int prio_lo = 0, prio_hi = 0;      // the reduction tail of an MSM is a few small kernels: its blocks go first when SMs free up
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
    if (cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&ctx->aux_stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithPriority(&ctx->post_stream, cudaStreamNonBlocking, prio_hi) != cudaSuccess ||
        cudaEventCreate(&ctx->ev0) != cudaSuccess || cudaEventCreate(&ctx->ev1) != cudaSuccess) {
        delete ctx;
        return B200_ERR_CUDA;
    }
    ctx->stream = ctx->own_stream;
    const char* c = getenv("B200_MSM_C");
    if (c && *c) { int v = atoi(c); if (v >= 4 && v <= 24) ctx->msm_c_override = v; }
    *out = ctx;
    return B200_OK;
} B2_ABI_CATCH_INT

void b200_destroy(b200_ctx* ctx) try {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    ntt_domain_free_all(ctx);
    if (ctx->scratch) cudaFree(ctx->scratch);
    if (ctx->poly_scratch) cudaFree(ctx->poly_scratch);
    for (auto& sp : ctx->prof_spans) { cudaEventDestroy(sp.e0); cudaEventDestroy(sp.e1); }
    for (auto e : ctx->prof_pool) cudaEventDestroy(e);
    cudaEventDestroy(ctx->ev0);
    cudaEventDestroy(ctx->ev1);
    cudaStreamDestroy(ctx->own_stream);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    if (ctx->aux_stream) cudaStreamDestroy(ctx->aux_stream);
    if (ctx->post_stream) cudaStreamDestroy(ctx->post_stream);
    if (ctx->h_stage) cudaFreeHost(ctx->h_stage);
    for (cudaEvent_t e : ctx->core_events) cudaEventDestroy(e);
    delete ctx;
} B2_ABI_CATCH_VOID

const char* b200_last_error(const b200_ctx* ctx) try { return ctx ? ctx->err.c_str() : g_init_error.c_str(); } B2_ABI_CATCH_NULL

void* b200_host_alloc(size_t bytes) try {
    void* p = nullptr;
    if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
} B2_ABI_CATCH_NULL
void b200_host_free(void* p) try { if (p) cudaFreeHost(p); } B2_ABI_CATCH_VOID

int b200_dev_alloc(b200_ctx* ctx, size_t bytes, void** dptr) try {
    B2_ENTER(ctx, BUILD_CURVE);
    if (!dptr) return ctx->fail(B200_ERR_ARG, "null out pointer");
    B2_CUDA(ctx, cudaMalloc(dptr, bytes ? bytes : 16));
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)
int b200_dev_free(b200_ctx* ctx, void* dptr) try {
    B2_ENTER(ctx, BUILD_CURVE);
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    B2_CUDA(ctx, cudaFree(dptr));
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)
int b200_h2d(b200_ctx* ctx, void* dst, const void* src, size_t bytes) try {
    B2_ENTER(ctx, BUILD_CURVE);
    B2_CUDA(ctx, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)
int b200_d2h(b200_ctx* ctx, void* dst, const void* src, size_t bytes) try {
    B2_ENTER(ctx, BUILD_CURVE);
    B2_CUDA(ctx, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)
int b200_sync(b200_ctx* ctx) try {
    B2_ENTER(ctx, BUILD_CURVE);
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)

// ---- element-wise ------------------------------------------------------------------------
static int vec_mul_host(b200_ctx* ctx, int fr, const uint64_t* a, const uint64_t* b, uint64_t* out, size_t n) {
    if (n == 0) return B200_OK;
    if (!a || !b || !out) return ctx->fail(B200_ERR_ARG, "null vector");
    DevBuf da, db;
    const size_t bytes = n * (fr ? sizeof(Fr) : sizeof(Fp));
    B2_CUDA(ctx, da.alloc(bytes));
    B2_CUDA(ctx, db.alloc(bytes));
    B2_CUDA(ctx, cudaMemcpyAsync(da.p, a, bytes, cudaMemcpyHostToDevice, ctx->stream));
    B2_CUDA(ctx, cudaMemcpyAsync(db.p, b, bytes, cudaMemcpyHostToDevice, ctx->stream));
    if (fr) B2_TRY(vec_mul_fr(ctx, da.as<Fr>(), db.as<Fr>(), da.as<Fr>(), n));
    else B2_TRY(vec_mul_fp(ctx, da.as<Fp>(), db.as<Fp>(), da.as<Fp>(), n));
    B2_CUDA(ctx, cudaMemcpyAsync(out, da.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B200_OK;
}
int b200_fr_vec_mul(b200_ctx* ctx, int curve, const uint64_t* a, const uint64_t* b, uint64_t* out, size_t n) try {
    B2_ENTER(ctx, curve);
    return vec_mul_host(ctx, 1, a, b, out, n);
} B2_ABI_CATCH_CTX(ctx)
int b200_fp_vec_mul(b200_ctx* ctx, int curve, const uint64_t* a, const uint64_t* b, uint64_t* out, size_t n) try {
    B2_ENTER(ctx, curve);
    return vec_mul_host(ctx, 0, a, b, out, n);
} B2_ABI_CATCH_CTX(ctx)
int b200_imad_peak(b200_ctx* ctx, double* mac32_per_s, double* imad32_per_s, double* sm_mhz) try {
    B2_ENTER(ctx, BUILD_CURVE);
    return imad_peak(ctx, mac32_per_s, imad32_per_s, sm_mhz);
} B2_ABI_CATCH_CTX(ctx)

// ---- NTT ---------------------------------------------------------------------------------
static int get_domain(b200_ctx* ctx, unsigned log_n, const uint64_t* generator, const uint64_t* coset_gen, NttDomain** d) {
    if (!generator) return ctx->fail(B200_ERR_ARG, "generator is required");
    Fr gen, coset;
    memcpy(gen.l, generator, 32);
    if (coset_gen) memcpy(coset.l, coset_gen, 32); else coset = Fr::one();
    return ntt_domain_get(ctx, log_n, gen, coset, d);
}

int b200_ntt_dev(b200_ctx* ctx, int curve, void* dev_inout, unsigned log_n, int inverse, int decimation, int on_coset,
                 const uint64_t* generator, const uint64_t* coset_gen, float* kernel_ms) try {
    B2_ENTER(ctx, curve);
    if (!dev_inout) return ctx->fail(B200_ERR_ARG, "null vector");
    if (decimation != B200_DIF && decimation != B200_DIT) return ctx->fail(B200_ERR_ARG, "bad decimation");
    if (on_coset && !coset_gen) return ctx->fail(B200_ERR_ARG, "coset_gen is required with on_coset");
    NttDomain* d;
    B2_TRY(get_domain(ctx, log_n, generator, coset_gen, &d));
    if (on_coset) {  // build the table outside the timed region
        const Fr* t;
        NttDomain::Table id = !inverse ? (decimation ? NttDomain::COSET_BR : NttDomain::COSET_NAT)
                                       : (decimation ? NttDomain::COSETINV_NAT : NttDomain::COSETINV_BR);
        B2_TRY(ntt_domain_table(ctx, d, id, &t));
    }
    B2_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    B2_TRY(ntt_gnark(ctx, d, (Fr*)dev_inout, inverse, decimation, on_coset));
    B2_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    B2_CUDA(ctx, cudaEventSynchronize(ctx->ev1));
    if (kernel_ms) B2_CUDA(ctx, cudaEventElapsedTime(kernel_ms, ctx->ev0, ctx->ev1));
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)

int b200_ntt(b200_ctx* ctx, int curve, uint64_t* host_inout, unsigned log_n, int inverse, int decimation, int on_coset,
             const uint64_t* generator, const uint64_t* coset_gen) try {
    B2_ENTER(ctx, curve);
    if (!host_inout) return ctx->fail(B200_ERR_ARG, "null vector");
    if (log_n > 27) return ctx->fail(B200_ERR_UNSUPPORTED, "NTT size above 2^27");
    const size_t bytes = ((size_t)1 << log_n) * sizeof(Fr);
    DevBuf d;
    B2_CUDA(ctx, d.alloc(bytes));
    B2_CUDA(ctx, cudaMemcpyAsync(d.p, host_inout, bytes, cudaMemcpyHostToDevice, ctx->stream));
    B2_TRY(b200_ntt_dev(ctx, curve, d.p, log_n, inverse, decimation, on_coset, generator, coset_gen, nullptr));
    B2_CUDA(ctx, cudaMemcpyAsync(host_inout, d.p, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)

int b200_groth16_compute_h(b200_ctx* ctx, int curve, const uint64_t* a, const uint64_t* b, const uint64_t* c,
                           size_t n_constraints, unsigned log_n, const uint64_t* generator, const uint64_t* coset_gen,
                           uint64_t* h_out) try {
    B2_ENTER(ctx, curve);
    if (!a || !b || !c || !h_out || !coset_gen) return ctx->fail(B200_ERR_ARG, "null argument");
    if (log_n > 27) return ctx->fail(B200_ERR_UNSUPPORTED, "NTT size above 2^27");
    const size_t n = (size_t)1 << log_n;
    if (n_constraints > n) return ctx->fail(B200_ERR_ARG, "more constraints than the domain holds");
    NttDomain* d;
    B2_TRY(get_domain(ctx, log_n, generator, coset_gen, &d));
    DevBuf buf[3];
    const uint64_t* src[3] = {a, b, c};
    for (int i = 0; i < 3; ++i) {
        B2_CUDA(ctx, buf[i].alloc(n * sizeof(Fr)));
        B2_CUDA(ctx, cudaMemcpyAsync(buf[i].p, src[i], n_constraints * sizeof(Fr), cudaMemcpyHostToDevice, ctx->stream));
        if (n > n_constraints)
            B2_CUDA(ctx, cudaMemsetAsync(buf[i].as<Fr>() + n_constraints, 0, (n - n_constraints) * sizeof(Fr), ctx->stream));
    }
    B2_TRY(ntt_compute_h(ctx, d, buf[0].as<Fr>(), buf[1].as<Fr>(), buf[2].as<Fr>()));
    B2_CUDA(ctx, cudaMemcpyAsync(h_out, buf[0].p, n * sizeof(Fr), cudaMemcpyDeviceToHost, ctx->stream));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)

// ---- MSM ---------------------------------------------------------------------------------
// H2D of n points (raw != 0: gnark-crypto RawEncoding bytes, converted in place on the device) into dst
static int copy_points(b200_ctx* ctx, DevBuf& dst, const void* host, size_t n, int group, int raw) {
    const size_t bytes = n * (group == 1 ? sizeof(G1Affine) : sizeof(G2Affine));
    B2_CUDA(ctx, dst.alloc(bytes));
    if (!bytes) return B200_OK;
    B2_CUDA(ctx, cudaMemcpyAsync(dst.p, host, bytes, cudaMemcpyHostToDevice, ctx->stream));
    if (raw) {
        DevBuf bad;
        B2_CUDA(ctx, bad.alloc(sizeof(uint32_t)));
        B2_CUDA(ctx, cudaMemsetAsync(bad.p, 0, sizeof(uint32_t), ctx->stream));
        B2_TRY(points_from_raw_inplace(ctx, dst.p, n, group, bad.as<uint32_t>()));
        uint32_t flag = 0;
        B2_CUDA(ctx, cudaMemcpyAsync(&flag, bad.p, sizeof flag, cudaMemcpyDeviceToHost, ctx->stream));
        B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (flag) return ctx->fail(B200_ERR_ARG, "not a raw-encoded point slice (compressed-form flag or coordinate >= p)");
    }
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B200_OK;
}

int b200_points_from_raw(b200_ctx* ctx, int curve, int group, const uint8_t* raw, size_t n, uint64_t* out_affine) try {
    B2_ENTER(ctx, curve);
    if (n && (!raw || !out_affine)) return ctx->fail(B200_ERR_ARG, "null argument");
    if (group != 1 && group != 2) return ctx->fail(B200_ERR_ARG, "group must be 1 or 2");
    DevBuf d;
    B2_TRY(copy_points(ctx, d, raw, n, group, 1));
    if (n) B2_CUDA(ctx, cudaMemcpyAsync(out_affine, d.p, d.bytes, cudaMemcpyDeviceToHost, ctx->stream));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)

static int bases_upload_impl(b200_ctx* ctx, int group, const void* host, size_t n, int raw, b200_bases** out);

int b200_bases_upload(b200_ctx* ctx, int curve, int group, const uint64_t* bases_affine, size_t n, b200_bases** out) try {
    B2_ENTER(ctx, curve);
    return bases_upload_impl(ctx, group, bases_affine, n, 0, out);
} B2_ABI_CATCH_CTX(ctx)
int b200_bases_upload_raw(b200_ctx* ctx, int curve, int group, const uint8_t* raw, size_t n, b200_bases** out) try {
    B2_ENTER(ctx, curve);
    return bases_upload_impl(ctx, group, raw, n, 1, out);
} B2_ABI_CATCH_CTX(ctx)

static int bases_upload_impl(b200_ctx* ctx, int group, const void* host, size_t n, int raw, b200_bases** out) {
    if (!out || (n && !host)) return ctx->fail(B200_ERR_ARG, "null argument");
    if (group != 1 && group != 2) return ctx->fail(B200_ERR_ARG, "group must be 1 or 2");
    b200_bases* b = new b200_bases();
    b->group = group;
    b->n = n;
    { const int st = copy_points(ctx, b->pts, host, n, group, raw); if (st != B200_OK) { delete b; return st; } }
    const int c = ctx->msm_fold_c ? ctx->msm_fold_c : msm_fold_window(n);
    if (ctx->msm_precompute && n >= 4096 && c != 0) {
        DevBuf table;
        const size_t reserve = msm_workspace_bytes(n, group, ctx->msm_mode(), c) + n * sizeof(Fr) + ((size_t)1 << 30);
        const int st = group == 1 ? msm_fold_build_g1(ctx, b->pts.as<G1Affine>(), n, c, reserve, table)
                                  : msm_fold_build_g2(ctx, b->pts.as<G2Affine>(), n, c, reserve, table);
        if (st == B200_OK) {
            b->pts.release();
            b->pts.p = table.p; b->pts.bytes = table.bytes;
            table.p = nullptr; table.bytes = 0;
            b->fold.c = c;
            b->fold.stride = n;
        } else if (st != B200_ERR_NOMEM) { delete b; return st; }
    }
    *out = b;
    return B200_OK;
}
int b200_bases_free(b200_ctx* ctx, b200_bases* bases) try {
    B2_ENTER(ctx, BUILD_CURVE);
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    delete bases;
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)

static void fill_set_info(b200_ctx* ctx, size_t n, size_t bytes, const MsmFold& f, b200_msm_set_info* o) {
    const MsmPlan p = f.c ? msm_plan_folded(f.c) : msm_plan(n, ctx->msm_c_override);
    o->n = n; o->device_bytes = bytes; o->c = p.c; o->windows = p.W; o->folded = p.fold;
    o->ba_rounds = msm_ba_rounds(n, ctx->msm_mode(), f.c);
}
int b200_bases_info(b200_ctx* ctx, const b200_bases* bases, b200_msm_set_info* out) try {
    if (!ctx) return B200_ERR_ARG;
    if (!bases || !out) return ctx->fail(B200_ERR_ARG, "null argument");
    fill_set_info(ctx, bases->n, bases->pts.bytes, bases->fold, out);
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)

int b200_msm_dev(b200_ctx* ctx, const b200_bases* bases, const void* scalars_dev, size_t n, uint64_t* out_jac, float* kernel_ms) try {
    B2_ENTER(ctx, BUILD_CURVE);
    if (!bases || !out_jac || (n && !scalars_dev)) return ctx->fail(B200_ERR_ARG, "null argument");
    if (n > bases->n) return ctx->fail(B200_ERR_ARG, "more scalars than resident bases");
    B2_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    if (bases->group == 1) {
        G1Jac r;
        B2_TRY(msm_g1_run(ctx, bases->pts.as<G1Affine>(), (const Fr*)scalars_dev, nullptr, n, &r, nullptr, &bases->fold));
        memcpy(out_jac, &r, sizeof r);
    } else {
        G2Jac r;
        B2_TRY(msm_g2_run(ctx, bases->pts.as<G2Affine>(), (const Fr*)scalars_dev, nullptr, n, &r, nullptr, &bases->fold));
        memcpy(out_jac, &r, sizeof r);
    }
    B2_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    B2_CUDA(ctx, cudaEventSynchronize(ctx->ev1));
    if (kernel_ms) B2_CUDA(ctx, cudaEventElapsedTime(kernel_ms, ctx->ev0, ctx->ev1));
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)

int b200_msm_resident(b200_ctx* ctx, const b200_bases* bases, const uint64_t* scalars, size_t n, uint64_t* out_jac) try {
    B2_ENTER(ctx, BUILD_CURVE);
    if (n && !scalars) return ctx->fail(B200_ERR_ARG, "null scalars");
    DevBuf ds;
    B2_CUDA(ctx, ds.alloc(n * sizeof(Fr)));
    if (n) B2_CUDA(ctx, cudaMemcpyAsync(ds.p, scalars, n * sizeof(Fr), cudaMemcpyHostToDevice, ctx->stream));
    return b200_msm_dev(ctx, bases, ds.p, n, out_jac, nullptr);
} B2_ABI_CATCH_CTX(ctx)

static int msm_host(b200_ctx* ctx, int curve, int group, const uint64_t* scalars, const uint64_t* bases_affine, size_t n, uint64_t* out_jac) {
    B2_ENTER(ctx, curve);
    if (!out_jac) return ctx->fail(B200_ERR_ARG, "null output");
    b200_bases* b = nullptr;
    const int saved = ctx->msm_precompute;
    ctx->msm_precompute = 0;   // a one-shot MSM cannot amortise the table
    const int up = b200_bases_upload(ctx, curve, group, bases_affine, n, &b);
    ctx->msm_precompute = saved;
    B2_TRY(up);
    const int st = b200_msm_resident(ctx, b, scalars, n, out_jac);
    cudaStreamSynchronize(ctx->stream);
    delete b;
    return st;
}
int b200_msm_g1(b200_ctx* ctx, int curve, const uint64_t* scalars, const uint64_t* bases_affine, size_t n, uint64_t* out_jac) try {
    return msm_host(ctx, curve, 1, scalars, bases_affine, n, out_jac);
} B2_ABI_CATCH_CTX(ctx)
int b200_msm_g2(b200_ctx* ctx, int curve, const uint64_t* scalars, const uint64_t* bases_affine, size_t n, uint64_t* out_jac) try {
    return msm_host(ctx, curve, 2, scalars, bases_affine, n, out_jac);
} B2_ABI_CATCH_CTX(ctx)

int b200_batch_scalar_mul(b200_ctx* ctx, int curve, int group, const uint64_t* base_affine, const uint64_t* scalars,
                          size_t n, uint64_t* out_affine) try {
    B2_ENTER(ctx, curve);
    if (!base_affine || (n && (!scalars || !out_affine))) return ctx->fail(B200_ERR_ARG, "null argument");
    if (group != 1 && group != 2) return ctx->fail(B200_ERR_ARG, "group must be 1 or 2");
    const size_t psz = group == 1 ? sizeof(G1Affine) : sizeof(G2Affine);
    DevBuf ds, dout;
    B2_CUDA(ctx, ds.alloc(n * sizeof(Fr)));
    B2_CUDA(ctx, dout.alloc(n * psz));
    if (n) B2_CUDA(ctx, cudaMemcpyAsync(ds.p, scalars, n * sizeof(Fr), cudaMemcpyHostToDevice, ctx->stream));
    if (group == 1) {
        G1Affine b; memcpy(&b, base_affine, sizeof b);
        B2_TRY(batch_mul_g1(ctx, b, ds.as<Fr>(), n, dout.as<G1Affine>()));
    } else {
        G2Affine b; memcpy(&b, base_affine, sizeof b);
        B2_TRY(batch_mul_g2(ctx, b, ds.as<Fr>(), n, dout.as<G2Affine>()));
    }
    if (n) B2_CUDA(ctx, cudaMemcpyAsync(out_affine, dout.p, n * psz, cudaMemcpyDeviceToHost, ctx->stream));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)

int b200_kzg_to_lagrange_g1(b200_ctx* ctx, int curve, const uint64_t* srs_g1_affine, unsigned log_n, const uint64_t* generator,
                            uint64_t* out_affine) try {
    B2_ENTER(ctx, curve);
    if (!srs_g1_affine || !generator || !out_affine) return ctx->fail(B200_ERR_ARG, "null argument");
    Fr gen;
    memcpy(gen.l, generator, sizeof gen.l);
    return kzg_to_lagrange_g1(ctx, srs_g1_affine, log_n, gen, out_affine);
} B2_ABI_CATCH_CTX(ctx)

// ---- Groth16 -----------------------------------------------------------------------------
int b200_groth16_pk_upload(b200_ctx* ctx, int curve, const b200_groth16_pk_desc* desc, b200_groth16_pk** out) try {
    B2_ENTER(ctx, curve);
    return groth16_pk_upload(ctx, desc, out);
} B2_ABI_CATCH_CTX(ctx)
int b200_groth16_setup(b200_ctx* ctx, int curve, const b200_groth16_setup_desc* desc, b200_groth16_pk** pk_out,
                       b200_groth16_setup_out* out) try {
    B2_ENTER(ctx, curve);
    if (pk_out) *pk_out = nullptr;
    return groth16_setup(ctx, desc, pk_out, out);
} B2_ABI_CATCH_CTX(ctx)
int b200_groth16_pk_free(b200_ctx* ctx, b200_groth16_pk* pk) try {
    B2_ENTER(ctx, BUILD_CURVE);
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    delete pk;
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)
int b200_groth16_pk_info(b200_ctx* ctx, const b200_groth16_pk* pk, b200_msm_set_info out[5]) try {
    if (!ctx) return B200_ERR_ARG;
    if (!pk || !out) return ctx->fail(B200_ERR_ARG, "null argument");
    fill_set_info(ctx, pk->n_a, pk->A.bytes, pk->fA, out + 0);
    fill_set_info(ctx, pk->n_b, pk->B.bytes, pk->fB, out + 1);
    fill_set_info(ctx, pk->n_k, pk->K.bytes, pk->fK, out + 2);
    fill_set_info(ctx, pk->n_z, pk->Z.bytes, pk->fZ, out + 3);
    fill_set_info(ctx, pk->n_b, pk->G2B.bytes, pk->fG2B, out + 4);
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)
int b200_groth16_pk_wire_range(b200_ctx* ctx, const b200_groth16_pk* pk, size_t* lo, size_t* hi) try {
    if (!ctx) return B200_ERR_ARG;
    if (!pk || !lo || !hi) return ctx->fail(B200_ERR_ARG, "null argument");
    *lo = pk->w_lo; *hi = pk->w_hi;
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)
int b200_groth16_prove_core(b200_ctx* ctx, const b200_groth16_pk* pk, const uint64_t* a, const uint64_t* b,
                            const uint64_t* c, size_t n_constraints, const uint64_t* w, const uint32_t* k_remove,
                            size_t n_remove, b200_groth16_out* out, b200_timings* timings) try {
    B2_ENTER(ctx, BUILD_CURVE);
    if (!pk || !out || !w || (n_constraints && (!a || !b || !c)) || (n_remove && !k_remove))
        return ctx->fail(B200_ERR_ARG, "null argument");
    return groth16_prove_core(ctx, pk, a, b, c, n_constraints, w, k_remove, n_remove, nullptr, out, timings);
} B2_ABI_CATCH_CTX(ctx)

int b200_groth16_prove_core_dev(b200_ctx* ctx, const b200_groth16_pk* pk, const void* a_dev, const void* b_dev,
                                const void* c_dev, size_t n_constraints, const void* w_dev, b200_groth16_out* out,
                                b200_timings* timings) try {
    B2_ENTER(ctx, BUILD_CURVE);
    if (!pk || !out || !w_dev || (n_constraints && (!a_dev || !b_dev || !c_dev)))
        return ctx->fail(B200_ERR_ARG, "null argument");
    return groth16_prove_core(ctx, pk, (const uint64_t*)a_dev, (const uint64_t*)b_dev, (const uint64_t*)c_dev,
                              n_constraints, (const uint64_t*)w_dev, nullptr, 0, nullptr, out, timings);
} B2_ABI_CATCH_CTX(ctx)

int b200_groth16_msms_dev(b200_ctx* ctx, const b200_groth16_pk* pk, const void* w_dev, const void* h_dev,
                          b200_groth16_out* out, b200_timings* timings) try {
    B2_ENTER(ctx, BUILD_CURVE);
    if (!pk || !out || !w_dev || !h_dev) return ctx->fail(B200_ERR_ARG, "null argument");
    return groth16_prove_core(ctx, pk, nullptr, nullptr, nullptr, 0, (const uint64_t*)w_dev, nullptr, 0,
                              (const uint64_t*)h_dev, out, timings);
} B2_ABI_CATCH_CTX(ctx)

// ---- distributed computeH phases ------------------------------------------------------------
int b200_ntt_dist_phase(b200_ctx* ctx, int curve, unsigned log_n, const uint64_t* generator, const uint64_t* coset_gen,
                        void* data_dev, void* tmp_dev, unsigned log_ranks, unsigned rank, int phase, int inverse, int scale) try {
    B2_ENTER(ctx, curve);
    if (!data_dev || !tmp_dev) return ctx->fail(B200_ERR_ARG, "null buffer");
    if (phase < B200_DIST_PACK || phase > B200_DIST_DIT_CROSS) return ctx->fail(B200_ERR_ARG, "bad phase");
    if (log_n > 27) return ctx->fail(B200_ERR_UNSUPPORTED, "NTT size above 2^27");
    NttDomain* d;
    B2_TRY(get_domain(ctx, log_n, generator, coset_gen, &d));
    const Fr* post = nullptr;
    const Fr* scalar = nullptr;
    switch (scale) {
        case B200_SCALE_NONE: break;
        case B200_SCALE_H_A: B2_TRY(ntt_domain_table(ctx, d, NttDomain::H_A, &post)); break;
        case B200_SCALE_H_BC: B2_TRY(ntt_domain_table(ctx, d, NttDomain::H_BC, &post)); break;
        case B200_SCALE_COSETINV_BR: B2_TRY(ntt_domain_table(ctx, d, NttDomain::COSETINV_BR, &post)); break;
        case B200_SCALE_NINV: scalar = &d->ninv; break;
        default: return ctx->fail(B200_ERR_ARG, "bad scale");
    }
    if (scale != B200_SCALE_NONE && phase != B200_DIST_DIF_CROSS && phase != B200_DIST_DIT_CROSS)
        return ctx->fail(B200_ERR_ARG, "a scale applies to the last (cross) phase of a transform");
    return ntt_dist_phase(ctx, d, (Fr*)data_dev, (Fr*)tmp_dev, log_ranks, rank, phase, inverse, post, scalar);
} B2_ABI_CATCH_CTX(ctx)

int b200_fr_mul_sub_dev(b200_ctx* ctx, int curve, void* a_dev, const void* b_dev, const void* c_dev, size_t n) try {
    B2_ENTER(ctx, curve);
    if (n && (!a_dev || !b_dev || !c_dev)) return ctx->fail(B200_ERR_ARG, "null buffer");
    return ntt_mul_sub(ctx, (Fr*)a_dev, (const Fr*)b_dev, (const Fr*)c_dev, n);
} B2_ABI_CATCH_CTX(ctx)

int b200_domain_cache_trim(b200_ctx* ctx, size_t* bytes_freed) try {
    if (!ctx) return B200_ERR_ARG;
    if (set_device(ctx) != B200_OK) return B200_ERR_CUDA;
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const size_t f = ntt_domain_trim(ctx);
    if (bytes_freed) *bytes_freed = f;
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)

int b200_use_stream(b200_ctx* ctx, void* cuda_stream) try {
    if (!ctx) return B200_ERR_ARG;
    if (set_device(ctx) != B200_OK) return B200_ERR_CUDA;
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->stream = cuda_stream ? (cudaStream_t)cuda_stream : ctx->own_stream;
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)

// ---- PlonK building blocks -----------------------------------------------------------------
int b200_fr_batch_inverse_dev(b200_ctx* ctx, int curve, const void* in_dev, void* out_dev, size_t n) try {
    B2_ENTER(ctx, curve);
    if (n && (!in_dev || !out_dev)) return ctx->fail(B200_ERR_ARG, "null buffer");
    if (in_dev == out_dev && n) return ctx->fail(B200_ERR_ARG, "batch inverse: in and out must not alias");
    B2_TRY(fr_batch_inverse(ctx, (const Fr*)in_dev, (Fr*)out_dev, n));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)
int b200_fr_prefix_product_dev(b200_ctx* ctx, int curve, const void* in_dev, void* out_dev, size_t n) try {
    B2_ENTER(ctx, curve);
    if (n && (!in_dev || !out_dev)) return ctx->fail(B200_ERR_ARG, "null buffer");
    B2_TRY(fr_prefix_product(ctx, (const Fr*)in_dev, (Fr*)out_dev, n));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)
int b200_fr_poly_eval_dev(b200_ctx* ctx, int curve, const void* coeffs_dev, size_t n, const uint64_t* x, uint64_t* value) try {
    B2_ENTER(ctx, curve);
    if (!x || !value || (n && !coeffs_dev)) return ctx->fail(B200_ERR_ARG, "null argument");
    Fr xx, out;
    memcpy(xx.l, x, 32);
    B2_TRY(fr_poly_eval(ctx, (const Fr*)coeffs_dev, n, xx, &out));
    memcpy(value, out.l, 32);
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)
int b200_fr_poly_div_linear_dev(b200_ctx* ctx, int curve, const void* coeffs_dev, size_t n, const uint64_t* z, void* quotient_dev) try {
    B2_ENTER(ctx, curve);
    if (!z || (n > 1 && (!coeffs_dev || !quotient_dev))) return ctx->fail(B200_ERR_ARG, "null argument");
    if (coeffs_dev == quotient_dev && n > 1) return ctx->fail(B200_ERR_ARG, "division: coefficients and quotient must not alias");
    Fr zz;
    memcpy(zz.l, z, 32);
    B2_TRY(fr_poly_div_linear(ctx, (const Fr*)coeffs_dev, n, zz, (Fr*)quotient_dev));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)
int b200_plonk_divide_by_xn_minus_one_dev(b200_ctx* ctx, int curve, void* vec_dev, unsigned log_big, unsigned log_small,
                                          const uint64_t* generator_big, const uint64_t* coset_gen) try {
    B2_ENTER(ctx, curve);
    if (!vec_dev || !generator_big || !coset_gen) return ctx->fail(B200_ERR_ARG, "null argument");
    if (log_small > log_big || log_big - log_small > 3 || log_big > 27) return ctx->fail(B200_ERR_ARG, "rho = 2^(log_big - log_small) must be 1, 2, 4 or 8");
    NttDomain* d;
    B2_TRY(get_domain(ctx, log_big, generator_big, coset_gen, &d));
    B2_TRY(plonk_divide_xn(ctx, d, log_small, (Fr*)vec_dev));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)
int b200_kzg_commit_dev(b200_ctx* ctx, const b200_bases* srs, const void* coeffs_dev, size_t n, uint64_t* out_affine) try {
    B2_ENTER(ctx, BUILD_CURVE);
    if (!srs || !out_affine || (n && !coeffs_dev)) return ctx->fail(B200_ERR_ARG, "null argument");
    if (srs->group != 1) return ctx->fail(B200_ERR_ARG, "the KZG SRS is a G1 base set");
    uint64_t jac[3 * B200_FP_WORDS];
    B2_TRY(b200_msm_dev(ctx, srs, coeffs_dev, n, jac, nullptr));
    return b200_g1_jac_to_affine(jac, out_affine);
} B2_ABI_CATCH_CTX(ctx)

int b200_plonk_build_ratio_dev(b200_ctx* ctx, int curve, const void* l_dev, const void* r_dev, const void* o_dev,
                               const void* perm_dev, const uint64_t* beta, const uint64_t* gamma, unsigned log_n,
                               const uint64_t* generator, const uint64_t* coset_gen, void* z_out_dev) try {
    B2_ENTER(ctx, curve);
    if (!l_dev || !r_dev || !o_dev || !perm_dev || !beta || !gamma || !generator || !coset_gen || !z_out_dev)
        return ctx->fail(B200_ERR_ARG, "null argument");
    if (log_n > 27) return ctx->fail(B200_ERR_UNSUPPORTED, "domain above 2^27");
    NttDomain* d;
    B2_TRY(get_domain(ctx, log_n, generator, coset_gen, &d));
    Fr b, g;
    memcpy(b.l, beta, 32);
    memcpy(g.l, gamma, 32);
    B2_TRY(plonk_build_ratio(ctx, d, (const Fr*)l_dev, (const Fr*)r_dev, (const Fr*)o_dev, (const int64_t*)perm_dev, b, g, (Fr*)z_out_dev));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)
int b200_plonk_quotient_rows_dev(b200_ctx* ctx, int curve, const b200_plonk_rows_desc* desc, void* cres_dev) try {
    B2_ENTER(ctx, curve);
    if (!desc || !cres_dev || !desc->generator_small) return ctx->fail(B200_ERR_ARG, "null argument");
    if (desc->nb_bsb > 8) return ctx->fail(B200_ERR_UNSUPPORTED, "more than 8 BSB22 commitments");
    if (desc->log_big < desc->log_n || desc->log_big > 27) return ctx->fail(B200_ERR_ARG, "bad domain sizes");
    PlonkRowArgs a;
    for (int k = 0; k < 15; ++k) {
        if (k != PID_ZS && !desc->x[k]) return ctx->fail(B200_ERR_ARG, "null polynomial");
        a.x[k] = (const Fr*)desc->x[k];
    }
    for (unsigned k = 0; k < 16; ++k) {
        if (k < 2 * desc->nb_bsb && !desc->qci[k]) return ctx->fail(B200_ERR_ARG, "null BSB22 polynomial");
        a.qci[k] = (const Fr*)desc->qci[k];
    }
    a.nb_bsb = desc->nb_bsb;
    memcpy(a.beta.l, desc->beta, 32); memcpy(a.gamma.l, desc->gamma, 32); memcpy(a.alpha.l, desc->alpha, 32);
    memcpy(a.cs.l, desc->cs, 32);
    a.css = a.cs * a.cs;
    for (int k = 0; k < 2; ++k) { memcpy(a.bl[k].l, desc->bl[k], 32); memcpy(a.br[k].l, desc->br[k], 32); memcpy(a.bo[k].l, desc->bo[k], 32); }
    for (int k = 0; k < 3; ++k) memcpy(a.bz[k].l, desc->bz[k], 32);
    NttDomain* d;
    B2_TRY(get_domain(ctx, desc->log_n, desc->generator_small, nullptr, &d));
    a.tw = d->tw_fwd.as<Fr>();
    a.n = 1u << desc->log_n;
    a.log_big = desc->log_big;
    a.rho = 1u << (desc->log_big - desc->log_n);
    a.coset_index = desc->coset_index;
    if (a.coset_index >= a.rho) return ctx->fail(B200_ERR_ARG, "coset index out of range");
    if (a.log_big == 0) return ctx->fail(B200_ERR_ARG, "big domain of one element");
    a.cres = (Fr*)cres_dev;
    B2_TRY(plonk_quotient_rows(ctx, a));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)

// ---- tuning knobs -------------------------------------------------------------------------
int b200_set_option(b200_ctx* ctx, const char* key, long value) try {
    if (!ctx || !key) return B200_ERR_ARG;
    const std::string k(key);
    if (k == "msm_c") { if (value != 0 && (value < 4 || value > 24)) return ctx->fail(B200_ERR_ARG, "msm_c out of range"); ctx->msm_c_override = (int)value; }
    else if (k == "ntt_tile_log") { if (value < 6 || value > 12) return ctx->fail(B200_ERR_ARG, "ntt_tile_log out of range"); ctx->ntt_tile_log = (int)value; }
    else if (k == "ntt_strided_stages") { if (value < 1 || value > 12) return ctx->fail(B200_ERR_ARG, "ntt_strided_stages out of range"); ctx->ntt_strided_stages = (int)value; }
    else if (k == "msm_ba") { if (value != 15 && (value < 0 || value > 9)) return ctx->fail(B200_ERR_ARG, "msm_ba out of range (0..9 rounds, 15 = automatic)"); ctx->msm_ba = (int)value; }
    else if (k == "msm_ba_force") { ctx->msm_ba_force = value ? 1 : 0; }
    else if (k == "msm_acc_smem") { ctx->msm_acc_smem = (int)(value & 3); }
    else if (k == "msm_overlap") { ctx->msm_overlap = value ? 1 : 0; }
    else if (k == "msm_precompute") { ctx->msm_precompute = value ? 1 : 0; }
    else if (k == "msm_fold_c") { if (value != 0 && (value < 8 || value > 24)) return ctx->fail(B200_ERR_ARG, "msm_fold_c out of range"); ctx->msm_fold_c = (int)value; }
    else if (k == "ntt_tma") ctx->ntt_tma = value ? 1 : 0;
    else if (k == "h_sort_overlap") ctx->h_sort_overlap = value ? 1 : 0;
    else if (k == "aux_persist_min") ctx->aux_persist_min = value < 0 ? -1 : (long long)value;
    else if (k == "h_pos") { if (value < -1 || value > 2) return ctx->fail(B200_ERR_ARG, "h_pos out of range (-1 automatic, 0..2)"); ctx->h_pos = (int)value; }
    else if (k == "l2_fetch") {   // device-wide hint: bytes DRAM -> L2 fetches at most per miss (32, 64 or 128); experiments on the table gathers
        if (value != 32 && value != 64 && value != 128) return ctx->fail(B200_ERR_ARG, "l2_fetch must be 32, 64 or 128");
        B2_CUDA(ctx, cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)value));
    }
    else if (k == "ntt_threads") { if (value < 32 || value > 1024 || (value & (value - 1))) return ctx->fail(B200_ERR_ARG, "ntt_threads must be a power of two in 32..1024"); ctx->ntt_threads = (int)value; }
    else return ctx->fail(B200_ERR_ARG, "unknown option", key);
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)

// ---- profiling ---------------------------------------------------------------------------
int b200_profile_start(b200_ctx* ctx) try {
    B2_ENTER(ctx, BUILD_CURVE);
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (auto& sp : ctx->prof_spans) { ctx->prof_pool.push_back(sp.e0); ctx->prof_pool.push_back(sp.e1); }
    ctx->prof_spans.clear();
    ctx->prof_on = true;
    ctx->n_launches = 0;
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)
int b200_profile_read(b200_ctx* ctx, b200_profile* out) try {
    B2_ENTER(ctx, BUILD_CURVE);
    if (!out) return ctx->fail(B200_ERR_ARG, "null output");
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    memset(out, 0, sizeof *out);
    for (auto& sp : ctx->prof_spans) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, sp.e0, sp.e1) == cudaSuccess && sp.slot >= 0 && sp.slot < B200_PROF_SLOTS) {
            out->ms[sp.slot] += ms;
            out->launches[sp.slot] += 1;
            out->units[sp.slot] += sp.units;
        }
        ctx->prof_pool.push_back(sp.e0);
        ctx->prof_pool.push_back(sp.e1);
    }
    cudaGetLastError();
    out->kernel_launches = ctx->n_launches;
    ctx->prof_spans.clear();
    ctx->prof_on = false;
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)

}  // extern "C"
