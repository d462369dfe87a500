// b200_fr_prefix_product_static: the prefix product of b200_fr_prefix_product_dev for callers that may not
// allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart
// in the reference.
//
// The sibling entry point takes its chunk-summary scratch from a grow-only cudaMalloc'd buffer of the ctx
// (ctx_poly_scratch).  Here the scratch is a file-scope `static __device__` array of a fixed size: every GPU that
// loads the library holds NA_SCRATCH_BYTES (~4.1 MiB) of it until the process exits.  Host code cannot use the array's
// name as an address (it is a host-side stand-in for the symbol): the device address is fetched ONCE PER DEVICE with
// cudaGetSymbolAddress, with that device current, and handed to the kernels as an ordinary pointer argument.
//
// The call path below performs no cudaMalloc / cudaFree / cudaMallocHost and calls no allocating library routine.
// All calls (of all contexts on one device) share the one array, which keeps the previous call's contents: every
// word the scan reads is written earlier in the same call (pass 1 writes the chunk products the recursion and pass 3
// read), so nothing has to start at zero and nothing is cleared.  Two calls that overlap on the GPU (different streams
// or contexts of one device) RACE on the array: serialise them.
#include <mutex>
#include "poly_host.cuh"

namespace b200 {
namespace {

constexpr int NA_CHUNK = 32;        // elements per thread (same split as poly.cu, so the results are bit-identical)
constexpr int NA_THREADS = 128;
constexpr size_t NA_MAX_N = (size_t)1 << 22;   // elements per call (BASELINE.json config 4: PlonK at 2^22 gates)

constexpr size_t na_scratch_elems(size_t n) {
    size_t tot = 0;
    while (n > 1) { n = (n + NA_CHUNK - 1) / NA_CHUNK; tot += n; }
    return tot + 4;
}
constexpr size_t NA_SCRATCH_ELEMS = na_scratch_elems(NA_MAX_N);       // 135 305 elements
constexpr size_t NA_SCRATCH_BYTES = NA_SCRATCH_ELEMS * 32;            // 4 329 760 B per device

// two uint4 per Fr element.  Device code in this file could name the array directly; host code needs its address.
static __device__ __align__(256) uint4 g_na_scratch[NA_SCRATCH_ELEMS * 2];

constexpr int NA_MAX_DEVICES = 64;
std::mutex g_na_mutex;
Fr* g_na_ptr[NA_MAX_DEVICES] = {nullptr};   // device address of g_na_scratch on device k (each device has its own copy)

// Set-up, once per device, with that device current.
int na_scratch_address(b200_ctx* ctx, Fr** out) {
    if (ctx->device < 0 || ctx->device >= NA_MAX_DEVICES) return ctx->fail(B200_ERR_UNSUPPORTED, "static scratch: device ordinal out of range");
    std::lock_guard<std::mutex> lock(g_na_mutex);
    if (!g_na_ptr[ctx->device]) {
        int cur = -1;
        cudaError_t e = cudaGetDevice(&cur);
        if (e != cudaSuccess || cur != ctx->device) return ctx->fail(B200_ERR_CUDA, "static scratch: the context's device is not current");
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_na_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            return ctx->fail(B200_ERR_CUDA, "cudaGetSymbolAddress(static scratch)", cudaGetErrorString(e));
        }
        g_na_ptr[ctx->device] = reinterpret_cast<Fr*>(p);
    }
    *out = g_na_ptr[ctx->device];
    return B200_OK;
}

__device__ __forceinline__ Fr na_ld(const Fr* p) {
    const uint4* q = reinterpret_cast<const uint4*>(p);
    const uint4 a = q[0], b = q[1];
    Fr f;
    f.l[0] = a.x; f.l[1] = a.y; f.l[2] = a.z; f.l[3] = a.w;
    f.l[4] = b.x; f.l[5] = b.y; f.l[6] = b.z; f.l[7] = b.w;
    return f;
}
__device__ __forceinline__ void na_st(Fr* p, const Fr& f) {
    uint4* q = reinterpret_cast<uint4*>(p);
    q[0] = make_uint4(f.l[0], f.l[1], f.l[2], f.l[3]);
    q[1] = make_uint4(f.l[4], f.l[5], f.l[6], f.l[7]);
}

// pass 1: prod[t] = product of chunk t  (prod points into the static scratch)
__global__ void __launch_bounds__(NA_THREADS) na_chunk_product_kernel(const Fr* __restrict__ in, size_t n, Fr* __restrict__ prod) {
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t lo = t * NA_CHUNK;
    if (lo >= n) return;
    const int cnt = (int)(n - lo < (size_t)NA_CHUNK ? n - lo : (size_t)NA_CHUNK);
    Fr acc = na_ld(in + lo);
    for (int i = 1; i < cnt; ++i) acc = acc * na_ld(in + lo + i);
    na_st(prod + t, acc);
}
// pass 3: out[i] = incl[t - 1] * in[lo] * ... * in[i]  (incl = inclusive prefix of the chunk products; may be null)
__global__ void __launch_bounds__(NA_THREADS) na_chunk_scan_kernel(const Fr* in, size_t n, const Fr* __restrict__ incl, Fr* out) {
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t lo = t * NA_CHUNK;
    if (lo >= n) return;
    const int cnt = (int)(n - lo < (size_t)NA_CHUNK ? n - lo : (size_t)NA_CHUNK);
    Fr acc = (t == 0 || incl == nullptr) ? Fr::one() : na_ld(incl + t - 1);
    for (int i = 0; i < cnt; ++i) {
        acc = acc * na_ld(in + lo + i);
        na_st(out + lo + i, acc);
    }
}

// inclusive prefix product, in -> out (may alias); levels of chunk products live in `scratch`
int na_prefix_rec(b200_ctx* ctx, const Fr* in, Fr* out, size_t n, Fr* scratch) {
    const size_t chunks = (n + NA_CHUNK - 1) / NA_CHUNK;
    const unsigned blocks = (unsigned)((chunks + NA_THREADS - 1) / NA_THREADS);
    if (chunks == 1) {
        na_chunk_scan_kernel<<<1, NA_THREADS, 0, ctx->stream>>>(in, n, nullptr, out);
        B2_CUDA(ctx, cudaGetLastError());
        ctx->n_launches += 1;
        return B200_OK;
    }
    Fr* prod = scratch;     // `chunks` elements, scanned in place by the next level
    na_chunk_product_kernel<<<blocks, NA_THREADS, 0, ctx->stream>>>(in, n, prod);
    B2_CUDA(ctx, cudaGetLastError());
    B2_TRY(na_prefix_rec(ctx, prod, prod, chunks, scratch + chunks));
    na_chunk_scan_kernel<<<blocks, NA_THREADS, 0, ctx->stream>>>(in, n, prod, out);
    B2_CUDA(ctx, cudaGetLastError());
    ctx->n_launches += 2;
    return B200_OK;
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" int b200_fr_prefix_product_static(b200_ctx* ctx, int curve, const void* in_dev, void* out_dev, size_t n) try {
    if (!ctx) return B200_ERR_ARG;
#if defined(B2_CURVE_BLS12_377)
    if (curve != B200_BLS12_377) return ctx->fail(B200_ERR_UNSUPPORTED, "this library is the BLS12-377 build");
#else
    if (curve != B200_BN254) return ctx->fail(B200_ERR_UNSUPPORTED, "this library is the BN254 build");
#endif
    if (n > NA_MAX_N) {
        char msg[160];
        snprintf(msg, sizeof msg, "b200_fr_prefix_product_static: n = %zu exceeds the static scratch (at most %zu elements per call); "
                                  "use b200_fr_prefix_product_dev", n, NA_MAX_N);
        return ctx->fail(B200_ERR_ARG, msg);
    }
    if (n == 0) return B200_OK;
    if (!in_dev || !out_dev) return ctx->fail(B200_ERR_ARG, "null buffer");
    B2_CUDA(ctx, cudaSetDevice(ctx->device));
    Fr* scratch = nullptr;
    B2_TRY(na_scratch_address(ctx, &scratch));
    return na_prefix_rec(ctx, (const Fr*)in_dev, (Fr*)out_dev, n, scratch);
} B2_ABI_CATCH_CTX(ctx)
