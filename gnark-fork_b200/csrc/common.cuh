// Shared host-side plumbing of the C ABI: context, error reporting, device buffers.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <new>
#include <stdexcept>
#include <string>
#include <vector>
#include "../../include/b200prover.h"
#include "field.cuh"
#include "ec.cuh"

namespace b200 {
struct NttDomain;
}

struct b200_ctx {
    int device = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;          // all work is ordered on this stream (own_stream unless b200_use_stream)
    cudaStream_t own_stream = nullptr;
    cudaStream_t copy_stream = nullptr;     // per-proof input copies, overlapped with the witness MSMs
    cudaStream_t aux_stream = nullptr;      // sort of the next MSM, overlapped with the current accumulation
    cudaStream_t post_stream = nullptr;     // bucket reduction of the previous MSM, overlapped with the current accumulation
    void* h_stage = nullptr;                // page-locked staging of MSM result points
    size_t h_stage_bytes = 0;
    long long aux_persist_min = -1;         // entries (windows x points) from which the persistent aux grids apply; -1 = default (msm_aux_persistent)
    int aux_persist = 0;                    // set by a caller that enqueues an MSM's sort beside another MSM's bucket stage: the sort's scatter and the round-0 denominator pass run as small persistent grids (msm_host.cuh)
    int h_pos = -1;                         // host-buffer proofs: computeH is enqueued after the accumulation of MSM h_pos (0 A, 1 B1, 2 K); -1 = automatic
    int h_sort_overlap = 0;                 // 1: with device-resident inputs the first MSM's sort runs under computeH instead of after it
    int msm_overlap = 1;                    // 0: the prove core runs its MSMs strictly one after the other
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    std::string err;
    std::vector<b200::NttDomain*> domains;  // cache keyed by (log_n, generator, coset_gen)
    void* scratch = nullptr;                // grow-only workspace shared by MSM / prove calls
    size_t scratch_bytes = 0;
    void* poly_scratch = nullptr;           // small persistent scratch of the polynomial primitives (poly.cu), stream-ordered
    size_t poly_scratch_bytes = 0;
    std::vector<cudaEvent_t> core_events;   // the prove core's events (groth16.cu)
    std::vector<uint32_t> kmap_host;        // K wire map of the proof in flight (api.Commit circuits, groth16.cu)
    int msm_c_override = 0;                 // tuning knobs (b200_set_option / B200_MSM_C env)
    int msm_ba = 15;                        // batched-affine pre-reduction rounds (msm_ba.cuh); 0 = off, 15 = chosen from the mean bucket load
    int msm_ba_force = 0;                   // tests: keep the rounds on MSMs too small to profit
    int msm_mode() const { return (msm_c_override & 0xff) | ((msm_ba & 0xf) << 8) | (msm_ba_force ? 0x1000 : 0); }   // what MsmLayout decodes
    int msm_precompute = 1;                 // build folded-window tables for resident base sets
    int msm_fold_c = 0;                     // 0 = automatic (msm_fold_window)
    int ntt_tma = 1;                        // 1: passes whose tile is a tensor box are staged by TMA (ntt_pass_tma_kernel); A/B in DESIGN.md
    int ntt_tile_log = 9, ntt_strided_stages = 8, ntt_threads = 128;   // measured best (scripts/ntt_sweep.py): many small CTAs overlap load, butterflies and store

    // optional per-kernel-class timing (b200_profile_*): CUDA event pairs on `stream`
    bool prof_on = false;
    uint64_t n_launches = 0;                // kernel launches since b200_profile_start
    struct ProfSpan { int slot; double units; cudaEvent_t e0, e1; };
    std::vector<ProfSpan> prof_spans;
    std::vector<cudaEvent_t> prof_pool;
    cudaEvent_t prof_event() {
        cudaEvent_t e = nullptr;
        if (!prof_pool.empty()) { e = prof_pool.back(); prof_pool.pop_back(); }
        else cudaEventCreate(&e);
        return e;
    }
    // returns an index to pass to prof_end (-1 when profiling is off)
    int prof_begin(int slot, double units, cudaStream_t on = nullptr) {
        if (!prof_on) return -1;
        ProfSpan sp{slot, units, prof_event(), prof_event()};
        cudaEventRecord(sp.e0, on ? on : stream);
        prof_spans.push_back(sp);
        return (int)prof_spans.size() - 1;
    }
    void prof_end(int idx, cudaStream_t on = nullptr) {
        if (idx >= 0) cudaEventRecord(prof_spans[idx].e1, on ? on : stream);
    }

    int fail(int status, const char* what, const char* detail = nullptr) {
        err = what;
        if (detail) { err += ": "; err += detail; }
        return status;
    }
    // bucket accumulation with the running sum in shared memory (msm_accumulate_smem_kernel): bit 0 = G1, bit 1 = G2.
    // BN254: G2 only (-10 % on the kernel; G1 already runs 4 blocks of 112 registers).  BLS12-377: off (measured, msm.cuh).
#if defined(B2_CURVE_BLS12_377)
    int msm_acc_smem = 0;
#else
    int msm_acc_smem = 2;
#endif
};

#define B2_CUDA(ctx, call)                                                                     \
    do {                                                                                       \
        cudaError_t e_ = (call);                                                               \
        if (e_ != cudaSuccess) {                                                               \
            char buf_[256];                                                                    \
            snprintf(buf_, sizeof buf_, "%s:%d %s -> %s", __FILE__, __LINE__, #call,           \
                     cudaGetErrorString(e_));                                                  \
            return (ctx)->fail(e_ == cudaErrorMemoryAllocation ? B200_ERR_NOMEM : B200_ERR_CUDA, \
                               "CUDA error", buf_);                                            \
        }                                                                                      \
    } while (0)

// No C++ exception may cross the C ABI: every extern "C" entry point is a function-try-block that ends
// in one of these handlers (scripts/add_exception_barrier.py put them there).
namespace b200 {
inline int abi_fail(b200_ctx* ctx, int status, const char* what) noexcept {
    try { if (ctx) ctx->fail(status, "exception stopped at the C ABI", what); } catch (...) {}
    return status;
}
}  // namespace b200
#define B2_ABI_CATCH_CTX(ctx)                                                                                    \
    catch (const std::bad_alloc&) { return b200::abi_fail(ctx, B200_ERR_NOMEM, "host allocation failed"); }       \
    catch (const std::exception& e_) { return b200::abi_fail(ctx, B200_ERR_INTERNAL, e_.what()); }                \
    catch (...) { return b200::abi_fail(ctx, B200_ERR_INTERNAL, "unknown exception"); }
#define B2_ABI_CATCH_INT                                           \
    catch (const std::bad_alloc&) { return B200_ERR_NOMEM; }       \
    catch (...) { return B200_ERR_INTERNAL; }
#define B2_ABI_CATCH_VOID catch (...) {}
#define B2_ABI_CATCH_NULL catch (...) { return nullptr; }

#define B2_TRY(expr)                    \
    do {                                \
        int st_ = (expr);               \
        if (st_ != B200_OK) return st_; \
    } while (0)

namespace b200 {

// RAII device allocation (freed on scope exit; all work is ordered on ctx->stream and the
// owning entry point synchronises before returning).
struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    DevBuf() {}
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    ~DevBuf() { release(); }
    cudaError_t alloc(size_t n) {
        release();
        if (n == 0) n = 16;
        cudaError_t e = cudaMalloc(&p, n);
        if (e == cudaSuccess) bytes = n; else p = nullptr;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; bytes = 0; }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

// Bump allocator over the ctx workspace (256-byte aligned slices).
struct Arena {
    char* base = nullptr;
    size_t off = 0;
    static size_t align(size_t x) { return (x + 255) & ~(size_t)255; }
    void* take(size_t bytes) {
        void* p = base ? base + off : nullptr;
        off += align(bytes ? bytes : 1);
        return p;
    }
};

// Make the ctx workspace at least `bytes` large (synchronises and reallocates when it grows).
static inline int ctx_reserve(b200_ctx* ctx, size_t bytes) {
    if (bytes <= ctx->scratch_bytes) return B200_OK;
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) return ctx->fail(B200_ERR_CUDA, "sync before workspace growth", cudaGetErrorString(e));
    if (ctx->scratch) cudaFree(ctx->scratch);
    ctx->scratch = nullptr;
    ctx->scratch_bytes = 0;
    e = cudaMalloc(&ctx->scratch, bytes);
    if (e != cudaSuccess) { ctx->scratch = nullptr; return ctx->fail(B200_ERR_NOMEM, "workspace", cudaGetErrorString(e)); }
    ctx->scratch_bytes = bytes;
    return B200_OK;
}

// Persistent small scratch of poly.cu (grow-only; work on it is ordered on ctx->stream).
static inline int ctx_poly_scratch(b200_ctx* ctx, size_t bytes, void** out) {
    if (bytes > ctx->poly_scratch_bytes) {
        cudaError_t e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) return ctx->fail(B200_ERR_CUDA, "sync before scratch growth", cudaGetErrorString(e));
        if (ctx->poly_scratch) cudaFree(ctx->poly_scratch);
        ctx->poly_scratch = nullptr;
        ctx->poly_scratch_bytes = 0;
        e = cudaMalloc(&ctx->poly_scratch, bytes);
        if (e != cudaSuccess) { ctx->poly_scratch = nullptr; return ctx->fail(B200_ERR_NOMEM, "polynomial scratch", cudaGetErrorString(e)); }
        ctx->poly_scratch_bytes = bytes;
    }
    *out = ctx->poly_scratch;
    return B200_OK;
}

static inline unsigned ceil_log2(size_t x) { unsigned l = 0; while (((size_t)1 << l) < x) ++l; return l; }

}  // namespace b200
