// Multi-GPU Groth16 prover in ONE process behind the C ABI: what a Go groth16.Prove calls to
// prove on all the GPUs of a box with one call, as /root/reference/backend/groth16/bn254/icicle/icicle.go:133-422 makes one
// call per proof on its single device.
//
//   * one b200_ctx per device, one host thread per device for the duration of a call (the launches of the ranks are
//     issued side by side);
//   * the five MultiExp (prove.go:201-290) run on point-range shards of the key, one shard per device; the 576-byte
//     partial sums are added on the host -- no data-path collective;
//   * computeH (prove.go:353-396) is distributed with ONE exchange per transform (csrc/ntt.cuh: layouts A and B), and the
//     exchange is not a collective call: it is fused into the kernels on either side over peer-mapped memory (NVLink /
//     NVSwitch P2P).  PACK scatters a rank's block straight into the peers' buffers (pack_push_kernel, coalesced peer
//     stores); every cross phase loads its input straight from the peers' send buffers (ntt_pass_kernel in pull mode,
//     2^(TL-lg) x 32 B contiguous per row).  What is left of the all-to-all is ordering: four points per computeH where
//     every rank's stream waits for an event of every other rank (cudaStreamWaitEvent across devices) -- no NCCL, no
//     copy engine, no host round trip of data;
//   * per rank and vector three buffers U, V, W of n / R elements: host block -> U; PACK: U -> peers' V; inverse DIF local
//     in V; cross (pull from the V's) -> U with the 1/n, coset and den scalings; forward DIT local U -> W; cross (pull from
//     the W's) -> V; a V <- a V * b V - c V; inverse DIF local in a's V; cross (pull) -> a's U = this rank's block of h in
//     the bit-reversed order pk.G1.Z is sharded in (groth16.cu: shard_range_z).
// The same device may appear several times in `devices` (virtual ranks: how the tests run this on one GPU).
#include <atomic>
#include <memory>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <chrono>
#include <cstdlib>
#include "groth16_host.cuh"
#include "host_group.cuh"

namespace b200 { int plonk_draw_blinding(uint64_t out[36]); }   // plonk.cu

using namespace b200;

namespace {

// All ranks meet here between "my kernels are enqueued, my event is recorded" and "wait for everybody's event".
struct HostBarrier {
    std::mutex m;
    std::condition_variable cv;
    int n = 1, count = 0;
    uint64_t gen = 0;
    bool broken = false;
    bool wait() {
        std::unique_lock<std::mutex> lk(m);
        if (broken) return false;
        const uint64_t g = gen;
        if (++count == n) { count = 0; ++gen; cv.notify_all(); return true; }
        cv.wait(lk, [&] { return gen != g || broken; });
        return !broken;
    }
    void abort() { std::lock_guard<std::mutex> lk(m); broken = true; cv.notify_all(); }
    void reset(int ranks) { std::lock_guard<std::mutex> lk(m); n = ranks; count = 0; broken = false; }
};

constexpr int SYNC_POINTS = 4;   // per computeH: after PACK, after the first local phase, after the DIT local phase, after the last local phase

}  // namespace

struct b200_group {
    std::vector<b200_ctx*> ctx;
    int R = 0;
    uint32_t lg = 0;
    bool pow2 = false;
    std::string err;
    HostBarrier barrier;
    std::vector<cudaEvent_t> ev;    // [rank * SYNC_POINTS + k], owned by the rank's device
    int fail(int st, const char* what, const char* detail = nullptr) {
        err = what;
        if (detail) { err += ": "; err += detail; }
        return st;
    }
};

struct b200_group_groth16_pk {
    std::vector<b200_groth16_pk*> pk;     // one shard per rank
    unsigned log_n = 0;
    bool distributed = false;             // computeH over all ranks (power-of-two group, 2 lg <= log_n)
    std::unique_ptr<DevBuf[]> buf;        // [rank * 9 + vector * 3 + {U, V, W}], n >> lg elements each
    size_t nbuf = 0;
    Fr* at(int rank, int v, int k) const { return buf[(size_t)rank * 9 + v * 3 + k].as<Fr>(); }
};

namespace {

template <class Fn>
int run_ranks(b200_group* g, Fn fn) {
    const int R = g->R;
    std::vector<int> st(R, B200_OK);
    g->barrier.reset(R);
    auto body = [&](int r) {
        try {
            if (cudaSetDevice(g->ctx[r]->device) != cudaSuccess) st[r] = g->ctx[r]->fail(B200_ERR_CUDA, "cudaSetDevice");
            else st[r] = fn(r);
        } catch (const std::bad_alloc&) {
            st[r] = abi_fail(g->ctx[r], B200_ERR_NOMEM, "host allocation failed");
        } catch (const std::exception& e) {
            st[r] = abi_fail(g->ctx[r], B200_ERR_INTERNAL, e.what());
        } catch (...) {
            st[r] = abi_fail(g->ctx[r], B200_ERR_INTERNAL, "unknown exception");
        }
        if (st[r] != B200_OK) g->barrier.abort();   // the others must not wait for this rank
    };
    if (R == 1) body(0);
    else {
        std::vector<std::thread> th;
        th.reserve(R);
        for (int r = 0; r < R; ++r) th.emplace_back(body, r);
        for (auto& t : th) t.join();
    }
    for (int r = 0; r < R; ++r)
        if (st[r] != B200_OK) {
            char where[32];
            snprintf(where, sizeof where, "rank %d", r);
            return g->fail(st[r], where, g->ctx[r]->err.c_str());
        }
    return B200_OK;
}

// Cross-device ordering point k: everything rank r has enqueued so far on its stream happens before anything any rank
// enqueues after the point.
int sync_point(b200_group* g, int r, int k) {
    b200_ctx* ctx = g->ctx[r];
    B2_CUDA(ctx, cudaEventRecord(g->ev[(size_t)r * SYNC_POINTS + k], ctx->stream));
    if (!g->barrier.wait()) return ctx->fail(B200_ERR_INTERNAL, "another rank failed");
    for (int q = 0; q < g->R; ++q)
        if (q != r) B2_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, g->ev[(size_t)q * SYNC_POINTS + k], 0));
    // event k is recorded again only by the next proof, i.e. after every rank's thread of this call has been joined
    return B200_OK;
}

// computeH over the group, enqueued on rank r's stream.  U of each vector holds the rank's natural-order block.
int dist_compute_h(b200_group* g, const b200_group_groth16_pk* gpk, int r) {
    b200_ctx* ctx = g->ctx[r];
    const b200_groth16_pk* pk = gpk->pk[r];
    NttDomain* d = pk->dom;
    const uint32_t lg = g->lg, rank = (uint32_t)r;
    const Fr *ha, *hbc, *cinv;
    B2_TRY(ntt_domain_table(ctx, d, NttDomain::H_A, &ha));
    B2_TRY(ntt_domain_table(ctx, d, NttDomain::H_BC, &hbc));
    B2_TRY(ntt_domain_table(ctx, d, NttDomain::COSETINV_BR, &cinv));
    const Fr* scales[3] = {ha, hbc, ha};
    Fr* peersV[3][NTT_MAX_RANKS];
    Fr* peersW[3][NTT_MAX_RANKS];
    for (int v = 0; v < 3; ++v)
        for (int q = 0; q < g->R; ++q) { peersV[v][q] = gpk->at(q, v, 1); peersW[v][q] = gpk->at(q, v, 2); }
    auto U = [&](int v) { return gpk->at(r, v, 0); };
    auto V = [&](int v) { return gpk->at(r, v, 1); };
    auto W = [&](int v) { return gpk->at(r, v, 2); };
    // block layout -> cyclic layout, written into the peers' V
    for (int v = 0; v < 3; ++v) B2_TRY(ntt_dist_pack_push(ctx, d, U(v), peersV[v], lg, rank));
    B2_TRY(sync_point(g, r, 0));
    // FFTInverse (DIF), prove.go:369-371: local stages in V
    for (int v = 0; v < 3; ++v) B2_TRY(ntt_dist_phase(ctx, d, V(v), nullptr, lg, rank, NTT_DIST_DIF_LOCAL, 1, nullptr, nullptr));
    B2_TRY(sync_point(g, r, 1));
    // ... cross stages pulled from the peers' V into U with 1/n, the coset shift and den; FFT (DIT) on the coset,
    // prove.go:373-375: local stages U -> W
    for (int v = 0; v < 3; ++v) {
        B2_TRY(ntt_dist_phase(ctx, d, U(v), nullptr, lg, rank, NTT_DIST_DIF_CROSS, 1, scales[v], nullptr, peersV[v]));
        B2_TRY(ntt_dist_phase(ctx, d, U(v), W(v), lg, rank, NTT_DIST_DIT_LOCAL, 0, nullptr, nullptr));
    }
    B2_TRY(sync_point(g, r, 2));
    // ... cross stages pulled from the peers' W into V (every peer has finished reading V: its DIT local phase is behind point 2)
    for (int v = 0; v < 3; ++v) B2_TRY(ntt_dist_phase(ctx, d, V(v), nullptr, lg, rank, NTT_DIST_DIT_CROSS, 0, nullptr, nullptr, peersW[v]));
    B2_TRY(ntt_mul_sub(ctx, V(0), V(1), V(2), d->n() >> lg));                    // prove.go:377-390 (den is in H_A)
    B2_TRY(ntt_dist_phase(ctx, d, V(0), nullptr, lg, rank, NTT_DIST_DIF_LOCAL, 1, nullptr, nullptr));   // prove.go:393
    B2_TRY(sync_point(g, r, 3));
    B2_TRY(ntt_dist_phase(ctx, d, U(0), nullptr, lg, rank, NTT_DIST_DIF_CROSS, 1, cinv, nullptr, peersV[0]));
    return B200_OK;
}

void combine(const std::vector<b200_groth16_out>& parts, b200_groth16_out* out) {
    *out = parts[0];
    for (size_t r = 1; r < parts.size(); ++r) {
        hostgrp::jac_add<Fp>(out->ar, parts[r].ar, out->ar);
        hostgrp::jac_add<Fp>(out->bs1, parts[r].bs1, out->bs1);
        hostgrp::jac_add<Fp>(out->krs, parts[r].krs, out->krs);
        hostgrp::jac_add<Fp>(out->krs2, parts[r].krs2, out->krs2);
        hostgrp::jac_add<Fp2>(out->bs2, parts[r].bs2, out->bs2);
    }
}

void max_timings(const std::vector<b200_timings>& t, b200_timings* out) {
    *out = t[0];
    for (size_t r = 1; r < t.size(); ++r) {
        const float* s = &t[r].h2d_ms;
        float* o = &out->h2d_ms;
        for (size_t k = 0; k < sizeof(b200_timings) / sizeof(float); ++k) o[k] = o[k] > s[k] ? o[k] : s[k];
    }
}

int finish_pk(b200_group* g, b200_group_groth16_pk* gpk, unsigned log_n) {
    gpk->log_n = log_n;
    gpk->distributed = g->pow2 && g->R > 1 && 2 * g->lg <= log_n && g->R <= NTT_MAX_RANKS;
    if (!gpk->distributed) return B200_OK;
    const size_t n_loc = ((size_t)1 << log_n) >> g->lg;
    gpk->nbuf = (size_t)g->R * 9;
    gpk->buf.reset(new DevBuf[gpk->nbuf]);
    return run_ranks(g, [&](int r) -> int {
        for (int k = 0; k < 9; ++k) B2_CUDA(g->ctx[r], gpk->buf[(size_t)r * 9 + k].alloc(n_loc * sizeof(Fr)));
        return B200_OK;
    });
}

int prove(b200_group* g, const b200_group_groth16_pk* gpk, const uint64_t* a, const uint64_t* b, const uint64_t* c,
          size_t n_constraints, const uint64_t* w, const void* const* a_dev, const void* const* b_dev, const void* const* c_dev,
          const void* const* w_dev, const uint32_t* k_remove, size_t n_remove, b200_groth16_out* out, b200_timings* timings) {
    const int R = g->R;
    const size_t n = (size_t)1 << gpk->log_n;
    if (n_constraints > n) return g->fail(B200_ERR_ARG, "more constraints than the domain holds");
    std::vector<b200_groth16_out> parts(R);
    std::vector<b200_timings> tms(R);
    const bool on_dev = a_dev != nullptr;
    // B200_TRACE=1: host-side timeline of the call on stderr (thread start, prove core per rank, join, combine)
    static const bool trace = getenv("B200_TRACE") != nullptr;
    auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t_enter = now();
    std::vector<double> t_start(R, 0.0), t_done(R, 0.0);
    int st = run_ranks(g, [&](int r) -> int {
        b200_ctx* ctx = g->ctx[r];
        t_start[r] = now();
        struct Done { double* t; decltype(now)* f; ~Done() { *t = (*f)(); } } done{&t_done[r], &now};
        const b200_groth16_pk* pk = gpk->pk[r];
        const uint64_t* wr = on_dev ? (const uint64_t*)w_dev[r] : w;
        if (!gpk->distributed) {      // every rank runs the whole computeH on its own copy of a, b, c
            if (on_dev) return ctx->fail(B200_ERR_UNSUPPORTED, "device-resident inputs need the distributed computeH (power-of-two group)");
            return groth16_prove_core(ctx, pk, a, b, c, n_constraints, wr, k_remove, n_remove, nullptr, &parts[r], &tms[r]);
        }
        const size_t n_loc = n >> g->lg, lo = n_loc * (size_t)r;
        ProveHHook hook;
        hook.h_out = gpk->at(r, 0, 0);
        hook.inputs_on_device = on_dev;
        const uint64_t* hsrc[3] = {a, b, c};
        const void* const* dsrc[3] = {a_dev, b_dev, c_dev};
        hook.copy_abc = [&, r](cudaStream_t cp) -> int {      // this rank's block of a, b, c -> U, zero padded
            for (int v = 0; v < 3; ++v) {
                Fr* U = gpk->at(r, v, 0);
                if (on_dev) {
                    B2_CUDA(ctx, cudaMemcpyAsync(U, dsrc[v][r], n_loc * sizeof(Fr), cudaMemcpyDeviceToDevice, cp));
                    continue;
                }
                const size_t have = n_constraints > lo ? std::min(n_loc, n_constraints - lo) : 0;
                if (have) B2_CUDA(ctx, cudaMemcpyAsync(U, hsrc[v] + 4 * lo, have * sizeof(Fr), cudaMemcpyDefault, cp));
                if (have < n_loc) B2_CUDA(ctx, cudaMemsetAsync(U + have, 0, (n_loc - have) * sizeof(Fr), cp));
            }
            return B200_OK;
        };
        hook.compute_h = [&, r]() -> int { return dist_compute_h(g, gpk, r); };
        return groth16_prove_core(ctx, pk, nullptr, nullptr, nullptr, 0, wr, k_remove, n_remove, nullptr, &parts[r], &tms[r], &hook);
    });
    if (st != B200_OK) return st;
    const double t_join = now();
    combine(parts, out);
    if (timings) max_timings(tms, timings);
    if (trace) {
        fprintf(stderr, "[b200 trace] group prove: join at %.3f ms, combine %.3f ms;", t_join - t_enter, now() - t_join);
        for (int r = 0; r < R; ++r) fprintf(stderr, " r%d start %.3f core %.3f (device %.3f)", r, t_start[r] - t_enter, t_done[r] - t_start[r], tms[r].total_ms);
        fprintf(stderr, "\n");
    }
    return B200_OK;
}

}  // namespace

extern "C" {

int b200_group_init(const int* devices, int ndev, b200_group** out) try {
    if (!out) return B200_ERR_ARG;
    *out = nullptr;
    if (!devices || ndev < 1 || ndev > 64) return B200_ERR_ARG;
    b200_group* g = new b200_group();
    g->R = ndev;
    g->pow2 = (ndev & (ndev - 1)) == 0;
    while ((1 << g->lg) < ndev) ++g->lg;
    int st = B200_OK;
    for (int r = 0; r < ndev && st == B200_OK; ++r) {
        b200_ctx* c = nullptr;
        st = b200_init(devices[r], &c);
        if (st == B200_OK) g->ctx.push_back(c);
    }
    // peer access between every pair of distinct devices (NVLink / NVSwitch): the kernels of one rank read and write the
    // buffers of the others
    for (int r = 0; r < ndev && st == B200_OK; ++r) {
        cudaSetDevice(devices[r]);
        for (int q = 0; q < ndev && st == B200_OK; ++q) {
            if (devices[q] == devices[r]) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, devices[r], devices[q]);
            if (!can) { g->pow2 = false; continue; }     // no P2P path: computeH stays replicated
            const cudaError_t e = cudaDeviceEnablePeerAccess(devices[q], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) st = B200_ERR_CUDA;
            cudaGetLastError();
        }
        for (int k = 0; k < SYNC_POINTS && st == B200_OK; ++k) {
            cudaEvent_t e;
            if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) st = B200_ERR_CUDA;
            else g->ev.push_back(e);
        }
    }
    if (st != B200_OK) {
        for (cudaEvent_t e : g->ev) cudaEventDestroy(e);
        for (b200_ctx* c : g->ctx) b200_destroy(c);
        delete g;
        return st;
    }
    *out = g;
    return B200_OK;
} B2_ABI_CATCH_INT

void b200_group_destroy(b200_group* g) try {
    if (!g) return;
    for (size_t i = 0; i < g->ev.size(); ++i) {
        cudaSetDevice(g->ctx[i / SYNC_POINTS]->device);
        cudaEventDestroy(g->ev[i]);
    }
    for (b200_ctx* c : g->ctx) b200_destroy(c);
    delete g;
} B2_ABI_CATCH_VOID

const char* b200_group_last_error(const b200_group* g) try { return g ? g->err.c_str() : ""; } B2_ABI_CATCH_NULL
int b200_group_size(const b200_group* g) try { return g ? g->R : 0; } B2_ABI_CATCH_INT
b200_ctx* b200_group_ctx(b200_group* g, int rank) try { return (g && rank >= 0 && rank < g->R) ? g->ctx[rank] : nullptr; } B2_ABI_CATCH_NULL

int b200_group_set_option(b200_group* g, const char* key, long value) try {
    if (!g) return B200_ERR_ARG;
    for (b200_ctx* c : g->ctx) {
        const int st = b200_set_option(c, key, value);
        if (st != B200_OK) return g->fail(st, "set_option", c->err.c_str());
    }
    return B200_OK;
} B2_ABI_CATCH_INT

int b200_group_groth16_pk_upload(b200_group* g, int curve, const b200_groth16_pk_desc* desc, b200_group_groth16_pk** out) try {
    if (!g || !desc || !out) return B200_ERR_ARG;
    *out = nullptr;
    if (curve != b200_build_curve()) return g->fail(B200_ERR_UNSUPPORTED, "curve not supported by this build");
    b200_group_groth16_pk* gpk = new b200_group_groth16_pk();
    gpk->pk.assign(g->R, nullptr);
    int st = run_ranks(g, [&](int r) -> int {
        b200_groth16_pk_desc d = *desc;
        d.shard_index = r;
        d.shard_count = g->R;
        d.presliced = 0;
        return groth16_pk_upload(g->ctx[r], &d, &gpk->pk[r]);
    });
    if (st == B200_OK) st = finish_pk(g, gpk, desc->log_n);
    if (st != B200_OK) { b200_group_groth16_pk_free(g, gpk); return st; }
    *out = gpk;
    return B200_OK;
} B2_ABI_CATCH_INT

int b200_group_groth16_setup(b200_group* g, int curve, const b200_groth16_setup_desc* desc, b200_group_groth16_pk** out,
                             b200_groth16_setup_out* so) try {
    if (!g || !desc || !out || !so) return B200_ERR_ARG;
    *out = nullptr;
    if (curve != b200_build_curve()) return g->fail(B200_ERR_UNSUPPORTED, "curve not supported by this build");
    b200_group_groth16_pk* gpk = new b200_group_groth16_pk();
    gpk->pk.assign(g->R, nullptr);
    // the scalar side once (rank 0's device for the batch inversion, host threads for the matrices), then every rank
    // generates its own point ranges; rank 0 also makes the verifying-key side
    SetupScalars S;
    cudaSetDevice(g->ctx[0]->device);
    int st = groth16_setup_scalars(g->ctx[0], desc, S);
    if (st != B200_OK) { delete gpk; return g->fail(st, "setup", g->ctx[0]->err.c_str()); }
    std::vector<b200_groth16_setup_out> outs(g->R, *so);
    st = run_ranks(g, [&](int r) -> int {
        b200_groth16_setup_out& local = outs[r];
        if (r != 0) { local.infinity_a = local.infinity_b = nullptr; local.vk_k = nullptr; local.ck_k = nullptr; }
        return groth16_setup_points(g->ctx[r], desc, S, r, g->R, &gpk->pk[r], &local, r == 0);
    });
    if (st == B200_OK) *so = outs[0];
    if (st == B200_OK) st = finish_pk(g, gpk, desc->log_n);
    if (st != B200_OK) { b200_group_groth16_pk_free(g, gpk); return st; }
    *out = gpk;
    return B200_OK;
} B2_ABI_CATCH_INT

int b200_group_groth16_pk_free(b200_group* g, b200_group_groth16_pk* gpk) try {
    if (!g) return B200_ERR_ARG;
    if (!gpk) return B200_OK;
    for (int r = 0; r < g->R; ++r) {
        cudaSetDevice(g->ctx[r]->device);
        cudaStreamSynchronize(g->ctx[r]->stream);
        if ((size_t)r < gpk->pk.size()) delete gpk->pk[r];
        for (int k = 0; k < 9 && (size_t)r * 9 + k < gpk->nbuf; ++k) gpk->buf[(size_t)r * 9 + k].release();
    }
    delete gpk;
    return B200_OK;
} B2_ABI_CATCH_INT

int b200_group_groth16_pk_info(b200_group* g, const b200_group_groth16_pk* gpk, int rank, b200_msm_set_info out[5], int* distributed_h) try {
    if (!g || !gpk || rank < 0 || rank >= g->R) return B200_ERR_ARG;
    if (distributed_h) *distributed_h = gpk->distributed ? 1 : 0;
    return out ? b200_groth16_pk_info(g->ctx[rank], gpk->pk[rank], out) : B200_OK;
} B2_ABI_CATCH_INT

int b200_group_groth16_prove_core(b200_group* g, const b200_group_groth16_pk* gpk, const uint64_t* a, const uint64_t* b,
                                  const uint64_t* c, size_t n_constraints, const uint64_t* w, const uint32_t* k_remove,
                                  size_t n_remove, b200_groth16_out* out, b200_timings* timings) try {
    if (!g || !gpk || !out || !w || (n_constraints && (!a || !b || !c))) return B200_ERR_ARG;
    return prove(g, gpk, a, b, c, n_constraints, w, nullptr, nullptr, nullptr, nullptr, k_remove, n_remove, out, timings);
} B2_ABI_CATCH_INT

int b200_group_groth16_prove_core_dev(b200_group* g, const b200_group_groth16_pk* gpk, const void* const* a_dev,
                                      const void* const* b_dev, const void* const* c_dev, const void* const* w_dev,
                                      b200_groth16_out* out, b200_timings* timings) try {
    if (!g || !gpk || !out || !a_dev || !b_dev || !c_dev || !w_dev) return B200_ERR_ARG;
    return prove(g, gpk, nullptr, nullptr, nullptr, 0, nullptr, a_dev, b_dev, c_dev, w_dev, nullptr, 0, out, timings);
} B2_ABI_CATCH_INT


// ---- PlonK over the group (P1 range-sharded commitments, P4 coset-parallel quotient) ------------------
// The session phases of plonk.cu, one session per rank, driven from the rank's host thread.  Between phases the partial
// commitments (one Jacobian point each) are added on the host by every thread from a shared table; after phase 2 every
// rank pulls the numerator's coset blocks it does not own from the owners' buffers (peer copies on its own stream).
struct b200_group_plonk_pk {
    std::vector<b200_plonk_pk*> pk;
};

int b200_group_plonk_pk_upload(b200_group* g, int curve, const b200_plonk_pk_desc* desc, b200_group_plonk_pk** out) try {
    if (!g || !desc || !out) return B200_ERR_ARG;
    *out = nullptr;
    b200_group_plonk_pk* gpk = new b200_group_plonk_pk();
    gpk->pk.assign(g->R, nullptr);
    const int st = run_ranks(g, [&](int r) -> int {
        b200_plonk_pk_desc d = *desc;
        d.shard_index = r;
        d.shard_count = g->R;
        return b200_plonk_pk_upload(g->ctx[r], curve, &d, &gpk->pk[r]);
    });
    if (st != B200_OK) { b200_group_plonk_pk_free(g, gpk); return st; }
    *out = gpk;
    return B200_OK;
} B2_ABI_CATCH_INT

int b200_group_plonk_pk_free(b200_group* g, b200_group_plonk_pk* gpk) try {
    if (!g) return B200_ERR_ARG;
    if (!gpk) return B200_OK;
    for (int r = 0; r < g->R && (size_t)r < gpk->pk.size(); ++r)
        if (gpk->pk[r]) b200_plonk_pk_free(g->ctx[r], gpk->pk[r]);
    delete gpk;
    return B200_OK;
} B2_ABI_CATCH_INT

int b200_group_plonk_prove(b200_group* g, const b200_group_plonk_pk* gpk, const uint64_t* l, const uint64_t* r_, const uint64_t* o,
                           const uint64_t* public_witness, const uint64_t* blinding, const b200_plonk_bsb22* bsb,
                           b200_plonk_proof* out, b200_plonk_timings* timings) try {
    if (!g || !gpk || !l || !r_ || !o || !out) return B200_ERR_ARG;
    const int R = g->R;
    uint64_t drawn[36];
    if (!blinding) {      // every rank must blind with the same polynomials
        if (plonk_draw_blinding(drawn) != B200_OK) return g->fail(B200_ERR_ARG, "no entropy source");
        blinding = drawn;
    }
    constexpr int JW = 3 * B200_FP_WORDS;                 // u64 words of a G1Jac
    std::vector<uint64_t> table((size_t)R * 3 * JW);       // [rank][up to 3 partials]
    std::vector<void*> num_ptr(R, nullptr);
    std::vector<b200_plonk_proof> proofs(R);
    std::vector<b200_plonk_timings> tms(R);
    size_t block_elems = 0;
    unsigned log_blocks = 0;
    auto totals = [&](int count, uint64_t* dst) {          // sum over the ranks, on this thread
        for (int k = 0; k < count; ++k) {
            memcpy(dst + JW * k, table.data() + JW * k, JW * 8);
            for (int q = 1; q < R; ++q) hostgrp::jac_add<Fp>(dst + JW * k, table.data() + ((size_t)q * 3 + k) * JW, dst + JW * k);
        }
    };
    const int st = run_ranks(g, [&](int r) -> int {
        b200_ctx* ctx = g->ctx[r];
        b200_plonk_session* S = nullptr;
        uint64_t part[3 * JW], tot[3 * JW];
        struct Guard { b200_ctx* c; b200_plonk_session** s; ~Guard() { if (*s) b200_plonk_session_finish(c, *s, nullptr, nullptr, nullptr); } } guard{ctx, &S};
        auto publish = [&](int count) -> int {             // partials -> table; all ranks; totals
            memcpy(table.data() + (size_t)r * 3 * JW, part, (size_t)count * JW * 8);
            if (!g->barrier.wait()) return ctx->fail(B200_ERR_INTERNAL, "another rank failed");
            totals(count, tot);
            if (!g->barrier.wait()) return ctx->fail(B200_ERR_INTERNAL, "another rank failed");   // the table may be overwritten again
            return B200_OK;
        };
        B2_TRY(b200_plonk_session_begin(ctx, gpk->pk[r], l, r_, o, public_witness, blinding, bsb, &S, part));
        B2_TRY(publish(3));
        B2_TRY(b200_plonk_session_step(ctx, S, 1, tot, part));
        B2_TRY(publish(1));
        B2_TRY(b200_plonk_session_step(ctx, S, 2, tot, nullptr));
        // the numerator: coset k was evaluated by rank k mod R into block bitrev(k) of its buffer
        size_t be = 0;
        unsigned lb = 0;
        B2_TRY(b200_plonk_session_numerator(ctx, S, &num_ptr[r], &be, &lb));
        B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (r == 0) { block_elems = be; log_blocks = lb; }
        if (!g->barrier.wait()) return ctx->fail(B200_ERR_INTERNAL, "another rank failed");
        const unsigned rho = 1u << log_blocks;
        for (unsigned k = 0; k < rho; ++k) {
            const int owner = (int)(k % (unsigned)R);
            if (owner == r) continue;
            unsigned blk = 0;
            for (unsigned b = 0; b < log_blocks; ++b) blk |= ((k >> b) & 1u) << (log_blocks - 1 - b);
            const size_t off = (size_t)blk * block_elems * sizeof(Fr);
            B2_CUDA(ctx, cudaMemcpyPeerAsync((char*)num_ptr[r] + off, ctx->device, (const char*)num_ptr[owner] + off, g->ctx[owner]->device,
                                             block_elems * sizeof(Fr), ctx->stream));
        }
        B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (!g->barrier.wait()) return ctx->fail(B200_ERR_INTERNAL, "another rank failed");   // nobody overwrites a block another rank still reads
        B2_TRY(b200_plonk_session_step(ctx, S, 3, nullptr, part));
        B2_TRY(publish(3));
        B2_TRY(b200_plonk_session_step(ctx, S, 4, tot, part));
        B2_TRY(publish(2));
        B2_TRY(b200_plonk_session_step(ctx, S, 5, tot, part));
        B2_TRY(publish(1));
        b200_plonk_session* fin = S;
        S = nullptr;                                        // finish frees the session
        return b200_plonk_session_finish(ctx, fin, tot, &proofs[r], &tms[r]);
    });
    if (st != B200_OK) return st;
    *out = proofs[0];
    if (timings) *timings = tms[0];
    return B200_OK;
} B2_ABI_CATCH_INT

}  // extern "C"
