// Groth16 prover core on the device: resident proving key, computeH, wire filtering and the five
// MultiExp of /root/reference/backend/groth16/bn254/prove.go:139-175,199-301 (device shape modelled
// on /root/reference/backend/groth16/bn254/icicle/icicle.go:31-130,196-393), plus the host-side
// fix-ups (prove.go:178-321) and Proof.WriteTo (marshal.go:41-66) for callers without Go.
#include <algorithm>
#include <cstdlib>
#include "groth16_host.cuh"
#include "misc_host.cuh"

namespace b200 {

static void shard_range(size_t len, int idx, int cnt, size_t* lo, size_t* hi) {
    *lo = len * (size_t)idx / (size_t)cnt;
    *hi = len * (size_t)(idx + 1) / (size_t)cnt;
}

// G1.Z is consumed against h in position order (bit-reversed), so with a power-of-two shard count
// its ranges follow the block layout of the distributed computeH (ntt.cuh, layout B): shard i owns
// positions [i n / count, (i+1) n / count) cut to len(Z).  Other counts: the balanced split.
static void shard_range_z(size_t n_z, unsigned log_n, int idx, int cnt, size_t* lo, size_t* hi) {
    const size_t n = (size_t)1 << log_n;
    if ((cnt & (cnt - 1)) == 0 && (size_t)cnt <= n) {
        const size_t block = n / (size_t)cnt;
        *lo = std::min(n_z, block * (size_t)idx);
        *hi = std::min(n_z, block * (size_t)(idx + 1));
    } else {
        shard_range(n_z, idx, cnt, lo, hi);
    }
}

// raw: the host bytes are gnark-crypto RawEncoding points (a section of a WriteRawTo file, marshal.go:246-304); same
// size per point as the Affine struct, converted in place after the copy.  *bad collects format violations.
template <class T>
static int upload_range(b200_ctx* ctx, DevBuf& dst, const uint64_t* host, size_t lo, size_t hi, bool presliced, bool raw = false,
                        uint32_t* d_bad = nullptr) {
    const size_t cnt = hi - lo;
    B2_CUDA(ctx, dst.alloc(cnt * sizeof(T)));
    if (cnt) {
        B2_CUDA(ctx, cudaMemcpyAsync(dst.p, reinterpret_cast<const char*>(host) + (presliced ? 0 : lo) * sizeof(T), cnt * sizeof(T),
                                     cudaMemcpyHostToDevice, ctx->stream));
        if (raw) B2_TRY(points_from_raw_inplace(ctx, dst.p, cnt, sizeof(T) == sizeof(G1Affine) ? 1 : 2, d_bad));
    }
    return B200_OK;
}

// Setup (setup.go:251,308): the same range as [scalar_i] * base from the device fixed-base kernel instead of a copy.
// scalars: host pointer (full array, indexed from lo) or, when on_device, a device pointer to the full array.
template <class T>
static int generate_range(b200_ctx* ctx, DevBuf& dst, const Fr* scalars, bool on_device, size_t lo, size_t hi, const T& base,
                          uint64_t* host_out) {
    const size_t cnt = hi - lo;
    B2_CUDA(ctx, dst.alloc(cnt * sizeof(T)));
    if (!cnt) return B200_OK;
    DevBuf sc;
    const Fr* d_sc = scalars + lo;
    if (!on_device) {
        B2_CUDA(ctx, sc.alloc(cnt * sizeof(Fr)));
        B2_CUDA(ctx, cudaMemcpyAsync(sc.p, scalars + lo, cnt * sizeof(Fr), cudaMemcpyHostToDevice, ctx->stream));
        d_sc = sc.as<Fr>();
    }
    if constexpr (sizeof(T) == sizeof(G1Affine)) B2_TRY(batch_mul_g1(ctx, base, d_sc, cnt, dst.as<G1Affine>()));
    else B2_TRY(batch_mul_g2(ctx, base, d_sc, cnt, dst.as<G2Affine>()));
    if (host_out) {
        B2_CUDA(ctx, cudaMemcpyAsync(reinterpret_cast<char*>(host_out) + lo * sizeof(T), dst.p, cnt * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream));
        B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    return B200_OK;
}

// Replace a resident base array by its folded-window table when enabled and it fits in memory.
template <class F>
static int maybe_fold(b200_ctx* ctx, DevBuf& bases, size_t n, MsmFold& fold, int group, size_t reserve_bytes) {
    fold = MsmFold();
    if (!ctx->msm_precompute || n < 4096) return B200_OK;
    const int c = ctx->msm_fold_c ? ctx->msm_fold_c : msm_fold_window(n);
    if (c == 0) return B200_OK;
    DevBuf table;
    int st = group == 1 ? msm_fold_build_g1(ctx, bases.as<G1Affine>(), n, c, reserve_bytes, table)
                        : msm_fold_build_g2(ctx, bases.as<G2Affine>(), n, c, reserve_bytes, table);
    if (st == B200_ERR_NOMEM) return B200_OK;   // keep the plain bases
    if (st != B200_OK) return st;
    bases.release();
    bases.p = table.p; bases.bytes = table.bytes;
    table.p = nullptr; table.bytes = 0;
    fold.c = c;
    fold.stride = n;
    return B200_OK;
}

// wire indices whose base is not the point at infinity, positions [lo, hi) of that list
static int wire_map(b200_ctx* ctx, std::vector<uint32_t>& map, const uint8_t* infinity, size_t n_wires, size_t expect) {
    map.clear();
    map.reserve(expect);
    for (size_t i = 0; i < n_wires; ++i)
        if (!infinity[i]) map.push_back((uint32_t)i);
    if (map.size() != expect) return ctx->fail(B200_ERR_ARG, "infinity bitmap does not match the base count");
    return B200_OK;
}
static int upload_wire_map(b200_ctx* ctx, DevBuf& dst, const std::vector<uint32_t>& map, size_t lo, size_t hi, uint32_t rebase) {
    std::vector<uint32_t> rel(map.begin() + lo, map.begin() + hi);
    for (auto& v : rel) v -= rebase;
    B2_CUDA(ctx, dst.alloc((hi - lo) * 4));
    if (hi > lo) B2_CUDA(ctx, cudaMemcpy(dst.p, rel.data(), (hi - lo) * 4, cudaMemcpyHostToDevice));
    return B200_OK;
}

int groth16_pk_upload(b200_ctx* ctx, const b200_groth16_pk_desc* d, b200_groth16_pk** out, const PkScalarSource* gen_src) {
    if (!d || !out || !d->generator || !d->coset_gen) return ctx->fail(B200_ERR_ARG, "null descriptor field");
    if (!d->infinity_a || !d->infinity_b) return ctx->fail(B200_ERR_ARG, "null infinity bitmap");
    if (!gen_src && ((d->n_a && !d->g1_a) || (d->n_b && (!d->g1_b || !d->g2_b)) || (d->n_k && !d->g1_k) || (d->n_z && !d->g1_z)))
        return ctx->fail(B200_ERR_ARG, "null base array");
    if (d->shard_count < 1 || d->shard_index < 0 || d->shard_index >= d->shard_count)
        return ctx->fail(B200_ERR_ARG, "bad shard");
    if (d->n_wires < d->nb_public || d->n_k > d->n_wires - d->nb_public)
        return ctx->fail(B200_ERR_ARG, "K longer than the private wires");
    if (d->n_z > ((size_t)1 << d->log_n)) return ctx->fail(B200_ERR_ARG, "Z longer than the domain");
    if (d->n_g2b != d->n_b) return ctx->fail(B200_ERR_ARG, "G1.B and G2.B differ in length");
    b200_groth16_pk* pk = new b200_groth16_pk();
    pk->log_n = d->log_n;
    pk->n_wires = d->n_wires;
    pk->nb_public = d->nb_public;
    pk->full_n_k = d->n_k;
    Fr gen, coset;
    memcpy(gen.l, d->generator, 32);
    memcpy(coset.l, d->coset_gen, 32);
    int st = ntt_domain_get(ctx, d->log_n, gen, coset, &pk->dom);
    if (st == B200_OK) ++pk->dom->refs;
    const int si = d->shard_index, sc = d->shard_count;
    const bool ps = d->presliced != 0;
    size_t lo, hi;
    // Only the wires this shard's A, B and K ranges reference are copied per proof: [w_lo, w_hi)
    std::vector<uint32_t> wa, wb;
    size_t a_lo = 0, a_hi = 0, b_lo = 0, b_hi = 0;
    if (st == B200_OK) st = wire_map(ctx, wa, d->infinity_a, d->n_wires, d->n_a);
    if (st == B200_OK) st = wire_map(ctx, wb, d->infinity_b, d->n_wires, d->n_b);
    if (st == B200_OK) {
        shard_range(d->n_a, si, sc, &a_lo, &a_hi); pk->n_a = a_hi - a_lo;
        shard_range(d->n_b, si, sc, &b_lo, &b_hi); pk->n_b = b_hi - b_lo;
        shard_range(d->n_k, si, sc, &lo, &hi); pk->n_k = hi - lo; pk->k_lo = lo;
        size_t w_lo = d->n_wires, w_hi = 0;
        if (a_hi > a_lo) { w_lo = std::min<size_t>(w_lo, wa[a_lo]); w_hi = std::max<size_t>(w_hi, (size_t)wa[a_hi - 1] + 1); }
        if (b_hi > b_lo) { w_lo = std::min<size_t>(w_lo, wb[b_lo]); w_hi = std::max<size_t>(w_hi, (size_t)wb[b_hi - 1] + 1); }
        if (pk->n_k) {
            // K scalars: W[nb_public + k_lo ...), shifted up by at most the committed wires removed per proof
            w_lo = std::min<size_t>(w_lo, d->nb_public + pk->k_lo);
            const size_t slack = d->n_wires - d->nb_public - d->n_k;   // = number of committed private wires
            w_hi = std::max<size_t>(w_hi, std::min<size_t>(d->n_wires, d->nb_public + pk->k_lo + pk->n_k + slack));
        }
        if (w_hi < w_lo) w_lo = w_hi = 0;
        pk->w_lo = w_lo; pk->w_hi = w_hi;
    }
    const bool raw = d->raw_points != 0;
    DevBuf bad;
    if (st == B200_OK && raw) {
        if (bad.alloc(sizeof(uint32_t)) != cudaSuccess || cudaMemsetAsync(bad.p, 0, sizeof(uint32_t), ctx->stream) != cudaSuccess)
            st = ctx->fail(B200_ERR_CUDA, "pk upload");
    }
    uint32_t* d_bad = raw ? bad.as<uint32_t>() : nullptr;
    if (gen_src) {
        // Setup: this shard's ranges come out of the fixed-base kernel (and, optionally, go back to the caller's arrays)
        const PkScalarSource& g = *gen_src;
        if (st == B200_OK) st = generate_range<G1Affine>(ctx, pk->A, g.a, false, a_lo, a_hi, g.g1, g.out_a);
        if (st == B200_OK) st = generate_range<G1Affine>(ctx, pk->B, g.b, false, b_lo, b_hi, g.g1, g.out_b);
        if (st == B200_OK) st = generate_range<G2Affine>(ctx, pk->G2B, g.b, false, b_lo, b_hi, g.g2, g.out_g2b);
        if (st == B200_OK) st = generate_range<G1Affine>(ctx, pk->K, g.k, false, pk->k_lo, pk->k_lo + pk->n_k, g.g1, g.out_k);
        if (st == B200_OK) { shard_range_z(d->n_z, d->log_n, si, sc, &lo, &hi); pk->n_z = hi - lo; pk->z_lo = lo;
            st = generate_range<G1Affine>(ctx, pk->Z, g.z_dev, true, lo, hi, g.g1, g.out_z); }
    } else {
    if (st == B200_OK) st = upload_range<G1Affine>(ctx, pk->A, d->g1_a, a_lo, a_hi, ps, raw, d_bad);
    if (st == B200_OK) st = upload_range<G1Affine>(ctx, pk->B, d->g1_b, b_lo, b_hi, ps, raw, d_bad);
    if (st == B200_OK) st = upload_range<G2Affine>(ctx, pk->G2B, d->g2_b, b_lo, b_hi, ps, raw, d_bad);
    if (st == B200_OK) st = upload_range<G1Affine>(ctx, pk->K, d->g1_k, pk->k_lo, pk->k_lo + pk->n_k, ps, raw, d_bad);
    if (st == B200_OK) { shard_range_z(d->n_z, d->log_n, si, sc, &lo, &hi); pk->n_z = hi - lo; pk->z_lo = lo;
        st = upload_range<G1Affine>(ctx, pk->Z, d->g1_z, lo, hi, ps, raw, d_bad); }
    }
    if (st == B200_OK) st = upload_wire_map(ctx, pk->mapA, wa, a_lo, a_hi, (uint32_t)pk->w_lo);
    if (st == B200_OK) st = upload_wire_map(ctx, pk->mapB, wb, b_lo, b_hi, (uint32_t)pk->w_lo);
    if (st == B200_OK && raw) {
        uint32_t flag = 0;
        if (cudaMemcpyAsync(&flag, bad.p, sizeof flag, cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess ||
            cudaStreamSynchronize(ctx->stream) != cudaSuccess) st = ctx->fail(B200_ERR_CUDA, "pk upload");
        else if (flag) st = ctx->fail(B200_ERR_ARG, "raw_points: a base array is not raw-encoded (compressed-form flag or coordinate >= p)");
    }
    if (st == B200_OK && cudaStreamSynchronize(ctx->stream) != cudaSuccess) st = ctx->fail(B200_ERR_CUDA, "pk upload");
    // Folded tables take what is left after the prover's own working set: the domain tables (built
    // below), a/b/c/W, and the largest MSM scratch.  G2 first: it gains the most per byte.
    const size_t n_dom = (size_t)1 << d->log_n;
    const size_t n_max = std::max(std::max(pk->n_a, pk->n_b), std::max(pk->n_k, pk->n_z));
    const int fc = ctx->msm_fold_c ? ctx->msm_fold_c : msm_fold_window(n_max);
    const size_t reserve = 9 * n_dom * sizeof(Fr) + d->n_wires * sizeof(Fr) + d->n_wires * 4
                         + msm_workspace_bytes(n_max, 2, ctx->msm_mode(), fc)
                         + msm_sizes(n_max, 2, ctx->msm_mode(), fc).sort_bytes + ((size_t)1 << 30);   // two sort groups
    if (st == B200_OK) st = maybe_fold<Fp2>(ctx, pk->G2B, pk->n_b, pk->fG2B, 2, reserve);
    if (st == B200_OK) st = maybe_fold<Fp>(ctx, pk->A, pk->n_a, pk->fA, 1, reserve);
    if (st == B200_OK) st = maybe_fold<Fp>(ctx, pk->B, pk->n_b, pk->fB, 1, reserve);
    if (st == B200_OK) st = maybe_fold<Fp>(ctx, pk->K, pk->n_k, pk->fK, 1, reserve);
    if (st == B200_OK) st = maybe_fold<Fp>(ctx, pk->Z, pk->n_z, pk->fZ, 1, reserve);
    // build the computeH tables now so the first proof does not pay for them
    const Fr* t;
    if (st == B200_OK) st = ntt_domain_table(ctx, pk->dom, NttDomain::H_A, &t);
    if (st == B200_OK) st = ntt_domain_table(ctx, pk->dom, NttDomain::H_BC, &t);
    if (st == B200_OK) st = ntt_domain_table(ctx, pk->dom, NttDomain::COSETINV_BR, &t);
    if (st == B200_OK && cudaStreamSynchronize(ctx->stream) != cudaSuccess)
        st = ctx->fail(B200_ERR_CUDA, "pk upload", cudaGetErrorString(cudaGetLastError()));
    if (st != B200_OK) { delete pk; return st; }
    *out = pk;
    return B200_OK;
}

static float ev_ms(cudaEvent_t a, cudaEvent_t b) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, a, b);
    return ms;
}

int groth16_prove_core(b200_ctx* ctx, const b200_groth16_pk* pk, const uint64_t* a, const uint64_t* b,
                       const uint64_t* c, size_t n_constraints, const uint64_t* w, const uint32_t* k_remove,
                       size_t n_remove, const uint64_t* h_local, b200_groth16_out* out, b200_timings* tm, const ProveHHook* hook) {
    const size_t n = (size_t)1 << pk->log_n;
    // hook: computeH is the caller's (the group's distributed one, group.cu), enqueued where this function would run its
    // own; h lands in hook->h_out.  Otherwise h_local != null means h was computed before the call.
    const bool own_h = !h_local && !hook;      // a, b, c are copied into the workspace and transformed here
    if (hook) h_local = reinterpret_cast<const uint64_t*>(hook->h_out);
    if (!own_h) n_constraints = 0;   // h comes from the distributed computeH: a, b, c are not read
    if (n_constraints > n) return ctx->fail(B200_ERR_ARG, "more constraints than the domain holds");
    if (n_remove && pk->full_n_k + n_remove != pk->n_wires - pk->nb_public)
        return ctx->fail(B200_ERR_ARG, "k_remove does not match len(pk.G1.K)");
    // The five MSMs: 0 A, 1 B1, 2 K, 3 B2 (G2), 4 Z.  Workspace: a, b, c (n Fr each), the wire slice
    // of W, optional K map, two sort groups (the sort of MSM k+1 runs on the aux stream while the
    // buckets of MSM k are accumulated on the main stream) and one accumulation group.
    const size_t n_w = pk->w_hi - pk->w_lo;   // wires this shard reads
    const size_t counts[5] = {pk->n_a, pk->n_b, pk->n_k, pk->n_b, pk->n_z};
    const int groups[5] = {1, 1, 1, 2, 1};
    const MsmFold* folds[5] = {&pk->fA, &pk->fB, &pk->fK, &pk->fG2B, &pk->fZ};
    MsmSizes sz[5];
    size_t sort_max = 0, acc_max = 0, ba_max = 0, res_off[6] = {0};
    for (int k = 0; k < 5; ++k) {
        sz[k] = msm_sizes(counts[k], groups[k], ctx->msm_mode(), folds[k]->c);
        sort_max = std::max(sort_max, sz[k].sort_bytes);
        acc_max = std::max(acc_max, sz[k].acc_bytes);
        ba_max = std::max(ba_max, sz[k].ba_bytes);
        res_off[k + 1] = res_off[k] + ((sz[k].result_bytes + 255) & ~(size_t)255);
    }
    const bool overlap = ctx->msm_overlap != 0;
    Arena ar;
    const size_t n_abc = own_h ? n : 0;
    const size_t o_a = ar.off; ar.take(n_abc * sizeof(Fr));
    const size_t o_b = ar.off; ar.take(n_abc * sizeof(Fr));
    const size_t o_c = ar.off; ar.take(n_abc * sizeof(Fr));
    const size_t o_w = ar.off; ar.take(n_w * sizeof(Fr));
    const size_t o_kmap = ar.off; ar.take(n_remove ? pk->n_k * 4 : 0);
    const size_t o_sort0 = ar.off; ar.take(sort_max);
    const size_t o_sort1 = overlap ? ar.off : o_sort0; if (overlap) ar.take(sort_max);
    const size_t o_acc = ar.off; ar.take(acc_max);
    const size_t o_acc1 = overlap ? ar.off : o_acc; if (overlap) ar.take(acc_max);   // second accumulation group: the reduction of MSM k runs beside the accumulation of k+1
    const size_t o_ba = ar.off; ar.take(ba_max);   // batched-affine round buffers: the accumulations run one after the other on the main stream
    B2_TRY(ctx_reserve(ctx, ar.off));
    if (ctx->h_stage_bytes < res_off[5]) {   // page-locked staging of the few result points per MSM
        if (ctx->h_stage) cudaFreeHost(ctx->h_stage);
        ctx->h_stage = nullptr; ctx->h_stage_bytes = 0;
        B2_CUDA(ctx, cudaMallocHost(&ctx->h_stage, res_off[5]));
        ctx->h_stage_bytes = res_off[5];
    }
    char* base = (char*)ctx->scratch;
    Fr* d_a = (Fr*)(base + o_a); Fr* d_b = (Fr*)(base + o_b); Fr* d_c = (Fr*)(base + o_c);
    Fr* d_w = (Fr*)(base + o_w);
    uint32_t* d_kmap = n_remove ? (uint32_t*)(base + o_kmap) : nullptr;
    char* sort_ws[2] = {base + o_sort0, base + o_sort1};
    char* acc_ws[2] = {base + o_acc, base + o_acc1};
    cudaStream_t st = ctx->stream;
    cudaStream_t sx = overlap ? ctx->aux_stream : st;
    cudaStream_t cp = ctx->copy_stream;
    cudaStream_t sp = overlap ? ctx->post_stream : nullptr;

    constexpr int NEV = 34;     // created once per ctx, reused by every proof (freed by b200_destroy)
    while (ctx->core_events.size() < NEV) {
        cudaEvent_t e;
        B2_CUDA(ctx, cudaEventCreate(&e));
        ctx->core_events.push_back(e);
    }
    cudaEvent_t* ev = ctx->core_events.data();
    cudaEvent_t ev_w = ev[9], ev_abc = ev[10], ev_h = ev[11];
    cudaEvent_t* ev_sort = ev + 12;   // [5]
    cudaEvent_t* ev_acc = ev + 17;    // [5]  accumulation + split-bucket merges done: the sort group is free
    cudaEvent_t* ev_split = ev + 22;  // [5]
    cudaEvent_t* ev_post = ev + 27;   // [5]  reduction + result copy done: the accumulation group is free

    // Copies run on their own stream: W first (the four witness MSMs need only W), then a, b, c,
    // which computeH picks up later -- the PCIe traffic hides behind the first MSM.
    B2_CUDA(ctx, cudaEventRecord(ev[0], st));
    B2_CUDA(ctx, cudaStreamWaitEvent(cp, ev[0], 0));
    if (overlap) B2_CUDA(ctx, cudaStreamWaitEvent(sx, ev[0], 0));
    const int span = ctx->prof_begin(B200_PROF_COPY_IN, (double)(3 * n_constraints + n_w) * sizeof(Fr), cp);
    if (n_w) B2_CUDA(ctx, cudaMemcpyAsync(d_w, w + 4 * pk->w_lo, n_w * sizeof(Fr), cudaMemcpyDefault, cp));
    if (n_remove) {
        // prove.go:238-242: K scalars are W[nb_public:] minus the committed wires.  The map lives in the ctx until the
        // next proof, so the copy needs no host synchronisation in the middle of the enqueue.
        std::vector<uint32_t>& map = ctx->kmap_host;
        map.clear();
        map.reserve(pk->full_n_k);
        size_t r = 0;
        for (size_t i = 0; i < pk->n_wires - pk->nb_public; ++i) {
            if (r < n_remove && k_remove[r] == i) { ++r; continue; }
            map.push_back((uint32_t)(pk->nb_public + i - pk->w_lo));
        }
        if (map.size() != pk->full_n_k || r != n_remove) return ctx->fail(B200_ERR_ARG, "k_remove must be sorted, unique and inside W[nb_public:]");
        B2_CUDA(ctx, cudaMemcpyAsync(d_kmap, map.data() + pk->k_lo, pk->n_k * 4, cudaMemcpyHostToDevice, cp));
    }
    B2_CUDA(ctx, cudaEventRecord(ev_w, cp));
    const uint64_t* src[3] = {a, b, c};
    Fr* dst[3] = {d_a, d_b, d_c};
    for (int i = 0; i < 3 && own_h; ++i) {
        if (n_constraints) B2_CUDA(ctx, cudaMemcpyAsync(dst[i], src[i], n_constraints * sizeof(Fr), cudaMemcpyDefault, cp));
        if (n > n_constraints) B2_CUDA(ctx, cudaMemsetAsync(dst[i] + n_constraints, 0, (n - n_constraints) * sizeof(Fr), cp));
    }
    if (hook) B2_TRY(hook->copy_abc(cp));
    ctx->prof_end(span, cp);
    B2_CUDA(ctx, cudaEventRecord(ev_abc, cp));

    // computeH goes first when a, b, c are already on the device (nothing to wait for), else after
    // the first MSM (by then the host copies have landed)
    bool h_first = hook ? hook->inputs_on_device : false;
    if (own_h && n_constraints) {
        cudaPointerAttributes pa;
        if (cudaPointerGetAttributes(&pa, a) == cudaSuccess) h_first = pa.type == cudaMemoryTypeDevice;
        else cudaGetLastError();
    }
    const Fr* h_z = h_local ? reinterpret_cast<const Fr*>(h_local) : d_a + pk->z_lo;
    const void* bases[5] = {pk->A.p, pk->B.p, pk->K.p, pk->G2B.p, pk->Z.p};
    const Fr* scal[5] = {d_w, d_w, d_kmap ? d_w : d_w + (pk->nb_public + pk->k_lo - pk->w_lo), d_w, h_z};
    const uint32_t* maps[5] = {pk->mapA.as<uint32_t>(), pk->mapB.as<uint32_t>(), d_kmap, pk->mapB.as<uint32_t>(), nullptr};
    void* jobs[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    struct JobGuard { void** j; const int* g; ~JobGuard() { for (int k = 0; k < 5; ++k) if (j[k]) { if (g[k] == 2) msm_g2_job_delete(j[k]); else msm_g1_job_delete(j[k]); } } } jguard{jobs, groups};
    // G1.B and G2.B carry the same wires (InfinityB, prove.go:160-175): when both run the same plan,
    // Bs2 reuses the sorted entries and task lists of Bs1 and its own sort is skipped.
    const bool share_b = overlap && pk->fB.c == pk->fG2B.c && pk->n_b > 0;   // needs Bs1's group kept while K sorts
    const int sort_slot[5] = {0, 1, 0, 1, 0};
    for (int k = 0; k < 5; ++k) {
        void* stage = (char*)ctx->h_stage + res_off[k];
        char* sw = sort_ws[overlap ? sort_slot[k] : 0];
        char* aw = acc_ws[overlap ? (k & 1) : 0];
        jobs[k] = groups[k] == 2 ? msm_g2_job_new(ctx, (const G2Affine*)bases[k], scal[k], maps[k], counts[k], folds[k], sw, aw, stage, base + o_ba)
                                 : msm_g1_job_new(ctx, (const G1Affine*)bases[k], scal[k], maps[k], counts[k], folds[k], sw, aw, stage, base + o_ba);
    }
    auto run_h = [&]() -> int {
        B2_CUDA(ctx, cudaStreamWaitEvent(st, ev_abc, 0));
        B2_CUDA(ctx, cudaEventRecord(ev[6], st));
        if (hook) B2_TRY(hook->compute_h());
        else B2_TRY(ntt_compute_h(ctx, pk->dom, d_a, d_b, d_c));
        B2_CUDA(ctx, cudaEventRecord(ev[7], st));
        B2_CUDA(ctx, cudaEventRecord(ev_h, st));
        return B200_OK;
    };
    // sort group of MSM k is free once the previous user of the same slot has been accumulated:
    // slots 0: A, K, Z;  1: B1, B2 -- so K waits for A, B2 for B1, Z for K
    auto run_sort = [&](int k) -> int {
        if (k < 4) B2_CUDA(ctx, cudaStreamWaitEvent(sx, ev_w, 0));
        else if (own_h || hook) B2_CUDA(ctx, cudaStreamWaitEvent(sx, ev_h, 0));
        if (k == 0 && h_first && !ctx->h_sort_overlap) B2_CUDA(ctx, cudaStreamWaitEvent(sx, ev_h, 0));  // the transforms and a sort only slow each other down
        if (k == 2) B2_CUDA(ctx, cudaStreamWaitEvent(sx, ev_acc[0], 0));
        if (k == 3) B2_CUDA(ctx, cudaStreamWaitEvent(sx, ev_acc[1], 0));
        if (k == 4) B2_CUDA(ctx, cudaStreamWaitEvent(sx, ev_acc[2], 0));
        // every sort but the first runs beside the bucket stage of an earlier MSM: small persistent grids (msm_host.cuh)
        struct Persist { b200_ctx* c; ~Persist() { c->aux_persist = 0; } } persist{ctx};
        ctx->aux_persist = (overlap && (k > 0 || (h_first && ctx->h_sort_overlap))) ? 1 : 0;   // k = 0: beside computeH when the inputs are resident
        if (k == 3 && share_b) B2_TRY(msm_g2_job_prep0(ctx, jobs[k], sx));   // batched-affine rounds: Bs2's own denominator pass over Bs1's entries
        else B2_TRY(groups[k] == 2 ? msm_g2_job_sort(ctx, jobs[k], sx) : msm_g1_job_sort(ctx, jobs[k], sx));
        B2_CUDA(ctx, cudaEventRecord(ev_sort[k], sx));
        return B200_OK;
    };
    auto run_acc = [&](int k) -> int {
        B2_CUDA(ctx, cudaStreamWaitEvent(st, ev_sort[k], 0));
        if (sp && k >= 2) B2_CUDA(ctx, cudaStreamWaitEvent(st, ev_post[k - 2], 0));   // the accumulation group of MSM k-2
        B2_TRY(groups[k] == 2 ? msm_g2_job_acc(ctx, jobs[k], st, sp, ev_split[k]) : msm_g1_job_acc(ctx, jobs[k], st, sp, ev_split[k]));
        B2_CUDA(ctx, cudaEventRecord(ev_acc[k], st));
        if (sp) B2_CUDA(ctx, cudaEventRecord(ev_post[k], sp));
        return B200_OK;
    };
    B2_CUDA(ctx, cudaEventRecord(ev[6], st));   // placeholders when computeH is skipped (h supplied)
    B2_CUDA(ctx, cudaEventRecord(ev[7], st));
    if (h_first) B2_TRY(run_h());
    // Host-buffer proofs: computeH waits for a, b, c to cross PCIe behind W.  On one GPU the first MSM covers that copy;
    // in a device group every rank's MSMs are 1/R as long while the ranks share the host's memory system for their
    // copies, so computeH moves behind more of them (h_pos).  The sort of Z, the only consumer of h, comes after all of them.
    const bool late_h = (own_h || hook) && !h_first;
    int h_pos = ctx->h_pos;
    {
        static const char* e = getenv("B200_H_POS");
        if (e && *e) h_pos = atoi(e);
    }
    if (h_pos < 0 || h_pos > 2) h_pos = 2;   // measured: e2e 0.2463 -> 0.2445 s on one GPU, 43.5 -> 42.7 ms on eight (positions 1 and 2 equal there)
    cudaEvent_t stage_start[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};   // != null: the stage starts at the end of computeH
    if (overlap) {
        B2_TRY(run_sort(0));
        B2_TRY(run_sort(1));
        B2_TRY(run_acc(0));
        if (late_h && h_pos == 0) { B2_TRY(run_h()); stage_start[1] = ev[7]; }
        B2_TRY(run_sort(2));
        B2_TRY(run_acc(1));
        if (late_h && h_pos == 1) { B2_TRY(run_h()); stage_start[2] = ev[7]; }
        B2_TRY(run_sort(3));
        B2_TRY(run_acc(2));
        if (late_h && h_pos == 2) { B2_TRY(run_h()); stage_start[3] = ev[7]; }
        B2_TRY(run_sort(4));
        B2_TRY(run_acc(3));
        B2_TRY(run_acc(4));
    } else {   // one stream, one sort group: strictly sort, accumulate, sort, ...
        for (int k = 0; k < 5; ++k) {
            if (k == 1 && late_h) { B2_TRY(run_h()); stage_start[1] = ev[7]; }
            B2_TRY(run_sort(k));
            B2_TRY(run_acc(k));
        }
    }
    if (sp) {    // the step ends when the last reductions have delivered their results
        B2_CUDA(ctx, cudaStreamWaitEvent(st, ev_post[3], 0));
        B2_CUDA(ctx, cudaStreamWaitEvent(st, ev_post[4], 0));
    }
    B2_CUDA(ctx, cudaEventRecord(ev[8], st));
    B2_CUDA(ctx, cudaStreamSynchronize(st));
    if (sp) B2_CUDA(ctx, cudaStreamSynchronize(sp));
    B2_CUDA(ctx, cudaStreamSynchronize(cp));
    if (overlap) B2_CUDA(ctx, cudaStreamSynchronize(sx));

    G1Jac ja, jb1, jk, jz;
    G2Jac jb2;
    msm_g1_job_finish(jobs[0], &ja);
    msm_g1_job_finish(jobs[1], &jb1);
    msm_g1_job_finish(jobs[2], &jk);
    msm_g2_job_finish(jobs[3], &jb2);
    msm_g1_job_finish(jobs[4], &jz);

    memcpy(out->ar, &ja, sizeof ja);
    memcpy(out->bs1, &jb1, sizeof jb1);
    memcpy(out->krs, &jk, sizeof jk);
    memcpy(out->krs2, &jz, sizeof jz);
    memcpy(out->bs2, &jb2, sizeof jb2);
    if (tm) {
        // Main-stream intervals: an MSM's figure is its bucket accumulation + reduction (its sort ran
        // on the aux stream under the previous MSM) plus any wait for that sort.
        tm->h2d_ms = ev_ms(ev[0], ev_abc);           // on the copy stream, overlapped
        tm->compute_h_ms = ev_ms(ev[6], ev[7]);
        const cudaEvent_t a_start = h_first ? ev[7] : ev[0];
        tm->msm_a_ms = ev_ms(a_start, ev_acc[0]);
        tm->msm_b1_ms = ev_ms(stage_start[1] ? stage_start[1] : ev_acc[0], ev_acc[1]);
        tm->msm_k_ms = ev_ms(stage_start[2] ? stage_start[2] : ev_acc[1], ev_acc[2]);
        tm->msm_b2_ms = ev_ms(stage_start[3] ? stage_start[3] : ev_acc[2], ev_acc[3]);
        tm->msm_z_ms = ev_ms(ev_acc[3], ev_acc[4]);
        tm->total_ms = ev_ms(ev[0], ev[8]);
    }
    return B200_OK;
}

}  // namespace b200
