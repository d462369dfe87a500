// PlonK prover (BN254): the host orchestration of /root/reference/backend/plonk/bn254/prove.go:116-1389 above
// the device kernels of poly.cu / ntt.cu / msm*.cu.  The reference runs eleven goroutines synchronised by
// channels; the data dependencies are a chain ordered by four Fiat-Shamir challenges
// (gamma,beta <- LRO; alpha <- Z; zeta <- H; KZG gamma <- openings), which is how the stages below are ordered.
//
// B200-first choices (results are the reference's, field arithmetic is exact):
//   * computeNumerator (prove.go:837-1079) transforms every polynomial ToCanonical -> scale -> ToLagrange on each
//     of the rho cosets (~2 rho 15 transforms).  Here the key's polynomials are evaluated on the rho cosets ONCE
//     at upload (the trade prove.go:995-999 declines for memory; 8 rho n elements = 4 GiB at n = 2^22), and the
//     five per-proof polynomials (L, R, O, Z, Qk) go to canonical form once and then take one forward transform
//     per coset with the coset scaling fused in the first pass: 5 + 5 rho transforms instead of ~30 rho.
//   * canonical coefficients are kept in bit-reversed order so that every coset evaluation is a DIT transform
//     (bit-reversed in, natural out) and no separate permutation pass exists in the quotient loop.
//   * commitments are the folded-window G1 MSM against two resident SRS tables.
// Fiat-Shamir (SHA-256) and the O(1) group arithmetic (blinding commitments, folded digests) run on the host,
// as in the reference.
#include <atomic>
#include <chrono>
#include <future>
#include <vector>
#include <memory>
#include <cstdio>
#include "host_group.cuh"
#include "msm_host.cuh"
#include "misc_host.cuh"
#include "poly_host.cuh"

using namespace b200;
using namespace b200::hostgrp;

struct b200_plonk_pk {
    unsigned log_n = 0, log_big = 0;
    size_t n = 0, big = 0, nb_public = 0;
    uint32_t rho = 0;
    NttDomain *d0 = nullptr, *d1 = nullptr;       // owned by the ctx cache
    DevBuf trace;        // ql qr qm qo qk s1 s2 s3, canonical / regular, n each
    DevBuf qk_lag;       // incomplete Qk in Lagrange / regular form (completeQk, prove.go:397-423)
    DevBuf perm;         // trace.S
    DevBuf coset_ev;     // [rho][8][n]: ql qr qm qo s1 s2 s3 LOne on coset k, Lagrange / regular
    DevBuf coset_pre;    // [rho][n]: shift_k^bitrev(pos) -- the coset scaling of a bit-reversed coefficient vector
    DevBuf work;         // per-proof workspace
    b200_bases *srs = nullptr, *srs_lag = nullptr;
    G1Affine srs_lo[3], srs_hi[3];   // pk.Kzg.G1[0..2], [n..n+2]  (commitBlindingFactor, prove.go:1159-1172)
    G1Affine vk_pts[8];              // S1 S2 S3 Ql Qr Qm Qo Qk
    mutable std::atomic<bool> in_flight{false};     // the per-proof workspace below is per KEY: one proof at a time
    uint32_t shard_index = 0, shard_count = 1;      // MSM point ranges and quotient cosets of this rank
    size_t srs_lo_idx = 0, srs_hi_idx = 0, lag_lo = 0, lag_hi = 0;
    // BSB22 (api.Commit circuits)
    uint32_t nb_bsb = 0;
    DevBuf qcp;          // [nb_bsb][n] trace.Qcp, canonical / regular
    DevBuf qcp_ev;       // [rho][nb_bsb][n] coset evaluations
    DevBuf bsb_work;     // per proof: [nb_bsb] x (n bit-reversed canonical | n regular canonical | n coset evaluation)
    G1Affine vk_qcp[B200_PLONK_MAX_BSB22];
    size_t commit_row[B200_PLONK_MAX_BSB22];   // nb_public + CommitmentIndex
    const Fr* qcp_c(uint32_t k) const { return qcp.as<Fr>() + (size_t)k * n; }
    const Fr* qcp_e(uint32_t coset, uint32_t k) const { return qcp_ev.as<Fr>() + ((size_t)coset * nb_bsb + k) * n; }
    const Fr* tr(int k) const { return trace.as<Fr>() + (size_t)k * n; }
    const Fr* cev(uint32_t coset, int k) const { return coset_ev.as<Fr>() + ((size_t)coset * 8 + k) * n; }
};
enum { TR_QL = 0, TR_QR, TR_QM, TR_QO, TR_QK, TR_S1, TR_S2, TR_S3 };
enum { CE_QL = 0, CE_QR, CE_QM, CE_QO, CE_S1, CE_S2, CE_S3, CE_LONE };

namespace {

// ---- SHA-256 (FIPS 180-4) ---------------------------------------------------------------------------
struct Sha256 {
    uint32_t h[8];
    uint8_t buf[64];
    uint64_t len = 0;
    size_t fill = 0;
    Sha256() {
        static const uint32_t iv[8] = {0x6a09e667u, 0xbb67ae85u, 0x3c6ef372u, 0xa54ff53au, 0x510e527fu, 0x9b05688cu, 0x1f83d9abu, 0x5be0cd19u};
        memcpy(h, iv, sizeof h);
    }
    static uint32_t rotr(uint32_t x, int k) { return (x >> k) | (x << (32 - k)); }
    void block(const uint8_t* p) {
        static const uint32_t K[64] = {
            0x428a2f98u, 0x71374491u, 0xb5c0fbcfu, 0xe9b5dba5u, 0x3956c25bu, 0x59f111f1u, 0x923f82a4u, 0xab1c5ed5u, 0xd807aa98u, 0x12835b01u,
            0x243185beu, 0x550c7dc3u, 0x72be5d74u, 0x80deb1feu, 0x9bdc06a7u, 0xc19bf174u, 0xe49b69c1u, 0xefbe4786u, 0x0fc19dc6u, 0x240ca1ccu,
            0x2de92c6fu, 0x4a7484aau, 0x5cb0a9dcu, 0x76f988dau, 0x983e5152u, 0xa831c66du, 0xb00327c8u, 0xbf597fc7u, 0xc6e00bf3u, 0xd5a79147u,
            0x06ca6351u, 0x14292967u, 0x27b70a85u, 0x2e1b2138u, 0x4d2c6dfcu, 0x53380d13u, 0x650a7354u, 0x766a0abbu, 0x81c2c92eu, 0x92722c85u,
            0xa2bfe8a1u, 0xa81a664bu, 0xc24b8b70u, 0xc76c51a3u, 0xd192e819u, 0xd6990624u, 0xf40e3585u, 0x106aa070u, 0x19a4c116u, 0x1e376c08u,
            0x2748774cu, 0x34b0bcb5u, 0x391c0cb3u, 0x4ed8aa4au, 0x5b9cca4fu, 0x682e6ff3u, 0x748f82eeu, 0x78a5636fu, 0x84c87814u, 0x8cc70208u,
            0x90befffau, 0xa4506cebu, 0xbef9a3f7u, 0xc67178f2u};
        uint32_t w[64];
        for (int i = 0; i < 16; ++i) w[i] = (uint32_t)p[4 * i] << 24 | (uint32_t)p[4 * i + 1] << 16 | (uint32_t)p[4 * i + 2] << 8 | p[4 * i + 3];
        for (int i = 16; i < 64; ++i) {
            const uint32_t s0 = rotr(w[i - 15], 7) ^ rotr(w[i - 15], 18) ^ (w[i - 15] >> 3);
            const uint32_t s1 = rotr(w[i - 2], 17) ^ rotr(w[i - 2], 19) ^ (w[i - 2] >> 10);
            w[i] = w[i - 16] + s0 + w[i - 7] + s1;
        }
        uint32_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4], f = h[5], g = h[6], hh = h[7];
        for (int i = 0; i < 64; ++i) {
            const uint32_t t1 = hh + (rotr(e, 6) ^ rotr(e, 11) ^ rotr(e, 25)) + ((e & f) ^ (~e & g)) + K[i] + w[i];
            const uint32_t t2 = (rotr(a, 2) ^ rotr(a, 13) ^ rotr(a, 22)) + ((a & b) ^ (a & c) ^ (b & c));
            hh = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
        }
        h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e; h[5] += f; h[6] += g; h[7] += hh;
    }
    void update(const uint8_t* p, size_t n) {
        len += n;
        while (n) {
            const size_t k = 64 - fill < n ? 64 - fill : n;
            memcpy(buf + fill, p, k);
            fill += k; p += k; n -= k;
            if (fill == 64) { block(buf); fill = 0; }
        }
    }
    void final(uint8_t out[32]) {
        const uint64_t bits = len * 8;
        const uint8_t one = 0x80, zero = 0;
        update(&one, 1);
        while (fill != 56) update(&zero, 1);
        uint8_t lb[8];
        for (int i = 0; i < 8; ++i) lb[i] = (uint8_t)(bits >> (56 - 8 * i));
        update(lb, 8);
        for (int i = 0; i < 8; ++i) { out[4 * i] = (uint8_t)(h[i] >> 24); out[4 * i + 1] = (uint8_t)(h[i] >> 16); out[4 * i + 2] = (uint8_t)(h[i] >> 8); out[4 * i + 3] = (uint8_t)h[i]; }
    }
};

// ---- gnark-crypto fiat-shamir.Transcript: H(name || previous challenge (raw 32 bytes) || bindings) ---------
// (byte layout as restated in-tree by backend/plonk/bn254/solidity.go:407-538)
struct Transcript {
    std::vector<std::string> names;
    std::vector<std::vector<uint8_t>> bound;
    std::vector<std::vector<uint8_t>> value;
    explicit Transcript(std::initializer_list<const char*> ns) {
        for (const char* s : ns) names.push_back(s);
        bound.resize(names.size());
        value.resize(names.size());
    }
    int idx(const char* name) const {
        for (size_t i = 0; i < names.size(); ++i) if (names[i] == name) return (int)i;
        return -1;
    }
    void bind(const char* name, const uint8_t* p, size_t n) { auto& b = bound[idx(name)]; b.insert(b.end(), p, p + n); }
    void challenge(const char* name, uint8_t out[32]) {
        const int i = idx(name);
        if (value[i].empty()) {
            Sha256 s;
            s.update((const uint8_t*)names[i].data(), names[i].size());
            if (i > 0) s.update(value[i - 1].data(), 32);
            s.update(bound[i].data(), bound[i].size());
            value[i].resize(32);
            s.final(value[i].data());
        }
        memcpy(out, value[i].data(), 32);
    }
};

// fr.Element.SetBytes: big-endian integer reduced mod r, to Montgomery form
Fr fr_from_be(const uint8_t b[32]) {
    uint32_t x[8];
    for (int i = 0; i < 8; ++i) x[7 - i] = (uint32_t)b[4 * i] << 24 | (uint32_t)b[4 * i + 1] << 16 | (uint32_t)b[4 * i + 2] << 8 | b[4 * i + 3];
    for (;;) {
        bool ge = true;
        for (int i = 7; i >= 0; --i) {
            if (x[i] != FrParams::p(i)) { ge = x[i] > FrParams::p(i); break; }
        }
        if (!ge) break;
        uint64_t borrow = 0;
        for (int i = 0; i < 8; ++i) {
            const uint64_t d = (uint64_t)x[i] - FrParams::p(i) - borrow;
            x[i] = (uint32_t)d;
            borrow = (d >> 63) & 1;
        }
    }
    Fr f;
    for (int i = 0; i < 8; ++i) f.l[i] = x[i];
    return f.to_mont();
}
// fr.Element.Marshal: canonical big-endian bytes
void fr_be(const Fr& m, uint8_t out[32]) {
    const Fr x = m.from_mont();
    for (int i = 0; i < 8; ++i) {
        const uint32_t v = x.l[7 - i];
        out[4 * i] = (uint8_t)(v >> 24); out[4 * i + 1] = (uint8_t)(v >> 16); out[4 * i + 2] = (uint8_t)(v >> 8); out[4 * i + 3] = (uint8_t)v;
    }
}
// G1Affine.RawBytes / Marshal: X || Y big-endian; infinity = 0x40 then zeros
constexpr int G1_RAW = 2 * FP_BYTES;   // 64 on BN254
constexpr size_t AFF_WORDS = sizeof(G1Affine) / 8;   // u64 words of an affine G1 point at the ABI: 8 on BN254, 12 on BLS12-377
void g1_raw(const G1Affine& p, uint8_t* out) {
    if (p.is_inf()) { memset(out, 0, G1_RAW); out[0] = 0x40; return; }
    fp_be(p.x, out);
    fp_be(p.y, out + FP_BYTES);
}
void bind_point(Transcript& fs, const char* name, const G1Affine& p) { uint8_t b[G1_RAW]; g1_raw(p, b); fs.bind(name, b, G1_RAW); }
void bind_fr(Transcript& fs, const char* name, const Fr& x) { uint8_t b[32]; fr_be(x, b); fs.bind(name, b, 32); }
Fr derive(Transcript& fs, const char* name) { uint8_t b[32]; fs.challenge(name, b); return fr_from_be(b); }

int os_random_fr(Fr* out, int count) {
    FILE* f = fopen("/dev/urandom", "rb");
    if (!f) return B200_ERR_ARG;
    for (int k = 0; k < count; ++k) {
        uint8_t b[64];
        if (fread(b, 1, 64, f) != 64) { fclose(f); return B200_ERR_ARG; }
        // 512 random bits reduced mod r: hi * 2^256 + lo  (bias 2^-258)
        const Fr lo = fr_from_be(b), hi = fr_from_be(b + 32);
        out[k] = hi * Fr::r2() + lo;      // r2() as a Montgomery element is 2^256 mod r
    }
    fclose(f);
    return B200_OK;
}

double now_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// kzg.Commit of several polynomials against one resident SRS: G1 MSMs software-pipelined like the Groth16
// prove core (groth16.cu) -- the counting sort of MSM k+1 runs on the aux stream under the bucket accumulation
// of MSM k.  wait[k] (optional): an event the sort of MSM k must wait for (its scalars' host copy).
int commit_many(b200_ctx* ctx, const b200_bases* srs, const Fr* const* coeffs, const size_t* counts, int m, G1XYZZ* out,
                const cudaEvent_t* wait = nullptr) {
    if (m < 1 || m > 4) return ctx->fail(B200_ERR_ARG, "commit_many: 1..4 polynomials");
    size_t sort_max = 0, acc_max = 0, ba_max = 0, res_off[5] = {0};
    for (int k = 0; k < m; ++k) {
        if (counts[k] > srs->n) return ctx->fail(B200_ERR_ARG, "polynomial longer than the SRS");
        const MsmSizes sz = msm_sizes(counts[k], 1, ctx->msm_mode(), srs->fold.c);
        sort_max = std::max(sort_max, sz.sort_bytes);
        acc_max = std::max(acc_max, sz.acc_bytes);
        ba_max = std::max(ba_max, sz.ba_bytes);
        res_off[k + 1] = res_off[k] + ((sz.result_bytes + 255) & ~(size_t)255);
    }
    const bool overlap = ctx->msm_overlap != 0 && m > 1;
    Arena ar;
    const size_t o_sort0 = ar.off; ar.take(sort_max);
    const size_t o_sort1 = overlap ? ar.off : o_sort0; if (overlap) ar.take(sort_max);
    const size_t o_acc = ar.off; ar.take(acc_max);
    const size_t o_ba = ar.off; ar.take(ba_max);
    B2_TRY(ctx_reserve(ctx, ar.off));
    if (ctx->h_stage_bytes < res_off[m]) {
        if (ctx->h_stage) cudaFreeHost(ctx->h_stage);
        ctx->h_stage = nullptr; ctx->h_stage_bytes = 0;
        B2_CUDA(ctx, cudaMallocHost(&ctx->h_stage, res_off[m]));
        ctx->h_stage_bytes = res_off[m];
    }
    char* base = (char*)ctx->scratch;
    char* sort_ws[2] = {base + o_sort0, base + o_sort1};
    cudaStream_t st = ctx->stream, sx = overlap ? ctx->aux_stream : st;
    cudaEvent_t ev[9];
    for (auto& e : ev) B2_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    struct EvGuard { cudaEvent_t* e; ~EvGuard() { for (int i = 0; i < 9; ++i) cudaEventDestroy(e[i]); } } guard{ev};
    cudaEvent_t *ev_sort = ev + 1, *ev_acc = ev + 5;
    void* jobs[4] = {nullptr, nullptr, nullptr, nullptr};
    struct JobGuard { void** j; ~JobGuard() { for (int k = 0; k < 4; ++k) if (j[k]) msm_g1_job_delete(j[k]); } } jguard{jobs};
    for (int k = 0; k < m; ++k)
        jobs[k] = msm_g1_job_new(ctx, srs->pts.as<G1Affine>(), coeffs[k], nullptr, counts[k], &srs->fold, sort_ws[overlap ? (k & 1) : 0],
                                 base + o_acc, (char*)ctx->h_stage + res_off[k], base + o_ba);
    B2_CUDA(ctx, cudaEventRecord(ev[0], st));
    if (overlap) B2_CUDA(ctx, cudaStreamWaitEvent(sx, ev[0], 0));
    auto run_sort = [&](int k) -> int {
        if (wait && wait[k]) B2_CUDA(ctx, cudaStreamWaitEvent(sx, wait[k], 0));
        if (k >= 2) B2_CUDA(ctx, cudaStreamWaitEvent(sx, ev_acc[k - 2], 0));       // its sort group is free again
        {   // every sort but the first runs beside the previous commitment's bucket stage (msm_aux_persistent, msm_host.cuh)
            struct Persist { b200_ctx* c; ~Persist() { c->aux_persist = 0; } } persist{ctx};
            ctx->aux_persist = (overlap && k > 0) ? 1 : 0;
            B2_TRY(msm_g1_job_sort(ctx, jobs[k], sx));
        }
        B2_CUDA(ctx, cudaEventRecord(ev_sort[k], sx));
        return B200_OK;
    };
    auto run_acc = [&](int k) -> int {
        B2_CUDA(ctx, cudaStreamWaitEvent(st, ev_sort[k], 0));
        B2_TRY(msm_g1_job_acc(ctx, jobs[k], st));
        B2_CUDA(ctx, cudaEventRecord(ev_acc[k], st));
        return B200_OK;
    };
    if (overlap) {
        B2_TRY(run_sort(0));
        for (int k = 0; k < m; ++k) {
            if (k + 1 < m) B2_TRY(run_sort(k + 1));
            B2_TRY(run_acc(k));
        }
    } else {
        for (int k = 0; k < m; ++k) { B2_TRY(run_sort(k)); B2_TRY(run_acc(k)); }
    }
    B2_CUDA(ctx, cudaStreamSynchronize(st));
    if (overlap) B2_CUDA(ctx, cudaStreamSynchronize(sx));
    for (int k = 0; k < m; ++k) {
        G1Jac r;
        msm_g1_job_finish(jobs[k], &r);
        out[k] = xyzz_from_jac(r);
    }
    return B200_OK;
}
int commit(b200_ctx* ctx, const b200_bases* srs, const Fr* d_coeffs, size_t m, G1XYZZ* out) {
    return commit_many(ctx, srs, &d_coeffs, &m, 1, out);
}
// commitBlindingFactor (prove.go:1159-1172): [b (X^n - 1)] = sum b_j (G1[n+j] - G1[j])
G1XYZZ commit_blinding(const b200_plonk_pk* pk, const Fr* b, int nb) {
    G1XYZZ hi = G1XYZZ::inf(), lo = G1XYZZ::inf();
    for (int j = 0; j < nb; ++j) {
        hi.add(scalar_mul(G1XYZZ::from_affine(pk->srs_hi[j]), b[j]));
        lo.add(scalar_mul(G1XYZZ::from_affine(pk->srs_lo[j]), b[j]));
    }
    G1Affine l = lo.to_affine();
    if (!l.is_inf()) l.y = l.y.neg();
    hi.add(G1XYZZ::from_affine(l));
    return hi;
}
void put_point(uint64_t* dst, const G1Affine& p) { memcpy(dst, &p, sizeof p); }
void put_fr(uint64_t* dst, const Fr& x) { memcpy(dst, x.l, 32); }

}  // namespace

namespace b200 {

// divideByXMinusOne (prove.go:1223-1250) on the big domain: r[i] *= 1/(X^n - 1) on the coset, then
// LagrangeCoset/BitReverse -> Canonical/Regular = FFTInverse(DIT, OnCoset)
int plonk_divide_xn(b200_ctx* ctx, NttDomain* d, unsigned log_small, Fr* vec) {
    const unsigned log_big = d->log_n;
    const unsigned lr = log_big - log_small;
    const uint32_t rho = 1u << lr;
    const uint64_t n_small = (uint64_t)1 << log_small;
    // evaluateXnMinusOneDomainBigCoset (prove.go:1253-1276): res[i] = (g * w^i)^n - 1, then inverted
    Fr res[8], f[8];
    const Fr step = d->gen.pow_u64(n_small);
    res[0] = d->coset.pow_u64(n_small);
    for (uint32_t i = 1; i < rho; ++i) res[i] = res[i - 1] * step;
    for (uint32_t i = 0; i < rho; ++i) {
        const Fr v = res[i] - Fr::one();
        if (v.is_zero()) return ctx->fail(B200_ERR_ARG, "X^n - 1 vanishes on the coset");
        res[i] = v.inv();
    }
    // r[i] *= res[bitrev(i) mod rho]  (prove.go:1240-1245): bitrev(i) mod rho = the top lr bits of i, reversed
    for (uint32_t k = 0; k < rho; ++k) {
        uint32_t rev = 0;
        for (unsigned b = 0; b < lr; ++b) rev |= ((k >> b) & 1u) << (lr - 1 - b);
        f[k] = res[rev];
    }
    B2_TRY(fr_scale_cyclic(ctx, vec, (size_t)1 << log_big, f, rho, log_big - lr));
    return ntt_gnark(ctx, d, vec, 1, B200_DIT, 1);
}

}  // namespace b200

static void plonk_pk_delete(b200_ctx* ctx, b200_plonk_pk* pk) {
    if (!pk) return;
    if (pk->srs) b200_bases_free(ctx, pk->srs);
    if (pk->srs_lag) b200_bases_free(ctx, pk->srs_lag);
    if (pk->d0) --pk->d0->refs;
    if (pk->d1) --pk->d1->refs;
    delete pk;
}

static int b200_bases_upload_raw_words(b200_ctx* ctx, int curve, int group, const uint64_t* raw, size_t n, b200_bases** out) {
    return b200_bases_upload_raw(ctx, curve, group, reinterpret_cast<const uint8_t*>(raw), n, out);
}

namespace b200 {
// the nine blinding coefficients of getRandomPolynomial (prove.go:295-301) from the OS entropy source, for the group
// prover (group.cu): a proof sharded over several ranks needs the same ones on every rank
int plonk_draw_blinding(uint64_t out[36]) {
    Fr b[9];
    const int st = os_random_fr(b, 9);
    if (st == B200_OK) memcpy(out, b, sizeof b);
    return st;
}
}  // namespace b200

extern "C" {

int b200_plonk_pk_upload(b200_ctx* ctx, int curve, const b200_plonk_pk_desc* d, b200_plonk_pk** out) try {
    if (!ctx) return B200_ERR_ARG;
    if (curve != b200_build_curve()) return ctx->fail(B200_ERR_UNSUPPORTED, "curve not supported by this build (one library per curve)");
    if (cudaSetDevice(ctx->device) != cudaSuccess) return ctx->fail(B200_ERR_CUDA, "cudaSetDevice");
    if (!d || !out || !d->generator || !d->generator_big || !d->coset_gen || !d->ql || !d->qr || !d->qm || !d->qo || !d->qk || !d->s1 ||
        !d->s2 || !d->s3 || !d->perm || !d->kzg_g1 || !d->kzg_lagrange_g1 || !d->vk_commitments)
        return ctx->fail(B200_ERR_ARG, "null argument");
    if (d->nb_bsb22 > B200_PLONK_MAX_BSB22) return ctx->fail(B200_ERR_UNSUPPORTED, "more BSB22 commitments than B200_PLONK_MAX_BSB22");
    if (d->nb_bsb22 && (!d->qcp || !d->vk_qcp || !d->commitment_index)) return ctx->fail(B200_ERR_ARG, "null BSB22 argument");
    if (d->log_n < 1 || d->log_big < d->log_n + 2 || d->log_big > d->log_n + 3 || d->log_big > 27)
        return ctx->fail(B200_ERR_ARG, "domains: need 2 <= n, big = 4n or 8n, big <= 2^27");
    const size_t n = (size_t)1 << d->log_n;
    if (d->nb_public > n) return ctx->fail(B200_ERR_ARG, "more public inputs than rows");
    b200_plonk_pk* pk = new b200_plonk_pk();
    pk->log_n = d->log_n; pk->log_big = d->log_big; pk->n = n; pk->big = (size_t)1 << d->log_big;
    pk->rho = 1u << (d->log_big - d->log_n);
    pk->nb_public = d->nb_public;
    pk->shard_count = d->shard_count > 1 ? (uint32_t)d->shard_count : 1u;
    pk->shard_index = d->shard_count > 1 ? (uint32_t)d->shard_index : 0u;
    if (pk->shard_index >= pk->shard_count || d->shard_index < 0) { delete pk; return ctx->fail(B200_ERR_ARG, "shard_index out of range"); }
    pk->srs_lo_idx = (n + 3) * pk->shard_index / pk->shard_count;
    pk->srs_hi_idx = (n + 3) * (pk->shard_index + 1) / pk->shard_count;
    pk->lag_lo = n * pk->shard_index / pk->shard_count;
    pk->lag_hi = n * (pk->shard_index + 1) / pk->shard_count;
    pk->nb_bsb = (uint32_t)d->nb_bsb22;
    for (uint32_t k = 0; k < pk->nb_bsb; ++k) {
        if (!d->qcp[k]) { delete pk; return ctx->fail(B200_ERR_ARG, "null Qcp"); }
        pk->commit_row[k] = d->nb_public + d->commitment_index[k];
        if (pk->commit_row[k] >= n) { delete pk; return ctx->fail(B200_ERR_ARG, "commitment index outside the domain"); }
    }
    if (pk->nb_bsb && !d->raw_file) memcpy(pk->vk_qcp, d->vk_qcp, pk->nb_bsb * sizeof(G1Affine));
    Fr gen, gen_big, cg;
    memcpy(gen.l, d->generator, 32); memcpy(gen_big.l, d->generator_big, 32); memcpy(cg.l, d->coset_gen, 32);
    int st = ntt_domain_get(ctx, d->log_n, gen, cg, &pk->d0);
    if (st == B200_OK) ++pk->d0->refs;
    if (st == B200_OK) st = ntt_domain_get(ctx, d->log_big, gen_big, cg, &pk->d1);
    if (st == B200_OK) ++pk->d1->refs;
    if (st != B200_OK) { plonk_pk_delete(ctx, pk); return st; }
    cudaStream_t s = ctx->stream;
    auto fail_cuda = [&](cudaError_t e, const char* what) {
        plonk_pk_delete(ctx, pk);
        return ctx->fail(e == cudaErrorMemoryAllocation ? B200_ERR_NOMEM : B200_ERR_CUDA, what, cudaGetErrorString(e));
    };
    cudaError_t e = pk->trace.alloc(8 * n * sizeof(Fr));
    if (e == cudaSuccess) e = pk->qk_lag.alloc(n * sizeof(Fr));
    if (e == cudaSuccess) e = pk->perm.alloc(3 * n * sizeof(int64_t));
    if (e == cudaSuccess) e = pk->coset_ev.alloc((size_t)pk->rho * 8 * n * sizeof(Fr));
    if (e == cudaSuccess) e = pk->coset_pre.alloc((size_t)pk->rho * n * sizeof(Fr));
    // workspace: 5 vectors of n, 9 of n + 4 (canonical blinded L R O Z; 5 coset evaluations reused by the openings), id, cres
    const size_t N4 = n + 4;
    if (e == cudaSuccess) e = pk->work.alloc((6 * n + 9 * N4 + pk->big) * sizeof(Fr));
    if (e == cudaSuccess && pk->nb_bsb) e = pk->qcp.alloc((size_t)pk->nb_bsb * n * sizeof(Fr));
    if (e == cudaSuccess && pk->nb_bsb) e = pk->qcp_ev.alloc((size_t)pk->rho * pk->nb_bsb * n * sizeof(Fr));
    if (e == cudaSuccess && pk->nb_bsb) e = pk->bsb_work.alloc((size_t)pk->nb_bsb * 3 * n * sizeof(Fr));
    if (e != cudaSuccess) return fail_cuda(e, "PlonK key allocation");
    for (uint32_t k = 0; k < pk->nb_bsb && e == cudaSuccess; ++k)
        e = cudaMemcpyAsync(pk->qcp.as<Fr>() + (size_t)k * n, d->qcp[k], n * sizeof(Fr), cudaMemcpyHostToDevice, s);
    const uint64_t* src[8] = {d->ql, d->qr, d->qm, d->qo, d->qk, d->s1, d->s2, d->s3};
    for (int k = 0; k < 8 && e == cudaSuccess; ++k)
        e = cudaMemcpyAsync(pk->trace.as<Fr>() + (size_t)k * n, src[k], n * sizeof(Fr), cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(pk->perm.p, d->perm, 3 * n * sizeof(int64_t), cudaMemcpyHostToDevice, s);
    if (e != cudaSuccess) return fail_cuda(e, "PlonK key upload");
    const bool raw = d->raw_file != 0;
    if (raw) {
        // The arrays are sections of a ProvingKey.WriteRawTo file (plonk/bn254/marshal.go:101-180): scalars big-endian
        // canonical, the permutation big-endian int64, points raw-encoded.  Converted on the device, in place.
        DevBuf bad;
        if (bad.alloc(4) != cudaSuccess || cudaMemsetAsync(bad.p, 0, 4, s) != cudaSuccess) return fail_cuda(cudaErrorMemoryAllocation, "PlonK key upload");
        st = fr_from_be_inplace(ctx, pk->trace.as<Fr>(), 8 * n, bad.as<uint32_t>());
        if (st == B200_OK && pk->nb_bsb) st = fr_from_be_inplace(ctx, pk->qcp.as<Fr>(), (size_t)pk->nb_bsb * n, bad.as<uint32_t>());
        if (st == B200_OK) st = i64_from_be_inplace(ctx, pk->perm.as<int64_t>(), 3 * n);
        // the few points the host keeps: 3 + 3 SRS points (commitBlindingFactor) and the 8 commitments of the key
        DevBuf pts;
        const size_t np = 3 + 3 + 8 + pk->nb_bsb;
        if (st == B200_OK && pts.alloc(np * sizeof(G1Affine)) != cudaSuccess) st = ctx->fail(B200_ERR_NOMEM, "PlonK key upload");
        if (st == B200_OK) {
            char* q = (char*)pts.p;
            cudaMemcpyAsync(q, d->kzg_g1, 3 * sizeof(G1Affine), cudaMemcpyHostToDevice, s);
            cudaMemcpyAsync(q + 3 * sizeof(G1Affine), d->kzg_g1 + AFF_WORDS * n, 3 * sizeof(G1Affine), cudaMemcpyHostToDevice, s);
            cudaMemcpyAsync(q + 6 * sizeof(G1Affine), d->vk_commitments, 8 * sizeof(G1Affine), cudaMemcpyHostToDevice, s);
            if (pk->nb_bsb) cudaMemcpyAsync(q + 14 * sizeof(G1Affine), d->vk_qcp, pk->nb_bsb * sizeof(G1Affine), cudaMemcpyHostToDevice, s);
            st = points_from_raw_inplace(ctx, pts.p, np, 1, bad.as<uint32_t>());
            std::vector<G1Affine> hp(np);
            uint32_t flag = 0;
            if (st == B200_OK && (cudaMemcpyAsync(hp.data(), pts.p, np * sizeof(G1Affine), cudaMemcpyDeviceToHost, s) != cudaSuccess ||
                                  cudaMemcpyAsync(&flag, bad.p, 4, cudaMemcpyDeviceToHost, s) != cudaSuccess || cudaStreamSynchronize(s) != cudaSuccess))
                st = ctx->fail(B200_ERR_CUDA, "PlonK key upload");
            if (st == B200_OK && flag) st = ctx->fail(B200_ERR_ARG, "raw_file: an element of the key file is not canonical (scalar >= r, coordinate >= p or compressed-form flag)");
            if (st == B200_OK) {
                memcpy(pk->srs_lo, hp.data(), 3 * sizeof(G1Affine));
                memcpy(pk->srs_hi, hp.data() + 3, 3 * sizeof(G1Affine));
                memcpy(pk->vk_pts, hp.data() + 6, 8 * sizeof(G1Affine));
                for (uint32_t k = 0; k < pk->nb_bsb; ++k) pk->vk_qcp[k] = hp[14 + k];
            }
        }
        if (st != B200_OK) { plonk_pk_delete(ctx, pk); return st; }
    } else {
    memcpy(pk->srs_lo, d->kzg_g1, 3 * sizeof(G1Affine));
    memcpy(pk->srs_hi, d->kzg_g1 + AFF_WORDS * n, 3 * sizeof(G1Affine));
    memcpy(pk->vk_pts, d->vk_commitments, 8 * sizeof(G1Affine));
    }
    // coset scaling tables: shift_k = g * w_big^k  (shifters, prove.go:903-907)
    Fr shift = cg;
    for (uint32_t k = 0; k < pk->rho && st == B200_OK; ++k) {
        st = gen_powers(ctx, pk->coset_pre.as<Fr>() + (size_t)k * n, shift, Fr::one(), n, (int)d->log_n);
        shift = shift * gen_big;
    }
    // coset evaluations of the key's polynomials and of LOne (canonical form of L_1 = (1/n) sum X^j)
    Fr* tmp = pk->work.as<Fr>();          // n elements: bit-reversed canonical coefficients
    const int which[7] = {TR_QL, TR_QR, TR_QM, TR_QO, TR_S1, TR_S2, TR_S3};
    for (int p = 0; p < 8 && st == B200_OK; ++p) {
        if (p < 7) st = fr_bitrev_copy(ctx, pk->tr(which[p]), tmp, d->log_n);
        else st = gen_powers(ctx, tmp, Fr::one(), pk->d0->ninv, n, 0);
        for (uint32_t k = 0; k < pk->rho && st == B200_OK; ++k) {
            Fr* dst = const_cast<Fr*>(pk->cev(k, p));
            if (cudaMemcpyAsync(dst, tmp, n * sizeof(Fr), cudaMemcpyDeviceToDevice, s) != cudaSuccess) { st = ctx->fail(B200_ERR_CUDA, "copy"); break; }
            st = ntt_run(ctx, pk->d0, dst, 0, 1, pk->coset_pre.as<Fr>() + (size_t)k * n, nullptr, nullptr);
        }
    }
    for (uint32_t p = 0; p < pk->nb_bsb && st == B200_OK; ++p) {
        st = fr_bitrev_copy(ctx, pk->qcp_c(p), tmp, d->log_n);
        for (uint32_t k = 0; k < pk->rho && st == B200_OK; ++k) {
            Fr* dst = const_cast<Fr*>(pk->qcp_e(k, p));
            if (cudaMemcpyAsync(dst, tmp, n * sizeof(Fr), cudaMemcpyDeviceToDevice, s) != cudaSuccess) { st = ctx->fail(B200_ERR_CUDA, "copy"); break; }
            st = ntt_run(ctx, pk->d0, dst, 0, 1, pk->coset_pre.as<Fr>() + (size_t)k * n, nullptr, nullptr);
        }
    }
    // Qk back to Lagrange form (Qk.Clone().ToLagrange().ToRegular(), prove.go:398)
    if (st == B200_OK) st = fr_bitrev_copy(ctx, pk->tr(TR_QK), pk->qk_lag.as<Fr>(), d->log_n);
    if (st == B200_OK) st = ntt_run(ctx, pk->d0, pk->qk_lag.as<Fr>(), 0, 1, nullptr, nullptr, nullptr);
    if (st == B200_OK && cudaStreamSynchronize(s) != cudaSuccess) st = ctx->fail(B200_ERR_CUDA, "PlonK key precomputation");
    // each rank keeps its point range of both SRS (the whole of them on one GPU)
    auto up = raw ? b200_bases_upload_raw_words : b200_bases_upload;
    if (st == B200_OK) st = up(ctx, curve, 1, d->kzg_g1 + AFF_WORDS * pk->srs_lo_idx, pk->srs_hi_idx - pk->srs_lo_idx, &pk->srs);
    if (st == B200_OK) st = up(ctx, curve, 1, d->kzg_lagrange_g1 + AFF_WORDS * pk->lag_lo, pk->lag_hi - pk->lag_lo, &pk->srs_lag);
    if (st != B200_OK) { plonk_pk_delete(ctx, pk); return st; }
    *out = pk;
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)

int b200_plonk_pk_free(b200_ctx* ctx, b200_plonk_pk* pk) try {
    if (!ctx) return B200_ERR_ARG;
    if (cudaSetDevice(ctx->device) != cudaSuccess) return ctx->fail(B200_ERR_CUDA, "cudaSetDevice");
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    plonk_pk_delete(ctx, pk);
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)

}  // extern "C"

// ---- the prover as a session of phases ------------------------------------------------------------------
// One proof = 7 phases separated by the points where commitments must be complete (Fiat-Shamir) or where the
// quotient's coset blocks must all be present.  On one GPU b200_plonk_prove_bsb22 runs them back to back; with
// the key sharded over R ranks (b200_plonk_pk_desc.shard_*), every rank runs the same phases, its MSMs cover its
// point range of the SRS, its quotient loop covers the cosets k = rank (mod R), and the caller sums the partial
// commitments (b200_g1_jac_add) and all-gathers the coset blocks between phases (PlonK P1, P4).
//   phase  consumes (totals)        produces (partials)
//   0      -                        [L] [R] [O]
//   1      [L] [R] [O]              [Z]
//   2      [Z]                      -            (coset blocks of the numerator written; exchange them now)
//   3      -                        [H1] [H2] [H3]
//   4      [H1] [H2] [H3]           [Z opening quotient] [linearised polynomial]
//   5      those two                [batch opening quotient]
//   6      that one                 -            (proof complete)
struct b200_plonk_session {
    const b200_plonk_pk* pk = nullptr;
    uint32_t nb = 0;
    int next_phase = 0;
    G1Affine bsb_com[B200_PLONK_MAX_BSB22];
    Fr bsb_val[B200_PLONK_MAX_BSB22];
    Fr *pi2_br[B200_PLONK_MAX_BSB22], *pi2_cn[B200_PLONK_MAX_BSB22], *pi2_ev[B200_PLONK_MAX_BSB22];
    Fr bp[9];
    std::vector<Fr> pub;
    Fr *v[4], *qk_br, *idv, *cn[4], *tmp[5], *cres;
    Fr *quot, *fh, *lin, *folded;
    Transcript fs{"gamma", "beta", "alpha", "zeta"};
    Fr gamma, beta, alpha, zeta, gk, zu, blz, brz, boz, s1z, s2z, qcpz[B200_PLONK_MAX_BSB22], claimed[7];
    G1Affine lro[3], z_com, h_com[3], zs_h, lin_digest, fh_digest, batched_h;
    G1XYZZ hx[3];
    b200_plonk_timings T;
    double t_start = 0, t0 = 0;
    cudaEvent_t ev_cp[4] = {nullptr, nullptr, nullptr, nullptr}, ev_t[2] = {nullptr, nullptr};
    // commitBlindingFactor of L, R, O, Z (prove.go:1159-1172): 2 x (2 + 2 + 2 + 3) host scalar multiplications that depend on
    // the blinding scalars only -- started as host tasks when the session begins, collected where the commitments are formed
    std::future<G1XYZZ> blind_com[4];
    std::future<G1Affine> fh_digest_task;       // the folded-H digest (2 scalar multiplications), phase 4 -> phase 5
    ~b200_plonk_session() {
        for (auto& f : blind_com) if (f.valid()) f.wait();
        if (fh_digest_task.valid()) fh_digest_task.wait();
        for (auto e : ev_cp) if (e) cudaEventDestroy(e);
        for (auto e : ev_t) if (e) cudaEventDestroy(e);
        if (pk) pk->in_flight.store(false);     // the key's workspace is free for the next proof
    }
    void lap(float& slot) { const double t = now_ms(); slot += (float)(t - t0); t0 = t; }
    const Fr* bl() const { return bp; }
    const Fr* br() const { return bp + 2; }
    const Fr* bo() const { return bp + 4; }
    const Fr* bz() const { return bp + 6; }
};

namespace {

// this shard's part of kzg.Commit for several polynomials against one SRS (lo..hi = the shard's point range)
int commit_shard(b200_ctx* ctx, const b200_bases* srs, size_t lo, size_t hi, const Fr* const* coeffs, const size_t* counts, int m,
                 G1XYZZ* out, const cudaEvent_t* wait = nullptr) {
    const Fr* sc[4];
    size_t cnt[4];
    cudaEvent_t wv[4];
    int idx[4], live = 0;
    for (int k = 0; k < m; ++k) {
        out[k] = G1XYZZ::inf();
        const size_t top = counts[k] < hi ? counts[k] : hi;
        if (top <= lo) continue;
        sc[live] = coeffs[k] + lo; cnt[live] = top - lo; wv[live] = wait ? wait[k] : nullptr; idx[live] = k;
        ++live;
    }
    if (!live) {
        if (wait) for (int k = 0; k < m; ++k) if (wait[k]) B2_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, wait[k], 0));
        return B200_OK;
    }
    G1XYZZ res[4];
    B2_TRY(commit_many(ctx, srs, sc, cnt, live, res, wait ? wv : nullptr));
    for (int k = 0; k < live; ++k) out[idx[k]] = res[k];
    return B200_OK;
}
constexpr int JAC_WORDS = (int)(sizeof(G1Jac) / 8);   // 12 u64 on BN254, 18 on BLS12-377
void put_jac(uint64_t* dst, const G1XYZZ& p) { const G1Jac j = p.to_jac(); memcpy(dst, &j, sizeof j); }
G1XYZZ get_jac(const uint64_t* src) { G1Jac j; memcpy(&j, src, sizeof j); return xyzz_from_jac(j); }

int phase0(b200_ctx* ctx, b200_plonk_session& S, const uint64_t* l, const uint64_t* r, const uint64_t* o, const b200_plonk_bsb22* bsb,
           uint64_t* partials) {
    const b200_plonk_pk* pk = S.pk;
    const size_t n = pk->n;
    cudaStream_t s = ctx->stream;
    // ---- solution to the device, [L] [R] [O]  (solveConstraints / commitToLRO, prove.go:365-452) ----------
    // the copies run on the copy stream; the sort of [L] starts when L has landed, R and O travel under it
    const uint64_t* src[3] = {l, r, o};
    for (auto& e : S.ev_cp) B2_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (auto& e : S.ev_t) B2_CUDA(ctx, cudaEventCreate(&e));
    B2_CUDA(ctx, cudaEventRecord(S.ev_cp[3], s));
    B2_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, S.ev_cp[3], 0));
    B2_CUDA(ctx, cudaEventRecord(S.ev_t[0], ctx->copy_stream));
    for (int k = 0; k < 3; ++k) {
        B2_CUDA(ctx, cudaMemcpyAsync(S.v[k], src[k], n * sizeof(Fr), cudaMemcpyDefault, ctx->copy_stream));
        B2_CUDA(ctx, cudaEventRecord(S.ev_cp[k], ctx->copy_stream));
    }
    for (uint32_t k = 0; k < S.nb; ++k)
        B2_CUDA(ctx, cudaMemcpyAsync(S.pi2_br[k], bsb->pi2[k], n * sizeof(Fr), cudaMemcpyDefault, ctx->copy_stream));
    B2_CUDA(ctx, cudaEventRecord(S.ev_t[1], ctx->copy_stream));
    const Fr* sc[3] = {S.v[0], S.v[1], S.v[2]};
    const size_t cnt[3] = {n, n, n};
    G1XYZZ c[3];
    if (ctx->msm_overlap) B2_TRY(commit_shard(ctx, pk->srs_lag, pk->lag_lo, pk->lag_hi, sc, cnt, 3, c, S.ev_cp));
    else {
        for (int k = 0; k < 3; ++k) B2_CUDA(ctx, cudaStreamWaitEvent(s, S.ev_cp[k], 0));
        B2_TRY(commit_shard(ctx, pk->srs_lag, pk->lag_lo, pk->lag_hi, sc, cnt, 3, c));
    }
    for (int k = 0; k < 3; ++k) B2_CUDA(ctx, cudaStreamWaitEvent(s, S.ev_cp[k], 0));
    B2_CUDA(ctx, cudaStreamWaitEvent(s, S.ev_t[1], 0));
    for (int k = 0; k < 3; ++k) put_jac(partials + JAC_WORDS * k, c[k]);
    S.lap(S.T.commit_lro_ms);
    B2_CUDA(ctx, cudaEventSynchronize(S.ev_t[1]));
    cudaEventElapsedTime(&S.T.h2d_ms, S.ev_t[0], S.ev_t[1]);   // on the copy stream, under the [L] [R] [O] commitments: not a summand of total_ms
    return B200_OK;
}

int phase1(b200_ctx* ctx, b200_plonk_session& S, const uint64_t* totals, uint64_t* partials) {
    const b200_plonk_pk* pk = S.pk;
    const size_t n = pk->n;
    const Fr* bps[3] = {S.bl(), S.br(), S.bo()};
    for (int k = 0; k < 3; ++k) {
        G1XYZZ c = get_jac(totals + JAC_WORDS * k);
        c.add(S.blind_com[k].valid() ? S.blind_com[k].get() : commit_blinding(pk, bps[k], 2));
        S.lro[k] = c.to_affine();
    }
    // ---- gamma, beta (deriveGammaAndBeta, prove.go:455-485; bindPublicData, verify.go:296-340) -----------
    for (int k = 0; k < 8; ++k) bind_point(S.fs, "gamma", pk->vk_pts[k]);
    for (uint32_t k = 0; k < S.nb; ++k) bind_point(S.fs, "gamma", pk->vk_qcp[k]);
    for (const Fr& x : S.pub) bind_fr(S.fs, "gamma", x);
    for (int k = 0; k < 3; ++k) bind_point(S.fs, "gamma", S.lro[k]);
    S.gamma = derive(S.fs, "gamma");
    S.beta = derive(S.fs, "beta");
    S.lap(S.T.commit_lro_ms);
    // ---- Z and [Z] (buildRatioCopyConstraint, prove.go:597-631) ------------------------------------------
    B2_TRY(plonk_build_ratio(ctx, pk->d0, S.v[0], S.v[1], S.v[2], pk->perm.as<int64_t>(), S.beta, S.gamma, S.v[3], S.tmp[0], S.tmp[1]));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    S.lap(S.T.ratio_ms);
    const Fr* sc[1] = {S.v[3]};
    const size_t cnt[1] = {n};
    G1XYZZ c;
    B2_TRY(commit_shard(ctx, pk->srs_lag, pk->lag_lo, pk->lag_hi, sc, cnt, 1, &c));
    put_jac(partials, c);
    S.lap(S.T.commit_z_ms);
    return B200_OK;
}

int phase2(b200_ctx* ctx, b200_plonk_session& S, const uint64_t* totals) {
    const b200_plonk_pk* pk = S.pk;
    const size_t n = pk->n;
    const unsigned log_n = pk->log_n;
    cudaStream_t s = ctx->stream;
    NttDomain *d0 = pk->d0, *d1 = pk->d1;
    {
        G1XYZZ c = get_jac(totals);
        c.add(S.blind_com[3].valid() ? S.blind_com[3].get() : commit_blinding(pk, S.bz(), 3));
        S.z_com = c.to_affine();
    }
    for (uint32_t k = 0; k < S.nb; ++k) bind_point(S.fs, "alpha", S.bsb_com[k]);     // deriveAlpha, prove.go:504-512
    bind_point(S.fs, "alpha", S.z_com);
    S.alpha = derive(S.fs, "alpha");
    S.lap(S.T.commit_z_ms);
    // ---- quotient (evaluateConstraints / computeNumerator, prove.go:521-595,837-1079) -----------------------
    // canonical forms: FFTInverse(DIF) leaves the coefficients bit-reversed, which is what the coset DIT wants
    B2_CUDA(ctx, cudaMemcpyAsync(S.qk_br, pk->qk_lag.p, n * sizeof(Fr), cudaMemcpyDeviceToDevice, s));
    if (pk->nb_public) B2_CUDA(ctx, cudaMemcpyAsync(S.qk_br, S.pub.data(), pk->nb_public * sizeof(Fr), cudaMemcpyHostToDevice, s));   // completeQk
    for (uint32_t k = 0; k < S.nb; ++k)      // qkCoeffs[nbPublic + CommitmentIndex] = commitmentVal, prove.go:415-417
        B2_CUDA(ctx, cudaMemcpyAsync(S.qk_br + pk->commit_row[k], S.bsb_val[k].l, sizeof(Fr), cudaMemcpyHostToDevice, s));
    for (int k = 0; k < 4; ++k) B2_TRY(ntt_gnark(ctx, d0, S.v[k], 1, B200_DIF, 0));
    B2_TRY(ntt_gnark(ctx, d0, S.qk_br, 1, B200_DIF, 0));
    for (int k = 0; k < 4; ++k) B2_TRY(fr_bitrev_copy(ctx, S.v[k], S.cn[k], log_n));   // regular layout, for the openings
    for (uint32_t k = 0; k < S.nb; ++k) {                                              // cCommitments[i].ToCanonical, prove.go:339
        B2_TRY(ntt_gnark(ctx, d0, S.pi2_br[k], 1, B200_DIF, 0));
        B2_TRY(fr_bitrev_copy(ctx, S.pi2_br[k], S.pi2_cn[k], log_n));
    }
    PlonkRowArgs a;
    memset(&a, 0, sizeof a);
    a.nb_bsb = S.nb;
    a.beta = S.beta; a.gamma = S.gamma; a.alpha = S.alpha;
    a.cs = d1->coset;
    a.css = a.cs * a.cs;
    a.tw = d0->tw_fwd.as<Fr>();
    a.n = (uint32_t)n; a.log_big = pk->log_big; a.rho = pk->rho;
    a.cres = S.cres;
    Fr coset = Fr::one();
    const Fr* srcs[5] = {S.v[0], S.v[1], S.v[2], S.v[3], S.qk_br};
    const Fr *bl = S.bl(), *br = S.br(), *bo = S.bo(), *bz = S.bz();
    for (uint32_t k = 0; k < pk->rho; ++k) {
        coset = coset * (k == 0 ? d1->coset : d1->gen);
        if (k % pk->shard_count != pk->shard_index) continue;            // another rank's coset
        // bl <- bl * ((s w^k)^n - 1) scaled by powers of the shift (prove.go:985-993)
        const Fr xn1 = coset.pow_u64(n) - Fr::one();
        Fr acc = xn1;
        for (int j = 0; j < 3; ++j) {
            if (j < 2) { a.bl[j] = bl[j] * acc; a.br[j] = br[j] * acc; a.bo[j] = bo[j] * acc; }
            a.bz[j] = bz[j] * acc;
            acc = acc * coset;
        }
        const Fr* pre = pk->coset_pre.as<Fr>() + (size_t)k * n;
        for (int p = 0; p < 5; ++p) {
            B2_CUDA(ctx, cudaMemcpyAsync(S.tmp[p], srcs[p], n * sizeof(Fr), cudaMemcpyDeviceToDevice, s));
            B2_TRY(ntt_run(ctx, d0, S.tmp[p], 0, 1, pre, nullptr, nullptr));
        }
        for (uint32_t p = 0; p < S.nb; ++p) {
            B2_CUDA(ctx, cudaMemcpyAsync(S.pi2_ev[p], S.pi2_br[p], n * sizeof(Fr), cudaMemcpyDeviceToDevice, s));
            B2_TRY(ntt_run(ctx, d0, S.pi2_ev[p], 0, 1, pre, nullptr, nullptr));
            a.qci[2 * p] = pk->qcp_e(k, p);
            a.qci[2 * p + 1] = S.pi2_ev[p];
        }
        B2_TRY(fr_scaled_omega(ctx, d0, S.beta * coset, S.idv));
        a.x[PID_L] = S.tmp[0]; a.x[PID_R] = S.tmp[1]; a.x[PID_O] = S.tmp[2]; a.x[PID_Z] = S.tmp[3]; a.x[PID_ZS] = nullptr;
        a.x[PID_QK] = S.tmp[4];
        a.x[PID_QL] = pk->cev(k, CE_QL); a.x[PID_QR] = pk->cev(k, CE_QR); a.x[PID_QM] = pk->cev(k, CE_QM); a.x[PID_QO] = pk->cev(k, CE_QO);
        a.x[PID_S1] = pk->cev(k, CE_S1); a.x[PID_S2] = pk->cev(k, CE_S2); a.x[PID_S3] = pk->cev(k, CE_S3);
        a.x[PID_LONE] = pk->cev(k, CE_LONE);
        a.x[PID_ID] = S.idv;
        a.coset_index = k;
        B2_TRY(plonk_quotient_rows(ctx, a));
    }
    B2_CUDA(ctx, cudaStreamSynchronize(s));
    S.lap(S.T.quotient_ms);
    return B200_OK;
}

int phase3(b200_ctx* ctx, b200_plonk_session& S, uint64_t* partials) {
    const b200_plonk_pk* pk = S.pk;
    const size_t n = pk->n;
    // divideByXMinusOne, prove.go:1223-1250 (every rank holds the whole numerator after the exchange)
    B2_TRY(plonk_divide_xn(ctx, pk->d1, pk->log_n, S.cres));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    S.lap(S.T.quotient_ms);
    const Fr* sc[3] = {S.cres, S.cres + (n + 2), S.cres + 2 * (n + 2)};          // h1 h2 h3, prove.go:652-666
    const size_t cnt[3] = {n + 2, n + 2, n + 2};
    G1XYZZ c[3];
    B2_TRY(commit_shard(ctx, pk->srs, pk->srs_lo_idx, pk->srs_hi_idx, sc, cnt, 3, c));   // commitToQuotient, prove.go:1198-1218
    for (int k = 0; k < 3; ++k) put_jac(partials + JAC_WORDS * k, c[k]);
    S.lap(S.T.commit_h_ms);
    return B200_OK;
}

int phase4(b200_ctx* ctx, b200_plonk_session& S, const uint64_t* totals, uint64_t* partials) {
    const b200_plonk_pk* pk = S.pk;
    const size_t n = pk->n;
    NttDomain* d0 = pk->d0;
    for (int k = 0; k < 3; ++k) {
        S.hx[k] = get_jac(totals + JAC_WORDS * k);
        S.h_com[k] = S.hx[k].to_affine();
        bind_point(S.fs, "zeta", S.h_com[k]);
    }
    S.zeta = derive(S.fs, "zeta");
    const Fr &zeta = S.zeta, &beta = S.beta, &gamma = S.gamma, &alpha = S.alpha;
    S.lap(S.T.commit_h_ms);
    // ---- open Z at w zeta (openZ, prove.go:634-650; kzg.Open) ------------------------------------------------
    B2_TRY(fr_blind(ctx, S.cn[0], n, S.bl(), 2));
    B2_TRY(fr_blind(ctx, S.cn[1], n, S.br(), 2));
    B2_TRY(fr_blind(ctx, S.cn[2], n, S.bo(), 2));
    B2_TRY(fr_blind(ctx, S.cn[3], n, S.bz(), 3));
    const Fr zeta_shifted = zeta * d0->gen;
    B2_TRY(fr_poly_eval(ctx, S.cn[3], n + 3, zeta_shifted, &S.zu));
    S.quot = S.tmp[0];
    B2_TRY(fr_poly_div_linear(ctx, S.cn[3], n + 3, zeta_shifted, S.quot));
    S.lap(S.T.open_z_ms);
    // ---- foldH (prove.go:670-705) ---------------------------------------------------------------------
    const Fr zn2 = zeta.pow_u64(n + 2);
    S.fh = S.tmp[1];
    B2_TRY(plonk_fold_h(ctx, S.cres, S.cres + (n + 2), S.cres + 2 * (n + 2), zn2, n + 2, S.fh));
    S.fh_digest_task = std::async(std::launch::async, [h0 = S.hx[0], h1 = S.hx[1], h2 = S.hx[2], zn2]() {   // needed in phase 5
        G1XYZZ fhd = scalar_mul(h2, zn2);
        fhd.add(h1);
        fhd = scalar_mul(fhd, zn2);
        fhd.add(h0);
        return fhd.to_affine();
    });
    // ---- linearised polynomial (computeLinearizedPolynomial, prove.go:708-777,1289-1389) --------------------
    // evaluateBlinded == the blinded polynomial at zeta
    B2_TRY(fr_poly_eval(ctx, S.cn[0], n + 2, zeta, &S.blz));
    B2_TRY(fr_poly_eval(ctx, S.cn[1], n + 2, zeta, &S.brz));
    B2_TRY(fr_poly_eval(ctx, S.cn[2], n + 2, zeta, &S.boz));
    B2_TRY(fr_poly_eval(ctx, pk->tr(TR_S1), n, zeta, &S.s1z));
    B2_TRY(fr_poly_eval(ctx, pk->tr(TR_S2), n, zeta, &S.s2z));
    for (uint32_t k = 0; k < S.nb; ++k) B2_TRY(fr_poly_eval(ctx, pk->qcp_c(k), n, zeta, &S.qcpz[k]));     // prove.go:724-729
    S.lin = S.tmp[2];
    Fr lin_s1, lin_s2;      // scalars of [S3] and [Z] in the linearised polynomial's digest
    {
        const Fr &blz = S.blz, &brz = S.brz, &boz = S.boz;
        PlonkLinArgs a;
        memset(&a, 0, sizeof a);
        a.bz = S.cn[3]; a.s3 = pk->tr(TR_S3); a.ql = pk->tr(TR_QL); a.qr = pk->tr(TR_QR); a.qm = pk->tr(TR_QM); a.qo = pk->tr(TR_QO);
        a.qk = pk->tr(TR_QK);
        a.nb_bsb = S.nb;
        for (uint32_t k = 0; k < S.nb; ++k) { a.pi2[k] = S.pi2_cn[k]; a.qcp_zeta[k] = S.qcpz[k]; }
        a.rl = brz * blz;
        a.c1 = (S.s1z * beta + blz + gamma) * (S.s2z * beta + brz + gamma) * S.zu * beta;
        const Fr u = d0->coset, uz = zeta * u, uuz = uz * u;
        a.c2 = ((beta * zeta + blz + gamma) * (beta * uz + brz + gamma) * (beta * uuz + boz + gamma)).neg();
        const Fr zn1 = zeta.pow_u64(n) - Fr::one();
        a.lag = zn1 * (zeta - Fr::one()).inv() * alpha * alpha * d0->ninv;
        a.alpha = alpha; a.l = blz; a.r = brz; a.o = boz;
        a.n = (uint32_t)n; a.len = (uint32_t)(n + 3);
        a.out = S.lin;
        B2_TRY(plonk_linearize(ctx, a));
        lin_s1 = a.c1 * alpha;
        lin_s2 = a.c2 * alpha + a.lag;
    }
    // The digest of the linearised polynomial.  The reference commits the polynomial with a full-size MultiExp
    // (kzg.Commit, prove.go:769); commitments are linear, so the same group element is the 7 + nb point combination
    // the VERIFIER forms from the key's and the proof's commitments (verify.go:210-250):
    //   l [Ql] + r [Qr] + r l [Qm] + o [Qo] + [Qk] + sum_i qcp_i(zeta) [Bsb22_i] + s1 [S3] + s2 [Z]
    // It runs on a host thread under the one MSM this phase still needs (the quotient of the Z opening); every
    // rank of a sharded proof computes it, rank 0 reports it as its partial sum.
    // one host task per scalar multiplication (6 + nb of them, ~0.4 ms each): at N > 1 the MSM they hide under is short
    std::vector<std::future<G1XYZZ>> terms;
    auto term = [&terms](const G1Affine& p, const Fr& k) {
        terms.push_back(std::async(std::launch::async, [p, k]() { return scalar_mul(G1XYZZ::from_affine(p), k); }));
    };
    {
        const G1Affine* vk = pk->vk_pts;   // S1 S2 S3 Ql Qr Qm Qo Qk
        term(vk[3], S.blz);
        term(vk[4], S.brz);
        term(vk[5], S.brz * S.blz);
        term(vk[6], S.boz);
        for (uint32_t k = 0; k < S.nb; ++k) term(S.bsb_com[k], S.qcpz[k]);
        term(vk[2], lin_s1);
        term(S.z_com, lin_s2);
    }
    const Fr* sc[1] = {S.quot};
    const size_t cnt[1] = {n + 2};
    G1XYZZ c[1];
    const int st_commit = commit_shard(ctx, pk->srs, pk->srs_lo_idx, pk->srs_hi_idx, sc, cnt, 1, c);
    G1XYZZ lin_digest = G1XYZZ::from_affine(pk->vk_pts[7]);      // + [Qk]
    for (auto& f : terms) lin_digest.add(f.get());                // joined before any early return
    B2_TRY(st_commit);
    put_jac(partials, c[0]);
    put_jac(partials + JAC_WORDS, pk->shard_index == 0 ? lin_digest : G1XYZZ::inf());
    S.lap(S.T.linearize_ms);
    return B200_OK;
}

int phase5(b200_ctx* ctx, b200_plonk_session& S, const uint64_t* totals, uint64_t* partials) {
    const b200_plonk_pk* pk = S.pk;
    const size_t n = pk->n;
    const uint32_t nb = S.nb;
    S.zs_h = get_jac(totals).to_affine();
    S.lin_digest = get_jac(totals + JAC_WORDS).to_affine();
    if (S.fh_digest_task.valid()) S.fh_digest = S.fh_digest_task.get();
    // ---- batch opening at zeta (batchOpening, prove.go:779-833; kzg.BatchOpenSinglePoint) --------------------
    B2_TRY(fr_poly_eval(ctx, S.fh, n + 2, S.zeta, &S.claimed[0]));
    B2_TRY(fr_poly_eval(ctx, S.lin, n + 3, S.zeta, &S.claimed[1]));
    S.claimed[2] = S.blz; S.claimed[3] = S.brz; S.claimed[4] = S.boz; S.claimed[5] = S.s1z; S.claimed[6] = S.s2z;
    {
        Transcript g({"gamma"});
        bind_fr(g, "gamma", S.zeta);
        const G1Affine* dg[7] = {&S.fh_digest, &S.lin_digest, &S.lro[0], &S.lro[1], &S.lro[2], &pk->vk_pts[0], &pk->vk_pts[1]};
        for (int k = 0; k < 7; ++k) bind_point(g, "gamma", *dg[k]);
        for (uint32_t k = 0; k < nb; ++k) bind_point(g, "gamma", pk->vk_qcp[k]);
        for (int k = 0; k < 7; ++k) bind_fr(g, "gamma", S.claimed[k]);
        for (uint32_t k = 0; k < nb; ++k) bind_fr(g, "gamma", S.qcpz[k]);
        bind_fr(g, "gamma", S.zu);
        S.gk = derive(g, "gamma");
    }
    S.folded = S.tmp[3];
    {
        PlonkFoldArgs a;
        memset(&a, 0, sizeof a);
        const Fr* ps[7] = {S.fh, S.lin, S.cn[0], S.cn[1], S.cn[2], pk->tr(TR_S1), pk->tr(TR_S2)};
        const size_t ls[7] = {n + 2, n + 3, n + 2, n + 2, n + 2, n, n};
        Fr acc = Fr::one();
        for (int k = 0; k < 7; ++k) { a.p[k] = ps[k]; a.len[k] = (uint32_t)ls[k]; a.g[k] = acc; acc = acc * S.gk; }
        for (uint32_t k = 0; k < nb; ++k) { a.p[7 + k] = pk->qcp_c(k); a.len[7 + k] = (uint32_t)n; a.g[7 + k] = acc; acc = acc * S.gk; }
        a.count = 7 + nb; a.width = (uint32_t)(n + 3);
        a.out = S.folded;
        B2_TRY(plonk_fold_polys(ctx, a));
    }
    B2_TRY(fr_poly_div_linear(ctx, S.folded, n + 3, S.zeta, S.quot));
    const Fr* sc[1] = {S.quot};
    const size_t cnt[1] = {n + 2};
    G1XYZZ c;
    B2_TRY(commit_shard(ctx, pk->srs, pk->srs_lo_idx, pk->srs_hi_idx, sc, cnt, 1, &c));
    put_jac(partials, c);
    S.lap(S.T.batch_open_ms);
    return B200_OK;
}

void phase6(b200_plonk_session& S, const uint64_t* totals, b200_plonk_proof* out) {
    S.batched_h = get_jac(totals).to_affine();
    S.T.total_ms = (float)(now_ms() - S.t_start);
    memset(out, 0, sizeof *out);
    for (int k = 0; k < 3; ++k) { put_point(out->lro[k], S.lro[k]); put_point(out->h[k], S.h_com[k]); }
    put_point(out->z, S.z_com);
    put_point(out->batched_h, S.batched_h);
    for (int k = 0; k < 7; ++k) put_fr(out->claimed[k], S.claimed[k]);
    put_point(out->zshift_h, S.zs_h);
    put_fr(out->zshift_value, S.zu);
    put_fr(out->challenges[0], S.gamma); put_fr(out->challenges[1], S.beta); put_fr(out->challenges[2], S.alpha);
    put_fr(out->challenges[3], S.zeta); put_fr(out->challenges[4], S.gk);
    out->nb_bsb22 = S.nb;
    for (uint32_t k = 0; k < S.nb; ++k) { put_fr(out->claimed_qcp[k], S.qcpz[k]); put_point(out->bsb22[k], S.bsb_com[k]); }
}

}  // namespace

extern "C" {

int b200_plonk_session_begin(b200_ctx* ctx, const b200_plonk_pk* pk, const uint64_t* l, const uint64_t* r, const uint64_t* o,
                             const uint64_t* public_witness, const uint64_t* blinding, const b200_plonk_bsb22* bsb,
                             b200_plonk_session** out, uint64_t* partials) try {
    if (!ctx) return B200_ERR_ARG;
    if (cudaSetDevice(ctx->device) != cudaSuccess) return ctx->fail(B200_ERR_CUDA, "cudaSetDevice");
    if (!pk || !l || !r || !o || !out || !partials || (pk->nb_public && !public_witness)) return ctx->fail(B200_ERR_ARG, "null argument");
    const uint32_t nb = pk->nb_bsb;
    if (nb && (!bsb || !bsb->pi2 || !bsb->commitments || !bsb->values)) return ctx->fail(B200_ERR_ARG, "the key has BSB22 commitments: bsb is required");
    if (pk->shard_count > 1 && !blinding) return ctx->fail(B200_ERR_ARG, "a sharded proof needs the same blinding on every rank: pass it");
    if (3 * (pk->n + 2) > pk->big) return ctx->fail(B200_ERR_ARG, "big domain below 3(n+2)");
    // the workspace of a proof lives in the key (and in the ctx): a second session on the same key before the first one
    // finished would silently corrupt both proofs -- refuse it
    if (pk->in_flight.exchange(true)) return ctx->fail(B200_ERR_ARG, "a proof is already in flight on this proving key (a key and a ctx prove one proof at a time)");
    std::unique_ptr<b200_plonk_session> S(new b200_plonk_session());
    S->pk = pk;          // ~b200_plonk_session releases the key
    S->nb = nb;
    const size_t n = pk->n, N4 = n + 4;
    for (uint32_t k = 0; k < nb; ++k) {
        if (!bsb->pi2[k]) return ctx->fail(B200_ERR_ARG, "null Pi2 column");
        memcpy(&S->bsb_com[k], bsb->commitments + AFF_WORDS * k, sizeof(G1Affine));
        memcpy(S->bsb_val[k].l, bsb->values + 4 * k, 32);
        S->pi2_br[k] = pk->bsb_work.as<Fr>() + (size_t)(3 * k) * n;
        S->pi2_cn[k] = S->pi2_br[k] + n;
        S->pi2_ev[k] = S->pi2_cn[k] + n;
    }
    memset(&S->T, 0, sizeof S->T);
    S->t_start = S->t0 = now_ms();
    // workspace
    Fr* w = pk->work.as<Fr>();
    for (int k = 0; k < 4; ++k) S->v[k] = w + (size_t)k * n;                   // L R O Z: Lagrange, then canonical bit-reversed
    S->qk_br = w + 4 * n;
    S->idv = w + 5 * n;
    for (int k = 0; k < 4; ++k) S->cn[k] = w + 6 * n + (size_t)k * N4;          // canonical / regular, blinded
    for (int k = 0; k < 5; ++k) S->tmp[k] = w + 6 * n + (size_t)(4 + k) * N4;   // coset evaluations of L R O Z Qk; reused after the quotient
    S->cres = w + 6 * n + 9 * N4;
    // blinding polynomials (initBlindingPolynomials, prove.go:295-301)
    if (blinding) memcpy(S->bp, blinding, sizeof S->bp);
    else if (os_random_fr(S->bp, 9) != B200_OK) return ctx->fail(B200_ERR_ARG, "no entropy source");
    for (int k = 0; k < 4; ++k) {
        const Fr* b = S->bp + 2 * k;
        const int nb = k < 3 ? 2 : 3;
        S->blind_com[k] = std::async(std::launch::async, [pk, b, nb]() { return commit_blinding(pk, b, nb); });
    }
    S->pub.resize(pk->nb_public);
    if (pk->nb_public) memcpy(S->pub.data(), public_witness, pk->nb_public * sizeof(Fr));
    B2_TRY(phase0(ctx, *S, l, r, o, bsb, partials));
    S->next_phase = 1;
    *out = S.release();
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)

int b200_plonk_session_step(b200_ctx* ctx, b200_plonk_session* S, int phase, const uint64_t* totals, uint64_t* partials) try {
    if (!ctx) return B200_ERR_ARG;
    if (cudaSetDevice(ctx->device) != cudaSuccess) return ctx->fail(B200_ERR_CUDA, "cudaSetDevice");
    if (!S || phase != S->next_phase || phase < 1 || phase > 5) return ctx->fail(B200_ERR_ARG, "phases run in order 1..5 after begin");
    if ((phase != 3 && !totals) || (phase != 2 && !partials)) return ctx->fail(B200_ERR_ARG, "null argument");
    S->t0 = now_ms();
    int st = B200_OK;
    switch (phase) {
        case 1: st = phase1(ctx, *S, totals, partials); break;
        case 2: st = phase2(ctx, *S, totals); break;
        case 3: st = phase3(ctx, *S, partials); break;
        case 4: st = phase4(ctx, *S, totals, partials); break;
        case 5: st = phase5(ctx, *S, totals, partials); break;
    }
    if (st == B200_OK) S->next_phase = phase + 1;
    return st;
} B2_ABI_CATCH_CTX(ctx)

int b200_plonk_session_numerator(b200_ctx* ctx, b200_plonk_session* S, void** cres_dev, size_t* block_elems, unsigned* log_blocks) try {
    if (!ctx) return B200_ERR_ARG;
    if (!S || !cres_dev || !block_elems || !log_blocks) return ctx->fail(B200_ERR_ARG, "null argument");
    *cres_dev = S->cres;
    *block_elems = S->pk->n;
    *log_blocks = S->pk->log_big - S->pk->log_n;
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)

int b200_plonk_session_finish(b200_ctx* ctx, b200_plonk_session* S, const uint64_t* totals, b200_plonk_proof* out, b200_plonk_timings* tm) try {
    if (!ctx) return B200_ERR_ARG;
    if (!S) return ctx->fail(B200_ERR_ARG, "null session");
    std::unique_ptr<b200_plonk_session> guard(S);
    if (S->next_phase != 6 && (out || totals)) return ctx->fail(B200_ERR_ARG, "finish comes after phase 5 (pass NULL totals and proof to abandon a session)");
    if (S->next_phase == 6) {
        if (!totals || !out) return ctx->fail(B200_ERR_ARG, "null argument");
        phase6(*S, totals, out);
        if (tm) *tm = S->T;
    }
    return B200_OK;
} B2_ABI_CATCH_CTX(ctx)

int b200_plonk_prove(b200_ctx* ctx, const b200_plonk_pk* pk, const uint64_t* l, const uint64_t* r, const uint64_t* o,
                     const uint64_t* public_witness, const uint64_t* blinding, b200_plonk_proof* out, b200_plonk_timings* tm) try {
    return b200_plonk_prove_bsb22(ctx, pk, l, r, o, public_witness, blinding, nullptr, out, tm);
} B2_ABI_CATCH_CTX(ctx)

int b200_plonk_prove_bsb22(b200_ctx* ctx, const b200_plonk_pk* pk, const uint64_t* l, const uint64_t* r, const uint64_t* o,
                           const uint64_t* public_witness, const uint64_t* blinding, const b200_plonk_bsb22* bsb,
                           b200_plonk_proof* out, b200_plonk_timings* tm) try {
    if (!ctx) return B200_ERR_ARG;
    if (!out) return ctx->fail(B200_ERR_ARG, "null argument");
    if (pk && pk->shard_count > 1) return ctx->fail(B200_ERR_ARG, "a sharded key proves through the session phases (b200_plonk_session_*)");
    b200_plonk_session* S = nullptr;
    uint64_t a[36], b[36];
    B2_TRY(b200_plonk_session_begin(ctx, pk, l, r, o, public_witness, blinding, bsb, &S, a));
    // one rank: the partial commitments are the totals
    int st = b200_plonk_session_step(ctx, S, 1, a, b);
    if (st == B200_OK) st = b200_plonk_session_step(ctx, S, 2, b, nullptr);
    if (st == B200_OK) st = b200_plonk_session_step(ctx, S, 3, nullptr, a);
    if (st == B200_OK) st = b200_plonk_session_step(ctx, S, 4, a, b);
    if (st == B200_OK) st = b200_plonk_session_step(ctx, S, 5, b, a);
    if (st != B200_OK) { b200_plonk_session_finish(ctx, S, nullptr, nullptr, nullptr); return st; }
    return b200_plonk_session_finish(ctx, S, a, out, tm);
} B2_ABI_CATCH_CTX(ctx)

int b200_plonk_derive_challenges(const uint64_t* vk_commitments, const uint64_t* vk_qcp, const uint64_t* public_witness, size_t nb_public,
                                 const b200_plonk_proof* p, uint64_t* out) try {
    if (!vk_commitments || !p || !out || (nb_public && !public_witness)) return B200_ERR_ARG;
    if (p->nb_bsb22 > B200_PLONK_MAX_BSB22 || (p->nb_bsb22 && !vk_qcp)) return B200_ERR_ARG;
    auto pt = [](const uint64_t* src) { G1Affine a; memcpy(&a, src, sizeof a); return a; };
    Transcript fs({"gamma", "beta", "alpha", "zeta"});
    for (int k = 0; k < 8; ++k) bind_point(fs, "gamma", pt(vk_commitments + AFF_WORDS * k));
    for (uint64_t k = 0; k < p->nb_bsb22; ++k) bind_point(fs, "gamma", pt(vk_qcp + AFF_WORDS * k));
    for (size_t k = 0; k < nb_public; ++k) { Fr x; memcpy(x.l, public_witness + 4 * k, 32); bind_fr(fs, "gamma", x); }
    for (int k = 0; k < 3; ++k) bind_point(fs, "gamma", pt(p->lro[k]));
    put_fr(out, derive(fs, "gamma"));
    put_fr(out + 4, derive(fs, "beta"));
    for (uint64_t k = 0; k < p->nb_bsb22; ++k) bind_point(fs, "alpha", pt(p->bsb22[k]));
    bind_point(fs, "alpha", pt(p->z));
    put_fr(out + 8, derive(fs, "alpha"));
    for (int k = 0; k < 3; ++k) bind_point(fs, "zeta", pt(p->h[k]));
    put_fr(out + 12, derive(fs, "zeta"));
    return B200_OK;
} B2_ABI_CATCH_INT

int b200_plonk_proof_bytes(const b200_plonk_proof* p, uint8_t* out, size_t cap) try {
    if (!p || !out || p->nb_bsb22 > B200_PLONK_MAX_BSB22) return B200_ERR_ARG;
    const uint32_t nb = (uint32_t)p->nb_bsb22;
    const size_t FB = FP_BYTES;                       // a compressed G1 point: 32 bytes on BN254, 48 on BLS12-377
    if (cap < 9 * FB + 8 + 32 * (8 + (size_t)nb) + FB * (size_t)nb) return B200_ERR_ARG;
    auto pt = [](const uint64_t* src) { G1Affine a; memcpy(&a, src, sizeof a); return a; };
    auto fr = [](const uint64_t* src) { Fr a; memcpy(a.l, src, 32); return a; };
    auto be32 = [](uint8_t* q, uint32_t v) { q[0] = (uint8_t)(v >> 24); q[1] = (uint8_t)(v >> 16); q[2] = (uint8_t)(v >> 8); q[3] = (uint8_t)v; };
    uint8_t* q = out;
    for (int k = 0; k < 3; ++k) { g1_compress(pt(p->lro[k]), q); q += FB; }
    g1_compress(pt(p->z), q); q += FB;
    for (int k = 0; k < 3; ++k) { g1_compress(pt(p->h[k]), q); q += FB; }
    g1_compress(pt(p->batched_h), q); q += FB;
    be32(q, 7 + nb); q += 4;                                                // []fr.Element: u32 BE length
    for (int k = 0; k < 7; ++k) { fr_be(fr(p->claimed[k]), q); q += 32; }
    for (uint32_t k = 0; k < nb; ++k) { fr_be(fr(p->claimed_qcp[k]), q); q += 32; }
    g1_compress(pt(p->zshift_h), q); q += FB;
    fr_be(fr(p->zshift_value), q); q += 32;
    be32(q, nb); q += 4;                                                    // Bsb22Commitments
    for (uint32_t k = 0; k < nb; ++k) { g1_compress(pt(p->bsb22[k]), q); q += FB; }
    return (int)(q - out);
} B2_ABI_CATCH_INT

}  // extern "C"
