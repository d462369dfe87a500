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
#include <chrono>
#include <cstdio>
#include "host_group.cuh"
#include "msm_host.cuh"
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
void g1_raw(const G1Affine& p, uint8_t out[64]) {
    if (p.is_inf()) { memset(out, 0, 64); out[0] = 0x40; return; }
    fp_be(p.x, out);
    fp_be(p.y, out + 32);
}
void bind_point(Transcript& fs, const char* name, const G1Affine& p) { uint8_t b[64]; g1_raw(p, b); fs.bind(name, b, 64); }
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
    size_t sort_max = 0, acc_max = 0, res_off[5] = {0};
    for (int k = 0; k < m; ++k) {
        if (counts[k] > srs->n) return ctx->fail(B200_ERR_ARG, "polynomial longer than the SRS");
        const MsmSizes sz = msm_sizes(counts[k], 1, ctx->msm_c_override, srs->fold.c);
        sort_max = std::max(sort_max, sz.sort_bytes);
        acc_max = std::max(acc_max, sz.acc_bytes);
        res_off[k + 1] = res_off[k] + ((sz.result_bytes + 255) & ~(size_t)255);
    }
    const bool overlap = ctx->msm_overlap != 0 && m > 1;
    Arena ar;
    const size_t o_sort0 = ar.off; ar.take(sort_max);
    const size_t o_sort1 = overlap ? ar.off : o_sort0; if (overlap) ar.take(sort_max);
    const size_t o_acc = ar.off; ar.take(acc_max);
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
                                 base + o_acc, (char*)ctx->h_stage + res_off[k]);
    B2_CUDA(ctx, cudaEventRecord(ev[0], st));
    if (overlap) B2_CUDA(ctx, cudaStreamWaitEvent(sx, ev[0], 0));
    auto run_sort = [&](int k) -> int {
        if (wait && wait[k]) B2_CUDA(ctx, cudaStreamWaitEvent(sx, wait[k], 0));
        if (k >= 2) B2_CUDA(ctx, cudaStreamWaitEvent(sx, ev_acc[k - 2], 0));       // its sort group is free again
        B2_TRY(msm_g1_job_sort(ctx, jobs[k], sx));
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
    delete pk;
}

extern "C" {

int b200_plonk_pk_upload(b200_ctx* ctx, int curve, const b200_plonk_pk_desc* d, b200_plonk_pk** out) {
    if (!ctx) return B200_ERR_ARG;
    if (curve != B200_BN254) return ctx->fail(B200_ERR_UNSUPPORTED, "only BN254 is built into this library");
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
    pk->nb_bsb = (uint32_t)d->nb_bsb22;
    for (uint32_t k = 0; k < pk->nb_bsb; ++k) {
        if (!d->qcp[k]) { delete pk; return ctx->fail(B200_ERR_ARG, "null Qcp"); }
        pk->commit_row[k] = d->nb_public + d->commitment_index[k];
        if (pk->commit_row[k] >= n) { delete pk; return ctx->fail(B200_ERR_ARG, "commitment index outside the domain"); }
    }
    if (pk->nb_bsb) memcpy(pk->vk_qcp, d->vk_qcp, pk->nb_bsb * sizeof(G1Affine));
    Fr gen, gen_big, cg;
    memcpy(gen.l, d->generator, 32); memcpy(gen_big.l, d->generator_big, 32); memcpy(cg.l, d->coset_gen, 32);
    int st = ntt_domain_get(ctx, d->log_n, gen, cg, &pk->d0);
    if (st == B200_OK) st = ntt_domain_get(ctx, d->log_big, gen_big, cg, &pk->d1);
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
    memcpy(pk->srs_lo, d->kzg_g1, 3 * sizeof(G1Affine));
    memcpy(pk->srs_hi, d->kzg_g1 + 8 * n, 3 * sizeof(G1Affine));
    memcpy(pk->vk_pts, d->vk_commitments, 8 * sizeof(G1Affine));
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
    if (st == B200_OK) st = b200_bases_upload(ctx, curve, 1, d->kzg_g1, n + 3, &pk->srs);
    if (st == B200_OK) st = b200_bases_upload(ctx, curve, 1, d->kzg_lagrange_g1, n, &pk->srs_lag);
    if (st != B200_OK) { plonk_pk_delete(ctx, pk); return st; }
    *out = pk;
    return B200_OK;
}

int b200_plonk_pk_free(b200_ctx* ctx, b200_plonk_pk* pk) {
    if (!ctx) return B200_ERR_ARG;
    if (cudaSetDevice(ctx->device) != cudaSuccess) return ctx->fail(B200_ERR_CUDA, "cudaSetDevice");
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    plonk_pk_delete(ctx, pk);
    return B200_OK;
}

int b200_plonk_prove(b200_ctx* ctx, const b200_plonk_pk* pk, const uint64_t* l, const uint64_t* r, const uint64_t* o,
                     const uint64_t* public_witness, const uint64_t* blinding, b200_plonk_proof* out, b200_plonk_timings* tm) {
    return b200_plonk_prove_bsb22(ctx, pk, l, r, o, public_witness, blinding, nullptr, out, tm);
}

int b200_plonk_prove_bsb22(b200_ctx* ctx, const b200_plonk_pk* pk, const uint64_t* l, const uint64_t* r, const uint64_t* o,
                           const uint64_t* public_witness, const uint64_t* blinding, const b200_plonk_bsb22* bsb,
                           b200_plonk_proof* out, b200_plonk_timings* tm) {
    if (!ctx) return B200_ERR_ARG;
    if (cudaSetDevice(ctx->device) != cudaSuccess) return ctx->fail(B200_ERR_CUDA, "cudaSetDevice");
    if (!pk || !l || !r || !o || !out || (pk->nb_public && !public_witness)) return ctx->fail(B200_ERR_ARG, "null argument");
    const uint32_t nb = pk->nb_bsb;
    if (nb && (!bsb || !bsb->pi2 || !bsb->commitments || !bsb->values)) return ctx->fail(B200_ERR_ARG, "the key has BSB22 commitments: bsb is required");
    G1Affine bsb_com[B200_PLONK_MAX_BSB22];
    Fr bsb_val[B200_PLONK_MAX_BSB22];
    Fr *pi2_br[B200_PLONK_MAX_BSB22], *pi2_cn[B200_PLONK_MAX_BSB22], *pi2_ev[B200_PLONK_MAX_BSB22];
    for (uint32_t k = 0; k < nb; ++k) {
        if (!bsb->pi2[k]) return ctx->fail(B200_ERR_ARG, "null Pi2 column");
        memcpy(&bsb_com[k], bsb->commitments + 8 * k, sizeof(G1Affine));
        memcpy(bsb_val[k].l, bsb->values + 4 * k, 32);
        pi2_br[k] = pk->bsb_work.as<Fr>() + (size_t)(3 * k) * pk->n;
        pi2_cn[k] = pi2_br[k] + pk->n;
        pi2_ev[k] = pi2_cn[k] + pk->n;
    }
    const size_t n = pk->n, N4 = n + 4, big = pk->big;
    const unsigned log_n = pk->log_n;
    cudaStream_t s = ctx->stream;
    NttDomain *d0 = pk->d0, *d1 = pk->d1;
    b200_plonk_timings T;
    memset(&T, 0, sizeof T);
    const double t_start = now_ms();
    double t0 = t_start;
    auto lap = [&](float& slot) { const double t = now_ms(); slot += (float)(t - t0); t0 = t; };

    // workspace
    Fr* w = pk->work.as<Fr>();
    Fr* v[4];                       // L R O Z: Lagrange, then canonical bit-reversed
    for (int k = 0; k < 4; ++k) v[k] = w + (size_t)k * n;
    Fr* qk_br = w + 4 * n;
    Fr* idv = w + 5 * n;
    Fr* cn[4];                      // canonical / regular, blinded: n + 2 (L R O) or n + 3 (Z) coefficients
    for (int k = 0; k < 4; ++k) cn[k] = w + 6 * n + (size_t)k * N4;
    Fr* tmp[5];                     // coset evaluations of L R O Z Qk; reused after the quotient
    for (int k = 0; k < 5; ++k) tmp[k] = w + 6 * n + (size_t)(4 + k) * N4;
    Fr* cres = w + 6 * n + 9 * N4;

    // blinding polynomials (initBlindingPolynomials, prove.go:295-301)
    Fr bp[9];
    if (blinding) memcpy(bp, blinding, sizeof bp);
    else if (os_random_fr(bp, 9) != B200_OK) return ctx->fail(B200_ERR_ARG, "no entropy source");
    const Fr *bl = bp, *br = bp + 2, *bo = bp + 4, *bz = bp + 6;
    std::vector<Fr> pub(pk->nb_public);
    if (pk->nb_public) memcpy(pub.data(), public_witness, pk->nb_public * sizeof(Fr));

    // ---- solution to the device, [L] [R] [O]  (solveConstraints / commitToLRO, prove.go:365-452) ----------
    // the copies run on the copy stream; the sort of [L] starts when L has landed, R and O travel under it
    const uint64_t* src[3] = {l, r, o};
    cudaEvent_t ev_cp[4];
    for (auto& e : ev_cp) B2_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    struct CpGuard { cudaEvent_t* e; ~CpGuard() { for (int i = 0; i < 4; ++i) cudaEventDestroy(e[i]); } } cpguard{ev_cp};
    cudaEvent_t ev_t[2];
    for (auto& e : ev_t) B2_CUDA(ctx, cudaEventCreate(&e));
    struct TGuard { cudaEvent_t* e; ~TGuard() { cudaEventDestroy(e[0]); cudaEventDestroy(e[1]); } } tguard{ev_t};
    B2_CUDA(ctx, cudaEventRecord(ev_cp[3], s));
    B2_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, ev_cp[3], 0));
    B2_CUDA(ctx, cudaEventRecord(ev_t[0], ctx->copy_stream));
    for (int k = 0; k < 3; ++k) {
        B2_CUDA(ctx, cudaMemcpyAsync(v[k], src[k], n * sizeof(Fr), cudaMemcpyDefault, ctx->copy_stream));
        B2_CUDA(ctx, cudaEventRecord(ev_cp[k], ctx->copy_stream));
    }
    for (uint32_t k = 0; k < nb; ++k)
        B2_CUDA(ctx, cudaMemcpyAsync(pi2_br[k], bsb->pi2[k], n * sizeof(Fr), cudaMemcpyDefault, ctx->copy_stream));
    B2_CUDA(ctx, cudaEventRecord(ev_t[1], ctx->copy_stream));
    G1Affine lro[3];
    const Fr* bps[3] = {bl, br, bo};
    {
        const Fr* sc[3] = {v[0], v[1], v[2]};
        const size_t cnt[3] = {n, n, n};
        G1XYZZ c[3];
        if (ctx->msm_overlap) B2_TRY(commit_many(ctx, pk->srs_lag, sc, cnt, 3, c, ev_cp));
        else {
            for (int k = 0; k < 3; ++k) B2_CUDA(ctx, cudaStreamWaitEvent(s, ev_cp[k], 0));
            B2_TRY(commit_many(ctx, pk->srs_lag, sc, cnt, 3, c));
        }
        for (int k = 0; k < 3; ++k) B2_CUDA(ctx, cudaStreamWaitEvent(s, ev_cp[k], 0));
        B2_CUDA(ctx, cudaStreamWaitEvent(s, ev_t[1], 0));
        for (int k = 0; k < 3; ++k) {
            c[k].add(commit_blinding(pk, bps[k], 2));
            lro[k] = c[k].to_affine();
        }
    }
    lap(T.commit_lro_ms);
    B2_CUDA(ctx, cudaEventSynchronize(ev_t[1]));
    cudaEventElapsedTime(&T.h2d_ms, ev_t[0], ev_t[1]);      // on the copy stream, under the [L] [R] [O] commitments: not a summand of total_ms

    // ---- gamma, beta (deriveGammaAndBeta, prove.go:455-485; bindPublicData, verify.go:296-340) -----------
    Transcript fs({"gamma", "beta", "alpha", "zeta"});
    for (int k = 0; k < 8; ++k) bind_point(fs, "gamma", pk->vk_pts[k]);
    for (uint32_t k = 0; k < nb; ++k) bind_point(fs, "gamma", pk->vk_qcp[k]);
    for (const Fr& x : pub) bind_fr(fs, "gamma", x);
    for (int k = 0; k < 3; ++k) bind_point(fs, "gamma", lro[k]);
    const Fr gamma = derive(fs, "gamma");
    const Fr beta = derive(fs, "beta");

    // ---- Z and [Z] (buildRatioCopyConstraint, prove.go:597-631) ------------------------------------------
    B2_TRY(plonk_build_ratio(ctx, d0, v[0], v[1], v[2], pk->perm.as<int64_t>(), beta, gamma, v[3], tmp[0], tmp[1]));
    B2_CUDA(ctx, cudaStreamSynchronize(s));
    lap(T.ratio_ms);
    G1Affine z_com;
    {
        G1XYZZ c;
        B2_TRY(commit(ctx, pk->srs_lag, v[3], n, &c));
        c.add(commit_blinding(pk, bz, 3));
        z_com = c.to_affine();
    }
    lap(T.commit_z_ms);
    for (uint32_t k = 0; k < nb; ++k) bind_point(fs, "alpha", bsb_com[k]);     // deriveAlpha, prove.go:504-512
    bind_point(fs, "alpha", z_com);
    const Fr alpha = derive(fs, "alpha");

    // ---- quotient (evaluateConstraints / computeNumerator / divideByXMinusOne, prove.go:521-595,837-1079,1223) ---
    // canonical forms: FFTInverse(DIF) leaves the coefficients bit-reversed, which is what the coset DIT wants
    B2_CUDA(ctx, cudaMemcpyAsync(qk_br, pk->qk_lag.p, n * sizeof(Fr), cudaMemcpyDeviceToDevice, s));
    if (pk->nb_public) B2_CUDA(ctx, cudaMemcpyAsync(qk_br, pub.data(), pk->nb_public * sizeof(Fr), cudaMemcpyHostToDevice, s));   // completeQk
    for (uint32_t k = 0; k < nb; ++k)      // qkCoeffs[nbPublic + CommitmentIndex] = commitmentVal, prove.go:415-417
        B2_CUDA(ctx, cudaMemcpyAsync(qk_br + pk->commit_row[k], bsb_val[k].l, sizeof(Fr), cudaMemcpyHostToDevice, s));
    for (int k = 0; k < 4; ++k) B2_TRY(ntt_gnark(ctx, d0, v[k], 1, B200_DIF, 0));
    B2_TRY(ntt_gnark(ctx, d0, qk_br, 1, B200_DIF, 0));
    for (int k = 0; k < 4; ++k) B2_TRY(fr_bitrev_copy(ctx, v[k], cn[k], log_n));   // regular layout, for the openings
    for (uint32_t k = 0; k < nb; ++k) {                                            // cCommitments[i].ToCanonical, prove.go:339
        B2_TRY(ntt_gnark(ctx, d0, pi2_br[k], 1, B200_DIF, 0));
        B2_TRY(fr_bitrev_copy(ctx, pi2_br[k], pi2_cn[k], log_n));
    }
    {
        PlonkRowArgs a;
        memset(&a, 0, sizeof a);
        a.nb_bsb = nb;
        a.beta = beta; a.gamma = gamma; a.alpha = alpha;
        a.cs = d1->coset;
        a.css = a.cs * a.cs;
        a.tw = d0->tw_fwd.as<Fr>();
        a.n = (uint32_t)n; a.log_big = pk->log_big; a.rho = pk->rho;
        a.cres = cres;
        Fr coset = Fr::one();
        const Fr* srcs[5] = {v[0], v[1], v[2], v[3], qk_br};
        for (uint32_t k = 0; k < pk->rho; ++k) {
            coset = coset * (k == 0 ? d1->coset : d1->gen);
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
                B2_CUDA(ctx, cudaMemcpyAsync(tmp[p], srcs[p], n * sizeof(Fr), cudaMemcpyDeviceToDevice, s));
                B2_TRY(ntt_run(ctx, d0, tmp[p], 0, 1, pre, nullptr, nullptr));
            }
            for (uint32_t p = 0; p < nb; ++p) {
                B2_CUDA(ctx, cudaMemcpyAsync(pi2_ev[p], pi2_br[p], n * sizeof(Fr), cudaMemcpyDeviceToDevice, s));
                B2_TRY(ntt_run(ctx, d0, pi2_ev[p], 0, 1, pre, nullptr, nullptr));
                a.qci[2 * p] = pk->qcp_e(k, p);
                a.qci[2 * p + 1] = pi2_ev[p];
            }
            B2_TRY(fr_scaled_omega(ctx, d0, beta * coset, idv));
            a.x[PID_L] = tmp[0]; a.x[PID_R] = tmp[1]; a.x[PID_O] = tmp[2]; a.x[PID_Z] = tmp[3]; a.x[PID_ZS] = nullptr;
            a.x[PID_QK] = tmp[4];
            a.x[PID_QL] = pk->cev(k, CE_QL); a.x[PID_QR] = pk->cev(k, CE_QR); a.x[PID_QM] = pk->cev(k, CE_QM); a.x[PID_QO] = pk->cev(k, CE_QO);
            a.x[PID_S1] = pk->cev(k, CE_S1); a.x[PID_S2] = pk->cev(k, CE_S2); a.x[PID_S3] = pk->cev(k, CE_S3);
            a.x[PID_LONE] = pk->cev(k, CE_LONE);
            a.x[PID_ID] = idv;
            a.coset_index = k;
            B2_TRY(plonk_quotient_rows(ctx, a));
        }
    }
    B2_TRY(plonk_divide_xn(ctx, d1, log_n, cres));
    B2_CUDA(ctx, cudaStreamSynchronize(s));
    lap(T.quotient_ms);
    if (3 * (n + 2) > big) return ctx->fail(B200_ERR_ARG, "big domain below 3(n+2)");
    const Fr *h1 = cres, *h2 = cres + (n + 2), *h3 = cres + 2 * (n + 2);      // prove.go:652-666
    G1XYZZ hx[3];
    G1Affine h_com[3];
    {                                                                          // commitToQuotient, prove.go:1198-1218
        const Fr* sc[3] = {h1, h2, h3};
        const size_t cnt[3] = {n + 2, n + 2, n + 2};
        B2_TRY(commit_many(ctx, pk->srs, sc, cnt, 3, hx));
        for (int k = 0; k < 3; ++k) h_com[k] = hx[k].to_affine();
    }
    lap(T.commit_h_ms);
    for (int k = 0; k < 3; ++k) bind_point(fs, "zeta", h_com[k]);
    const Fr zeta = derive(fs, "zeta");

    // ---- open Z at w zeta (openZ, prove.go:634-650; kzg.Open) ------------------------------------------------
    B2_TRY(fr_blind(ctx, cn[0], n, bl, 2));
    B2_TRY(fr_blind(ctx, cn[1], n, br, 2));
    B2_TRY(fr_blind(ctx, cn[2], n, bo, 2));
    B2_TRY(fr_blind(ctx, cn[3], n, bz, 3));
    const Fr zeta_shifted = zeta * d0->gen;
    Fr zu;
    B2_TRY(fr_poly_eval(ctx, cn[3], n + 3, zeta_shifted, &zu));
    Fr* quot = tmp[0];
    B2_TRY(fr_poly_div_linear(ctx, cn[3], n + 3, zeta_shifted, quot));
    lap(T.open_z_ms);

    // ---- foldH (prove.go:670-705) ---------------------------------------------------------------------
    const Fr zn2 = zeta.pow_u64(n + 2);
    Fr* fh = tmp[1];
    B2_TRY(plonk_fold_h(ctx, h1, h2, h3, zn2, n + 2, fh));
    G1XYZZ fhd = scalar_mul(hx[2], zn2);
    fhd.add(hx[1]);
    fhd = scalar_mul(fhd, zn2);
    fhd.add(hx[0]);
    const G1Affine fh_digest = fhd.to_affine();

    // ---- linearised polynomial (computeLinearizedPolynomial, prove.go:708-777,1289-1389) --------------------
    Fr blz, brz, boz, s1z, s2z;                // evaluateBlinded == the blinded polynomial at zeta
    B2_TRY(fr_poly_eval(ctx, cn[0], n + 2, zeta, &blz));
    B2_TRY(fr_poly_eval(ctx, cn[1], n + 2, zeta, &brz));
    B2_TRY(fr_poly_eval(ctx, cn[2], n + 2, zeta, &boz));
    B2_TRY(fr_poly_eval(ctx, pk->tr(TR_S1), n, zeta, &s1z));
    B2_TRY(fr_poly_eval(ctx, pk->tr(TR_S2), n, zeta, &s2z));
    Fr qcpz[B200_PLONK_MAX_BSB22];
    for (uint32_t k = 0; k < nb; ++k) B2_TRY(fr_poly_eval(ctx, pk->qcp_c(k), n, zeta, &qcpz[k]));     // prove.go:724-729
    Fr* lin = tmp[2];
    {
        PlonkLinArgs a;
        memset(&a, 0, sizeof a);
        a.bz = cn[3]; a.s3 = pk->tr(TR_S3); a.ql = pk->tr(TR_QL); a.qr = pk->tr(TR_QR); a.qm = pk->tr(TR_QM); a.qo = pk->tr(TR_QO);
        a.qk = pk->tr(TR_QK);
        a.nb_bsb = nb;
        for (uint32_t k = 0; k < nb; ++k) { a.pi2[k] = pi2_cn[k]; a.qcp_zeta[k] = qcpz[k]; }
        a.rl = brz * blz;
        a.c1 = (s1z * beta + blz + gamma) * (s2z * beta + brz + gamma) * zu * beta;
        const Fr u = d0->coset, uz = zeta * u, uuz = uz * u;
        a.c2 = ((beta * zeta + blz + gamma) * (beta * uz + brz + gamma) * (beta * uuz + boz + gamma)).neg();
        const Fr zn1 = zeta.pow_u64(n) - Fr::one();
        a.lag = zn1 * (zeta - Fr::one()).inv() * alpha * alpha * d0->ninv;
        a.alpha = alpha; a.l = blz; a.r = brz; a.o = boz;
        a.n = (uint32_t)n; a.len = (uint32_t)(n + 3);
        a.out = lin;
        B2_TRY(plonk_linearize(ctx, a));
    }
    G1Affine lin_digest, zs_h;
    {   // [quotient of the Z opening] and [linearised polynomial]: independent, committed as one pipelined pair
        const Fr* sc[2] = {quot, lin};
        const size_t cnt[2] = {n + 2, n + 3};
        G1XYZZ c[2];
        B2_TRY(commit_many(ctx, pk->srs, sc, cnt, 2, c));
        zs_h = c[0].to_affine();
        lin_digest = c[1].to_affine();
    }
    lap(T.linearize_ms);

    // ---- batch opening at zeta (batchOpening, prove.go:779-833; kzg.BatchOpenSinglePoint) --------------------
    Fr claimed[7];
    B2_TRY(fr_poly_eval(ctx, fh, n + 2, zeta, &claimed[0]));
    B2_TRY(fr_poly_eval(ctx, lin, n + 3, zeta, &claimed[1]));
    claimed[2] = blz; claimed[3] = brz; claimed[4] = boz; claimed[5] = s1z; claimed[6] = s2z;
    Fr gk;
    {
        Transcript g({"gamma"});
        bind_fr(g, "gamma", zeta);
        const G1Affine* dg[7] = {&fh_digest, &lin_digest, &lro[0], &lro[1], &lro[2], &pk->vk_pts[0], &pk->vk_pts[1]};
        for (int k = 0; k < 7; ++k) bind_point(g, "gamma", *dg[k]);
        for (uint32_t k = 0; k < nb; ++k) bind_point(g, "gamma", pk->vk_qcp[k]);
        for (int k = 0; k < 7; ++k) bind_fr(g, "gamma", claimed[k]);
        for (uint32_t k = 0; k < nb; ++k) bind_fr(g, "gamma", qcpz[k]);
        bind_fr(g, "gamma", zu);
        gk = derive(g, "gamma");
    }
    Fr* folded = tmp[3];
    {
        PlonkFoldArgs a;
        memset(&a, 0, sizeof a);
        const Fr* ps[7] = {fh, lin, cn[0], cn[1], cn[2], pk->tr(TR_S1), pk->tr(TR_S2)};
        const size_t ls[7] = {n + 2, n + 3, n + 2, n + 2, n + 2, n, n};
        Fr acc = Fr::one();
        for (int k = 0; k < 7; ++k) { a.p[k] = ps[k]; a.len[k] = (uint32_t)ls[k]; a.g[k] = acc; acc = acc * gk; }
        for (uint32_t k = 0; k < nb; ++k) { a.p[7 + k] = pk->qcp_c(k); a.len[7 + k] = (uint32_t)n; a.g[7 + k] = acc; acc = acc * gk; }
        a.count = 7 + nb; a.width = (uint32_t)(n + 3);
        a.out = folded;
        B2_TRY(plonk_fold_polys(ctx, a));
    }
    B2_TRY(fr_poly_div_linear(ctx, folded, n + 3, zeta, quot));
    G1Affine batched_h;
    {
        G1XYZZ c;
        B2_TRY(commit(ctx, pk->srs, quot, n + 2, &c));
        batched_h = c.to_affine();
    }
    lap(T.batch_open_ms);
    T.total_ms = (float)(now_ms() - t_start);

    memset(out, 0, sizeof *out);
    for (int k = 0; k < 3; ++k) { put_point(out->lro[k], lro[k]); put_point(out->h[k], h_com[k]); }
    put_point(out->z, z_com);
    put_point(out->batched_h, batched_h);
    for (int k = 0; k < 7; ++k) put_fr(out->claimed[k], claimed[k]);
    put_point(out->zshift_h, zs_h);
    put_fr(out->zshift_value, zu);
    put_fr(out->challenges[0], gamma); put_fr(out->challenges[1], beta); put_fr(out->challenges[2], alpha);
    put_fr(out->challenges[3], zeta); put_fr(out->challenges[4], gk);
    out->nb_bsb22 = nb;
    for (uint32_t k = 0; k < nb; ++k) { put_fr(out->claimed_qcp[k], qcpz[k]); put_point(out->bsb22[k], bsb_com[k]); }
    if (tm) *tm = T;
    return B200_OK;
}

int b200_plonk_derive_challenges(const uint64_t* vk_commitments, const uint64_t* vk_qcp, const uint64_t* public_witness, size_t nb_public,
                                 const b200_plonk_proof* p, uint64_t* out) {
    if (!vk_commitments || !p || !out || (nb_public && !public_witness)) return B200_ERR_ARG;
    if (p->nb_bsb22 > B200_PLONK_MAX_BSB22 || (p->nb_bsb22 && !vk_qcp)) return B200_ERR_ARG;
    auto pt = [](const uint64_t* src) { G1Affine a; memcpy(&a, src, sizeof a); return a; };
    Transcript fs({"gamma", "beta", "alpha", "zeta"});
    for (int k = 0; k < 8; ++k) bind_point(fs, "gamma", pt(vk_commitments + 8 * k));
    for (uint64_t k = 0; k < p->nb_bsb22; ++k) bind_point(fs, "gamma", pt(vk_qcp + 8 * k));
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
}

int b200_plonk_proof_bytes(const b200_plonk_proof* p, uint8_t* out, size_t cap) {
    if (!p || !out || p->nb_bsb22 > B200_PLONK_MAX_BSB22) return B200_ERR_ARG;
    const uint32_t nb = (uint32_t)p->nb_bsb22;
    if (cap < 552 + 64 * (size_t)nb) return B200_ERR_ARG;
    auto pt = [](const uint64_t* src) { G1Affine a; memcpy(&a, src, sizeof a); return a; };
    auto fr = [](const uint64_t* src) { Fr a; memcpy(a.l, src, 32); return a; };
    auto be32 = [](uint8_t* q, uint32_t v) { q[0] = (uint8_t)(v >> 24); q[1] = (uint8_t)(v >> 16); q[2] = (uint8_t)(v >> 8); q[3] = (uint8_t)v; };
    uint8_t* q = out;
    for (int k = 0; k < 3; ++k) { g1_compress(pt(p->lro[k]), q); q += 32; }
    g1_compress(pt(p->z), q); q += 32;
    for (int k = 0; k < 3; ++k) { g1_compress(pt(p->h[k]), q); q += 32; }
    g1_compress(pt(p->batched_h), q); q += 32;
    be32(q, 7 + nb); q += 4;                                                // []fr.Element: u32 BE length
    for (int k = 0; k < 7; ++k) { fr_be(fr(p->claimed[k]), q); q += 32; }
    for (uint32_t k = 0; k < nb; ++k) { fr_be(fr(p->claimed_qcp[k]), q); q += 32; }
    g1_compress(pt(p->zshift_h), q); q += 32;
    fr_be(fr(p->zshift_value), q); q += 32;
    be32(q, nb); q += 4;                                                    // Bsb22Commitments
    for (uint32_t k = 0; k < nb; ++k) { g1_compress(pt(p->bsb22[k]), q); q += 32; }
    return (int)(q - out);
}

}  // extern "C"
