// Groth16 Setup with the proving key generated in place on the device.
// Restates /root/reference/backend/groth16/bn254/setup.go:85-337 (Setup) and :352-433 (setupABC) for a caller-
// supplied toxic waste.  Where the reference walks the constraints with the sequential recurrence
// L_{j+1} = w L_j (t - w^j) / (t - w^{j+1}) (:424-427), all L_j(t) = (t^n - 1)/n * w^j / (t - w^j) come from three
// device kernels around one batch inversion; the accumulation over the sparse matrices runs on host threads over a
// wire-major (CSC) view; the two BatchScalarMultiplication calls (:251,308) are the device fixed-base kernel writing
// straight into the resident key (groth16.cu: generate_range).
#include <algorithm>
#include <thread>
#include "groth16_host.cuh"
#include "poly_host.cuh"
#include "misc_host.cuh"

namespace b200 {
namespace {

__global__ void setup_tau_minus_kernel(Fr* __restrict__ d, const Fr* __restrict__ w, const Fr tau, size_t m) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m) st_vec(d + i, tau - ld_vec(w + i));
}
// L[j] = inv[j] * w[j] * c
__global__ void setup_lagrange_kernel(Fr* __restrict__ inv_io, const Fr* __restrict__ w, const Fr c, size_t m) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m) st_vec(inv_io + i, ld_vec(inv_io + i) * ld_vec(w + i) * c);
}

template <class Fn>
void parallel_chunks(size_t n, size_t grain, Fn fn) {
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const size_t chunks = std::max<size_t>(1, std::min<size_t>(hw, (n + grain - 1) / grain));
    if (chunks == 1) { fn((size_t)0, n); return; }
    std::vector<std::thread> th;
    th.reserve(chunks);
    for (size_t c = 0; c < chunks; ++c) th.emplace_back([=]() { fn(n * c / chunks, n * (c + 1) / chunks); });
    for (auto& t : th) t.join();
}

inline Fr load_fr(const uint64_t* p) { Fr f; memcpy(f.l, p, sizeof f.l); return f; }

// res[wire] = sum over the terms (j, cid, wire) of coeffs[cid] * L[j]   (setupABC's accumulate, setup.go:384-422)
int accumulate_matrix(b200_ctx* ctx, const b200_r1cs_matrix& mtx, size_t nb_constraints, size_t nb_wires, const Fr* coeffs,
                      size_t nb_coeffs, const Fr* L, std::vector<Fr>& res) {
    res.assign(nb_wires, Fr::zero());
    if (!mtx.row_ptr || (!mtx.terms && mtx.row_ptr[nb_constraints])) return ctx->fail(B200_ERR_ARG, "setup: null constraint matrix");
    const size_t nnz = mtx.row_ptr[nb_constraints];
    // wire-major view: counting sort of the terms by wire id
    std::vector<uint64_t> col_ptr(nb_wires + 1, 0);
    for (size_t j = 0; j < nb_constraints; ++j) {
        if (mtx.row_ptr[j + 1] < mtx.row_ptr[j] || mtx.row_ptr[j + 1] > nnz) return ctx->fail(B200_ERR_ARG, "setup: row_ptr is not monotone");
        for (uint64_t e = mtx.row_ptr[j]; e < mtx.row_ptr[j + 1]; ++e) {
            const uint32_t cid = mtx.terms[2 * e], vid = mtx.terms[2 * e + 1];
            if (vid >= nb_wires || cid >= nb_coeffs) return ctx->fail(B200_ERR_ARG, "setup: term refers to a wire or coefficient out of range");
            ++col_ptr[vid + 1];
        }
    }
    for (size_t i = 0; i < nb_wires; ++i) col_ptr[i + 1] += col_ptr[i];
    struct Ent { uint32_t row, cid; };
    std::vector<Ent> ent(nnz);
    {
        std::vector<uint64_t> pos(col_ptr.begin(), col_ptr.end() - 1);
        for (size_t j = 0; j < nb_constraints; ++j)
            for (uint64_t e = mtx.row_ptr[j]; e < mtx.row_ptr[j + 1]; ++e)
                ent[pos[mtx.terms[2 * e + 1]]++] = {(uint32_t)j, mtx.terms[2 * e]};
    }
    const Fr one = Fr::one();
    parallel_chunks(nb_wires, 4096, [&](size_t lo, size_t hi) {
        for (size_t i = lo; i < hi; ++i) {
            Fr acc = Fr::zero();
            for (uint64_t e = col_ptr[i]; e < col_ptr[i + 1]; ++e) {
                const Fr& c = coeffs[ent[e].cid];
                acc = acc + (c == one ? L[ent[e].row] : c * L[ent[e].row]);
            }
            res[i] = acc;
        }
    });
    return B200_OK;
}

template <class A>
int mul_points(b200_ctx* ctx, const A& base, const std::vector<Fr>& sc, uint64_t* host_out) {
    if (sc.empty() || !host_out) return B200_OK;
    DevBuf d_sc, d_pt;
    B2_CUDA(ctx, d_sc.alloc(sc.size() * sizeof(Fr)));
    B2_CUDA(ctx, d_pt.alloc(sc.size() * sizeof(A)));
    B2_CUDA(ctx, cudaMemcpyAsync(d_sc.p, sc.data(), sc.size() * sizeof(Fr), cudaMemcpyHostToDevice, ctx->stream));
    if constexpr (sizeof(A) == sizeof(G1Affine)) B2_TRY(batch_mul_g1(ctx, base, d_sc.as<Fr>(), sc.size(), d_pt.as<G1Affine>()));
    else B2_TRY(batch_mul_g2(ctx, base, d_sc.as<Fr>(), sc.size(), d_pt.as<G2Affine>()));
    B2_CUDA(ctx, cudaMemcpyAsync(host_out, d_pt.p, sc.size() * sizeof(A), cudaMemcpyDeviceToHost, ctx->stream));
    B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return B200_OK;
}

}  // namespace

int groth16_setup_scalars(b200_ctx* ctx, const b200_groth16_setup_desc* d, SetupScalars& S) {
    if (!d || !d->generator || !d->coset_gen || !d->g1_gen || !d->g2_gen) return ctx->fail(B200_ERR_ARG, "setup: null argument");
    if (d->log_n > 27) return ctx->fail(B200_ERR_UNSUPPORTED, "setup: domain above 2^27");
    const size_t n = (size_t)1 << d->log_n, m = d->nb_constraints, nw = d->nb_wires;
    if (m > n || (d->log_n && m <= n / 2)) return ctx->fail(B200_ERR_ARG, "setup: the domain is not fft.NewDomain(nb_constraints)");
    if (d->nb_public > nw || nw >= ((size_t)1 << 32)) return ctx->fail(B200_ERR_ARG, "setup: bad wire counts");
    if (d->nb_coeffs && !d->coeffs) return ctx->fail(B200_ERR_ARG, "setup: null coefficient table");
    if (d->nb_commitments && !d->wire_class) return ctx->fail(B200_ERR_ARG, "setup: commitments without wire classes");
    const Fr tau = load_fr(d->tau), alpha = load_fr(d->alpha), beta = load_fr(d->beta), gamma = load_fr(d->gamma), delta = load_fr(d->delta);
    const Fr gen = load_fr(d->generator);
    if (gamma.is_zero() || delta.is_zero()) return ctx->fail(B200_ERR_ARG, "setup: gamma and delta must be invertible");
    const Fr zt = tau.pow_u64((uint64_t)n) - Fr::one();     // t^n - 1
    if (zt.is_zero()) return ctx->fail(B200_ERR_ARG, "setup: tau lies in the evaluation domain");
    const Fr gamma_inv = gamma.inv(), delta_inv = delta.inv();
    const Fr ninv = Fr::from_u32(1).dbl().pow_u64(d->log_n).inv();

    // ---- L_j(tau), j < nb_constraints, on the device ------------------------------------------------------
    std::vector<Fr> L(m);
    if (m) {
        DevBuf d_w, d_d, d_inv;
        B2_CUDA(ctx, d_w.alloc(m * sizeof(Fr)));
        B2_CUDA(ctx, d_d.alloc(m * sizeof(Fr)));
        B2_CUDA(ctx, d_inv.alloc(m * sizeof(Fr)));
        B2_TRY(gen_powers(ctx, d_w.as<Fr>(), gen, Fr::one(), m, 0));
        const unsigned blocks = (unsigned)((m + 255) / 256);
        setup_tau_minus_kernel<<<blocks, 256, 0, ctx->stream>>>(d_d.as<Fr>(), d_w.as<Fr>(), tau, m);
        B2_CUDA(ctx, cudaGetLastError());
        B2_TRY(fr_batch_inverse(ctx, d_d.as<Fr>(), d_inv.as<Fr>(), m));
        setup_lagrange_kernel<<<blocks, 256, 0, ctx->stream>>>(d_inv.as<Fr>(), d_w.as<Fr>(), zt * ninv, m);
        B2_CUDA(ctx, cudaGetLastError());
        ctx->n_launches += 2;
        B2_CUDA(ctx, cudaMemcpyAsync(L.data(), d_inv.p, m * sizeof(Fr), cudaMemcpyDeviceToHost, ctx->stream));
        B2_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }

    // ---- A_i(tau), B_i(tau), C_i(tau) (setupABC) ------------------------------------------------------------
    const Fr* coeffs = reinterpret_cast<const Fr*>(d->coeffs);
    std::vector<Fr> A, B, C;
    B2_TRY(accumulate_matrix(ctx, d->l, m, nw, coeffs, d->nb_coeffs, L.data(), A));
    B2_TRY(accumulate_matrix(ctx, d->r, m, nw, coeffs, d->nb_coeffs, L.data(), B));
    B2_TRY(accumulate_matrix(ctx, d->o, m, nw, coeffs, d->nb_coeffs, L.data(), C));
    std::vector<Fr>().swap(L);

    // ---- K scalars by wire class (setup.go:144-200) ---------------------------------------------------------
    // class 0 below nb_public and class 1: vk.G1.K (in wire order); class 2 + j: commitment basis j; else pk.G1.K
    auto klass = [&](size_t i) -> int { return i < d->nb_public ? -1 : (d->wire_class ? (int)d->wire_class[i] : 0); };
    std::vector<Fr> kall(nw);
    parallel_chunks(nw, 4096, [&](size_t lo, size_t hi) {
        for (size_t i = lo; i < hi; ++i) {
            const Fr t = A[i] * beta + B[i] * alpha + C[i];
            kall[i] = t * (klass(i) == 0 ? delta_inv : gamma_inv);
        }
    });
    std::vector<Fr>& vkK = S.vkK;
    std::vector<Fr>& pkK = S.pkK;
    std::vector<std::vector<Fr>>& ckK = S.ckK;
    vkK.clear(); pkK.clear();
    ckK.assign(d->nb_commitments, std::vector<Fr>());
    for (size_t i = 0; i < nw; ++i) {
        const int k = klass(i);
        if (k == -1 || k == 1) vkK.push_back(kall[i]);
        else if (k == 0) pkK.push_back(kall[i]);
        else if ((size_t)(k - 2) < d->nb_commitments) ckK[k - 2].push_back(kall[i]);
        else return ctx->fail(B200_ERR_ARG, "setup: wire class beyond nb_commitments");
    }
    std::vector<Fr>().swap(kall);
    std::vector<Fr>().swap(C);
    if (vkK.size() != d->nb_public + d->nb_commitments) return ctx->fail(B200_ERR_ARG, "setup: the number of commitment wires differs from nb_commitments");

    // ---- infinity filters (setup.go:217-243) -------------------------------------------------------------------
    std::vector<uint8_t>& infA = S.infA;
    std::vector<uint8_t>& infB = S.infB;
    infA.assign(nw, 0); infB.assign(nw, 0);
    size_t na = 0, nb = 0;
    for (size_t i = 0; i < nw; ++i) {
        infA[i] = A[i].is_zero();
        if (!infA[i]) A[na++] = A[i];
        infB[i] = B[i].is_zero();
        if (!infB[i]) B[nb++] = B[i];
    }
    A.resize(na);
    B.resize(nb);
    S.A.swap(A);
    S.B.swap(B);
    S.z_mult = zt * delta_inv;
    S.tau = tau; S.alpha = alpha; S.beta = beta; S.gamma = gamma; S.delta = delta;
    return B200_OK;
}

// The point side of Setup for one shard: this device's ranges of the key from the scalars, the verifying-key side and the
// Pedersen bases when `out` asks for them.
int groth16_setup_points(b200_ctx* ctx, const b200_groth16_setup_desc* d, const SetupScalars& S, int shard_index, int shard_count,
                         b200_groth16_pk** pk_out, b200_groth16_setup_out* out, bool vk_side) {
    if (!pk_out || !out) return ctx->fail(B200_ERR_ARG, "setup: null argument");
    const size_t n = (size_t)1 << d->log_n, nw = d->nb_wires;
    const std::vector<Fr>&A = S.A, &B = S.B, &pkK = S.pkK, &vkK = S.vkK;
    const std::vector<std::vector<Fr>>& ckK = S.ckK;
    const size_t na = A.size(), nb = B.size();
    const Fr alpha = S.alpha, beta = S.beta, gamma = S.gamma, delta = S.delta;
    if (out->infinity_a) memcpy(out->infinity_a, S.infA.data(), nw);
    if (out->infinity_b) memcpy(out->infinity_b, S.infB.data(), nw);

    // ---- Z scalars (setup.go:203-214), bit-reversed (:265), on the device ---------------------------------------
    DevBuf d_z;
    B2_CUDA(ctx, d_z.alloc(n * sizeof(Fr)));
    B2_TRY(gen_powers(ctx, d_z.as<Fr>(), S.tau, S.z_mult, n, d->log_n ? (int)d->log_n : 0));

    // ---- the resident key: this shard's ranges of [A]1 [B]1 [K]1 [Z]1 [B]2 -----------------------------------
    PkScalarSource src;
    src.a = A.data(); src.b = B.data(); src.k = pkK.data(); src.z_dev = d_z.as<Fr>();
    memcpy(&src.g1, d->g1_gen, sizeof src.g1);
    memcpy(&src.g2, d->g2_gen, sizeof src.g2);
    src.out_a = out->g1_a; src.out_b = out->g1_b; src.out_k = out->g1_k; src.out_z = out->g1_z; src.out_g2b = out->g2_b;
    b200_groth16_pk_desc pd;
    memset(&pd, 0, sizeof pd);
    pd.log_n = d->log_n; pd.generator = d->generator; pd.coset_gen = d->coset_gen;
    pd.n_a = na; pd.n_b = nb; pd.n_g2b = nb; pd.n_k = pkK.size(); pd.n_z = n - 1;
    pd.infinity_a = S.infA.data(); pd.infinity_b = S.infB.data();
    pd.n_wires = nw; pd.nb_public = d->nb_public;
    pd.shard_index = shard_index; pd.shard_count = shard_count ? shard_count : 1;
    B2_TRY(groth16_pk_upload(ctx, &pd, pk_out, &src));
    d_z.release();

    // ---- verifying-key side and the Pedersen bases -----------------------------------------------------------------
    int st = B200_OK;
    if (vk_side) {
        std::vector<Fr> s1 = {alpha, beta, delta}, s2 = {beta, delta, gamma};
        std::vector<uint64_t> p1(3 * sizeof(G1Affine) / 8), p2(3 * sizeof(G2Affine) / 8);
        st = mul_points<G1Affine>(ctx, src.g1, s1, p1.data());
        if (st == B200_OK) st = mul_points<G2Affine>(ctx, src.g2, s2, p2.data());
        if (st == B200_OK) {
            memcpy(out->g1_alpha, p1.data(), sizeof(G1Affine));
            memcpy(out->g1_beta, p1.data() + sizeof(G1Affine) / 8, sizeof(G1Affine));
            memcpy(out->g1_delta, p1.data() + 2 * sizeof(G1Affine) / 8, sizeof(G1Affine));
            memcpy(out->g2_beta, p2.data(), sizeof(G2Affine));
            memcpy(out->g2_delta, p2.data() + sizeof(G2Affine) / 8, sizeof(G2Affine));
            memcpy(out->g2_gamma, p2.data() + 2 * sizeof(G2Affine) / 8, sizeof(G2Affine));
        }
    }
    if (st == B200_OK && vk_side) st = mul_points<G1Affine>(ctx, src.g1, vkK, out->vk_k);
    for (size_t j = 0; j < d->nb_commitments && st == B200_OK && vk_side; ++j)
        if (out->ck_k && out->ck_k[j]) st = mul_points<G1Affine>(ctx, src.g1, ckK[j], out->ck_k[j]);
    if (st != B200_OK) {
        delete *pk_out;
        *pk_out = nullptr;
        return st;
    }
    out->n_a = na; out->n_b = nb; out->n_k = pkK.size(); out->n_z = n - 1; out->n_vk_k = vkK.size();
    return B200_OK;
}

int groth16_setup(b200_ctx* ctx, const b200_groth16_setup_desc* d, b200_groth16_pk** pk_out, b200_groth16_setup_out* out) {
    SetupScalars S;
    B2_TRY(groth16_setup_scalars(ctx, d, S));
    return groth16_setup_points(ctx, d, S, d->shard_index, d->shard_count, pk_out, out, true);
}

}  // namespace b200
