// CPU replay of the device algorithms' index logic (TEST INFRASTRUCTURE).
// Compiled with g++ from the very headers the kernels use (ntt.cuh, msm.cuh, field.cuh, ec.cuh,
// host path), so the tile geometry, twiddle exponents, signed-digit recoding, reduction tree and
// group formulas are checked against oracle/ on a machine without a GPU.
#include <vector>
#include <cstring>
#include "../../gnark-fork_b200/csrc/ntt.cuh"
#include "../../gnark-fork_b200/csrc/msm.cuh"

using namespace b200;

extern "C" {

// sequential replay of ntt_pass_kernel over the plan ntt_plan() produces
void model_ntt(uint64_t* data, unsigned log_n, int inverse, int dit, unsigned tl_max, unsigned s_max,
               const uint64_t* gen_mont) {
    Fr* x = reinterpret_cast<Fr*>(data);
    const size_t n = (size_t)1 << log_n;
    Fr gen; memcpy(gen.l, gen_mont, 32);
    if (inverse) gen = gen.inv();
    std::vector<Fr> tw(n / 2 ? n / 2 : 1);
    tw[0] = Fr::one();
    for (size_t j = 1; j < tw.size(); ++j) tw[j] = tw[j - 1] * gen;
    if (log_n == 0) return;
    NttPassGeom plan[8];
    const int np = ntt_plan(log_n, dit, tl_max, s_max < tl_max ? s_max : tl_max, plan);
    for (int p = 0; p < np; ++p) {
        const NttPassGeom g = plan[p];
        const uint32_t tile_n = 1u << g.TL, tiles = (uint32_t)(n >> g.TL);
        std::vector<Fr> sm(tile_n);
        for (uint32_t tile = 0; tile < tiles; ++tile) {
            for (uint32_t loc = 0; loc < tile_n; ++loc) sm[loc] = x[ntt_gidx(g, tile, loc)];
            for (uint32_t k = 0; k < g.S; ++k) {
                const uint32_t bit = ntt_stage_bit(g, dit, k), s = g.s0 + k;
                for (uint32_t bf = 0; bf < tile_n / 2; ++bf) {
                    const uint32_t i0 = ntt_bf_lower(bf, bit), i1 = i0 | (1u << bit);
                    const uint32_t e = ntt_twiddle_exp(g.log_n, dit, s, ntt_gidx(g, tile, i0));
                    const Fr u = sm[i0], v = sm[i1];
                    if (dit) { const Fr t = v * tw[e]; sm[i0] = u + t; sm[i1] = u - t; }
                    else { sm[i0] = u + v; sm[i1] = (u - v) * tw[e]; }
                }
            }
            for (uint32_t loc = 0; loc < tile_n; ++loc) x[ntt_gidx(g, tile, loc)] = sm[loc];
        }
    }
}

// sequential replay of one phase of the distributed transform (ntt.cu: ntt_dist_phase + the index maps of
// ntt_pass_kernel).  data / tmp: this rank's two buffers of n >> lg elements; post: optional table
// of n elements indexed by global position, applied with the last store.
void model_ntt_dist_phase(uint64_t* data_, uint64_t* tmp_, unsigned log_n, unsigned lg, unsigned rank, int kind,
                          int inverse, unsigned tl_max, unsigned s_max, const uint64_t* gen_mont, const uint64_t* post_) {
    Fr* data = reinterpret_cast<Fr*>(data_);
    Fr* tmp = reinterpret_cast<Fr*>(tmp_);
    const Fr* post = reinterpret_cast<const Fr*>(post_);
    const size_t n = (size_t)1 << log_n;
    const uint32_t log_loc = log_n - lg;
    const size_t n_loc = (size_t)1 << log_loc;
    const NttDistPhase ph = ntt_dist_phase_desc(log_n, lg, rank, kind);
    if (kind == NTT_DIST_PACK) {
        for (uint32_t i = 0; i < n_loc; ++i) tmp[ntt_map(ph.smap, i)] = data[i];
        return;
    }
    Fr gen; memcpy(gen.l, gen_mont, 32);
    if (inverse) gen = gen.inv();
    std::vector<Fr> tw(n / 2 ? n / 2 : 1);
    tw[0] = Fr::one();
    for (size_t j = 1; j < tw.size(); ++j) tw[j] = tw[j - 1] * gen;
    NttPassGeom plan[8];
    int np;
    if (ph.cross) { plan[0] = ntt_dist_cross_geom(log_n, lg, ph.dit, tl_max); np = 1; }
    else np = ntt_plan(log_loc, ph.dit, tl_max, s_max < tl_max ? s_max : tl_max, plan);
    for (int p = 0; p < np; ++p) {
        const bool last = p == np - 1;
        const NttPassGeom g = plan[p];
        const Fr* src = (kind == NTT_DIST_DIF_CROSS) ? tmp : data;
        Fr* dst = (kind == NTT_DIST_DIT_LOCAL && last) ? tmp : data;
        const NttIndexMap smap = last ? ph.smap : ntt_map_identity();
        const uint32_t tile_n = 1u << g.TL, tiles = (uint32_t)(n_loc >> g.TL);
        std::vector<Fr> sm(tile_n);
        for (uint32_t tile = 0; tile < tiles; ++tile) {
            for (uint32_t loc = 0; loc < tile_n; ++loc) sm[loc] = src[ntt_gidx(g, tile, loc)];
            for (uint32_t k = 0; k < g.S; ++k) {
                const uint32_t bit = ntt_stage_bit(g, ph.dit, k), s = ph.s_off + g.s0 + k;
                for (uint32_t bf = 0; bf < tile_n / 2; ++bf) {
                    const uint32_t i0 = ntt_bf_lower(bf, bit), i1 = i0 | (1u << bit);
                    const uint32_t e = ntt_twiddle_exp(log_n, ph.dit, s, ntt_map(ph.gmap, ntt_gidx(g, tile, i0)));
                    const Fr u = sm[i0], v = sm[i1];
                    if (ph.dit) { const Fr t = v * tw[e]; sm[i0] = u + t; sm[i1] = u - t; }
                    else { sm[i0] = u + v; sm[i1] = (u - v) * tw[e]; }
                }
            }
            for (uint32_t loc = 0; loc < tile_n; ++loc) {
                const uint32_t bi = ntt_gidx(g, tile, loc);
                Fr v = sm[loc];
                if (last && post) v = v * post[ntt_map(ph.gmap, bi)];
                dst[ntt_map(smap, bi)] = v;
            }
        }
    }
}

void model_fr_mul_sub(uint64_t* a_, const uint64_t* b_, const uint64_t* c_, size_t n) {
    Fr* a = reinterpret_cast<Fr*>(a_);
    const Fr* b = reinterpret_cast<const Fr*>(b_);
    const Fr* c = reinterpret_cast<const Fr*>(c_);
    for (size_t i = 0; i < n; ++i) a[i] = a[i] * b[i] - c[i];
}

}  // extern "C"

// the whole MSM pipeline replayed on the host with the device formulas (XYZZ, reduction tree)
template <class F>
static void model_msm(const uint64_t* scalars, const uint64_t* bases, size_t n, int c_override, uint64_t* out_jac) {
    const Fr* sc = reinterpret_cast<const Fr*>(scalars);
    const Affine<F>* pts = reinterpret_cast<const Affine<F>*>(bases);
    Jac<F> res = Jac<F>::inf();
    if (n) {
        const MsmPlan pl = msm_plan(n, c_override);
        std::vector<int32_t> digits((size_t)pl.W * n);
        for (size_t i = 0; i < n; ++i) {
            const Fr s = sc[i].from_mont();
            msm_signed_digits(s.l, pl.c, pl.W, &digits[i], n);
        }
        const size_t M = (size_t)pl.W * pl.nb;
        std::vector<XYZZ<F>> buckets(M, XYZZ<F>::inf());
        for (int w = 0; w < pl.W; ++w)
            for (size_t i = 0; i < n; ++i) {
                const int32_t d = digits[(size_t)w * n + i];
                if (!d) continue;
                Affine<F> p = pts[i];
                if (d < 0) p.y = p.y.neg();
                buckets[(size_t)w * pl.nb + (size_t)((d < 0 ? -d : d) - 1)].add_mixed(p);
            }
        const uint32_t T = pl.nb / pl.L0;
        std::vector<XYZZ<F>> S((size_t)pl.W * T), Wt((size_t)pl.W * T);
        for (size_t t = 0; t < S.size(); ++t) {
            XYZZ<F> run = XYZZ<F>::inf(), acc = XYZZ<F>::inf();
            for (int j = (int)pl.L0 - 1; j >= 0; --j) { run.add(buckets[t * pl.L0 + j]); acc.add(run); }
            S[t] = run; Wt[t] = acc;
        }
        // upper part exactly as msm_tree_* + the host Horner of msm_run_impl:
        //   sum_k k B_k = V_0 + L0 * sum_b 2^b V_{1+b}
        uint32_t logT = 0; while ((1u << logT) < T) ++logT;
        const uint32_t Q = 1 + logT;
        uint32_t logL0 = 0; while ((1u << logL0) < pl.L0) ++logL0;
        XYZZ<F> acc = XYZZ<F>::inf();
        for (int w = pl.WB - 1; w >= 0; --w) {
            for (int k = 0; k < pl.c; ++k) acc = acc.dbl();
            std::vector<XYZZ<F>> v(Q, XYZZ<F>::inf());
            for (uint32_t t = 0; t < T; ++t) {
                v[0].add(Wt[(size_t)w * T + t]);
                for (uint32_t bit = 0; bit < logT; ++bit) if ((t >> bit) & 1u) v[1 + bit].add(S[(size_t)w * T + t]);
            }
            XYZZ<F> hi = XYZZ<F>::inf();
            for (int bit = (int)Q - 2; bit >= 0; --bit) { hi = hi.dbl(); hi.add(v[1 + bit]); }
            for (uint32_t k = 0; k < logL0; ++k) hi = hi.dbl();
            hi.add(v[0]);
            acc.add(hi);
        }
        res = acc.to_jac();
    }
    memcpy(out_jac, &res, sizeof res);
}
extern "C" {
void model_msm_g1(const uint64_t* s, const uint64_t* b, size_t n, int c, uint64_t* out) { model_msm<Fp>(s, b, n, c, out); }
void model_msm_g2(const uint64_t* s, const uint64_t* b, size_t n, int c, uint64_t* out) { model_msm<Fp2>(s, b, n, c, out); }

void model_fr_mul(const uint64_t* a, const uint64_t* b, uint64_t* o, size_t n) {
    for (size_t i = 0; i < n; ++i) reinterpret_cast<Fr*>(o)[i] = reinterpret_cast<const Fr*>(a)[i] * reinterpret_cast<const Fr*>(b)[i];
}
void model_fr_inv(const uint64_t* a, uint64_t* o, size_t n) {
    for (size_t i = 0; i < n; ++i) reinterpret_cast<Fr*>(o)[i] = reinterpret_cast<const Fr*>(a)[i].inv();
}
void model_fp_inv(const uint64_t* a, uint64_t* o, size_t n) {
    for (size_t i = 0; i < n; ++i) reinterpret_cast<Fp*>(o)[i] = reinterpret_cast<const Fp*>(a)[i].inv();
}
void model_fp_mul(const uint64_t* a, const uint64_t* b, uint64_t* o, size_t n) {
    for (size_t i = 0; i < n; ++i) reinterpret_cast<Fp*>(o)[i] = reinterpret_cast<const Fp*>(a)[i] * reinterpret_cast<const Fp*>(b)[i];
}
}
