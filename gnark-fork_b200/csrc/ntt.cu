// Fr NTT kernels (sm_100a) and the device-side fft.Domain.
// Reference semantics: gnark-crypto fft.Domain.FFT / FFTInverse as called from
// /root/reference/backend/groth16/bn254/prove.go:369-375,393 (replaces iciclegnark.NttOnDevice /
// INttOnDevice, /root/reference/backend/groth16/bn254/icicle/icicle.go:489-502).
#include <cstdlib>
#include <cuda.h>
#include "ntt_host.cuh"

namespace b200 {

// ---------------------------------------------------------------------------------------
// shared-memory tile: element i is split into its low and high 16-byte halves, kept in two
// planes so that a warp touching 32 consecutive elements issues conflict-free 128-bit accesses
// (plane 1 is shifted by 64 B = 16 banks).
// ---------------------------------------------------------------------------------------
struct SmemTile {
    uint4* lo;
    uint4* hi;
    __device__ __forceinline__ SmemTile(void* base, uint32_t tl) {
        lo = reinterpret_cast<uint4*>(base);
        hi = lo + ((size_t)1 << tl) + 4;
    }
    __device__ __forceinline__ Fr get(uint32_t i) const {
        const uint4 a = lo[i], b = hi[i];
        Fr f;
        f.l[0] = a.x; f.l[1] = a.y; f.l[2] = a.z; f.l[3] = a.w;
        f.l[4] = b.x; f.l[5] = b.y; f.l[6] = b.z; f.l[7] = b.w;
        return f;
    }
    __device__ __forceinline__ void put(uint32_t i, const Fr& f) {
        lo[i] = make_uint4(f.l[0], f.l[1], f.l[2], f.l[3]);
        hi[i] = make_uint4(f.l[4], f.l[5], f.l[6], f.l[7]);
    }
};
static inline size_t smem_tile_bytes(uint32_t tl) { return (((size_t)2 << tl) + 4) * sizeof(uint4); }

__device__ __forceinline__ Fr ld_fr(const Fr* p) {
    const uint4* q = reinterpret_cast<const uint4*>(p);
    const uint4 a = q[0], b = q[1];
    Fr f;
    f.l[0] = a.x; f.l[1] = a.y; f.l[2] = a.z; f.l[3] = a.w;
    f.l[4] = b.x; f.l[5] = b.y; f.l[6] = b.z; f.l[7] = b.w;
    return f;
}
__device__ __forceinline__ Fr ldg_fr(const Fr* p) {  // read-only path (tables)
    const uint4* q = reinterpret_cast<const uint4*>(p);
    const uint4 a = __ldg(q), b = __ldg(q + 1);
    Fr f;
    f.l[0] = a.x; f.l[1] = a.y; f.l[2] = a.z; f.l[3] = a.w;
    f.l[4] = b.x; f.l[5] = b.y; f.l[6] = b.z; f.l[7] = b.w;
    return f;
}
__device__ __forceinline__ void st_fr(Fr* p, const Fr& f) {
    uint4* q = reinterpret_cast<uint4*>(p);
    q[0] = make_uint4(f.l[0], f.l[1], f.l[2], f.l[3]);
    q[1] = make_uint4(f.l[4], f.l[5], f.l[6], f.l[7]);
}

// Pull mode of a cross phase (multi-GPU in one process, group.cu): the pass does not read a received buffer but loads
// straight from the ranks' send buffers through peer-mapped pointers (NVLink P2P loads; the same device for virtual
// ranks).  Buffer index bi = (source rank r << shift) | j is read from src[r] + off + j, off = this rank's chunk in r's
// send buffer: the all-to-all is fused into the first load of the cross pass, 2^(TL - lg) * 32 B contiguous per row.
struct NttPeerPull {
    const Fr* src[NTT_MAX_RANKS];
    uint32_t shift, off;
    int on;
};

struct NttPassArgs {
    const Fr* src;   // loads   (== dst for an in-place pass)
    Fr* dst;         // stores, at smap(buffer index)
    const Fr* tw;    // w^j (or w^-j for an inverse transform), j < n/2 of the GLOBAL domain
    const Fr* pre;   // optional scale applied on load  (first pass), indexed by global position
    const Fr* post;  // optional scale applied on store (last pass), indexed by global position
    NttPassGeom g;   // geometry over the buffer
    NttIndexMap gmap, smap;   // buffer index -> global position / store index (identity on one GPU)
    uint32_t glog_n, s_off;   // global transform size and stage offset (distributed phases)
    int has_post_scalar;
    Fr post_scalar;
    NttPeerPull pull;
};

constexpr int NTT_THREADS = 1024;

template <int DIT>
__global__ void __launch_bounds__(NTT_THREADS) ntt_pass_kernel(const NttPassArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SmemTile sm(smem_raw, a.g.TL);
    const NttPassGeom g = a.g;
    const uint32_t tile = blockIdx.x;
    const uint32_t tile_n = 1u << g.TL;

    for (uint32_t loc = threadIdx.x; loc < tile_n; loc += blockDim.x) {
        const uint32_t bi = ntt_gidx(g, tile, loc);
        Fr v = a.pull.on ? ld_fr(a.pull.src[bi >> a.pull.shift] + a.pull.off + (bi & ((1u << a.pull.shift) - 1u))) : ld_fr(a.src + bi);
        if (a.pre) v = v * ldg_fr(a.pre + ntt_map(a.gmap, bi));
        sm.put(loc, v);
    }
    __syncthreads();

    const uint32_t n_bf = tile_n >> 1;
    for (uint32_t k = 0; k < g.S; ++k) {
        const uint32_t bit = ntt_stage_bit(g, DIT, k);
        const uint32_t s = a.s_off + g.s0 + k;
        for (uint32_t bf = threadIdx.x; bf < n_bf; bf += blockDim.x) {
            const uint32_t i0 = ntt_bf_lower(bf, bit), i1 = i0 | (1u << bit);
            const uint32_t g0 = ntt_map(a.gmap, ntt_gidx(g, tile, i0));
            const uint32_t e = ntt_twiddle_exp(a.glog_n, DIT, s, g0);
            const Fr u = sm.get(i0), v = sm.get(i1);
            if (DIT) {
                const Fr t = (e == 0) ? v : v * ldg_fr(a.tw + e);
                sm.put(i0, u + t);
                sm.put(i1, u - t);
            } else {
                const Fr d = u - v;
                sm.put(i0, u + v);
                sm.put(i1, (e == 0) ? d : d * ldg_fr(a.tw + e));
            }
        }
        __syncthreads();
    }

    for (uint32_t loc = threadIdx.x; loc < tile_n; loc += blockDim.x) {
        const uint32_t bi = ntt_gidx(g, tile, loc);
        Fr v = sm.get(loc);
        if (a.post) v = v * ldg_fr(a.post + ntt_map(a.gmap, bi));
        if (a.has_post_scalar) v = v * a.post_scalar;
        st_fr(a.dst + ntt_map(a.smap, bi), v);
    }
}

// ---------------------------------------------------------------------------------------
// TMA-staged variant of the pass (north_star: "TMA-staged shared-memory butterflies").  The tile of a pass is a box
// of a tensor view of the vector -- [top][2^S coupled rows, stride 2^L elements][2^LB contiguous elements] for a
// strided pass, [rows of 64 elements] for the contiguous one -- so ONE cp.async.bulk.tensor (UTMALDG) brings it into
// shared memory, signalled on an mbarrier, and one UTMASTG writes it back: no per-element index arithmetic, no LSU
// instructions and no registers on the load / store path; the butterflies are the same radix-2 network on the dense
// tile.  Used for passes whose store index is the buffer index (all single-GPU passes and the in-place passes of the
// distributed local phases); option "ntt_tma" (b200_set_option), A/B numbers in DESIGN.md.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

struct DenseTile {
    uint4* base;
    __device__ __forceinline__ Fr get(uint32_t i) const {
        const uint4 a = base[2 * i], b = base[2 * i + 1];
        Fr f;
        f.l[0] = a.x; f.l[1] = a.y; f.l[2] = a.z; f.l[3] = a.w;
        f.l[4] = b.x; f.l[5] = b.y; f.l[6] = b.z; f.l[7] = b.w;
        return f;
    }
    __device__ __forceinline__ void put(uint32_t i, const Fr& f) {
        base[2 * i] = make_uint4(f.l[0], f.l[1], f.l[2], f.l[3]);
        base[2 * i + 1] = make_uint4(f.l[4], f.l[5], f.l[6], f.l[7]);
    }
};

template <int DIT>
__global__ void __launch_bounds__(NTT_THREADS) ntt_pass_tma_kernel(const __grid_constant__ CUtensorMap tmap, const NttPassArgs a,
                                                                   const int contiguous) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t mbar;
    DenseTile sm{reinterpret_cast<uint4*>(smem_raw)};
    const NttPassGeom g = a.g;
    const uint32_t tile = blockIdx.x;
    const uint32_t tile_n = 1u << g.TL;
    // box coordinates (u64 units along dim 0)
    int32_t c0, c1, c2;
    if (contiguous) { c0 = 0; c1 = (int32_t)(tile << (g.TL - 6u)); c2 = 0; }
    else {
        const uint32_t midbits = g.L - g.LB;
        c0 = (int32_t)((tile & ((1u << midbits) - 1u)) << (g.LB + 2u)); c1 = 0; c2 = (int32_t)(tile >> midbits);
    }
    const uint32_t bar = smem_u32(&mbar), dst = smem_u32(smem_raw);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(tile_n * 32u) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                     ::"r"(dst), "l"(reinterpret_cast<uint64_t>(&tmap)), "r"(bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
    }
    __syncthreads();     // the barrier is initialised for everybody
    {
        uint32_t done = 0;
        while (!done) {
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(bar) : "memory");
        }
    }
    if (a.pre) {         // coset scaling fused in the first pass: one sweep over the tile as it sits in shared memory
        for (uint32_t loc = threadIdx.x; loc < tile_n; loc += blockDim.x)
            sm.put(loc, sm.get(loc) * ldg_fr(a.pre + ntt_map(a.gmap, ntt_gidx(g, tile, loc))));
        __syncthreads();
    }
    const uint32_t n_bf = tile_n >> 1;
    for (uint32_t k = 0; k < g.S; ++k) {
        const uint32_t bit = ntt_stage_bit(g, DIT, k);
        const uint32_t s = a.s_off + g.s0 + k;
        const bool last = (k + 1 == g.S) && (a.post || a.has_post_scalar);
        for (uint32_t bf = threadIdx.x; bf < n_bf; bf += blockDim.x) {
            const uint32_t i0 = ntt_bf_lower(bf, bit), i1 = i0 | (1u << bit);
            const uint32_t b0 = ntt_gidx(g, tile, i0);
            const uint32_t g0 = ntt_map(a.gmap, b0);
            const uint32_t e = ntt_twiddle_exp(a.glog_n, DIT, s, g0);
            const Fr u = sm.get(i0), v = sm.get(i1);
            Fr x, y;
            if (DIT) {
                const Fr t = (e == 0) ? v : v * ldg_fr(a.tw + e);
                x = u + t; y = u - t;
            } else {
                const Fr d = u - v;
                x = u + v; y = (e == 0) ? d : d * ldg_fr(a.tw + e);
            }
            if (last) {      // scaling fused in the last stage of the last pass (1/n, coset, den tables)
                if (a.post) {
                    x = x * ldg_fr(a.post + g0);
                    y = y * ldg_fr(a.post + ntt_map(a.gmap, ntt_gidx(g, tile, i1)));
                }
                if (a.has_post_scalar) { x = x * a.post_scalar; y = y * a.post_scalar; }
            }
            sm.put(i0, x);
            sm.put(i1, y);
        }
        __syncthreads();
    }
    // generic-proxy writes to shared memory must be visible to the async proxy before the bulk store reads them
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
                     ::"l"(reinterpret_cast<uint64_t>(&tmap)), "r"(dst), "r"(c0), "r"(c1), "r"(c2) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");     // the tile must outlive the read of the store
    }
}

// out[pos(k)] = mult * base^k for k < count; pos = k or bitrev_{log_n}(k).  Each thread owns a
// run of 16 consecutive exponents: one square-and-multiply, then 15 products.
constexpr int POW_RUN = 16;
__global__ void powers_kernel(Fr* out, const Fr base, const Fr mult, uint32_t count, int bitrev_log_n) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t k0 = (uint64_t)t * POW_RUN;
    if (k0 >= count) return;
    Fr x = base.pow_u64(k0) * mult;
    for (int j = 0; j < POW_RUN && k0 + j < count; ++j) {
        const uint32_t k = (uint32_t)(k0 + j);
        const uint32_t pos = bitrev_log_n ? (__brev(k) >> (32 - bitrev_log_n)) : k;
        st_fr(out + pos, x);
        x = x * base;
    }
}

// dst[smap(i)] = src[i]
__global__ void permute_kernel(const Fr* __restrict__ src, Fr* __restrict__ dst, const NttIndexMap smap, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    st_fr(dst + ntt_map(smap, i), ld_fr(src + i));
}

// PACK fused with its all-to-all (multi-GPU in one process): element i of this rank's natural-order block goes to rank
// q = i mod R, slot (rank << m) | (i >> lg) of q's buffer -- written straight into the peer's memory.  blockIdx.y = q;
// consecutive threads take consecutive slots: 1 KiB contiguous peer stores per warp, the strided local reads of the R
// blocks of one row share their cache lines in L2.
struct NttPeerPush {
    Fr* dst[NTT_MAX_RANKS];
};
__global__ void pack_push_kernel(const Fr* __restrict__ src, const NttPeerPush peers, uint32_t lg, uint32_t m, uint32_t rank) {
    const uint32_t off = blockIdx.x * blockDim.x + threadIdx.x, q = blockIdx.y;
    if (off >> m) return;
    st_fr(peers.dst[q] + ((size_t)rank << m) + off, ld_fr(src + (((size_t)off << lg) | q)));
}

__global__ void mul_sub_kernel(Fr* a, const Fr* b, const Fr* c, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    st_fr(a + i, ld_fr(a + i) * ld_fr(b + i) - ld_fr(c + i));
}

// ---------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------
int gen_powers(b200_ctx* ctx, Fr* out, const Fr& base, const Fr& mult, size_t count, int bitrev_log_n) {
    if (count == 0) return B200_OK;
    const size_t threads = (count + POW_RUN - 1) / POW_RUN;
    const int bs = 128;
    powers_kernel<<<(unsigned)((threads + bs - 1) / bs), bs, 0, ctx->stream>>>(out, base, mult, (uint32_t)count, bitrev_log_n);
    B2_CUDA(ctx, cudaGetLastError());
    return B200_OK;
}

int ntt_domain_get(b200_ctx* ctx, unsigned log_n, const Fr& gen, const Fr& coset, NttDomain** out) {
    if (log_n > 27) return ctx->fail(B200_ERR_UNSUPPORTED, "NTT size above 2^27");
    for (NttDomain* d : ctx->domains)
        if (d->log_n == log_n && d->gen == gen && d->coset == coset) { *out = d; return B200_OK; }
    NttDomain* d = new NttDomain();
    d->log_n = log_n;
    d->gen = gen;
    d->coset = coset;
    d->gen_inv = gen.inv();
    d->coset_inv = coset.inv();
    d->ninv = Fr::from_u32(1).dbl().pow_u64(log_n).inv();  // (2^log_n)^-1
    d->den = (coset.pow_u64((uint64_t)1 << log_n) - Fr::one()).inv();
    const size_t half = log_n ? ((size_t)1 << (log_n - 1)) : 1;
    cudaError_t e = d->tw_fwd.alloc(half * sizeof(Fr));
    if (e == cudaSuccess) e = d->tw_inv.alloc(half * sizeof(Fr));
    if (e != cudaSuccess) { delete d; return ctx->fail(B200_ERR_NOMEM, "twiddle tables", cudaGetErrorString(e)); }
    int st = gen_powers(ctx, d->tw_fwd.as<Fr>(), d->gen, Fr::one(), half, 0);
    if (st == B200_OK) st = gen_powers(ctx, d->tw_inv.as<Fr>(), d->gen_inv, Fr::one(), half, 0);
    if (st != B200_OK) { delete d; return st; }
    ctx->domains.push_back(d);
    *out = d;
    return B200_OK;
}

int ntt_domain_table(b200_ctx* ctx, NttDomain* d, NttDomain::Table t, const Fr** out) {
    if (!d->tab[t].p) {
        const size_t n = d->n();
        cudaError_t e = d->tab[t].alloc(n * sizeof(Fr));
        if (e != cudaSuccess) return ctx->fail(B200_ERR_NOMEM, "coset table", cudaGetErrorString(e));
        Fr base, mult = Fr::one();
        int br = 0;
        switch (t) {
            case NttDomain::COSET_NAT: base = d->coset; break;
            case NttDomain::COSET_BR: base = d->coset; br = 1; break;
            case NttDomain::COSETINV_NAT: base = d->coset_inv; mult = d->ninv; break;
            case NttDomain::COSETINV_BR: base = d->coset_inv; mult = d->ninv; br = 1; break;
            case NttDomain::H_BC: base = d->coset; mult = d->ninv; br = 1; break;
            case NttDomain::H_A: base = d->coset; mult = d->ninv * d->den; br = 1; break;
            default: return ctx->fail(B200_ERR_ARG, "bad table id");
        }
        int st = gen_powers(ctx, d->tab[t].as<Fr>(), base, mult, n, (br && d->log_n) ? (int)d->log_n : 0);
        if (st != B200_OK) { d->tab[t].release(); return st; }
    }
    *out = d->tab[t].as<Fr>();
    return B200_OK;
}

size_t ntt_domain_trim(b200_ctx* ctx) {
    size_t freed = 0;
    std::vector<NttDomain*> keep;
    for (NttDomain* d : ctx->domains) {
        if (d->refs > 0) { keep.push_back(d); continue; }
        freed += d->tw_fwd.bytes + d->tw_inv.bytes;
        for (auto& t : d->tab) freed += t.bytes;
        delete d;
    }
    ctx->domains.swap(keep);
    return freed;
}

void ntt_domain_free_all(b200_ctx* ctx) {
    for (NttDomain* d : ctx->domains) delete d;
    ctx->domains.clear();
}

// ---- host side of the TMA variant ------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn tma_encoder() {     // the driver entry point through the runtime: no link-time dependency on libcuda
    static EncodeTiledFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) p = nullptr;
        cudaGetLastError();
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}

// can this pass run as one tensor box per tile?  (box dimensions are at most 256 elements each)
static bool ntt_tma_ok(const NttPassArgs& a, size_t n_buf, int* contiguous) {
    const NttPassGeom& g = a.g;
    if (a.pull.on || a.src != a.dst) return false;
    if (a.smap.m != 31u || a.smap.sh_hi || a.smap.sh_lo || a.smap.add) return false;     // store index = buffer index
    if (n_buf < ((size_t)1 << g.TL) || g.TL < 6 || g.TL > 12) return false;
    *contiguous = (g.L == 0 && g.LB == 0) ? 1 : 0;
    if (*contiguous) return g.S == g.TL && g.TL - 6u <= 8u && g.log_n >= g.TL;
    return g.LB <= 6 && g.S <= 8 && g.L >= g.LB && g.L + g.S <= g.log_n;
}

static int ntt_tma_map(b200_ctx* ctx, const NttPassArgs& a, size_t n_buf, int contiguous, CUtensorMap* out) {
    EncodeTiledFn enc = tma_encoder();
    if (!enc) return ctx->fail(B200_ERR_UNSUPPORTED, "cuTensorMapEncodeTiled is not available in this driver");
    const NttPassGeom& g = a.g;
    cuuint64_t dim[3], stride[2];
    cuuint32_t box[3], est[3] = {1, 1, 1};
    if (contiguous) {          // rows of 64 elements (256 u64); a tile is 2^(TL-6) consecutive rows
        dim[0] = 256; dim[1] = n_buf >> 6; dim[2] = 1;
        stride[0] = 2048; stride[1] = (cuuint64_t)n_buf * 32;
        box[0] = 256; box[1] = 1u << (g.TL - 6u); box[2] = 1;
    } else {                   // [top][2^S rows, stride 2^L elements][2^L elements]; a tile takes 2^LB of the inner ones
        dim[0] = (cuuint64_t)4 << g.L; dim[1] = (cuuint64_t)1 << g.S; dim[2] = n_buf >> (g.L + g.S);
        stride[0] = (cuuint64_t)32 << g.L; stride[1] = (cuuint64_t)32 << (g.L + g.S);
        box[0] = 4u << g.LB; box[1] = 1u << g.S; box[2] = 1;
    }
    const CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_UINT64, 3, const_cast<Fr*>(a.src), dim, stride, box, est, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char msg[64];
        snprintf(msg, sizeof msg, "CUresult %d", (int)r);
        return ctx->fail(B200_ERR_CUDA, "cuTensorMapEncodeTiled", msg);
    }
    return B200_OK;
}

static int ntt_launch_pass(b200_ctx* ctx, const NttPassArgs& a, int dit, size_t n_buf) {
    const uint32_t tiles = (uint32_t)(n_buf >> a.g.TL);
    uint32_t threads = 1u << (a.g.TL > 1 ? a.g.TL - 1 : 0);
    if (threads > (uint32_t)ctx->ntt_threads) threads = (uint32_t)ctx->ntt_threads;
    if (threads > NTT_THREADS) threads = NTT_THREADS;
    if (threads < 32) threads = 32;
    const size_t smem = smem_tile_bytes(a.g.TL);
    int contiguous = 0;
    CUtensorMap tmap;
    const bool tma = ctx->ntt_tma && tma_encoder() && ntt_tma_ok(a, n_buf, &contiguous);   // else the per-thread-load pass (same GPU path)
    if (tma) B2_TRY(ntt_tma_map(ctx, a, n_buf, contiguous, &tmap));
    const int span = ctx->prof_begin(B200_PROF_NTT_PASS, (double)n_buf * a.g.S * 0.5);
    if (tma) {
        const size_t dense = (size_t)32 << a.g.TL;
        if (dit) ntt_pass_tma_kernel<1><<<tiles, threads, dense, ctx->stream>>>(tmap, a, contiguous);
        else ntt_pass_tma_kernel<0><<<tiles, threads, dense, ctx->stream>>>(tmap, a, contiguous);
    } else if (dit) ntt_pass_kernel<1><<<tiles, threads, smem, ctx->stream>>>(a);
    else ntt_pass_kernel<0><<<tiles, threads, smem, ctx->stream>>>(a);
    ctx->prof_end(span);
    ctx->n_launches += 1;
    B2_CUDA(ctx, cudaGetLastError());
    return B200_OK;
}

static int ntt_prepare(b200_ctx* ctx) {
    B2_CUDA(ctx, cudaFuncSetAttribute(ntt_pass_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_tile_bytes(12)));
    B2_CUDA(ctx, cudaFuncSetAttribute(ntt_pass_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_tile_bytes(12)));
    B2_CUDA(ctx, cudaFuncSetAttribute(ntt_pass_tma_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32 << 12));
    B2_CUDA(ctx, cudaFuncSetAttribute(ntt_pass_tma_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32 << 12));
    return B200_OK;
}

int ntt_run(b200_ctx* ctx, NttDomain* d, Fr* data, int inverse, int dit, const Fr* pre, const Fr* post,
            const Fr* post_scalar) {
    B2_TRY(ntt_prepare(ctx));
    if (d->log_n == 0) {  // n = 1: every variant is the identity (1/n = 1, g^0 = 1)
        return B200_OK;
    }
    const uint32_t tl_max = (uint32_t)ctx->ntt_tile_log;
    const uint32_t s_max = (uint32_t)ctx->ntt_strided_stages;
    NttPassGeom plan[8];
    const int np = ntt_plan(d->log_n, dit, tl_max, s_max < tl_max ? s_max : tl_max, plan);
    for (int i = 0; i < np; ++i) {
        NttPassArgs a;
        a.src = data;
        a.dst = data;
        a.tw = inverse ? d->tw_inv.as<Fr>() : d->tw_fwd.as<Fr>();
        a.pre = (i == 0) ? pre : nullptr;
        a.post = (i == np - 1) ? post : nullptr;
        a.g = plan[i];
        a.gmap = ntt_map_identity();
        a.smap = ntt_map_identity();
        a.glog_n = d->log_n;
        a.s_off = 0;
        a.has_post_scalar = (i == np - 1 && post_scalar) ? 1 : 0;
        a.post_scalar = post_scalar ? *post_scalar : Fr::one();
        a.pull.on = 0;
        B2_TRY(ntt_launch_pass(ctx, a, dit, d->n()));
    }
    return B200_OK;
}

// One phase of a transform distributed over 2^lg ranks (ntt.cuh).  `data` and `tmp` are this
// rank's two buffers of n >> lg elements; see NttDistPhase for which one a phase reads and writes.
int ntt_dist_phase(b200_ctx* ctx, NttDomain* d, Fr* data, Fr* tmp, uint32_t lg, uint32_t rank, int kind, int inverse,
                   const Fr* post, const Fr* post_scalar, const Fr* const* peer_src) {
    B2_TRY(ntt_prepare(ctx));
    if (peer_src && (kind != NTT_DIST_DIF_CROSS && kind != NTT_DIST_DIT_CROSS)) return ctx->fail(B200_ERR_ARG, "peer pull applies to a cross phase");
    if (peer_src && lg > 4) return ctx->fail(B200_ERR_UNSUPPORTED, "more than 16 ranks");
    if (lg == 0 || 2 * lg > d->log_n) return ctx->fail(B200_ERR_ARG, "distributed NTT needs 1 <= log2(ranks) <= log_n / 2");
    if (rank >> lg) return ctx->fail(B200_ERR_ARG, "rank out of range");
    const uint32_t log_loc = d->log_n - lg;
    const size_t n_loc = (size_t)1 << log_loc;
    const NttDistPhase ph = ntt_dist_phase_desc(d->log_n, lg, rank, kind);
    if (kind == NTT_DIST_PACK) {
        const int bs = 256;
        permute_kernel<<<(unsigned)((n_loc + bs - 1) / bs), bs, 0, ctx->stream>>>(data, tmp, ph.smap, (uint32_t)n_loc);
        ctx->n_launches += 1;
        B2_CUDA(ctx, cudaGetLastError());
        return B200_OK;
    }
    const uint32_t tl_max = (uint32_t)ctx->ntt_tile_log;
    const uint32_t s_max = (uint32_t)ctx->ntt_strided_stages;
    NttPassGeom plan[8];
    int np;
    if (ph.cross) {
        if (lg > tl_max) return ctx->fail(B200_ERR_ARG, "more ranks than a tile holds");
        plan[0] = ntt_dist_cross_geom(d->log_n, lg, ph.dit, tl_max);
        np = 1;
    } else {
        np = ntt_plan(log_loc, ph.dit, tl_max, s_max < tl_max ? s_max : tl_max, plan);
    }
    for (int i = 0; i < np; ++i) {
        const bool last = i == np - 1;
        NttPassArgs a;
        // DIF cross reads the received buffer (tmp) and writes data; DIT local writes its last pass to tmp
        a.src = (kind == NTT_DIST_DIF_CROSS) ? tmp : data;
        a.dst = (kind == NTT_DIST_DIT_LOCAL && last) ? tmp : data;
        a.tw = inverse ? d->tw_inv.as<Fr>() : d->tw_fwd.as<Fr>();
        a.pre = nullptr;
        a.post = last ? post : nullptr;
        a.g = plan[i];
        a.gmap = ph.gmap;
        a.smap = last ? ph.smap : ntt_map_identity();
        a.glog_n = d->log_n;
        a.s_off = ph.s_off;
        a.has_post_scalar = (last && post_scalar) ? 1 : 0;
        a.post_scalar = post_scalar ? *post_scalar : Fr::one();
        a.pull.on = 0;
        if (peer_src) {     // cross phase fused with its exchange: load from the ranks' send buffers, store to `data`
            a.pull.on = 1;
            a.pull.shift = d->log_n - 2u * lg;
            a.pull.off = rank << a.pull.shift;
            for (uint32_t r = 0; r < (1u << lg); ++r) a.pull.src[r] = peer_src[r];
            a.src = nullptr;
            a.dst = data;
        }
        B2_TRY(ntt_launch_pass(ctx, a, ph.dit, n_loc));
    }
    return B200_OK;
}

int ntt_dist_pack_push(b200_ctx* ctx, NttDomain* d, const Fr* block, Fr* const* peer_dst, uint32_t lg, uint32_t rank) {
    if (lg == 0 || lg > 4 || 2 * lg > d->log_n) return ctx->fail(B200_ERR_ARG, "distributed NTT needs 1 <= log2(ranks) <= min(4, log_n / 2)");
    const uint32_t m = d->log_n - 2u * lg;
    NttPeerPush peers;
    for (uint32_t r = 0; r < (1u << lg); ++r) peers.dst[r] = peer_dst[r];
    const uint32_t bs = 256, chunk = 1u << m;
    pack_push_kernel<<<dim3((chunk + bs - 1) / bs, 1u << lg), bs, 0, ctx->stream>>>(block, peers, lg, m, rank);
    ctx->n_launches += 1;
    B2_CUDA(ctx, cudaGetLastError());
    return B200_OK;
}

int ntt_gnark(b200_ctx* ctx, NttDomain* d, Fr* data, int inverse, int dit, int on_coset) {
    const Fr* pre = nullptr;
    const Fr* post = nullptr;
    const Fr* scalar = nullptr;
    if (!inverse) {
        if (on_coset) B2_TRY(ntt_domain_table(ctx, d, dit ? NttDomain::COSET_BR : NttDomain::COSET_NAT, &pre));
    } else {
        if (on_coset) B2_TRY(ntt_domain_table(ctx, d, dit ? NttDomain::COSETINV_NAT : NttDomain::COSETINV_BR, &post));
        else scalar = &d->ninv;
    }
    return ntt_run(ctx, d, data, inverse, dit, pre, post, scalar);
}

int ntt_mul_sub(b200_ctx* ctx, Fr* a, const Fr* b, const Fr* c, size_t n) {
    if (n == 0) return B200_OK;
    const int bs = 256;
    const int span = ctx->prof_begin(B200_PROF_POINTWISE, (double)n);
    mul_sub_kernel<<<(unsigned)((n + bs - 1) / bs), bs, 0, ctx->stream>>>(a, b, c, n);
    ctx->prof_end(span);
    ctx->n_launches += 1;
    B2_CUDA(ctx, cudaGetLastError());
    return B200_OK;
}

// computeH (prove.go:353-396) with the scalings folded:
//   x <- FFTInverse(x, DIF)            (prove.go:369-371)  1/n   } one table product on the last
//   x <- FFT(x, DIT, OnCoset)          (prove.go:373-375)  g^br(i)} store of the inverse transform
//   a <- (a*b - c) * den               (prove.go:377-390)  den folded into a's and c's table
//   a <- FFTInverse(a, DIF, OnCoset)   (prove.go:393)      g^-br(i)/n on the last store
// Field elements are canonical (fully reduced) so re-associating the products is bit-exact.
int ntt_compute_h(b200_ctx* ctx, NttDomain* d, Fr* a, Fr* b, Fr* c) {
    const Fr *ha, *hbc, *cinv;
    B2_TRY(ntt_domain_table(ctx, d, NttDomain::H_A, &ha));
    B2_TRY(ntt_domain_table(ctx, d, NttDomain::H_BC, &hbc));
    B2_TRY(ntt_domain_table(ctx, d, NttDomain::COSETINV_BR, &cinv));
    B2_TRY(ntt_run(ctx, d, a, 1, 0, nullptr, ha, nullptr));
    B2_TRY(ntt_run(ctx, d, b, 1, 0, nullptr, hbc, nullptr));
    B2_TRY(ntt_run(ctx, d, c, 1, 0, nullptr, ha, nullptr));  // (a*b - c)*den = (a*den)*b - c*den
    B2_TRY(ntt_run(ctx, d, a, 0, 1, nullptr, nullptr, nullptr));
    B2_TRY(ntt_run(ctx, d, b, 0, 1, nullptr, nullptr, nullptr));
    B2_TRY(ntt_run(ctx, d, c, 0, 1, nullptr, nullptr, nullptr));
    B2_TRY(ntt_mul_sub(ctx, a, b, c, d->n()));
    B2_TRY(ntt_run(ctx, d, a, 1, 0, nullptr, cinv, nullptr));
    return B200_OK;
}

}  // namespace b200
