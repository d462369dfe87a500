// CPU oracle (TEST INFRASTRUCTURE ONLY) -- C++ restatement of the BN254 hot path.
//
// Restates, from their published algorithms, the gnark-crypto routines that
// /root/reference/backend/groth16/bn254/prove.go calls (gnark-crypto
// v0.12.2-0.20231117165148-e77308824822, /root/reference/go.mod:10, NOT vendored):
//   fr/fp Montgomery arithmetic (4 x u64 limbs, R = 2^256)
//   fft.Domain.FFT / FFTInverse with DIF/DIT and OnCoset   (prove.go:369-375,393)
//   computeH                                               (prove.go:353-396)
//   G1Jac.MultiExp / G2Jac.MultiExp, signed-digit buckets  (prove.go:201,214,234,244,290)
//   BatchScalarMultiplicationG1/G2 of a fixed base          (setup.go:251,308)
// The structural witness for the FFT recursion that IS in the tree is
// /root/reference/backend/groth16/bn254/mpcsetup/lagrange.go:143-179.
//
// PARITY UNPINNED: the reference has no golden vectors for this path and cannot be built
// here (no Go). This file is cross-checked against oracle/*.py (big-int definition) in
// tests/test_oracle_*.py. It is the bit-exact checker for mid sizes and the timed
// "CPU restatement, not gnark-crypto (no amd64 asm)" baseline. Product code never links it.
//
// Build: see oracle/Makefile (g++ -O3 -fopenmp -shared -fPIC).
#include <cstdint>
#include <cstring>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include <omp.h>

typedef uint64_t u64;
typedef unsigned __int128 u128;

// ------------------------------------------------------------------------------------
// Montgomery field, N x u64 limbs (N = 4: BN254 Fp/Fr and BLS12-377 Fr; N = 6: BLS12-377 Fp).
// Parameters derived at start-up from the modulus only.
// ------------------------------------------------------------------------------------
template <int N>
struct FieldParams {
    u64 m[N];     // modulus
    u64 inv;      // -m^-1 mod 2^64
    u64 one[N];   // R mod m
    u64 r2[N];    // R^2 mod m
};

template <int N> static inline bool geq(const u64* a, const u64* b) {
    for (int i = N - 1; i >= 0; --i) {
        if (a[i] != b[i]) return a[i] > b[i];
    }
    return true;
}
template <int N> static inline u64 addn(u64* r, const u64* a, const u64* b) {
    u128 c = 0;
    for (int i = 0; i < N; ++i) { c += (u128)a[i] + b[i]; r[i] = (u64)c; c >>= 64; }
    return (u64)c;
}
template <int N> static inline u64 subn(u64* r, const u64* a, const u64* b) {
    u64 br = 0;
    for (int i = 0; i < N; ++i) {
        u128 d = (u128)a[i] - b[i] - br;
        r[i] = (u64)d; br = (u64)(d >> 64) & 1;
    }
    return br;
}

template <int N, const FieldParams<N>& PRM>
struct Fe {
    u64 v[N];
    static constexpr int LIMBS = N;
    static Fe zero() { Fe r; memset(r.v, 0, sizeof r.v); return r; }
    static Fe one() { Fe r; memcpy(r.v, PRM.one, sizeof r.v); return r; }
    bool is_zero() const { u64 o = 0; for (int i = 0; i < N; ++i) o |= v[i]; return o == 0; }
    bool operator==(const Fe& o) const { return memcmp(v, o.v, sizeof v) == 0; }
    Fe operator+(const Fe& o) const {
        Fe r; u64 c = addn<N>(r.v, v, o.v);
        if (c || geq<N>(r.v, PRM.m)) subn<N>(r.v, r.v, PRM.m);
        return r;
    }
    Fe operator-(const Fe& o) const {
        Fe r; if (subn<N>(r.v, v, o.v)) addn<N>(r.v, r.v, PRM.m);
        return r;
    }
    Fe neg() const { if (is_zero()) return *this; Fe r; subn<N>(r.v, PRM.m, v); return r; }
    Fe dbl() const { return *this + *this; }
    // CIOS Montgomery product
    Fe operator*(const Fe& o) const {
        u64 t[N + 2];
        memset(t, 0, sizeof t);
        for (int i = 0; i < N; ++i) {
            u128 c = 0;
            for (int j = 0; j < N; ++j) {
                c += (u128)v[j] * o.v[i] + t[j];
                t[j] = (u64)c; c >>= 64;
            }
            c += t[N]; t[N] = (u64)c; t[N + 1] = (u64)(c >> 64);
            u64 q = t[0] * PRM.inv;
            c = (u128)q * PRM.m[0] + t[0]; c >>= 64;
            for (int j = 1; j < N; ++j) {
                c += (u128)q * PRM.m[j] + t[j];
                t[j - 1] = (u64)c; c >>= 64;
            }
            c += t[N]; t[N - 1] = (u64)c; t[N] = t[N + 1] + (u64)(c >> 64);
        }
        Fe r; memcpy(r.v, t, sizeof r.v);
        if (t[N] || geq<N>(r.v, PRM.m)) subn<N>(r.v, r.v, PRM.m);
        return r;
    }
    Fe sqr() const { return *this * *this; }
    Fe pow(const u64* e, int nlimbs) const {
        Fe acc = one();
        for (int i = nlimbs * 64 - 1; i >= 0; --i) {
            acc = acc.sqr();
            if ((e[i / 64] >> (i % 64)) & 1) acc = acc * *this;
        }
        return acc;
    }
    Fe inv() const {  // Fermat
        u64 e[N]; u64 two[N] = {2};
        subn<N>(e, PRM.m, two);
        return pow(e, N);
    }
    Fe to_mont() const { Fe r2; memcpy(r2.v, PRM.r2, sizeof r2.v); return *this * r2; }
    Fe from_mont() const { Fe o = zero(); o.v[0] = 1; return *this * o; }
    static Fe from_u64(u64 x) { Fe r = zero(); r.v[0] = x; return r.to_mont(); }
};

template <int N>
static void init_params(FieldParams<N>& p, const u64* m) {
    memcpy(p.m, m, sizeof p.m);
    u64 x = 1;  // Newton: x = m^-1 mod 2^64
    for (int i = 0; i < 6; ++i) x *= 2 - m[0] * x;
    p.inv = (u64)0 - x;
    // one = 2^(64 N) mod m by modular doublings of 1; r2 after twice as many
    u64 t[N] = {1};
    for (int i = 0; i < 128 * N; ++i) {
        u64 c = addn<N>(t, t, t);
        if (c || geq<N>(t, m)) subn<N>(t, t, m);
        if (i == 64 * N - 1) memcpy(p.one, t, sizeof p.one);
    }
    memcpy(p.r2, t, sizeof p.r2);
}

// Curve selection: the default build is BN254; -DORACLE_BLS12_377 builds liboracle377.so (config 5) from the same
// file -- p, r from /root/reference/std/math/emulated/emparams/emparams.go:123,137, Fp2 = Fp[u]/(u^2 + 5)
// (/root/reference/std/algebra/native/fields_bls12377/e12.go:79), as the reference generates both curves from one template.
#if defined(ORACLE_BLS12_377)
constexpr int FP_LIMBS = 6;
static const u64 MOD_P[6] = {0x8508c00000000001ULL, 0x170b5d4430000000ULL, 0x1ef3622fba094800ULL,
                             0x1a22d9f300f5138fULL, 0xc63b05c06ca1493bULL, 0x01ae3a4617c510eaULL};
static const u64 MOD_R[4] = {0x0a11800000000001ULL, 0x59aa76fed0000001ULL, 0x60b44d1e5c37b001ULL, 0x12ab655e9a2ca556ULL};
constexpr u64 FP2_NONRESIDUE_NEG = 5;   // u^2 = -5
constexpr int FR_BITS = 253;
#else
constexpr int FP_LIMBS = 4;
static const u64 MOD_P[4] = {0x3c208c16d87cfd47ULL, 0x97816a916871ca8dULL, 0xb85045b68181585dULL, 0x30644e72e131a029ULL};
static const u64 MOD_R[4] = {0x43e1f593f0000001ULL, 0x2833e84879b97091ULL, 0xb85045b68181585dULL, 0x30644e72e131a029ULL};
constexpr u64 FP2_NONRESIDUE_NEG = 1;   // u^2 = -1
constexpr int FR_BITS = 254;
#endif

static FieldParams<FP_LIMBS> FP;
static FieldParams<4> FR;

struct Init { Init() { init_params<FP_LIMBS>(FP, MOD_P); init_params<4>(FR, MOD_R); } };
static Init g_init;

typedef Fe<FP_LIMBS, FP> Fp;
typedef Fe<4, FR> Fr;

// Fp2 = Fp[u]/(u^2 + k), k = FP2_NONRESIDUE_NEG
struct Fp2 {
    Fp a0, a1;
    static Fp mul_k(const Fp& x) {   // k * x
        if (FP2_NONRESIDUE_NEG == 1) return x;
        Fp x2 = x.dbl(), x4 = x2.dbl();
        return x4 + x;               // k = 5
    }
    static Fp2 zero() { return {Fp::zero(), Fp::zero()}; }
    static Fp2 one() { return {Fp::one(), Fp::zero()}; }
    bool is_zero() const { return a0.is_zero() && a1.is_zero(); }
    bool operator==(const Fp2& o) const { return a0 == o.a0 && a1 == o.a1; }
    Fp2 operator+(const Fp2& o) const { return {a0 + o.a0, a1 + o.a1}; }
    Fp2 operator-(const Fp2& o) const { return {a0 - o.a0, a1 - o.a1}; }
    Fp2 neg() const { return {a0.neg(), a1.neg()}; }
    Fp2 dbl() const { return {a0.dbl(), a1.dbl()}; }
    Fp2 operator*(const Fp2& o) const {  // Karatsuba
        Fp t0 = a0 * o.a0, t1 = a1 * o.a1;
        Fp t2 = (a0 + a1) * (o.a0 + o.a1);
        return {t0 - mul_k(t1), t2 - t0 - t1};
    }
    Fp2 sqr() const { return *this * *this; }
    Fp2 inv() const {
        Fp d = (a0.sqr() + mul_k(a1.sqr())).inv();
        return {a0 * d, a1.neg() * d};
    }
};

// ------------------------------------------------------------------------------------
// Short Weierstrass y^2 = x^3 + b (a = 0), Jacobian coordinates, generic over the field.
// Affine infinity is encoded (0,0) like gnark-crypto's G1Affine/G2Affine; Jacobian Z = 0.
// ------------------------------------------------------------------------------------
template <class F>
struct Aff {
    F x, y;
    bool is_inf() const { return x.is_zero() && y.is_zero(); }
};

template <class F>
struct Jac {
    F X, Y, Z;
    static Jac inf() { return {F::one(), F::one(), F::zero()}; }
    bool is_inf() const { return Z.is_zero(); }
    static Jac from_aff(const Aff<F>& a) {
        if (a.is_inf()) return inf();
        return {a.x, a.y, F::one()};
    }
    Jac dbl() const {  // dbl-2009-l
        if (is_inf()) return *this;
        F A = X.sqr(), B = Y.sqr(), C = B.sqr();
        F D = ((X + B).sqr() - A - C).dbl();
        F E = A.dbl() + A, Fq = E.sqr();
        Jac r;
        r.X = Fq - D.dbl();
        r.Y = E * (D - r.X) - C.dbl().dbl().dbl();
        r.Z = (Y * Z).dbl();
        return r;
    }
    Jac add_mixed(const Aff<F>& q) const {  // madd-2007-bl
        if (q.is_inf()) return *this;
        if (is_inf()) return from_aff(q);
        F Z1Z1 = Z.sqr(), U2 = q.x * Z1Z1, S2 = q.y * Z * Z1Z1;
        F H = U2 - X, rr = (S2 - Y).dbl();
        if (H.is_zero()) {
            if (rr.is_zero()) return dbl();
            return inf();
        }
        F HH = H.sqr(), I = HH.dbl().dbl(), J = H * I, V = X * I;
        Jac r;
        r.X = rr.sqr() - J - V.dbl();
        r.Y = rr * (V - r.X) - (Y * J).dbl();
        r.Z = (Z + H).sqr() - Z1Z1 - HH;
        return r;
    }
    Jac add(const Jac& q) const {  // add-2007-bl
        if (q.is_inf()) return *this;
        if (is_inf()) return q;
        F Z1Z1 = Z.sqr(), Z2Z2 = q.Z.sqr();
        F U1 = X * Z2Z2, U2 = q.X * Z1Z1;
        F S1 = Y * q.Z * Z2Z2, S2 = q.Y * Z * Z1Z1;
        F H = U2 - U1, rr = (S2 - S1).dbl();
        if (H.is_zero()) {
            if (rr.is_zero()) return dbl();
            return inf();
        }
        F I = H.dbl().sqr(), J = H * I, V = U1 * I;
        Jac r;
        r.X = rr.sqr() - J - V.dbl();
        r.Y = rr * (V - r.X) - (S1 * J).dbl();
        r.Z = ((Z + q.Z).sqr() - Z1Z1 - Z2Z2) * H;
        return r;
    }
    Jac neg() const { return {X, Y.neg(), Z}; }
    Aff<F> to_aff() const {
        if (is_inf()) return {F::zero(), F::zero()};
        F zi = Z.inv(), zi2 = zi.sqr();
        return {X * zi2, Y * zi2 * zi};
    }
};

// ------------------------------------------------------------------------------------
// MultiExp: signed-digit bucket method (the algorithm family gnark-crypto's MultiExp uses;
// the result is compared after affine normalisation, so window size / schedule are free).
// ------------------------------------------------------------------------------------
static inline int msm_window(size_t n) {
    int c = 4;
    while (c < 16 && ((size_t)1 << (c + 3)) < n) ++c;
    return c;
}

template <class F>
static Jac<F> msm(const Fr* scalars_mont, const Aff<F>* bases, size_t n) {
    if (n == 0) return Jac<F>::inf();
    const int c = msm_window(n);
    const int W = (FR_BITS / c) + 1;  // signed digits: the carry of the top window is absorbed (r < 2^FR_BITS)
    // signed digits, digit[w*n + i]
    std::vector<int32_t> digits((size_t)W * n);
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < n; ++i) {
        Fr s = scalars_mont[i].from_mont();
        int carry = 0;
        for (int w = 0; w < W; ++w) {
            int bit = w * c;
            u64 lo = s.v[bit / 64] >> (bit % 64);
            if (bit % 64 + c > 64 && bit / 64 + 1 < 4) lo |= s.v[bit / 64 + 1] << (64 - bit % 64);
            int d = (int)(lo & (((u64)1 << c) - 1)) + carry;
            carry = 0;
            if (d > (1 << (c - 1))) { d -= (1 << c); carry = 1; }
            digits[(size_t)w * n + i] = d;
        }
        // r < 2^FR_BITS and W*c > FR_BITS bits => final carry is always absorbed
    }
    int nthreads = omp_get_max_threads();
    int chunks = std::max(1, (nthreads + W - 1) / W);
    if (n < 4096) chunks = 1;
    const size_t nb = (size_t)1 << (c - 1);
    std::vector<Jac<F>> partial((size_t)W * chunks);
#pragma omp parallel for schedule(dynamic, 1)
    for (int t = 0; t < W * chunks; ++t) {
        int w = t / chunks, ch = t % chunks;
        size_t lo = n * ch / chunks, hi = n * (ch + 1) / chunks;
        std::vector<Jac<F>> bucket(nb, Jac<F>::inf());
        const int32_t* dg = &digits[(size_t)w * n];
        for (size_t i = lo; i < hi; ++i) {
            int d = dg[i];
            if (d == 0) continue;
            if (d > 0) bucket[d - 1] = bucket[d - 1].add_mixed(bases[i]);
            else { Aff<F> m = {bases[i].x, bases[i].y.neg()}; if (bases[i].is_inf()) m = bases[i]; bucket[-d - 1] = bucket[-d - 1].add_mixed(m); }
        }
        Jac<F> run = Jac<F>::inf(), sum = Jac<F>::inf();
        for (size_t k = nb; k-- > 0;) { run = run.add(bucket[k]); sum = sum.add(run); }
        partial[t] = sum;
    }
    Jac<F> acc = Jac<F>::inf();
    for (int w = W - 1; w >= 0; --w) {
        for (int k = 0; k < c; ++k) acc = acc.dbl();
        for (int ch = 0; ch < chunks; ++ch) acc = acc.add(partial[(size_t)w * chunks + ch]);
    }
    return acc;
}

// Fixed-base batch scalar multiplication (BatchScalarMultiplicationG1/G2, setup.go:251,308):
// out[i] = scalars[i] * base, affine.
template <class F>
static void batch_mul(const Aff<F>& base, const Fr* scalars_mont, size_t n, Aff<F>* out) {
    const int c = 8, W = 32;
    std::vector<Aff<F>> table((size_t)W * 255);
    Jac<F> b = Jac<F>::from_aff(base);
    for (int w = 0; w < W; ++w) {
        Jac<F> acc = b;
        for (int k = 1; k <= 255; ++k) {
            table[(size_t)w * 255 + k - 1] = acc.to_aff();
            acc = acc.add(b);
        }
        b = acc;  // 256 * previous base
    }
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < n; ++i) {
        Fr s = scalars_mont[i].from_mont();
        Jac<F> acc = Jac<F>::inf();
        for (int w = 0; w < W; ++w) {
            int d = (int)((s.v[w / 8] >> ((w % 8) * c)) & 0xff);
            if (d) acc = acc.add_mixed(table[(size_t)w * 255 + d - 1]);
        }
        out[i] = acc.to_aff();
    }
}

// ------------------------------------------------------------------------------------
// fft.Domain transforms. In-place radix-2; DIF: natural -> bit-reversed, DIT: bit-reversed
// -> natural (same butterflies as mpcsetup/lagrange.go:75-79,158-163 but on Fr).
// ------------------------------------------------------------------------------------
static inline size_t bitrev(size_t i, unsigned log_n) {
    size_t r = 0;
    for (unsigned k = 0; k < log_n; ++k) { r = (r << 1) | (i & 1); i >>= 1; }
    return r;
}

static std::vector<Fr> powers(const Fr& w, size_t count) {
    std::vector<Fr> t(count);
    if (count == 0) return t;
    // blocked so that it parallelises: t[k*B + j] = w^(k*B) * w^j
    const size_t B = 1024;
    size_t nb = std::min(B, count);
    t[0] = Fr::one();
    for (size_t j = 1; j < nb; ++j) t[j] = t[j - 1] * w;
    if (count > B) {
        Fr wB = t[B - 1] * w;
        size_t blocks = (count + B - 1) / B;
        std::vector<Fr> heads(blocks);
        heads[0] = Fr::one();
        for (size_t k = 1; k < blocks; ++k) heads[k] = heads[k - 1] * wB;
#pragma omp parallel for schedule(static)
        for (size_t k = 1; k < blocks; ++k)
            for (size_t j = 0; j < B && k * B + j < count; ++j) t[k * B + j] = heads[k] * t[j];
    }
    return t;
}

static void dif_inplace(Fr* a, unsigned log_n, const std::vector<Fr>& tw) {
    size_t n = (size_t)1 << log_n;
    for (unsigned s = 0; s < log_n; ++s) {
        size_t m = n >> s, h = m >> 1;
#pragma omp parallel for schedule(static)
        for (size_t k = 0; k < n / 2; ++k) {
            size_t j = k & (h - 1), base = (k >> (log_n - s - 1)) * m;
            Fr u = a[base + j], v = a[base + j + h];
            a[base + j] = u + v;
            a[base + j + h] = (u - v) * tw[j << s];
        }
    }
}

static void dit_inplace(Fr* a, unsigned log_n, const std::vector<Fr>& tw) {
    size_t n = (size_t)1 << log_n;
    for (unsigned s = 0; s < log_n; ++s) {
        size_t m = (size_t)2 << s, h = m >> 1;
        unsigned shift = log_n - s - 1;  // twiddle = w^(j * n/m)
#pragma omp parallel for schedule(static)
        for (size_t k = 0; k < n / 2; ++k) {
            size_t j = k & (h - 1), base = (k >> s) * m;
            Fr u = a[base + j], v = a[base + j + h] * tw[j << shift];
            a[base + j] = u + v;
            a[base + j + h] = u - v;
        }
    }
}

// Domain.FFT / Domain.FFTInverse. gen = Domain.Generator, coset = FrMultiplicativeGen.
static void ntt(Fr* a, unsigned log_n, int inverse, int decimation, int on_coset, Fr gen, Fr coset) {
    size_t n = (size_t)1 << log_n;
    Fr w = inverse ? gen.inv() : gen;
    std::vector<Fr> tw = powers(w, n / 2 ? n / 2 : 1);
    if (!inverse && on_coset) {
        std::vector<Fr> ct = powers(coset, n);
#pragma omp parallel for schedule(static)
        for (size_t i = 0; i < n; ++i) a[i] = a[i] * ct[decimation == 0 ? i : bitrev(i, log_n)];
    }
    if (decimation == 0) dif_inplace(a, log_n, tw); else dit_inplace(a, log_n, tw);
    if (inverse) {
        Fr ninv = Fr::from_u64((u64)n).inv();
        if (on_coset) {
            std::vector<Fr> ct = powers(coset.inv(), n);
#pragma omp parallel for schedule(static)
            for (size_t i = 0; i < n; ++i) a[i] = a[i] * ct[decimation == 0 ? bitrev(i, log_n) : i] * ninv;
        } else {
#pragma omp parallel for schedule(static)
            for (size_t i = 0; i < n; ++i) a[i] = a[i] * ninv;
        }
    }
}

// ------------------------------------------------------------------------------------
// C entry points (ctypes). All field elements: 4 x u64 LE limbs, Montgomery form.
// ------------------------------------------------------------------------------------
extern "C" {

void oracle_set_threads(int n) { if (n > 0) omp_set_num_threads(n); }
int oracle_max_threads() { return omp_get_max_threads(); }

void oracle_fr_mul(const u64* a, const u64* b, u64* out, size_t n) {
    const Fr* A = (const Fr*)a; const Fr* B = (const Fr*)b; Fr* O = (Fr*)out;
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < n; ++i) O[i] = A[i] * B[i];
}
void oracle_fp_mul(const u64* a, const u64* b, u64* out, size_t n) {
    const Fp* A = (const Fp*)a; const Fp* B = (const Fp*)b; Fp* O = (Fp*)out;
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < n; ++i) O[i] = A[i] * B[i];
}

void oracle_ntt(u64* data, unsigned log_n, int inverse, int decimation, int on_coset,
                const u64* gen, const u64* coset_gen) {
    Fr g, c; memcpy(g.v, gen, 32); memcpy(c.v, coset_gen, 32);
    ntt((Fr*)data, log_n, inverse, decimation, on_coset, g, c);
}

// computeH (prove.go:353-396). a, b, c each hold n = 2^log_n elements (already zero-padded);
// result overwrites a, in bit-reversed order.
void oracle_compute_h(u64* a, u64* b, u64* c, unsigned log_n, const u64* gen, const u64* coset_gen) {
    size_t n = (size_t)1 << log_n;
    Fr g, cg; memcpy(g.v, gen, 32); memcpy(cg.v, coset_gen, 32);
    Fr* A = (Fr*)a; Fr* B = (Fr*)b; Fr* C = (Fr*)c;
    ntt(A, log_n, 1, 0, 0, g, cg); ntt(B, log_n, 1, 0, 0, g, cg); ntt(C, log_n, 1, 0, 0, g, cg);
    ntt(A, log_n, 0, 1, 1, g, cg); ntt(B, log_n, 0, 1, 1, g, cg); ntt(C, log_n, 0, 1, 1, g, cg);
    u64 e[1] = {(u64)n};
    Fr den = (cg.pow(e, 1) - Fr::one()).inv();
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < n; ++i) A[i] = (A[i] * B[i] - C[i]) * den;
    ntt(A, log_n, 1, 0, 1, g, cg);
}

// MultiExp. bases: n x (X,Y) affine Montgomery, (0,0) = infinity. out: Jacobian X,Y,Z.
void oracle_msm_g1(const u64* scalars, const u64* bases, size_t n, u64* out_jac) {
    Jac<Fp> r = msm<Fp>((const Fr*)scalars, (const Aff<Fp>*)bases, n);
    memcpy(out_jac, &r, sizeof(r));
}
void oracle_msm_g2(const u64* scalars, const u64* bases, size_t n, u64* out_jac) {
    Jac<Fp2> r = msm<Fp2>((const Fr*)scalars, (const Aff<Fp2>*)bases, n);
    memcpy(out_jac, &r, sizeof(r));
}
void oracle_g1_batch_mul(const u64* base, const u64* scalars, size_t n, u64* out_aff) {
    Aff<Fp> b; memcpy(&b, base, sizeof(b));
    batch_mul<Fp>(b, (const Fr*)scalars, n, (Aff<Fp>*)out_aff);
}
void oracle_g2_batch_mul(const u64* base, const u64* scalars, size_t n, u64* out_aff) {
    Aff<Fp2> b; memcpy(&b, base, sizeof(b));
    batch_mul<Fp2>(b, (const Fr*)scalars, n, (Aff<Fp2>*)out_aff);
}
void oracle_g1_jac_to_aff(const u64* jac, u64* aff) {
    Jac<Fp> j; memcpy(&j, jac, sizeof(j)); Aff<Fp> a = j.to_aff(); memcpy(aff, &a, sizeof(a));
}
void oracle_g2_jac_to_aff(const u64* jac, u64* aff) {
    Jac<Fp2> j; memcpy(&j, jac, sizeof(j)); Aff<Fp2> a = j.to_aff(); memcpy(aff, &a, sizeof(a));
}

}  // extern "C"
