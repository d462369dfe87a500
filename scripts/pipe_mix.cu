// Development probe: do DFMA, IMAD.WIDE and IADD3 overlap on sm_100?  Times of mixed loops.
#include <cstdio>
#include <cuda_runtime.h>
constexpr int ITERS = 2048;
template <int ND, int NW, int NA>
__global__ void __launch_bounds__(256) mix_kernel(const double* in, double* out) {
    double d[8]; unsigned long long w[8]; unsigned a[16];
    const double b = in[threadIdx.x & 31], c = in[32 + (threadIdx.x & 31)];
    const unsigned ub = (unsigned)threadIdx.x | 1u, uc = (unsigned)blockIdx.x;
#pragma unroll
    for (int j = 0; j < 8; ++j) { d[j] = c + j; w[j] = j + threadIdx.x; }
#pragma unroll
    for (int j = 0; j < 16; ++j) a[j] = j;
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int r = 0; r < 2; ++r) {
#pragma unroll
            for (int j = 0; j < ND; ++j) d[j] = __fma_rz(d[j], b, c);
#pragma unroll
            for (int j = 0; j < NW; ++j) { const unsigned lo = (unsigned)w[j]; asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(w[j]) : "r"(lo), "r"(ub)); }
#pragma unroll
            for (int j = 0; j < NA; ++j) asm volatile("xor.b32 %0, %0, %1;\n\tadd.u32 %0, %0, %2;" : "+r"(a[j]) : "r"(ub), "r"(uc));
        }
    }
    double s = 0; unsigned long long t = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) { s += d[j]; t ^= w[j]; }
#pragma unroll
    for (int j = 0; j < 16; ++j) t ^= a[j];
    if (s == 1.2345 || t == 0x123456789ull) out[0] = s;
}
template <int ND, int NW, int NA>
static void run(double* d, int sms, cudaEvent_t e0, cudaEvent_t e1) {
    float best = 1e9f;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0);
        mix_kernel<ND, NW, NA><<<sms * 8, 256>>>(d, d);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep && ms < best) best = ms;
    }
    // SMSP cycles per loop iteration per warp
    const double cyc = best * 1e-3 * 1.965e9 / (8.0 * 256 / 32 / 4) / ITERS / 2;
    printf("DFMA %d  IMAD.WIDE %d  IADD3 %d : %.3f ms  -> %.1f SMSP cycles per (warp, round)\n", ND, NW, NA, best, cyc);
}
int main() {
    double* d; cudaMalloc(&d, 1 << 16); cudaMemset(d, 0, 1 << 16);
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    run<8, 0, 0>(d, sms, e0, e1);
    run<0, 8, 0>(d, sms, e0, e1);
    run<0, 0, 8>(d, sms, e0, e1);
    run<8, 8, 0>(d, sms, e0, e1);
    run<8, 0, 8>(d, sms, e0, e1);
    run<0, 8, 8>(d, sms, e0, e1);
    run<8, 8, 8>(d, sms, e0, e1);
    run<8, 4, 8>(d, sms, e0, e1);
    run<8, 4, 16>(d, sms, e0, e1);
    run<4, 8, 16>(d, sms, e0, e1);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
