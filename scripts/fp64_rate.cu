// Development probe: DFMA / DADD / IADD3 issue rates per SM.
#include <cstdio>
#include <cuda_runtime.h>
constexpr int ITERS = 4096;
__global__ void __launch_bounds__(256) dfma_kernel(const double* in, double* out) {
    double a[16]; const double b = in[threadIdx.x & 31], c = in[32 + (threadIdx.x & 31)];
#pragma unroll
    for (int j = 0; j < 16; ++j) a[j] = c + j;
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int j = 0; j < 16; ++j) a[j] = __fma_rz(a[j], b, c);
    }
    double s = 0;
#pragma unroll
    for (int j = 0; j < 16; ++j) s += a[j];
    if (s == 1.2345) out[0] = s;
}
__global__ void __launch_bounds__(256) dadd_kernel(const double* in, double* out) {
    double a[16]; const double c = in[32 + (threadIdx.x & 31)];
#pragma unroll
    for (int j = 0; j < 16; ++j) a[j] = c + j;
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int j = 0; j < 16; ++j) a[j] = a[j] + c;
    }
    double s = 0;
#pragma unroll
    for (int j = 0; j < 16; ++j) s += a[j];
    if (s == 1.2345) out[0] = s;
}
__global__ void __launch_bounds__(256) iadd_kernel(const unsigned* in, unsigned* out) {
    unsigned a[16]; const unsigned b = in[threadIdx.x & 31], c = in[32 + (threadIdx.x & 31)];
#pragma unroll
    for (int j = 0; j < 16; ++j) a[j] = c + j;
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int j = 0; j < 16; ++j) asm volatile("add.u32 %0, %0, %1;\n\tadd.u32 %0, %0, %2;" : "+r"(a[j]) : "r"(b), "r"(c));
    }
    unsigned s = 0;
#pragma unroll
    for (int j = 0; j < 16; ++j) s ^= a[j];
    if (s == 0x12345678u) out[0] = s;
}
int main() {
    double* d; cudaMalloc(&d, 1 << 16); cudaMemset(d, 0, 1 << 16);
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int which = 0; which < 3; ++which) {
        float best = 1e9f;
        for (int rep = 0; rep < 4; ++rep) {
            cudaEventRecord(e0);
            if (which == 0) dfma_kernel<<<sms * 8, 256>>>(d, d);
            if (which == 1) dadd_kernel<<<sms * 8, 256>>>(d, d);
            if (which == 2) iadd_kernel<<<sms * 8, 256>>>((unsigned*)d, (unsigned*)d);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            if (rep && ms < best) best = ms;
        }
        const double ops = (double)sms * 8 * 256 * ITERS * 16 * (which == 2 ? 2 : 1);
        printf("%s: %.3e /s = %.1f per clk per SM at %.0f MHz nominal (%s)\n", which == 0 ? "DFMA.RZ" : which == 1 ? "DADD" : "IADD",
               ops / (best * 1e-3), ops / (best * 1e-3) / sms / (khz * 1e3), khz / 1e3, cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
