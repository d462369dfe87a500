// Which load fetches only the 64 bytes it asks for?  Random 64-byte gathers from an 8 GiB table (the access pattern of the
// folded-window MSM: every table point is read once, at a random time), one variant per launch; run under
//   ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum ./gather_probe
// and compare dram__bytes_read with the algorithmic n * 64 bytes.  (r01: __ldg pulls 128-byte lines: 2.1x.)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gather_probe gather_probe.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t hash32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}

template <int V>
__device__ __forceinline__ uint4 load16(const uint4* p) {
    uint4 r;
    if (V == 0) r = __ldg(p);
    else if (V == 1) asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    else if (V == 2) asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    else if (V == 3) asm volatile("ld.global.cs.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    else if (V == 4) asm volatile("ld.global.nc.L2::64B.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    else if (V == 5) asm volatile("ld.global.L1::no_allocate.L2::64B.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    else if (V == 6) asm volatile("ld.global.lu.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    else if (V == 7) asm volatile("ld.global.cv.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    else if (V == 8) asm volatile("ld.global.nc.L1::no_allocate.L2::128B.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

// bytes: how much of the 64-byte record a thread reads (64 = x and y, 32 = x only)
template <int V, int BYTES>
__global__ void gather_kernel(const uint4* __restrict__ table, uint32_t mask, uint32_t n, uint32_t* __restrict__ out) {
    uint32_t acc = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint4* p = table + (size_t)(hash32(i) & mask) * 4;
#pragma unroll
        for (int k = 0; k < BYTES / 16; ++k) {
            const uint4 v = load16<V>(p + k);
            acc += v.x ^ v.y ^ v.z ^ v.w;
        }
    }
    if (acc == 0x12345678u) out[0] = acc;
}

// bulk-copy engine: every warp brings 32 records of 64 bytes into shared memory with cp.async.bulk, one per lane
__global__ void gather_bulk_kernel(const uint4* __restrict__ table, uint32_t mask, uint32_t n, uint32_t* __restrict__ out) {
    __shared__ __align__(128) uint4 buf[128 * 4];
    __shared__ __align__(8) uint64_t bar;
    const uint32_t bar_addr = (uint32_t)__cvta_generic_to_shared(&bar);
    if (threadIdx.x == 0) asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar_addr), "r"(1));
    __syncthreads();
    uint32_t acc = 0, phase = 0;
    for (uint32_t base = blockIdx.x * blockDim.x; base < n; base += gridDim.x * blockDim.x) {
        const uint32_t i = base + threadIdx.x;
        if (threadIdx.x == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_addr), "r"(blockDim.x * 64u));
        __syncthreads();
        const uint4* p = table + (size_t)(hash32(i) & mask) * 4;
        const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&buf[threadIdx.x * 4]);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 64, [%2];" ::"r"(dst), "l"(p), "r"(bar_addr) : "memory");
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(bar_addr), "r"(phase) : "memory");
        phase ^= 1;
#pragma unroll
        for (int k = 0; k < 4; ++k) { const uint4 v = buf[threadIdx.x * 4 + k]; acc += v.x ^ v.y ^ v.z ^ v.w; }
        __syncthreads();
    }
    if (acc == 0x12345678u) out[0] = acc;
}

template <int V, int BYTES>
static void run(const char* name, const uint4* table, uint32_t mask, uint32_t n, uint32_t* out) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    gather_kernel<V, BYTES><<<148 * 16, 256>>>(table, mask, n, out);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    printf("%-44s %2d B/record  %8.3f ms  %7.1f GB/s algorithmic  (%s)\n", name, BYTES, ms, (double)n * BYTES / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
}

int main(int argc, char** argv) {
    const uint32_t log_records = 27;                 // 2^27 x 64 B = 8 GiB
    const uint32_t n = 1u << 26;                     // gathers per launch (4 GiB algorithmic at 64 B)
    if (argc > 1) cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)atoi(argv[1]));
    size_t gran = 0;
    cudaDeviceGetLimit(&gran, cudaLimitMaxL2FetchGranularity);
    printf("cudaLimitMaxL2FetchGranularity = %zu\n", gran);
    uint4* table;
    uint32_t* out;
    cudaMalloc(&table, (size_t)64 << log_records);
    cudaMalloc(&out, 4);
    cudaMemset(table, 1, (size_t)64 << log_records);
    const uint32_t mask = (1u << log_records) - 1u;
    run<0, 64>("ld.global.nc (__ldg)", table, mask, n, out);
    run<1, 64>("ld.global", table, mask, n, out);
    run<2, 64>("ld.global.cg", table, mask, n, out);
    run<3, 64>("ld.global.cs", table, mask, n, out);
    run<4, 64>("ld.global.nc.L2::64B", table, mask, n, out);
    run<5, 64>("ld.global.L1::no_allocate.L2::64B", table, mask, n, out);
    run<6, 64>("ld.global.lu", table, mask, n, out);
    run<7, 64>("ld.global.cv", table, mask, n, out);
    run<8, 64>("ld.global.nc.L1::no_allocate.L2::128B", table, mask, n, out);
    run<0, 32>("ld.global.nc (__ldg), x only", table, mask, n, out);
    run<4, 32>("ld.global.nc.L2::64B, x only", table, mask, n, out);
    run<2, 32>("ld.global.cg, x only", table, mask, n, out);
    {
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        cudaEventRecord(e0);
        gather_bulk_kernel<<<148 * 16, 128>>>(table, mask, n, out);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        printf("%-44s %2d B/record  %8.3f ms  %7.1f GB/s algorithmic  (%s)\n", "cp.async.bulk 64 B -> shared", 64, ms, (double)n * 64 / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
    }
    cudaDeviceSynchronize();
    printf("end: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
