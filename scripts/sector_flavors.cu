// Microbenchmark (B200): what one isolated 32-byte access costs the memory system, by load flavour.  Every thread reads
// random 32-byte sectors of a 16 GB buffer; run under ncu to see DRAM bytes and L2 sectors per access:
//   ncu --metrics dram__bytes_read.sum,lts__t_sectors_srcunit_tex_op_read.sum,gpu__time_duration.sum /tmp/sector_flavors
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint64_t mix(uint64_t x) {
    x ^= x >> 33;
    x *= 0xff51afd7ed558ccdULL;
    x ^= x >> 33;
    x *= 0xc4ceb9fe1a85ec53ULL;
    x ^= x >> 33;
    return x;
}

template <int MODE>
__device__ __forceinline__ uint32_t load32(const uint4 *p) {
    uint4 a = make_uint4(0, 0, 0, 0), b = make_uint4(0, 0, 0, 0);
    if (MODE == 0) {
        a = __ldcs(p);
        b = __ldcs(p + 1);
    } else if (MODE == 1) {
        a = p[0];
        b = p[1];
    } else if (MODE == 2) {
        a = __ldcg(p);
        b = __ldcg(p + 1);
    } else if (MODE == 3) {
        a = __ldg(p);
        b = __ldg(p + 1);
    } else if (MODE == 4) {
        a = __ldcv(p);
        b = __ldcv(p + 1);
    } else if (MODE == 5) {
        asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w) : "l"(p));
        asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w) : "l"(p + 1));
    } else if (MODE == 6) {
        a = __ldcs(p); // 16 bytes only
    } else if (MODE == 7) {
        a.x = __ldcs(reinterpret_cast<const uint32_t *>(p)); // 4 bytes only
    } else if (MODE == 8) {
        unsigned long long q0, q1, q2, q3;
        asm volatile("ld.global.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(q0), "=l"(q1), "=l"(q2), "=l"(q3) : "l"(p));
        a.x = (uint32_t)q0;
        b.w = (uint32_t)(q3 >> 32) ^ (uint32_t)q1 ^ (uint32_t)q2;
    } else if (MODE == 9) {
        asm volatile("ld.global.cs.L2::64B.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w) : "l"(p));
        asm volatile("ld.global.cs.L2::64B.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w) : "l"(p + 1));
    }
    return a.x ^ b.w;
}

template <int MODE>
__global__ void __launch_bounds__(256) flavorKernel(const uint4 *buf, const uint64_t nsectors, const int iters, uint32_t *sink) {
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t acc = 0;
    for (int it = 0; it < iters; it += 4) {
        uint64_t s[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) s[k] = mix(tid * 0x9E3779B97F4A7C15ULL + (uint64_t)(it + k)) % nsectors;
#pragma unroll
        for (int k = 0; k < 4; ++k) acc ^= load32<MODE>(buf + 2 * s[k]);
    }
    if (acc == 0x12345678u) *sink = acc;
}

template <int MODE>
static void run(const char *name, const uint4 *buf, uint64_t bytes, uint32_t *sink) {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int blocks = sms * 4, iters = 1024;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0);
    flavorKernel<MODE><<<blocks, 256>>>(buf, bytes / 32, iters, sink);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    std::printf("mode %d %-28s: %6.1f G accesses/s (%.2f ms, %lld accesses) %s\n", MODE, name, (double)blocks * 256 * iters / ms / 1e6, ms,
                (long long)blocks * 256 * iters, cudaGetErrorString(cudaGetLastError()));
}

int main(int argc, char **argv) {
    if (argc > 1) { // fetch granularity asked for before anything else touches the device
        cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)atoi(argv[1]));
    }
    size_t got = 0;
    cudaDeviceGetLimit(&got, cudaLimitMaxL2FetchGranularity);
    std::printf("cudaLimitMaxL2FetchGranularity = %zu\n", got);
    uint32_t *sink;
    cudaMalloc(&sink, 4);
    const uint64_t bytes = 16ull << 30;
    uint4 *buf = nullptr;
    if (cudaMalloc(&buf, bytes) != cudaSuccess) return 1;
    cudaMemset(buf, 0, bytes);
    run<0>("ld.cs 2x16", buf, bytes, sink);
    run<1>("ld (default) 2x16", buf, bytes, sink);
    run<2>("ld.cg 2x16", buf, bytes, sink);
    run<3>("ld.nc (ldg) 2x16", buf, bytes, sink);
    run<4>("ld.cv 2x16", buf, bytes, sink);
    run<5>("ld.L1::no_allocate 2x16", buf, bytes, sink);
    run<6>("ld.cs 1x16", buf, bytes, sink);
    run<7>("ld.cs 1x4", buf, bytes, sink);
    run<8>("ld.v4.u64 1x32", buf, bytes, sink);
    run<9>("ld.cs.L2::64B 2x16", buf, bytes, sink);
    cudaDeviceSynchronize();
    std::printf("last error: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
