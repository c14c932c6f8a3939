// Microbenchmark (B200): throughput of ISOLATED 32-byte sector accesses to a buffer far larger than L2 -- the access pattern
// of the shortest-path searches (label entries, queue entries: one random sector each).  Reports useful GB/s for
//   read32   two 16-byte loads of one random sector          (label probe)
//   write32  two 16-byte stores covering one random sector   (label / queue-entry store)
//   write4   one 4-byte store into a random sector           (predecessor word, tombstone: a partial-sector write)
// at several footprints and numbers of independent accesses in flight per thread.
// Build + run: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/sector_bw scripts/sector_bw.cu && /tmp/sector_bw
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

template <int MODE, int ILP>
__global__ void __launch_bounds__(256) sectorKernel(uint4 *buf, const uint64_t nsectors, const int iters, uint32_t *sink) {
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t acc = 0;
    for (int it = 0; it < iters; it += ILP) {
        uint64_t s[ILP];
#pragma unroll
        for (int k = 0; k < ILP; ++k) s[k] = mix(tid * 0x9E3779B97F4A7C15ULL + (uint64_t)(it + k)) % nsectors;
        if (MODE == 0) {
            uint4 a[ILP], b[ILP];
#pragma unroll
            for (int k = 0; k < ILP; ++k) {
                a[k] = __ldcs(buf + 2 * s[k]);
                b[k] = __ldcs(buf + 2 * s[k] + 1);
            }
#pragma unroll
            for (int k = 0; k < ILP; ++k) acc ^= a[k].x ^ b[k].w;
        } else if (MODE == 1) {
#pragma unroll
            for (int k = 0; k < ILP; ++k) {
                __stcs(buf + 2 * s[k], make_uint4((uint32_t)tid, it, k, 1u));
                __stcs(buf + 2 * s[k] + 1, make_uint4((uint32_t)tid, it, k, 2u));
            }
        } else {
#pragma unroll
            for (int k = 0; k < ILP; ++k) __stcs(reinterpret_cast<uint32_t *>(buf + 2 * s[k]) + (it & 7), (uint32_t)tid);
        }
    }
    if (acc == 0x12345678u) *sink = acc;
}

template <int MODE, int ILP>
static void run(const char *name, uint4 *buf, uint64_t bytes, uint32_t *sink, int threadsPerSm) {
    const uint64_t nsectors = bytes / 32;
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int blocks = sms * threadsPerSm / 256, iters = 2048;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    sectorKernel<MODE, ILP><<<blocks, 256>>>(buf, nsectors, 256, sink); // warm-up
    cudaEventRecord(e0);
    sectorKernel<MODE, ILP><<<blocks, 256>>>(buf, nsectors, iters, sink);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double accesses = (double)blocks * 256 * iters;
    const double useful = accesses * (MODE == 2 ? 4 : 32);
    std::printf("%-8s footprint %5.1f GB  %4d thr/SM  ilp %d : %7.1f G accesses/s  %7.1f GB/s useful (%.2f ms)\n", name, bytes / 1e9, threadsPerSm,
                ILP, accesses / ms / 1e6, useful / ms / 1e6, ms);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
}

int main() {
    uint32_t *sink;
    cudaMalloc(&sink, 4);
    const uint64_t sizes[3] = {2ull << 30, 16ull << 30, 64ull << 30};
    uint4 *buf = nullptr;
    if (cudaMalloc(&buf, sizes[2]) != cudaSuccess) {
        std::printf("cudaMalloc failed\n");
        return 1;
    }
    cudaMemset(buf, 0, sizes[2]);
    for (int g = 0; g < 2; ++g) {
        if (g == 1) {
            cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, 32);
            size_t got = 0;
            cudaDeviceGetLimit(&got, cudaLimitMaxL2FetchGranularity);
            std::printf("--- cudaLimitMaxL2FetchGranularity asked 32, now %zu\n", got);
        }
        for (int f = 0; f < 3; ++f) {
            run<0, 1>("read32", buf, sizes[f], sink, 1024);
            run<0, 4>("read32", buf, sizes[f], sink, 1024);
            run<0, 8>("read32", buf, sizes[f], sink, 2048);
            run<1, 4>("write32", buf, sizes[f], sink, 2048);
            run<2, 4>("write4", buf, sizes[f], sink, 2048);
        }
    }
    // the searches of the pairs engine: 6 warps per scheduler, one or two dependent accesses in flight per lane
    run<0, 1>("read32", buf, sizes[1], sink, 768);
    run<0, 2>("read32", buf, sizes[1], sink, 768);
    cudaFree(buf);
    std::printf("last error: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
