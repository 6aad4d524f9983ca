// tools/gather_bench.cu — micro-benchmark: what random-gather bandwidth can a B200 deliver from a multi-GiB table?
// It bounds the IBF query kernel from above: the kernel's traffic is h random aligned blocks of W*8 bytes per
// k-mer with nothing else to do.  Sweeps line size (bytes fetched contiguously per random address), loads in
// flight per thread (U) and resident warps per SM.  Prints one CSV line per configuration.
//
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o gather_bench tools/gather_bench.cu && ./gather_bench [GiB]
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint64_t mix(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}

__device__ __forceinline__ uint4 ldg_na(const uint8_t* p)
{
    uint4 c;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(c.x), "=r"(c.y), "=r"(c.z), "=r"(c.w) : "l"(p));
    return c;
}

// LPL = lanes per line (line bytes = 16*LPL); each group of LPL lanes fetches one random line per load slot.
template <int LPL, int U>
__global__ void gather(const uint8_t* __restrict__ table, uint64_t n_lines, uint32_t iters, uint32_t* __restrict__ sink)
{
    const uint32_t lane = threadIdx.x & 31, li = lane % LPL;
    const uint64_t grp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) / LPL;
    uint32_t acc = 0;
    for (uint32_t it = 0; it < iters; ++it) {
        uint4 v[U];
        #pragma unroll
        for (int u = 0; u < U; ++u) {
            uint64_t line = mix(grp * 0x100000001B3ULL + (uint64_t)it * U + u) & (n_lines - 1);   // n_lines is a power of two
            v[u] = ldg_na(table + line * (16ull * LPL) + li * 16);
        }
        #pragma unroll
        for (int u = 0; u < U; ++u) acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
    }
    if (acc == 0x12345678u) sink[0] = acc;
}

template <int LPL, int U>
static void run(const uint8_t* table, uint64_t bytes, uint32_t* sink, int blocks_per_sm, int threads, int sms)
{
    const uint64_t line_bytes = 16ull * LPL, n_lines = bytes / line_bytes;
    const int grid = sms * blocks_per_sm;
    const uint64_t groups = (uint64_t)grid * threads / LPL;
    // ~64 GB of traffic per measurement
    uint32_t iters = (uint32_t)((64ull << 30) / (groups * U * line_bytes));
    if (iters < 4) iters = 4;
    int max_blocks = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&max_blocks, gather<LPL, U>, threads, 0));
    if (max_blocks < blocks_per_sm) { printf("%llu,%d,%d,%d,skipped (occupancy %d)\n", (unsigned long long)line_bytes, U, blocks_per_sm, threads, max_blocks); return; }
    gather<LPL, U><<<grid, threads>>>(table, n_lines, iters / 8 + 1, sink);   // warm-up
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    CK(cudaEventRecord(a));
    gather<LPL, U><<<grid, threads>>>(table, n_lines, iters, sink);
    CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
    float ms = 0; CK(cudaEventElapsedTime(&ms, a, b));
    double gb = (double)groups * U * iters * line_bytes / 1e9;
    printf("%llu,%d,%d,%d,%.1f,%.2f\n", (unsigned long long)line_bytes, U, blocks_per_sm * threads / 32, threads, gb / (ms / 1e3), ms);
    fflush(stdout);
}

int main(int argc, char** argv)
{
    uint64_t gib = argc > 1 ? strtoull(argv[1], 0, 10) : 16;
    uint64_t bytes = gib << 30;
    uint8_t* table; uint32_t* sink;
    CK(cudaMalloc(&table, bytes)); CK(cudaMemset(table, 1, bytes)); CK(cudaMalloc(&sink, 4));
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int sms = p.multiProcessorCount;
    printf("# %s, %d SMs, table %llu GiB\nline_bytes,U,warps_per_sm,threads_per_block,GBps,ms\n", p.name, sms, (unsigned long long)gib);
    // 128-byte lines (1024-bin IBF block): U x occupancy
    for (int bps : {2, 4, 8, 16}) {
        run<8, 1>(table, bytes, sink, bps, 128, sms);
        run<8, 4>(table, bytes, sink, bps, 128, sms);
        run<8, 8>(table, bytes, sink, bps, 128, sms);
        run<8, 12>(table, bytes, sink, bps, 128, sms);
        run<8, 16>(table, bytes, sink, bps, 128, sms);
        run<8, 24>(table, bytes, sink, bps, 128, sms);
    }
    // other line sizes at a good operating point
    for (int bps : {4, 8, 16}) {
        run<2, 12>(table, bytes, sink, bps, 128, sms);    // 32 B  (<= 256-bin IBF, W even... sector-sized blocks)
        run<4, 12>(table, bytes, sink, bps, 128, sms);    // 64 B
        run<16, 12>(table, bytes, sink, bps, 128, sms);   // 256 B
        run<32, 12>(table, bytes, sink, bps, 128, sms);   // 512 B
    }
    return 0;
}
