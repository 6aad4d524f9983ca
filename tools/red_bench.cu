// tools/red_bench.cu — micro-benchmark: how many fully divergent 32-bit RED.OR operations per second does a B200 retire
// when their targets stay in the L2?  It is the upper bound of the sliced build's scatter (csrc/build_sliced.cu): every lane
// of a warp ORs one bit into its own random 32-byte sector of a window of one or a few bins' bit vectors (16 MiB each at
// cfg2), i.e. 32 sectors per request, no coalescing, nothing else to do.  The kernel below does only that, with the hash
// replaced by a cheap integer mix, so what it reaches is the memory system's RED rate, not the ALUs'.
// Sweeps the window (1 / 2 / 4 / 8 vectors of 16 MiB, and 256 MiB = beyond the L2), resident warps per SM and the REDs
// in flight per thread.  One CSV line per configuration; the best line per window is the peak bench_support/extras.py
// divides the scatter's measured rate by (profiles/l2_red_peak.json).
//
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o red_bench tools/red_bench.cu && ./red_bench
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t mix32(uint32_t x)
{
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
    return x;
}

// every thread: iters x U REDs, each to word (random & mask) of the window; result unused -> RED, not ATOM
template <int U>
__global__ void __launch_bounds__(256) red_or(uint32_t* __restrict__ window, uint32_t word_mask, uint32_t iters)
{
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t s = mix32(tid * 2654435761U + 12345U);
    for (uint32_t it = 0; it < iters; ++it) {
        #pragma unroll
        for (int u = 0; u < U; ++u) {
            s = mix32(s + 0x9E3779B9U * (u + 1));
            atomicOr(window + (s & word_mask), 1u << (s >> 27));
        }
    }
}

template <int U>
static void run(uint32_t* window, uint64_t window_bytes, int blocks_per_sm, int sms)
{
    const int threads = 256, grid = sms * blocks_per_sm;
    int max_blocks = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&max_blocks, red_or<U>, threads, 0));
    if (max_blocks < blocks_per_sm) return;
    const uint64_t per_iter = (uint64_t)grid * threads * U;
    uint32_t iters = (uint32_t)((8ull << 30) / per_iter);             // ~8.6 G REDs per measurement
    if (iters < 8) iters = 8;
    const uint32_t word_mask = (uint32_t)(window_bytes / 4 - 1);
    red_or<U><<<grid, threads>>>(window, word_mask, iters / 8 + 1);    // warm-up, fills the L2 with the window
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        CK(cudaEventRecord(a));
        red_or<U><<<grid, threads>>>(window, word_mask, iters);
        CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
        float ms = 0; CK(cudaEventElapsedTime(&ms, a, b));
        if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    const double reds = (double)per_iter * iters;
    printf("%llu,%d,%d,%.2f,%.3f,%.3f\n", (unsigned long long)(window_bytes >> 20), blocks_per_sm * threads / 32, U,
           reds / (best / 1e3) / 1e9, best, reds / (best / 1e3) / 1e9 / sms);
    fflush(stdout);
    CK(cudaEventDestroy(a)); CK(cudaEventDestroy(b));
}

int main()
{
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    const int sms = p.multiProcessorCount;
    const uint64_t max_bytes = 256ull << 20;
    uint32_t* window; CK(cudaMalloc(&window, max_bytes)); CK(cudaMemset(window, 0, max_bytes));
    int clk_khz = 0; CK(cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0));
    printf("# %s, %d SMs, L2 %d MiB, max SM clock %d MHz; fully divergent RED.OR.b32 (32 sectors per warp request)\n", p.name, sms,
           p.l2CacheSize >> 20, clk_khz / 1000);
    printf("window_MiB,warps_per_sm,reds_in_flight_per_thread,G_red_per_s,ms,G_red_per_s_per_sm\n");
    for (uint64_t mib : {16ull, 32ull, 64ull, 128ull, 256ull}) {
        for (int bps : {1, 2, 4, 6, 8}) {
            run<1>(window, mib << 20, bps, sms);
            run<4>(window, mib << 20, bps, sms);
            run<8>(window, mib << 20, bps, sms);
        }
    }
    CK(cudaFree(window));
    return 0;
}
