// compact_kernel.cu — threshold compaction: per-read selection bitmasks -> per-bin ascending read-id lists.
//
// Reference code replaced: the tail of the per-read loop in clasifyLoadedReads, src/d_mapper.h:360-368
// (`origReadIdMap[binNo].push_back(readID)` under a global mutex).  With --threads 1 the reference fills
// every bin's list in ascending read id; that order is what this produces (as CSR: bin_offsets + read_ids),
// deterministically and without atomics.
//
// Scheme: the rows are cut into chunks of `chunk_rows` consecutive rows and the bins into column tiles of
// 32 u32 columns (1024 bins).  A warp owns one (chunk, column tile): lane c owns column word c — it reads that word of
// every row of the chunk (a row = one coalesced 128-byte request of the warp) and keeps the running counters of its 32
// bins in shared memory, so no votes, atomics or synchronisation are needed and ids come out ascending.
//   pass 1 (count): counters -> ws[bin * n_chunks + chunk]
//   scan          : exclusive prefix sum over ws in that bin-major order = where (bin, chunk) starts
//   pass 2 (fill) : same walk, lanes whose row has the bit write their row id at base + rank.
#include "dyibf_internal.h"

namespace dyibf {

static const int CW = 4;     // warps per block in count/fill (1024 u64 counters each = 32 KB shared per block)

struct CompactGeom { uint32_t chunk_rows, n_chunks, n_ctiles, ncols; };

static CompactGeom compact_geom(uint32_t bins, uint32_t n_rows)
{
    CompactGeom g;
    g.ncols = ((bins + 63) / 64) * 2;
    g.n_ctiles = (g.ncols + 31) / 32;
    // aim for >= ~768 warps of work (148 SMs x 4 blocks and some slack), chunk a multiple of 32 rows in [32, COMPACT_CHUNK_ROWS].
    // Every chunk costs bins counters in the workspace (written, scanned, read back) and a counter init / flush per warp:
    // at 4096 warps a 100 000-row batch spent 0.24 ms here (13 % of its query time) on 25 MB of counters; large batches
    // reach the 2048-row cap either way.
    uint64_t want = ((uint64_t)n_rows * g.n_ctiles + 767) / 768;
    uint32_t cr = (uint32_t)((want + 31) / 32 * 32);
    if (cr < 32) cr = 32;
    if (cr > COMPACT_CHUNK_ROWS) cr = COMPACT_CHUNK_ROWS;
    g.chunk_rows = cr;
    g.n_chunks = n_rows ? (n_rows + cr - 1) / cr : 0;
    return g;
}

static const uint32_t SCAN_TILE = 4096;   // elements per scan block (1024 threads x 4)

size_t compact_workspace_bytes(uint32_t bins, uint32_t n_rows)
{
    CompactGeom g = compact_geom(bins, n_rows);
    uint64_t n = (uint64_t)bins * g.n_chunks;
    uint64_t n_blocks = (n + SCAN_TILE - 1) / SCAN_TILE;
    return (size_t)((n + 1 + n_blocks + 1) * sizeof(uint64_t));
}

template <bool FILL>
__global__ void __launch_bounds__(CW * 32) compact_walk(const uint32_t* __restrict__ rows32, uint32_t n_rows, uint32_t bins,
                                                        CompactGeom g, uint64_t* __restrict__ ws,
                                                        uint32_t* __restrict__ read_ids, uint64_t capacity, uint32_t id_base)
{
    __shared__ uint64_t counters[CW][1024];
    const uint32_t lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const uint64_t unit = (uint64_t)blockIdx.x * CW + wib;          // (chunk, ctile)
    const uint64_t n_units = (uint64_t)g.n_chunks * g.n_ctiles;
    if (unit >= n_units) return;
    const uint32_t chunk = (uint32_t)(unit / g.n_ctiles), ctile = (uint32_t)(unit - (uint64_t)chunk * g.n_ctiles);
    uint64_t* cnt = counters[wib];
    const uint32_t col0 = ctile * 32;
    const uint32_t ncol = g.ncols - col0 < 32 ? g.ncols - col0 : 32;

    // counter of bin (col0 + c) * 32 + b lives at cnt[b * 32 + c]: lane c only ever touches column c, conflict-free
    for (uint32_t i = lane; i < 1024; i += 32) {
        uint64_t v = 0;
        if (FILL) {
            const uint32_t bin = (col0 + (i & 31u)) * 32 + (i >> 5);
            if (bin < bins) v = ws[(uint64_t)bin * g.n_chunks + chunk];
        }
        cnt[i] = v;
    }
    __syncwarp();

    // lane c reads word c of every row: a row is one coalesced 128-byte request, 32 of them in flight; the set bits of the
    // lane's words (the matrix is sparse when the filter works) bump the lane's own counters — rows in ascending order,
    // no votes, no atomics, nothing shared between lanes
    const uint32_t first_bin = (col0 + lane) * 32;
    uint32_t lane_mask = 0;                                            // bins beyond `bins` never select
    if (lane < ncol && first_bin < bins) lane_mask = bins - first_bin < 32 ? (1u << (bins - first_bin)) - 1u : 0xffffffffu;
    const uint32_t row_begin = chunk * g.chunk_rows;
    const uint32_t row_end = (row_begin + g.chunk_rows < n_rows) ? row_begin + g.chunk_rows : n_rows;
    for (uint32_t r0 = row_begin; r0 < row_end; r0 += 32) {
        uint32_t x[32];
        #pragma unroll
        for (int i = 0; i < 32; ++i) {
            const uint32_t r = r0 + i;
            x[i] = (lane_mask && r < row_end) ? __ldg(rows32 + (uint64_t)r * g.ncols + col0 + lane) : 0u;
        }
        #pragma unroll
        for (int i = 0; i < 32; ++i) {
            uint32_t v = x[i] & lane_mask;
            while (v) {
                const uint32_t b = __ffs(v) - 1;
                v &= v - 1;
                const uint64_t c = cnt[b * 32 + lane];
                if (FILL && c < capacity) read_ids[c] = id_base + r0 + i;
                cnt[b * 32 + lane] = c + 1;
            }
        }
    }
    if (!FILL) {
        __syncwarp();
        for (uint32_t i = lane; i < 1024; i += 32) {
            const uint32_t bin = (col0 + (i & 31u)) * 32 + (i >> 5);
            if (bin < bins) ws[(uint64_t)bin * g.n_chunks + chunk] = cnt[i];
        }
    }
}

// ---- exclusive scan of u64[n] in place (three small kernels) ---------------------------------------------
__global__ void __launch_bounds__(1024) scan_local(uint64_t* __restrict__ a, uint64_t n, uint64_t* __restrict__ block_sums)
{
    __shared__ uint64_t warp_sums[32];
    const uint64_t base = (uint64_t)blockIdx.x * SCAN_TILE + (uint64_t)threadIdx.x * 4;
    uint64_t v[4], tot = 0;
    #pragma unroll
    for (int i = 0; i < 4; ++i) { v[i] = (base + i < n) ? a[base + i] : 0; tot += v[i]; }
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint64_t inc = tot;
    #pragma unroll
    for (int d = 1; d < 32; d <<= 1) { uint64_t t = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += t; }
    if (lane == 31) warp_sums[w] = inc;
    __syncthreads();
    if (w == 0) {
        uint64_t s = warp_sums[lane], si = s;
        #pragma unroll
        for (int d = 1; d < 32; d <<= 1) { uint64_t t = __shfl_up_sync(0xffffffffu, si, d); if (lane >= d) si += t; }
        warp_sums[lane] = si - s;                                    // exclusive
        if (lane == 31) block_sums[blockIdx.x] = si;
    }
    __syncthreads();
    uint64_t run = warp_sums[w] + inc - tot;
    #pragma unroll
    for (int i = 0; i < 4; ++i) { if (base + i < n) a[base + i] = run; run += v[i]; }
}

__global__ void __launch_bounds__(1024) scan_block_sums(uint64_t* __restrict__ block_sums, uint64_t n_blocks, uint64_t* __restrict__ total_out)
{
    __shared__ uint64_t warp_sums[32];
    __shared__ uint64_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (uint64_t b0 = 0; b0 < n_blocks; b0 += 1024) {
        const uint64_t i = b0 + threadIdx.x;
        const uint64_t v = i < n_blocks ? block_sums[i] : 0;
        uint64_t inc = v;
        #pragma unroll
        for (int d = 1; d < 32; d <<= 1) { uint64_t t = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += t; }
        if (lane == 31) warp_sums[w] = inc;
        __syncthreads();
        if (w == 0) {
            uint64_t s = warp_sums[lane], si = s;
            #pragma unroll
            for (int d = 1; d < 32; d <<= 1) { uint64_t t = __shfl_up_sync(0xffffffffu, si, d); if (lane >= d) si += t; }
            warp_sums[lane] = si - s;
        }
        __syncthreads();
        const uint64_t excl = carry + warp_sums[w] + inc - v;
        if (i < n_blocks) block_sums[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total_out = carry;
}

__global__ void __launch_bounds__(1024) scan_add(uint64_t* __restrict__ a, uint64_t n, const uint64_t* __restrict__ block_sums)
{
    const uint64_t base = (uint64_t)blockIdx.x * SCAN_TILE + (uint64_t)threadIdx.x * 4;
    const uint64_t add = block_sums[blockIdx.x];
    #pragma unroll
    for (int i = 0; i < 4; ++i) if (base + i < n) a[base + i] += add;
}

__global__ void gather_bin_offsets(const uint64_t* __restrict__ ws, uint32_t bins, uint32_t n_chunks, uint64_t n,
                                   uint64_t* __restrict__ bin_offsets)
{
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < bins) bin_offsets[b] = n_chunks ? ws[(uint64_t)b * n_chunks] : 0;
    else if (b == bins) bin_offsets[b] = ws[n];                      // total
}

cudaError_t launch_compact_count(const uint64_t* rows, uint32_t n_rows, uint32_t bins, uint64_t* ws, cudaStream_t st)
{
    CompactGeom g = compact_geom(bins, n_rows);
    uint64_t units = (uint64_t)g.n_chunks * g.n_ctiles;
    if (units == 0) return cudaSuccess;
    compact_walk<false><<<(unsigned)((units + CW - 1) / CW), CW * 32, 0, st>>>(
        reinterpret_cast<const uint32_t*>(rows), n_rows, bins, g, ws, nullptr, 0, 0);
    return cudaGetLastError();
}

size_t scan_workspace_elems(uint64_t n)
{
    return (size_t)(n + 1 + (n + SCAN_TILE - 1) / SCAN_TILE + 1);
}

// a[0..n) -> exclusive prefix sums in place, a[n] = total; a[n+1..] is scratch (scan_workspace_elems(n) elements in all)
cudaError_t launch_scan_u64(uint64_t* a, uint64_t n, cudaStream_t st)
{
    const uint64_t n_blocks = (n + SCAN_TILE - 1) / SCAN_TILE;
    uint64_t* block_sums = a + n + 1;
    if (n == 0) return cudaMemsetAsync(a, 0, sizeof(uint64_t), st);
    scan_local<<<(unsigned)n_blocks, 1024, 0, st>>>(a, n, block_sums);
    scan_block_sums<<<1, 1024, 0, st>>>(block_sums, n_blocks, a + n);
    scan_add<<<(unsigned)n_blocks, 1024, 0, st>>>(a, n, block_sums);
    return cudaGetLastError();
}

cudaError_t launch_compact_scan(uint64_t* ws, uint32_t bins, uint32_t n_rows, uint64_t* bin_offsets, cudaStream_t st)
{
    CompactGeom g = compact_geom(bins, n_rows);
    const uint64_t n = (uint64_t)bins * g.n_chunks;
    cudaError_t e = launch_scan_u64(ws, n, st);
    if (e != cudaSuccess) return e;
    gather_bin_offsets<<<(bins + 1 + 255) / 256, 256, 0, st>>>(ws, bins, g.n_chunks, n, bin_offsets);
    return cudaGetLastError();
}

cudaError_t launch_compact_fill(const uint64_t* rows, uint32_t n_rows, uint32_t bins, const uint64_t* ws,
                                uint32_t* read_ids, uint64_t capacity, uint32_t id_base, cudaStream_t st)
{
    CompactGeom g = compact_geom(bins, n_rows);
    uint64_t units = (uint64_t)g.n_chunks * g.n_ctiles;
    if (units == 0) return cudaSuccess;
    compact_walk<true><<<(unsigned)((units + CW - 1) / CW), CW * 32, 0, st>>>(
        reinterpret_cast<const uint32_t*>(rows), n_rows, bins, g, const_cast<uint64_t*>(ws), read_ids, capacity, id_base);
    return cudaGetLastError();
}

} // namespace dyibf
