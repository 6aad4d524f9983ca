// build_kernel.cu — IBF construction and bin clearing.
//
// Reference loops replaced:
//   _addKmers / _insertKmer, src/d_bloom_filter.h:293-315: for every k-mer, for every hash function, set bit
//   (block*S + bin).  The reference does a non-atomic `word |= mask` from several threads
//   (sdsl int_vector.hpp:721-728); here it is a fire-and-forget 64-bit atomic OR (RED.OR at the L2), which is
//   race-free and, OR being commutative, equals the single-threaded reference result bit for bit.
//   clearBins, src/d_bloom_filter.h:99-123: clear bit (block*S + bin) in every hash block.
//
// Work decomposition (build): one thread per packed word of the forward strand = the 32 k-mers that start in
// that word; a warp owns 32 consecutive packed words and locates its first sequence with one binary search
// over SeqMeta::word_off (monotone), the same scheme as pack_kernel.cu — so 150 bp reads and chromosome-sized
// contigs run on one path with every lane busy.
#include "dyibf_internal.h"
#include "ibf_hash.cuh"

namespace dyibf {

// Calls f(word_index, bit) for every (k-mer, hash function) of the 32 k-mers that start in packed word t0 + lane.
// A warp owns the 32 consecutive packed words [t0, t0+32).
template <int H, typename F>
__device__ __forceinline__ void for_each_update(const HashParams& hp, const uint64_t* __restrict__ packed,
                                                const SeqMeta* __restrict__ meta, const uint32_t* __restrict__ bin_ids,
                                                uint32_t n_seqs, uint64_t t0, uint64_t total_words, F&& f)
{
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t k = hp.k;
    const uint32_t h = H ? H : hp.h;
    const uint64_t t = t0 + lane;
    uint64_t lo = 0, hi = n_seqs;                  // word_off(lo) <= t0 < word_off(hi)
    while (hi - lo > 1) {
        uint64_t mid = (lo + hi) >> 1;
        if (meta[mid].word_off <= t0) lo = mid; else hi = mid;
    }
    uint64_t s = lo;
    if (t >= total_words) return;
    while (s + 1 < n_seqs && meta[s + 1].word_off <= t) ++s;
    const SeqMeta m = meta[s];
    if (m.len < k) return;                         // addFastaFile skips records shorter than k (:142)
    const uint64_t j = t - m.word_off;
    const uint64_t nk = (uint64_t)m.len - k + 1;
    const uint64_t i0 = j * 32;
    if (i0 >= nk) return;
    const uint32_t cnt = (uint32_t)((nk - i0) < 32 ? (nk - i0) : 32);
    const uint64_t hi_w = packed[t], lo_w = packed[t + 1];
    const uint32_t firstN = m.flags & 1u;
    const uint32_t bin = bin_ids[s];
    const uint32_t bit = bin & 63u;
    const uint64_t col = bin >> 6;
    #pragma unroll 4
    for (uint32_t x = 0; x < cnt; ++x) {
        const uint64_t i = i0 + x;
        // i can exceed 32 bits only for sequences > 4 Gbp; the leading-N term needs i < 32 anyway
        const uint64_t kv = kmer_value2(hi_w, lo_w, (uint32_t)(i & 31u) | (i >= 32 ? 32u : 0u), k, firstN);
        #pragma unroll
        for (int jh = 0; jh < (H ? H : DYIBF_MAX_HASH); ++jh) {
            if (jh < (int)h) f(block_index(kv, jh, hp) * hp.W + col, bit);
        }
    }
}

// ---- direct build: one fire-and-forget atomic OR per (k-mer, hash function) ------------------------------------
template <int H>
__global__ void __launch_bounds__(256) build_kernel(const HashParams hp, uint64_t* __restrict__ filter,
                                                    const uint64_t* __restrict__ packed, const SeqMeta* __restrict__ meta,
                                                    const uint32_t* __restrict__ bin_ids, uint32_t n_seqs, uint64_t total_words)
{
    const uint64_t n_warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    for (uint64_t tile = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; tile * 32 < total_words; tile += n_warps)
        for_each_update<H>(hp, packed, meta, bin_ids, n_seqs, tile * 32, total_words, [&](uint64_t widx, uint32_t bit) {
            atomicOr(reinterpret_cast<unsigned long long*>(filter + widx), 1ull << bit);
        });
}

cudaError_t launch_build(const HashParams& hp, uint64_t* filter, const uint64_t* packed, const SeqMeta* meta,
                         const uint32_t* bin_ids, uint32_t n_seqs, uint64_t total_words, cudaStream_t st)
{
    if (n_seqs == 0 || total_words == 0) return cudaSuccess;
    uint64_t tiles = (total_words + 31) / 32;
    uint64_t blocks = (tiles + 7) / 8;
    if (blocks > 148ull * 32) blocks = 148ull * 32;
    dim3 grid((unsigned)blocks), block(256);
    switch (hp.h) {
        case 2:  build_kernel<2><<<grid, block, 0, st>>>(hp, filter, packed, meta, bin_ids, n_seqs, total_words); break;
        case 3:  build_kernel<3><<<grid, block, 0, st>>>(hp, filter, packed, meta, bin_ids, n_seqs, total_words); break;
        case 4:  build_kernel<4><<<grid, block, 0, st>>>(hp, filter, packed, meta, bin_ids, n_seqs, total_words); break;
        case 5:  build_kernel<5><<<grid, block, 0, st>>>(hp, filter, packed, meta, bin_ids, n_seqs, total_words); break;
        default: build_kernel<0><<<grid, block, 0, st>>>(hp, filter, packed, meta, bin_ids, n_seqs, total_words); break;
    }
    return cudaGetLastError();
}

// ---- bucketed build -------------------------------------------------------------------------------------------
// The direct kernel's updates are single 64-bit words at uniformly random addresses of a multi-GiB table: every one
// costs a 32-byte DRAM sector read and, later, a sector write-back, and the rate of such random row accesses — not
// bytes — is what HBM limits (ncu: ~21 G updates/s = ~42 G random sector accesses/s, the same access rate the query
// kernel's 128-byte gathers reach).  The bucketed build turns the random accesses into streaming ones:
//   scatter : the filter's word range is cut into <= 1024 regions of 2^R words (16 MiB for a 16 GiB filter).  Every
//             update becomes a 32-bit record (word-in-region << 6 | bit) appended to its region's bucket in HBM.  A
//             block first histograms the updates of a tile of 1024 packed words in shared memory, reserves one run per
//             bucket with a single global atomic, then recomputes the hashes and writes the records into the runs
//             (~100 records = ~400 contiguous bytes per bucket per tile, merged into full sectors by the L2).
//   apply   : bucket by bucket, records stream in and are OR-ed into their region, which — touched by thousands of
//             updates within microseconds — lives in the 126 MB L2: DRAM sees each region sector once in, once out.
// Updates that do not fit a bucket (the hash is uniform, buckets get 6 % slack) fall back to the direct atomic, so
// the result never depends on the bucket capacity.  OR is commutative and idempotent: same filter, bit for bit.
static const uint32_t MAX_REGIONS  = 1024;
static const uint32_t SCATTER_TILE = 1024;      // packed words per block iteration (8 warps x 4 groups x 32 words)
static const uint32_t APPLY_CHUNK  = 8192;      // records per apply block

template <int H>
__global__ void __launch_bounds__(256) build_scatter_kernel(const HashParams hp, uint64_t* __restrict__ filter,
        const uint64_t* __restrict__ packed, const SeqMeta* __restrict__ meta, const uint32_t* __restrict__ bin_ids,
        uint32_t n_seqs, uint64_t word_begin, uint64_t total_words, uint32_t R, uint32_t n_regions, uint64_t cap,
        uint32_t* __restrict__ records, unsigned long long* __restrict__ cursors)
{
    // packed words [word_begin, total_words) of the batch; word_begin is a multiple of SCATTER_TILE
    __shared__ uint32_t hist[MAX_REGIONS];
    __shared__ unsigned long long base[MAX_REGIONS];
    const uint32_t warp = threadIdx.x >> 5;
    const uint64_t rmask = (1ull << R) - 1;
    for (uint64_t tile0 = word_begin + (uint64_t)blockIdx.x * SCATTER_TILE; tile0 < total_words; tile0 += (uint64_t)gridDim.x * SCATTER_TILE) {
        for (uint32_t i = threadIdx.x; i < n_regions; i += 256) hist[i] = 0;
        __syncthreads();
        for (uint32_t g = 0; g < SCATTER_TILE / 256; ++g) {
            const uint64_t t0 = tile0 + (uint64_t)(g * 8 + warp) * 32;
            if (t0 < total_words)
                for_each_update<H>(hp, packed, meta, bin_ids, n_seqs, t0, total_words, [&](uint64_t widx, uint32_t) {
                    atomicAdd(&hist[widx >> R], 1u);
                });
        }
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < n_regions; i += 256) {
            const uint32_t c = hist[i];
            base[i] = c ? atomicAdd(&cursors[i], (unsigned long long)c) : 0ull;
            hist[i] = 0;
        }
        __syncthreads();
        for (uint32_t g = 0; g < SCATTER_TILE / 256; ++g) {
            const uint64_t t0 = tile0 + (uint64_t)(g * 8 + warp) * 32;
            if (t0 < total_words)
                for_each_update<H>(hp, packed, meta, bin_ids, n_seqs, t0, total_words, [&](uint64_t widx, uint32_t bit) {
                    const uint32_t r = (uint32_t)(widx >> R);
                    const unsigned long long pos = base[r] + atomicAdd(&hist[r], 1u);
                    if (pos < cap) records[(uint64_t)r * cap + pos] = (uint32_t)((widx & rmask) << 6) | bit;
                    else atomicOr(reinterpret_cast<unsigned long long*>(filter + widx), 1ull << bit);
                });
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256) build_apply_kernel(uint64_t* __restrict__ filter, const uint32_t* __restrict__ records,
        const unsigned long long* __restrict__ cursors, uint32_t R, uint64_t cap, uint32_t blocks_per_region)
{
    const uint32_t r = blockIdx.x / blocks_per_region, j = blockIdx.x - r * blocks_per_region;
    unsigned long long cnt = cursors[r];
    if (cnt > cap) cnt = cap;
    const uint64_t begin = (uint64_t)j * APPLY_CHUNK;
    if (begin >= cnt) return;
    const uint64_t end = begin + APPLY_CHUNK < cnt ? begin + APPLY_CHUNK : cnt;
    unsigned long long* region = reinterpret_cast<unsigned long long*>(filter + ((uint64_t)r << R));
    const uint32_t* rec = records + (uint64_t)r * cap;
    // four records per load (buckets are 128-byte aligned), streaming: the records are read once
    const uint64_t end4 = begin + ((end - begin) & ~3ull);
    for (uint64_t i = begin + 4ull * threadIdx.x; i < end4; i += 4 * 256) {
        const uint4 v = __ldcs(reinterpret_cast<const uint4*>(rec + i));
        atomicOr(region + (v.x >> 6), 1ull << (v.x & 63u));
        atomicOr(region + (v.y >> 6), 1ull << (v.y & 63u));
        atomicOr(region + (v.z >> 6), 1ull << (v.z & 63u));
        atomicOr(region + (v.w >> 6), 1ull << (v.w & 63u));
    }
    for (uint64_t i = end4 + threadIdx.x; i < end; i += 256) {
        const uint32_t v = __ldcs(rec + i);
        atomicOr(region + (v >> 6), 1ull << (v & 63u));
    }
}

// geometry of the bucketed build for a filter of data_words words and a record budget of budget_bytes
bool bucket_plan(uint64_t data_words, uint64_t budget_bytes, uint32_t* R_out, uint32_t* n_regions_out, uint64_t* cap_out)
{
    uint32_t R = 21;                                              // 2^21 words = 16 MiB regions
    while (((data_words + (1ull << R) - 1) >> R) > MAX_REGIONS) ++R;
    if (R > 26) return false;                                     // word-in-region must fit 26 bits (+ 6 bits of bit index)
    const uint64_t n_regions = (data_words + (1ull << R) - 1) >> R;
    if (n_regions < 8) return false;                              // fits the L2 anyway: direct atomics
    uint64_t cap = budget_bytes / 4 / n_regions;
    cap -= cap % 32;                                              // 128-byte aligned buckets
    if (cap < 65536) return false;
    *R_out = R; *n_regions_out = (uint32_t)n_regions; *cap_out = cap;
    return true;
}

cudaError_t launch_build_scatter(const HashParams& hp, uint64_t* filter, const uint64_t* packed, const SeqMeta* meta,
                                 const uint32_t* bin_ids, uint32_t n_seqs, uint64_t word_begin, uint64_t total_words, uint32_t R,
                                 uint32_t n_regions, uint64_t cap, uint32_t* records, unsigned long long* cursors, cudaStream_t st)
{
    if (n_seqs == 0 || total_words <= word_begin) return cudaSuccess;
    if (word_begin % SCATTER_TILE) return cudaErrorInvalidValue;
    uint64_t blocks = (total_words - word_begin + SCATTER_TILE - 1) / SCATTER_TILE;
    if (blocks > 148ull * 8) blocks = 148ull * 8;
    dim3 grid((unsigned)blocks), block(256);
#define DYIBF_SC(HH) build_scatter_kernel<HH><<<grid, block, 0, st>>>(hp, filter, packed, meta, bin_ids, n_seqs, word_begin, total_words, R, n_regions, cap, records, cursors)
    switch (hp.h) {
        case 2: DYIBF_SC(2); break;
        case 3: DYIBF_SC(3); break;
        case 4: DYIBF_SC(4); break;
        case 5: DYIBF_SC(5); break;
        default: DYIBF_SC(0); break;
    }
#undef DYIBF_SC
    return cudaGetLastError();
}

cudaError_t launch_build_apply(uint64_t* filter, const uint32_t* records, unsigned long long* cursors, uint32_t R,
                               uint32_t n_regions, uint64_t cap, uint64_t max_fill, cudaStream_t st)
{
    // max_fill: upper bound of any bucket's fill known to the host (<= cap); bounds the grid
    if (max_fill > cap) max_fill = cap;
    if (max_fill == 0) return cudaSuccess;
    const uint32_t bpr = (uint32_t)((max_fill + APPLY_CHUNK - 1) / APPLY_CHUNK);
    build_apply_kernel<<<n_regions * bpr, 256, 0, st>>>(filter, records, cursors, R, cap, bpr);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    return cudaMemsetAsync(cursors, 0, (size_t)n_regions * sizeof(unsigned long long), st);
}

// one thread per (hash block, touched word column)
__global__ void __launch_bounds__(256) clear_kernel(uint64_t* __restrict__ filter, uint64_t P, uint32_t W,
                                                    const uint32_t* __restrict__ cols, const uint64_t* __restrict__ masks,
                                                    uint32_t n_cols)
{
    const uint64_t total = P * n_cols;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) {
        const uint64_t p = t / n_cols;
        const uint32_t c = (uint32_t)(t - p * n_cols);
        uint64_t* w = filter + p * W + cols[c];
        *w &= ~masks[c];
    }
}

cudaError_t launch_clear(const HashParams& hp, uint64_t* filter, const uint32_t* cols, const uint64_t* masks, uint32_t n_cols,
                         cudaStream_t st)
{
    if (n_cols == 0) return cudaSuccess;
    uint64_t total = hp.P * n_cols;
    uint64_t blocks = (total + 255) / 256;
    if (blocks > 148ull * 16) blocks = 148ull * 16;
    clear_kernel<<<(unsigned)blocks, 256, 0, st>>>(filter, hp.P, hp.W, cols, masks, n_cols);
    return cudaGetLastError();
}

} // namespace dyibf
