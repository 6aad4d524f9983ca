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
    if (bin >= hp.bins) {                          // device-resident bin ids the host could not check: flag, add nothing
        if (hp.err) atomicOr(hp.err, DYIBF_DEVERR_BIN_RANGE);
        return;
    }
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
    if (blocks > sm_count_current() * 32) blocks = sm_count_current() * 32;
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

// (the bucketed build — scatter to per-region buckets, apply region by region — lives in build_bucket.cu)

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
    if (blocks > sm_count_current() * 16) blocks = sm_count_current() * 16;
    clear_kernel<<<(unsigned)blocks, 256, 0, st>>>(filter, hp.P, hp.W, cols, masks, n_cols);
    return cudaGetLastError();
}

} // namespace dyibf
