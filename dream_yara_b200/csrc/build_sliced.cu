// build_sliced.cu — the bit-sliced IBF build: _insertKmer's single-bit ORs (d_bloom_filter.h:293-302) go into a per-bin
// Bloom bit vector that lives in the L2, and the vectors are then interleaved into the reference's word layout.
// Same filter, bit for bit, as the direct atomics of build_kernel.cu (OR is commutative and idempotent).
//
// Why: an IBF is B plain Bloom filters of P bits each, interleaved — bit `bin` of hash block p.  The updates of ONE bin hit
// the same bit of the same word column in blocks p chosen by the hash: in the interleaved table that is one word out of
// every touched 8W-byte block, 16 GiB apart in total (cfg2), every update its own DRAM sector; in the bin's own bit vector
// the same updates land in P/8 bytes = 16 MiB, which the 126 MB L2 holds.  Genomes arrive bin by bin (build_filter adds
// one FASTA file per bin, d_build_filter.cpp:227-260), so at any moment the resident warps work on two or three bins and
// their vectors stay in the L2 while they fill: the REDs are L2 hits, DRAM only sees each vector stream in and out once.
//
//   pass, for a group of word columns (all W of them when the scratch fits: 16 GiB at 1024 bins / P = 2^27; else 4, 2 or 1):
//     1. zero the shadow
//     2. sliced_scatter_kernel: every k-mer of every sequence whose bin is in the group -> h REDs (32-bit OR) into the
//        bin's bit vector
//     3. sliced_interleave_kernel: 64 bins x 32 blocks bit matrices are transposed with 32 x 32 butterflies (shuffles),
//        collected in shared memory and OR-ed into the table as whole rows of the group's columns
//   and the next group.  Sequences may come in any order (a pass picks its own by bin); the speed-up needs them clustered
//   by bin, which the host driver checks before it chooses this path over the bucketed one (build_bucket.cu).
//
// Shadow layout: [bin of the group][P / 32 words], every bin's vector contiguous (16 MiB at cfg2) and a whole number of
// 128-byte lines.  The interleave takes tiles of 1024 blocks: one line from each bin's vector (rows 16 MiB apart, but whole
// lines), transposed through shared memory into whole table rows.  (A tile-major layout — [tile][bin][32 words], everything
// a tile needs contiguous — was tried and dropped: the scatter then spreads one bin's updates over the whole 16 GiB scratch
// and ran four times slower, 258 instead of 66 ms per 4 Gbp, with an odd line stride as well: the REDs need the bin's
// vector contiguous, i.e. few pages, not only few lines.  The first interleave read one 32-byte sector per bin and updated
// 32 of the 128 bytes of every table row: ncu showed 26.9 GB of DRAM reads per 4-column pass where 8.6 GB were needed.)
//
// Against the bucketed build (two radix-split levels + apply: 58 + 49 + 69 ms per 4 Gbp step on a B200) the two split
// levels disappear: what remains is one RED per update — the same L1TEX-bound rate as the bucketed build's apply — plus
// streaming passes for zero / interleave.
#include "dyibf_internal.h"
#include <cstdlib>
#include "ibf_hash.cuh"

namespace dyibf {

static const uint32_t FULLW = 0xffffffffu;

// ---- scatter: k-mers -> REDs into the group's bit vectors ---------------------------------------------------------
// A warp owns the 32 consecutive packed words [t0, t0+32) (= 1024 k-mer start positions).  Tiles are CLAIMED in batch order
// from a global counter, so the tiles in flight are always the next n_warps ones = a window of two or three bins, however
// unevenly the warps progress.  (A static grid-stride walk lets warps drift apart by whole iterations, i.e. by multiples of
// n_warps tiles: measured 30 G updates/s, the vectors of hundreds of bins in flight, against 130 G updates/s claimed.)
template <int H>
__global__ void __launch_bounds__(256) sliced_scatter_kernel(const HashParams hp, uint32_t* __restrict__ shadow, uint64_t pw,
        uint32_t first_bin, uint32_t group_bins, const uint64_t* __restrict__ packed, const SeqMeta* __restrict__ meta,
        const uint32_t* __restrict__ bin_ids, uint32_t n_seqs, uint64_t total_words, unsigned long long* __restrict__ work_counter)
{
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t k = hp.k;
    const uint32_t h = H ? H : hp.h;
    uint64_t hint = 0;
    for (;;) {
        unsigned long long tile = 0;
        if (lane == 0) tile = atomicAdd(work_counter, 1ull);
        tile = __shfl_sync(FULLW, tile, 0);
        if (tile * 32 >= total_words) break;
        const uint64_t t0 = tile * 32, t = t0 + lane;
        // sequence of the warp's first word: the previous tile's answer is usually right or one off
        uint64_t s = hint;
        if (!(s < n_seqs && __ldg(&meta[s].word_off) <= t0 && (s + 1 == n_seqs || __ldg(&meta[s + 1].word_off) > t0))) {
            uint64_t lo = 0, hi = n_seqs;                           // word_off(lo) <= t0 < word_off(hi)
            while (hi - lo > 1) {
                const uint64_t mid = (lo + hi) >> 1;
                if (__ldg(&meta[mid].word_off) <= t0) lo = mid; else hi = mid;
            }
            s = lo;
        }
        hint = s;
        if (t >= total_words) continue;
        while (s + 1 < n_seqs && __ldg(&meta[s + 1].word_off) <= t) ++s;
        const uint32_t bin = __ldg(bin_ids + s);
        if (bin >= hp.bins) {                                       // device-resident bin ids the host could not check
            if (hp.err) atomicOr(hp.err, DYIBF_DEVERR_BIN_RANGE);
            continue;
        }
        if (bin - first_bin >= group_bins) continue;                // another pass's bin (unsigned: also bin < first_bin)
        const uint64_t word_off = __ldg(&meta[s].word_off);
        const uint32_t len = __ldg(&meta[s].len), flags = __ldg(&meta[s].flags);
        if (len < k) continue;                                      // addFastaFile skips records shorter than k (:142)
        const uint64_t nk = (uint64_t)len - k + 1;
        const uint64_t i0 = (t - word_off) * 32;
        if (i0 >= nk) continue;
        const uint32_t cnt = (uint32_t)((nk - i0) < 32 ? (nk - i0) : 32);
        const uint64_t hi_w = __ldg(packed + t), lo_w = __ldg(packed + t + 1);
        const uint32_t firstN = flags & 1u, i0_hi = i0 >= 32 ? 32u : 0u;
        uint32_t* vec = shadow + (uint64_t)(bin - first_bin) * pw;
        #pragma unroll 4
        for (uint32_t x = 0; x < cnt; ++x) {
            const uint64_t kv = kmer_value2(hi_w, lo_w, x | i0_hi, k, firstN);
            #pragma unroll
            for (int j = 0; j < (H ? H : DYIBF_MAX_HASH); ++j) {
                if (j < (int)h) {
                    const uint64_t blk = block_index(kv, j, hp);
                    atomicOr(vec + (blk >> 5), 1u << ((uint32_t)blk & 31u));
                }
            }
        }
    }
}

// ---- interleave: shadow[bin][block] -> table[block * W + bin / 64] bit bin % 64 ---------------------------------------
// in-place transpose of a 32 x 32 bit matrix held one row per lane: afterwards lane j holds column j (bit i = old row i's
// bit j).  Five butterfly stages; a lane keeps or swaps half of its bits with lane ^ d.
__device__ __forceinline__ uint32_t transpose32(uint32_t x, uint32_t lane)
{
    #pragma unroll
    for (int s = 4; s >= 0; --s) {
        const uint32_t d = 1u << s;
        // mask of bit positions whose bit s is 0
        const uint32_t m = s == 4 ? 0x0000ffffu : s == 3 ? 0x00ff00ffu : s == 2 ? 0x0f0f0f0fu : s == 1 ? 0x33333333u : 0x55555555u;
        const uint32_t y = __shfl_xor_sync(FULLW, x, d);
        // rows with bit s clear keep their low-d columns and take the partner's low-d columns into their high ones; rows
        // with bit s set the other way round
        x = (lane & d) ? ((x & ~m) | ((y >> d) & m)) : ((x & m) | ((y << d) & ~m));
    }
    return x;
}

// A block handles one shadow tile (1024 blocks x the group's bins) in sub-tiles of SW * 32 blocks.  Warp w takes the word
// columns w, w + 8, ...: lane l loads SW words (SW * 4 contiguous bytes) of bins c*64 + l and c*64 + 32 + l; after the
// transposes lane j holds, for each of the SW words, the 64-bit table word of block (word * 32 + j) in column c, which goes to
// shared memory [block][column].  The block then ORs the sub-tile into the table: SW * 32 consecutive rows of n_cols words —
// whole 128-byte rows, coalesced, when the group spans all W columns.
template <int SW>
__global__ void __launch_bounds__(256, SW <= 8 ? 4 : 3) sliced_interleave_kernel(uint64_t* __restrict__ filter, const uint32_t* __restrict__ shadow,
        uint64_t pw, uint64_t n_tiles, uint64_t P, uint32_t W, uint32_t first_col, uint32_t n_cols)
{
    extern __shared__ uint64_t il_sm[];                              // [SW * 32 rows][stride]
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    const uint32_t stride = n_cols | 1u;                             // odd: the transposed stores of a warp spread over the banks
    const uint32_t rows = SW * 32;
    for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const uint32_t* base = shadow + tile * 32;                    // word 0 of the tile in bin 0's vector
        for (uint32_t sub = 0; sub < 32 / SW; ++sub) {
            bool any = false;
            for (uint32_t c = warp; c < n_cols; c += n_warps) {
                uint32_t lo[SW], hi[SW];
                const uint4* plo = reinterpret_cast<const uint4*>(base + (uint64_t)(c * 64 + lane) * pw + sub * SW);
                const uint4* phi = reinterpret_cast<const uint4*>(base + (uint64_t)(c * 64 + 32 + lane) * pw + sub * SW);
                #pragma unroll
                for (int q = 0; q < SW / 4; ++q) {
                    const uint4 a = __ldcs(plo + q), b = __ldcs(phi + q);
                    lo[4 * q] = a.x; lo[4 * q + 1] = a.y; lo[4 * q + 2] = a.z; lo[4 * q + 3] = a.w;
                    hi[4 * q] = b.x; hi[4 * q + 1] = b.y; hi[4 * q + 2] = b.z; hi[4 * q + 3] = b.w;
                }
                #pragma unroll
                for (int w = 0; w < SW; ++w) {
                    // sparse passes: skip the butterflies when nobody has a bit
                    const bool some = __any_sync(FULLW, (lo[w] | hi[w]) != 0);
                    uint64_t v = 0;
                    if (some) {
                        v = (uint64_t)transpose32(lo[w], lane) | ((uint64_t)transpose32(hi[w], lane) << 32);
                        any = true;
                    }
                    il_sm[(uint32_t)(w * 32 + lane) * stride + c] = v;
                }
            }
            const int nonzero = __syncthreads_or(any);
            if (nonzero) {
                const uint64_t blk0 = tile * 1024 + (uint64_t)sub * rows;
                const uint32_t rows_here = blk0 >= P ? 0u : (uint32_t)(P - blk0 < rows ? P - blk0 : rows);
                if (n_cols == W && (W & 1u) == 0) {
                    // the group spans whole table rows: the sub-tile is rows_here * W consecutive words.  16 bytes per thread
                    // and access, four accesses in flight per thread before the first store.
                    ulonglong2* T = reinterpret_cast<ulonglong2*>(filter + blk0 * W);
                    const uint32_t pairs = rows_here * (W >> 1), half = W >> 1;
                    for (uint32_t i0 = threadIdx.x; i0 < pairs; i0 += 4 * blockDim.x) {
                        ulonglong2 v[4], o[4];
                        #pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const uint32_t i = i0 + j * blockDim.x;
                            v[j] = make_ulonglong2(0, 0);
                            if (i < pairs) {
                                const uint32_t r = i / half, c = (i - r * half) << 1;
                                v[j].x = il_sm[r * stride + c]; v[j].y = il_sm[r * stride + c + 1];
                            }
                        }
                        #pragma unroll
                        for (int j = 0; j < 4; ++j)
                            if (v[j].x | v[j].y) {
                                const longlong2 t = __ldcs(reinterpret_cast<const longlong2*>(T + i0 + j * blockDim.x));
                                o[j] = make_ulonglong2((unsigned long long)t.x, (unsigned long long)t.y);
                            }
                        #pragma unroll
                        for (int j = 0; j < 4; ++j)
                            if (v[j].x | v[j].y)
                                __stcs(reinterpret_cast<longlong2*>(T + i0 + j * blockDim.x),
                                       make_longlong2((long long)(o[j].x | v[j].x), (long long)(o[j].y | v[j].y)));
                    }
                } else {
                    // a narrower group: n_cols words of every row.  Batches of 4 words per thread, loads before stores.
                    const uint32_t total = rows_here * n_cols;
                    for (uint32_t i0 = threadIdx.x; i0 < total; i0 += 4 * blockDim.x) {
                        uint64_t v[4], old[4];
                        #pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const uint32_t i = i0 + j * blockDim.x;
                            v[j] = 0;
                            if (i < total) { const uint32_t r = i / n_cols; v[j] = il_sm[r * stride + (i - r * n_cols)]; }
                        }
                        #pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const uint32_t i = i0 + j * blockDim.x, r = i / n_cols;
                            old[j] = v[j] ? __ldcs(filter + (blk0 + r) * W + first_col + (i - r * n_cols)) : 0;
                        }
                        #pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const uint32_t i = i0 + j * blockDim.x, r = i / n_cols;
                            if (v[j]) __stcs(filter + (blk0 + r) * W + first_col + (i - r * n_cols), old[j] | v[j]);
                        }
                    }
                }
            }
            __syncthreads();
        }
    }
}

// which column groups of `cols_per_group` word columns have at least one sequence: mask[g] = 1
__global__ void sliced_presence_kernel(const uint32_t* __restrict__ bin_ids, uint32_t n_seqs, uint32_t bins, uint32_t group_bins,
                                       uint32_t* __restrict__ mask, unsigned long long* __restrict__ switches)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_seqs) return;
    const uint32_t b = bin_ids[i];
    if (b >= bins) return;
    mask[b / group_bins] = 1u;
    if (i > 0) {
        const uint32_t pb = bin_ids[i - 1];
        if (pb < bins && pb != b) atomicAdd(switches, 1ull);       // bin changes along the batch: how clustered it is
    }
}

// ---- launchers ------------------------------------------------------------------------------------------------------
bool sliced_plan(const HashParams& hp, uint64_t budget_bytes, SlicedPlan* p)
{
    if (hp.direct) return false;
    const uint64_t pw = ((hp.P + 31) / 32 + 31) & ~31ull;           // 32-bit words of one bin's vector: whole 128-byte lines
    if (pw * 4 > (32ull << 20)) return false;                       // a bin's vector must sit comfortably in the L2
    if (pw * 4 < (64ull << 10)) return false;                       // tiny tables: direct atomics are L2 hits anyway
    const uint64_t n_tiles = pw / 32;                               // tiles of 1024 blocks = one line per bin
    // One group of all W columns when the scratch allows (the interleave then streams whole table rows: 16 GiB at cfg2).  Narrower
    // groups (W > 16, small budgets, DYIBF_SLICED_MAX_COLS) get TWO shadow buffers: while group g+1 is scattered, group g is
    // interleaved into the table and zeroed again on a second stream.  That overlap does not pay by itself — measured on
    // 1024 bins x 4 Mbp: four groups of 4 columns 105.8 ms per step against 83.0 ms for the one whole-table group (the
    // partial-row interleave is slow and contends with the scatter for the L2) — it only hides what narrow groups cost anyway.
    const char* env_cols = getenv("DYIBF_SLICED_MAX_COLS");   // per plan (one getenv per build call): the tests switch it inside one process
    uint32_t cols = env_cols && atoi(env_cols) > 0 ? (uint32_t)atoi(env_cols) : 16u;
    if (cols > 16) cols = 16;
    if (cols > hp.W) cols = hp.W;
    auto bytes_for = [&](uint32_t c) { return (uint64_t)(c < hp.W ? 2 : 1) * c * 64 * pw * 4; };
    while (cols > 1 && bytes_for(cols) > budget_bytes) cols >>= 1;
    if (bytes_for(cols) > budget_bytes) return false;
    p->pw = pw; p->n_tiles = n_tiles; p->cols_per_group = cols; p->group_bins = cols * 64;
    p->n_groups = (hp.W + cols - 1) / cols;
    p->shadow_bytes = (uint64_t)cols * 64 * pw * 4;
    p->n_buffers = p->n_groups > 1 ? 2 : 1;
    p->scatter_blocks = SLICED_MAX_BLOCKS;
    return true;
}

cudaError_t launch_sliced_presence(const uint32_t* bin_ids, uint32_t n_seqs, uint32_t bins, const SlicedPlan& p, uint32_t* mask,
                                   unsigned long long* switches, cudaStream_t st)
{
    cudaError_t e = cudaMemsetAsync(mask, 0, (size_t)p.n_groups * 4, st);
    if (e == cudaSuccess) e = cudaMemsetAsync(switches, 0, 8, st);
    if (e != cudaSuccess || n_seqs == 0) return e;
    sliced_presence_kernel<<<(n_seqs + 255) / 256, 256, 0, st>>>(bin_ids, n_seqs, bins, p.group_bins, mask, switches);
    return cudaGetLastError();
}

cudaError_t launch_sliced_zero(void* shadow, const SlicedPlan& p, cudaStream_t st)
{
    return cudaMemsetAsync(shadow, 0, p.shadow_bytes, st);
}

cudaError_t launch_sliced_scatter(const HashParams& hp, void* shadow, const SlicedPlan& p, uint32_t group, const uint64_t* packed,
                                  const SeqMeta* meta, const uint32_t* bin_ids, uint32_t n_seqs, uint64_t total_words,
                                  unsigned long long* work_counter, cudaStream_t st)
{
    if (n_seqs == 0 || total_words == 0) return cudaSuccess;
    cudaError_t ce = cudaMemsetAsync(work_counter, 0, sizeof(unsigned long long), st);
    if (ce != cudaSuccess) return ce;
    const uint64_t tiles = (total_words + 31) / 32;
    uint64_t blocks = (tiles + 7) / 8;
    // The window of bins in flight must stay small (plan.scatter_blocks, set from the batch's average run per bin): measured on
    // 1024 bins x 1 Mbp into 16 GiB, 1184 / 592 / 296 / 148 blocks scatter in 60 / 46 / 32 / 23 ms (the RED rate does not
    // need more than 8 warps per SM, the L2 needs the vectors of the bins in flight to fit); 4 Mbp per bin: 66 / 66 / 67 / 69 ms.
    const uint64_t max_blocks = p.scatter_blocks ? p.scatter_blocks : SLICED_MAX_BLOCKS;
    if (blocks > max_blocks) blocks = max_blocks;
    static const char* env_blocks = getenv("DYIBF_SLICED_BLOCKS");  // experiments
    if (env_blocks && atoi(env_blocks) > 0 && blocks > (uint64_t)atoi(env_blocks)) blocks = (uint64_t)atoi(env_blocks);
    const uint32_t first_bin = group * p.group_bins;
    uint32_t* sh = static_cast<uint32_t*>(shadow);
    dim3 grid((unsigned)blocks), block(256);
    switch (hp.h) {
        case 2:  sliced_scatter_kernel<2><<<grid, block, 0, st>>>(hp, sh, p.pw, first_bin, p.group_bins, packed, meta, bin_ids, n_seqs, total_words, work_counter); break;
        case 3:  sliced_scatter_kernel<3><<<grid, block, 0, st>>>(hp, sh, p.pw, first_bin, p.group_bins, packed, meta, bin_ids, n_seqs, total_words, work_counter); break;
        case 4:  sliced_scatter_kernel<4><<<grid, block, 0, st>>>(hp, sh, p.pw, first_bin, p.group_bins, packed, meta, bin_ids, n_seqs, total_words, work_counter); break;
        default: sliced_scatter_kernel<0><<<grid, block, 0, st>>>(hp, sh, p.pw, first_bin, p.group_bins, packed, meta, bin_ids, n_seqs, total_words, work_counter); break;
    }
    return cudaGetLastError();
}

cudaError_t launch_sliced_interleave(const HashParams& hp, uint64_t* filter, const void* shadow, const SlicedPlan& p, uint32_t group,
                                     cudaStream_t st)
{
    const uint32_t first_col = group * p.cols_per_group;
    const uint32_t n_cols = hp.W - first_col < p.cols_per_group ? hp.W - first_col : p.cols_per_group;
    // words per lane and sub-tile: 8 (32 bytes per lane and bin, 256 table rows per sub-tile: measured 18.6 / 23.4 / 29.1 ms per 16 GiB
    // pass for 8 / 16 / 32 — the smaller the shared-memory tile, the more blocks per SM) unless DYIBF_SLICED_SW says 16 or 32
    static const int sw = getenv("DYIBF_SLICED_SW") ? atoi(getenv("DYIBF_SLICED_SW")) : 8;
    const uint32_t* sh = static_cast<const uint32_t*>(shadow);
    uint64_t blocks = p.n_tiles;
    if (blocks > sm_count_current() * 16) blocks = sm_count_current() * 16;
#define DYIBF_IL(SWV) do { \
        const size_t smem = (size_t)SWV * 32 * (n_cols | 1u) * 8; \
        if (smem > 48 * 1024) {                                     /* per device, so per launch (a group spans several GPUs) */ \
            cudaError_t e = cudaFuncSetAttribute(sliced_interleave_kernel<SWV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
            if (e != cudaSuccess) return e; \
        } \
        sliced_interleave_kernel<SWV><<<(unsigned)blocks, 256, smem, st>>>(filter, sh, p.pw, p.n_tiles, hp.P, hp.W, first_col, n_cols); \
    } while (0)
    if (sw == 16) DYIBF_IL(16); else if (sw == 32) DYIBF_IL(32); else DYIBF_IL(8);
#undef DYIBF_IL
    return cudaGetLastError();
}

} // namespace dyibf
