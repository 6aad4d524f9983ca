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
//   pass, for a group of up to 4 word columns (256 bins; shadow = 256 x P bits = 4 GiB at P = 2^27):
//     1. zero the shadow
//     2. sliced_scatter_kernel: every k-mer of every sequence whose bin is in the group -> h REDs (32-bit OR) into
//        shadow[bin - first][block / 32]
//     3. sliced_interleave_kernel: 64 bins x 256 blocks tiles are transposed with 32 x 32 bit-matrix butterflies (shuffles)
//        and OR-ed into the table, 4 columns = one full 32-byte sector per block and lane
//   and the next group.  Sequences may come in any order (a pass picks its own by bin); the speed-up needs them clustered
//   by bin, which the host driver checks before it chooses this path over the bucketed one (build_bucket.cu).
//
// Against the bucketed build (two radix-split levels + apply: 58 + 49 + 69 ms per 4 Gbp step on a B200) the two split
// levels disappear: what remains is one RED per update — the same L1TEX-bound rate as the bucketed build's apply — plus
// ~50 GiB of streaming for the zero / interleave passes.
#include "dyibf_internal.h"
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

// A warp handles one tile: `NC` word columns x 256 blocks (8 shadow words per bin).  Lane l first holds the 8 words of bins
// c*64 + l and c*64 + 32 + l of column c (two 32-byte loads: full sectors); after the transposes lane j holds, for each of
// the 8 sub-blocks w, the 64-bit table word of block q0*32 + w*32 + j in column c.  With NC = 4 a lane then owns 4 adjacent
// words = one 32-byte sector of the table per block: one read, OR, one write.
template <int NC>
__global__ void __launch_bounds__(256) sliced_interleave_kernel(uint64_t* __restrict__ filter, const uint32_t* __restrict__ shadow,
        uint64_t pw, uint64_t P, uint32_t W, uint32_t first_col, uint32_t n_cols, uint32_t group_bins)
{
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t n_tiles = (pw + 7) / 8;                          // tiles of 8 shadow words = 256 blocks
    const uint64_t n_warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    for (uint64_t tile = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; tile < n_tiles; tile += n_warps) {
        const uint64_t q0 = tile * 8;
        uint64_t out[NC][8];
        bool any = false;
        #pragma unroll
        for (int c = 0; c < NC; ++c) {
            uint32_t lo[8], hi[8];
            #pragma unroll
            for (int w = 0; w < 8; ++w) { lo[w] = 0; hi[w] = 0; }
            if ((uint32_t)c < n_cols) {
                const uint32_t b_lo = c * 64 + lane, b_hi = b_lo + 32;
                // pw is a multiple of 8, shadow rows are 32-byte aligned
                if (b_lo < group_bins) {
                    const uint4* p = reinterpret_cast<const uint4*>(shadow + (uint64_t)b_lo * pw + q0);
                    const uint4 a = __ldcs(p), b = __ldcs(p + 1);
                    lo[0] = a.x; lo[1] = a.y; lo[2] = a.z; lo[3] = a.w; lo[4] = b.x; lo[5] = b.y; lo[6] = b.z; lo[7] = b.w;
                }
                if (b_hi < group_bins) {
                    const uint4* p = reinterpret_cast<const uint4*>(shadow + (uint64_t)b_hi * pw + q0);
                    const uint4 a = __ldcs(p), b = __ldcs(p + 1);
                    hi[0] = a.x; hi[1] = a.y; hi[2] = a.z; hi[3] = a.w; hi[4] = b.x; hi[5] = b.y; hi[6] = b.z; hi[7] = b.w;
                }
            }
            #pragma unroll
            for (int w = 0; w < 8; ++w) {
                // most tiles of a sparse pass are empty: skip the butterflies when nobody has a bit
                const bool some = __any_sync(FULLW, (lo[w] | hi[w]) != 0);
                uint64_t v = 0;
                if (some) {
                    v = (uint64_t)transpose32(lo[w], lane) | ((uint64_t)transpose32(hi[w], lane) << 32);
                    any = true;
                }
                out[c][w] = v;
            }
        }
        if (!any) continue;
        #pragma unroll
        for (int w = 0; w < 8; ++w) {
            const uint64_t blk = (q0 + w) * 32 + lane;
            if (blk >= P) continue;
            uint64_t* dst = filter + blk * W + first_col;
            if (NC == 4 && ((W | first_col) & 3u) == 0) {           // one aligned 32-byte sector: two 16-byte accesses
                if ((out[0][w] | out[1][w] | out[NC - 2][w] | out[NC - 1][w]) == 0) continue;
                ulonglong2* d2 = reinterpret_cast<ulonglong2*>(dst);
                ulonglong2 a = d2[0], b = d2[1];
                a.x |= out[0][w]; a.y |= out[1][w]; b.x |= out[NC - 2][w]; b.y |= out[NC - 1][w];
                d2[0] = a; d2[1] = b;
            } else {
                #pragma unroll
                for (int c = 0; c < NC; ++c)
                    if ((uint32_t)c < n_cols && out[c][w]) dst[c] |= out[c][w];
            }
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
    const uint64_t pw = ((hp.P + 31) / 32 + 7) & ~7ull;             // 32-bit words per bin vector, 32-byte aligned rows
    if (pw * 4 > (32ull << 20)) return false;                       // a bin's vector must sit comfortably in the L2
    if (pw * 4 < (64ull << 10)) return false;                       // tiny tables: direct atomics are L2 hits anyway
    uint32_t cols = hp.W >= 4 ? 4 : hp.W >= 2 ? 2 : 1;              // the interleave kernel comes in these widths
    while (cols > 1 && (uint64_t)cols * 64 * pw * 4 > budget_bytes) cols >>= 1;
    if ((uint64_t)cols * 64 * pw * 4 > budget_bytes) return false;
    p->pw = pw; p->cols_per_group = cols; p->group_bins = cols * 64;
    p->n_groups = (hp.W + cols - 1) / cols;
    p->shadow_bytes = (uint64_t)cols * 64 * pw * 4;
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
    if (blocks > 148ull * 8) blocks = 148ull * 8;                   // 64 warps per SM: the window of bins in flight stays small
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
    const uint32_t first_bin = group * p.group_bins;
    const uint32_t gbins = hp.bins - first_bin < p.group_bins ? hp.bins - first_bin : p.group_bins;
    const uint64_t tiles = (p.pw + 7) / 8;
    uint64_t blocks = (tiles + 7) / 8;
    if (blocks > 148ull * 8) blocks = 148ull * 8;
    const uint32_t* sh = static_cast<const uint32_t*>(shadow);
    if (p.cols_per_group == 4)      sliced_interleave_kernel<4><<<(unsigned)blocks, 256, 0, st>>>(filter, sh, p.pw, hp.P, hp.W, first_col, n_cols, gbins);
    else if (p.cols_per_group == 2) sliced_interleave_kernel<2><<<(unsigned)blocks, 256, 0, st>>>(filter, sh, p.pw, hp.P, hp.W, first_col, n_cols, gbins);
    else                            sliced_interleave_kernel<1><<<(unsigned)blocks, 256, 0, st>>>(filter, sh, p.pw, hp.P, hp.W, first_col, n_cols, gbins);
    return cudaGetLastError();
}

} // namespace dyibf
