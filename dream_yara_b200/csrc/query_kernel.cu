// query_kernel.cu — the IBF lookup: k-mer values -> h hash blocks -> AND -> per-bin counts -> threshold.
//
// Reference loop replaced: SeqAnBloomFilter::whichBins, src/d_bloom_filter.h:161-224, as driven per read by
// clasifyLoadedReads, src/d_mapper.h:338-359 (fwd, rc [, mate fwd, mate rc] OR-ed into one selection).
//
// Mapping to the hardware (B200, sm_100a):
//   * A hash block is W u64 words = the filter bits of all bins for one hash position.  It is cut into
//     VEC-byte chunks (16 B when W is even -> LDG.128, else 8 B); a sub-group of T = 2^TL2 lanes owns one
//     sequence and lane `li` owns chunk `li` of every block, so one warp-wide load instruction fetches
//     32/T different blocks (for 1024 bins: T = 8, four 128-byte blocks = four full cache lines per LDG.128).
//     Blocks wider than 32 chunks are processed in column tiles of 32 chunks.
//   * k-mer values need no rolling state (ibf_hash.cuh), so the T lanes of a sub-group compute the hash
//     positions of T different k-mers at once and broadcast them with width-T shuffles.
//   * Loads are issued in batches of U k-mers x h blocks before anything is consumed (memory-level
//     parallelism is what the random 128-byte gathers from a multi-GiB table need), through the
//     non-coherent path with L1 no-allocate: there is no reuse to cache.
//   * Counting is bit-sliced: lane-private vertical counters, PLANES bit-planes per 32 bins, fed by
//     carry-save adders (LOP3 xor3/majority).  No shared memory, no atomics, no popc/ctz loop; the
//     reference's `counts[bin]` is the number whose bit p is bit `bin%32` of plane p.
//   * `count >= threshold` is evaluated bit-sliced as well (borrow chain of count - threshold).
#include "dyibf_internal.h"
#include "ibf_hash.cuh"
#include <type_traits>

namespace dyibf {

template <int VEC> struct Chunk;
template <> struct Chunk<8>  { uint32_t w[2]; };
template <> struct Chunk<16> { uint32_t w[4]; };

template <int VEC>
__device__ __forceinline__ Chunk<VEC> ld_chunk(const uint8_t* p);
template <>
__device__ __forceinline__ Chunk<8> ld_chunk<8>(const uint8_t* p)
{
    Chunk<8> c;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(c.w[0]), "=r"(c.w[1]) : "l"(p));
    return c;
}
template <>
__device__ __forceinline__ Chunk<16> ld_chunk<16>(const uint8_t* p)
{
    Chunk<16> c;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(c.w[0]), "=r"(c.w[1]), "=r"(c.w[2]), "=r"(c.w[3]) : "l"(p));
    return c;
}

__device__ __forceinline__ uint32_t xor3(uint32_t a, uint32_t b, uint32_t c) { return a ^ b ^ c; }
__device__ __forceinline__ uint32_t maj3(uint32_t a, uint32_t b, uint32_t c) { return (a & b) | (c & (a | b)); }

// add the carry word `c` (weight 2^from) into planes[from..]
template <int PLANES>
__device__ __forceinline__ void ripple(uint32_t (&pl)[PLANES], int from, uint32_t c)
{
    #pragma unroll
    for (int p = 0; p < PLANES; ++p) {
        if (p >= from) { uint32_t t = pl[p] & c; pl[p] ^= c; c = t; }
    }
}

// count >= thr per bit position (thr < 2^PLANES): borrow of (count - thr), LSB first
template <int PLANES>
__device__ __forceinline__ uint32_t ge_threshold(const uint32_t (&pl)[PLANES], uint32_t thr)
{
    uint32_t borrow = 0;
    #pragma unroll
    for (int p = 0; p < PLANES; ++p)
        borrow = ((thr >> p) & 1u) ? (~pl[p] | borrow) : (~pl[p] & borrow);
    return ~borrow;
}

// H > 0: compile-time number of hash functions, 32-bit block indices, U = 4 k-mers in flight.
// H == 0: run-time hp.h (<= DYIBF_MAX_HASH), 64-bit block indices, U = 2.
template <int VEC, int TL2, int H, int PLANES>
__global__ void __launch_bounds__(128, 4)
query_kernel(const HashParams hp, const uint8_t* __restrict__ filter, const uint64_t* __restrict__ packed,
             const SeqMeta* __restrict__ meta, uint32_t n_rows, uint32_t group, uint32_t thr,
             uint64_t* __restrict__ out_rows, uint16_t* __restrict__ out_counts, uint32_t* __restrict__ work_counter,
             const uint32_t flags, unsigned long long* __restrict__ lookup_stat)
{
    constexpr int T  = 1 << TL2;            // lanes per sequence
    constexpr int SG = 32 / T;              // sequences per warp
    constexpr int NW = VEC / 4;             // u32 words per lane
    constexpr int U  = H ? 4 : 2;           // k-mers per load batch
    constexpr int KR = T > U ? T : U;       // k-mers per hash round
    constexpr int M  = KR / T;              // hash computations per lane per round
    constexpr int HM = H ? H : DYIBF_MAX_HASH;
    typedef typename std::conditional<H != 0, uint32_t, uint64_t>::type idx_t;

    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t li   = lane & (T - 1);
    const uint32_t sub  = lane >> TL2;
    const uint32_t h    = H ? H : hp.h;
    const uint32_t k    = hp.k;
    const uint32_t block_bytes = hp.W * 8u;
    const uint32_t chunks = block_bytes / VEC;
    const uint32_t n_tiles = (chunks + T - 1) / T;
    // Work items (SG rows each) are claimed from a global counter: SMs do not all see the same HBM latency
    // (two dies, near/far L2 slices), so a static split leaves the fast ones idle at the end.  The next claim
    // is issued before the current item is processed, so its round trip is off the critical path.
    const uint32_t n_items = (n_rows + SG - 1) / SG;
    const bool early_exit = (flags & QUERY_EARLY_EXIT) != 0 && out_counts == nullptr;
    const uint32_t sub_mask = (T == 32 ? 0xffffffffu : ((1u << T) - 1u)) << (sub * T);
    uint32_t n_lookups = 0;                 // k-mer lookups this lane took part in (statistics only)
    uint32_t next_item = 0;
    if (lane == 0) next_item = atomicAdd(work_counter, 1u);
    for (;;) {
        const uint32_t item = __shfl_sync(0xffffffffu, next_item, 0);
        if (item >= n_items) break;
        if (lane == 0) next_item = atomicAdd(work_counter, 1u);
        const uint64_t r0 = (uint64_t)item * SG;
        const uint64_t r = r0 + sub;
        const bool row_ok = r < n_rows;
        for (uint32_t tile = 0; tile < n_tiles; ++tile) {
            const uint32_t chunk = tile * T + li;
            const bool lane_on = row_ok && chunk < chunks;
            const uint8_t* lane_base = filter + (size_t)chunk * VEC;
            uint32_t sel[NW];
            #pragma unroll
            for (int w = 0; w < NW; ++w) sel[w] = 0;

            for (uint32_t g = 0; g < group; ++g) {
                const uint64_t s = r + (uint64_t)g * n_rows;
                uint32_t nk = 0, firstN = 0;
                const uint64_t* wp = packed;
                if (row_ok) {
                    const SeqMeta m = meta[s];
                    wp = packed + m.word_off;
                    firstN = m.flags & 1u;
                    nk = m.len >= k ? m.len - k + 1 : 0;
                    if (PLANES < 16 && nk >= (1u << PLANES)) {
                        // the caller's max_seq_len did not bound this sequence (device mode: nothing could check it): its
                        // counts would wrap in the 8-plane counters.  Flag it and let it select nothing.
                        if (li == 0 && tile == 0 && hp.err) atomicOr(hp.err, DYIBF_DEVERR_SEQ_TOO_LONG);
                        nk = 0;
                    }
                }
                const uint32_t nk_max = __reduce_max_sync(0xffffffffu, nk);

                uint32_t pl[NW][PLANES];
                #pragma unroll
                for (int w = 0; w < NW; ++w)
                    #pragma unroll
                    for (int p = 0; p < PLANES; ++p) pl[w][p] = 0;

                // Early exit (optional): the result of a sequence is decided as soon as every bin either reached the
                // threshold, was selected by an earlier strand/mate of the row, or cannot reach it any more with the
                // k-mers that are left; the remaining lookups cannot change `sel` and are skipped.  `open_mask` holds
                // the lanes of this warp whose sequence is still undecided.
                bool seq_open = nk > 0;
                if (early_exit && (thr == 0 || (PLANES < 32 && thr >= (1u << PLANES)))) seq_open = false;
                uint32_t open_mask = 0xffffffffu;
                for (uint32_t base = 0; base < nk_max && open_mask; base += KR) {
                    // --- hash round: lane li computes the block indices of k-mers base + li + m*T
                    idx_t blk[M][HM];
                    #pragma unroll
                    for (int m = 0; m < M; ++m) {
                        const uint32_t i = base + li + m * T;
                        #pragma unroll
                        for (int j = 0; j < HM; ++j) blk[m][j] = 0;
                        if (i < nk) {
                            const uint64_t kv = kmer_value(wp, i, k, firstN);
                            #pragma unroll
                            for (int j = 0; j < HM; ++j)
                                if (j < (int)h) blk[m][j] = (idx_t)block_index(kv, j, hp);
                        }
                    }
                    // --- load batches of U k-mers
                    #pragma unroll 1
                    for (uint32_t ub = 0; ub < KR; ub += U) {
                        if (base + ub >= nk_max || !open_mask) break;
                        uint32_t x[U][NW];
                        if constexpr (H != 0) {
                            Chunk<VEC> v[U][HM];
                            #pragma unroll
                            for (int u = 0; u < U; ++u) {
                                const uint32_t step = ub + u;
                                const bool on = lane_on && seq_open && (base + step < nk);
                                n_lookups += on;
                                #pragma unroll
                                for (int j = 0; j < HM; ++j) {
                                    const idx_t b = __shfl_sync(0xffffffffu, blk[M == 1 ? 0 : (u / T)][j],
                                                                M == 1 ? step : (u % T), T);
                                    if (on) v[u][j] = ld_chunk<VEC>(lane_base + (size_t)b * block_bytes);
                                    else {
                                        #pragma unroll
                                        for (int w = 0; w < NW; ++w) v[u][j].w[w] = 0;
                                    }
                                }
                            }
                            #pragma unroll
                            for (int u = 0; u < U; ++u)
                                #pragma unroll
                                for (int w = 0; w < NW; ++w) {
                                    uint32_t a = v[u][0].w[w];
                                    #pragma unroll
                                    for (int j = 1; j < HM; ++j) a &= v[u][j].w[w];
                                    x[u][w] = a;
                                }
                        } else {
                            bool on[U];
                            #pragma unroll
                            for (int u = 0; u < U; ++u) {
                                on[u] = lane_on && seq_open && (base + ub + u < nk);
                                n_lookups += on[u];
                                #pragma unroll
                                for (int w = 0; w < NW; ++w) x[u][w] = on[u] ? 0xffffffffu : 0u;
                            }
                            for (uint32_t j = 0; j < h; ++j) {
                                Chunk<VEC> v[U];
                                #pragma unroll
                                for (int u = 0; u < U; ++u) {
                                    const uint32_t step = ub + u;
                                    idx_t mine = 0;
                                    #pragma unroll
                                    for (int jj = 0; jj < HM; ++jj)          // register array: select, don't index
                                        if (jj == (int)j) mine = blk[M == 1 ? 0 : (u / T)][jj];
                                    const idx_t b = __shfl_sync(0xffffffffu, mine, M == 1 ? step : (u % T), T);
                                    if (on[u]) v[u] = ld_chunk<VEC>(lane_base + (size_t)b * block_bytes);
                                    else {
                                        #pragma unroll
                                        for (int w = 0; w < NW; ++w) v[u].w[w] = 0;
                                    }
                                }
                                #pragma unroll
                                for (int u = 0; u < U; ++u)
                                    #pragma unroll
                                    for (int w = 0; w < NW; ++w) x[u][w] &= v[u].w[w];
                            }
                        }
                        // --- bit-sliced accumulate
                        #pragma unroll
                        for (int w = 0; w < NW; ++w) {
                            if constexpr (U == 4) {
                                uint32_t ca = maj3(pl[w][0], x[0][w], x[1][w]);
                                pl[w][0]    = xor3(pl[w][0], x[0][w], x[1][w]);
                                uint32_t cb = maj3(pl[w][0], x[2][w], x[3][w]);
                                pl[w][0]    = xor3(pl[w][0], x[2][w], x[3][w]);
                                uint32_t c4 = maj3(pl[w][1], ca, cb);
                                pl[w][1]    = xor3(pl[w][1], ca, cb);
                                ripple<PLANES>(pl[w], 2, c4);
                            } else {
                                uint32_t c2 = maj3(pl[w][0], x[0][w], x[1][w]);
                                pl[w][0]    = xor3(pl[w][0], x[0][w], x[1][w]);
                                ripple<PLANES>(pl[w], 1, c2);
                            }
                        }
                        if (early_exit) {
                            const uint32_t done_k = base + ub + U;                 // k-mers of this sequence consumed
                            bool undecided = false;
                            if (seq_open && done_k < nk) {
                                const uint32_t left = nk - done_k;
                                undecided = true;
                                if (left < thr) {                                  // bins below thr - left are dead
                                    uint32_t u_any = 0;
                                    #pragma unroll
                                    for (int w = 0; w < NW; ++w)
                                        u_any |= ge_threshold<PLANES>(pl[w], thr - left) & ~ge_threshold<PLANES>(pl[w], thr) & ~sel[w];
                                    undecided = lane_on && u_any != 0;
                                }
                            }
                            open_mask = __ballot_sync(0xffffffffu, undecided);
                            seq_open = seq_open && (open_mask & sub_mask) != 0;
                        }
                    }
                }
                // --- threshold (whichBins :219-223: OR-accumulates; a sequence shorter than k selects nothing)
                if (nk > 0 && (PLANES >= 32 || thr < (1u << PLANES))) {
                    #pragma unroll
                    for (int w = 0; w < NW; ++w) sel[w] |= ge_threshold<PLANES>(pl[w], thr);
                }
                if (out_counts != nullptr && lane_on) {
                    #pragma unroll
                    for (int w = 0; w < NW; ++w)
                        for (uint32_t b = 0; b < 32; ++b) {
                            const uint32_t bin = chunk * (VEC * 8) + w * 32 + b;
                            if (bin < hp.bins) {
                                uint32_t c = 0;
                                #pragma unroll
                                for (int p = 0; p < PLANES; ++p) c |= ((pl[w][p] >> b) & 1u) << p;
                                out_counts[s * hp.bins + bin] = (uint16_t)c;
                            }
                        }
                }
            }
            if (lane_on) {
                uint32_t* dst = reinterpret_cast<uint32_t*>(out_rows + r * hp.W) + chunk * NW;
                #pragma unroll
                for (int w = 0; w < NW; ++w) {
                    const uint32_t first_bin = chunk * (VEC * 8) + w * 32;
                    uint32_t mask = 0xffffffffu;                         // bins beyond `bins` never select
                    if (first_bin >= hp.bins) mask = 0;
                    else if (hp.bins - first_bin < 32) mask = (1u << (hp.bins - first_bin)) - 1u;
                    sel[w] &= mask;
                }
                if (NW == 4) *reinterpret_cast<uint4*>(dst) = make_uint4(sel[0], sel[1], sel[NW - 2], sel[NW - 1]);
                else         *reinterpret_cast<uint2*>(dst) = make_uint2(sel[0], sel[1]);
            }
        }
    }
    if (lookup_stat != nullptr) {                                   // lane 0 of every sub-group holds its sequence's count
        const uint32_t tot = __reduce_add_sync(0xffffffffu, li == 0 ? n_lookups : 0u);
        if (lane == 0 && tot) atomicAdd(lookup_stat, (unsigned long long)tot);
    }
}

template <int VEC, int TL2, int H, int PLANES>
static cudaError_t launch_one(const HashParams& hp, const uint64_t* filter, const uint64_t* packed, const SeqMeta* meta,
                              uint32_t n_rows, uint32_t group, uint32_t thr, uint64_t* out_rows, uint16_t* out_counts,
                              uint32_t* work_counter, uint32_t flags, unsigned long long* lookup_stat, int sm_count, cudaStream_t st)
{
    auto kern = query_kernel<VEC, TL2, H, PLANES>;
    int per_sm = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 128, 0);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) per_sm = 1;
    constexpr int SG = 32 >> TL2;
    uint64_t warps_needed = ((uint64_t)n_rows + SG - 1) / SG;
    uint64_t blocks = (warps_needed + 3) / 4;
    uint64_t cap = (uint64_t)sm_count * per_sm;
    if (blocks > cap) blocks = cap;
    if (blocks == 0) return cudaSuccess;
    e = cudaMemsetAsync(work_counter, 0, sizeof(uint32_t), st);
    if (e != cudaSuccess) return e;
    kern<<<(unsigned)blocks, 128, 0, st>>>(hp, reinterpret_cast<const uint8_t*>(filter), packed, meta, n_rows, group,
                                           thr, out_rows, out_counts, work_counter, flags, lookup_stat);
    return cudaGetLastError();
}

template <int VEC, int H, int PLANES>
static cudaError_t launch_tl2(int tl2, const HashParams& hp, const uint64_t* filter, const uint64_t* packed,
                              const SeqMeta* meta, uint32_t n_rows, uint32_t group, uint32_t thr, uint64_t* out_rows,
                              uint16_t* out_counts, uint32_t* work_counter, uint32_t flags, unsigned long long* lookup_stat, int sm_count, cudaStream_t st)
{
    switch (tl2) {
        case 0: return launch_one<VEC, 0, H, PLANES>(hp, filter, packed, meta, n_rows, group, thr, out_rows, out_counts, work_counter, flags, lookup_stat, sm_count, st);
        case 1: return launch_one<VEC, 1, H, PLANES>(hp, filter, packed, meta, n_rows, group, thr, out_rows, out_counts, work_counter, flags, lookup_stat, sm_count, st);
        case 2: return launch_one<VEC, 2, H, PLANES>(hp, filter, packed, meta, n_rows, group, thr, out_rows, out_counts, work_counter, flags, lookup_stat, sm_count, st);
        case 3: return launch_one<VEC, 3, H, PLANES>(hp, filter, packed, meta, n_rows, group, thr, out_rows, out_counts, work_counter, flags, lookup_stat, sm_count, st);
        case 4: return launch_one<VEC, 4, H, PLANES>(hp, filter, packed, meta, n_rows, group, thr, out_rows, out_counts, work_counter, flags, lookup_stat, sm_count, st);
        default: return launch_one<VEC, 5, H, PLANES>(hp, filter, packed, meta, n_rows, group, thr, out_rows, out_counts, work_counter, flags, lookup_stat, sm_count, st);
    }
}

cudaError_t launch_query(const HashParams& hp, const uint64_t* filter, const uint64_t* packed, const SeqMeta* meta,
                         uint32_t n_rows, uint32_t group, uint32_t threshold, uint32_t max_kmers,
                         uint64_t* out_rows, uint16_t* out_counts, uint32_t* work_counter, uint32_t flags, unsigned long long* lookup_stat, int sm_count, cudaStream_t st)
{
    if (n_rows == 0) return cudaSuccess;
    const bool vec16 = (hp.W % 2u) == 0;
    const uint32_t chunks = vec16 ? hp.W / 2 : hp.W;
    int tl2 = 0;
    while (tl2 < 5 && (1u << tl2) < chunks) ++tl2;
    const bool fast = (hp.h == 2 || hp.h == 3) && max_kmers != 0 && max_kmers <= 255 && hp.P <= (1ull << 32);
#define DYIBF_GO(VEC, H, PLANES) \
    return launch_tl2<VEC, H, PLANES>(tl2, hp, filter, packed, meta, n_rows, group, threshold, out_rows, out_counts, work_counter, flags, lookup_stat, sm_count, st)
    if (fast) {
        if (vec16) { if (hp.h == 2) DYIBF_GO(16, 2, 8); else DYIBF_GO(16, 3, 8); }
        else       { if (hp.h == 2) DYIBF_GO(8, 2, 8);  else DYIBF_GO(8, 3, 8); }
    }
    if (vec16) DYIBF_GO(16, 0, 16);
    DYIBF_GO(8, 0, 16);
#undef DYIBF_GO
}

} // namespace dyibf
