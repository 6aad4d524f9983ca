// build_bucket.cu — the bucketed IBF build: turns _insertKmer's random single-bit ORs (d_bloom_filter.h:293-302) into
// streaming passes.  Same filter, bit for bit, as build_kernel.cu's direct atomics (OR is commutative and idempotent).
//
// Why: the direct kernel's updates are single 64-bit words at uniformly random addresses of a multi-GiB table: every one
// costs a 32-byte DRAM sector read and, later, a sector write-back, and the rate of such random accesses — not bytes —
// is what HBM limits (B200, 16 GiB table: 21.5 G updates/s).  Here the table is cut into <= 1024 regions of 2^R words
// (16 MiB for a 16 GiB table); every update becomes a 32-bit record (word-in-region << 6 | bit) in its region's bucket,
// and the buckets are then OR-ed in region by region, each region living in the 126 MB L2 while its updates arrive.
//
// How the records get to their buckets (history with numbers: profiles/r01_build_kernels_ncu_*.md; all times per
// 12.3 G records = one 4 Gbp step of bench.py --workload cfg5 on a B200):
//   * v1 ranked records with shared-memory atomics on a block-wide 1024-entry histogram and stored every record with
//     its own 4-byte store: bounded by the ATOMS rate (~2 cycles per lane, twice per record), 303 ms.
//   * v2 gave every warp a private 1024-entry histogram: no atomics, but 7 000 resident warps x 1024 buckets keep 7 M
//     partially written 32-byte sectors open — far more than the L2 holds — so nearly every record store became a DRAM
//     read-modify-write: 457 ms.
//   * v3 split in two levels of 32 buckets with register-resident counters (below) but still one store per record, in a
//     counting walk and a writing walk: 324 ms for both levels — ncu showed nearly every sector leaving the L2 half
//     written (DRAM reads = 0.9 x record bytes, DRAM writes = 2.2 x).
//   * v4 (this file): two-level radix split, 32 buckets per level, ONE walk.  A warp splits its stream 32 ways; lane b
//     keeps bucket b's state in registers; the 32 records a warp handles at once are ranked with five warp votes (bit i
//     of the key -> one ballot; a lane's peers are the lanes whose five bits agree with its own) — no atomics, no
//     histogram; records are staged in one 128-byte line per bucket in shared memory and a line that fills up is
//     written to its bucket as one coalesced 128-byte store (`Splitter` below).  Level 1 splits the hashed updates by
//     the upper five region bits into 32 super-buckets, level 2 splits each super-bucket into its 32 regions: 61 + 58 ms.
//   Tables of at most 32 regions (<= 512 MiB) need level 1 only.
//   Updates that do not fit a bucket (the hash is uniform; buckets get 10 % slack) fall back to the direct atomic, so
//   the result never depends on the bucket capacity.
#include "dyibf_internal.h"
#include "ibf_hash.cuh"

#include <cstdlib>

namespace dyibf {

static const uint32_t FULL          = 0xffffffffu;
static const uint32_t MAX_REGIONS   = 1024;
static const uint32_t SPLIT1_ITEM   = 128;      // packed words per warp work item of level 1 (4 groups of 32 words)
static const uint32_t SPLIT2_ITEM   = 8192;     // records per warp work item of level 2
static const uint32_t APPLY_CHUNK   = 8192;     // records per apply block
static const uint64_t MAX_BUCKET_CAP = 1ull << 26;   // records per final bucket; 32 x this (a level-1 bucket) fits 32 bits

// ---- locating the sequence of a packed word ---------------------------------------------------------------------
// The sequence that contains packed word t (meta[s].word_off <= t < meta[s+1].word_off), starting from the warp-uniform
// answer s0 for the warp's first word t0 <= t.
__device__ __forceinline__ uint64_t seq_of_word(const SeqMeta* __restrict__ meta, uint32_t n_seqs, uint64_t s0, uint64_t t)
{
    uint64_t s = s0;
    while (s + 1 < n_seqs && __ldg(&meta[s + 1].word_off) <= t) ++s;
    return s;
}

__device__ __forceinline__ uint64_t seq_of_first_word(const SeqMeta* __restrict__ meta, uint32_t n_seqs, uint64_t hint, uint64_t t0)
{
    if (hint < n_seqs && __ldg(&meta[hint].word_off) <= t0 && (hint + 1 == n_seqs || __ldg(&meta[hint + 1].word_off) > t0))
        return hint;
    uint64_t lo = 0, hi = n_seqs;                  // word_off(lo) <= t0 < word_off(hi)
    while (hi - lo > 1) {
        const uint64_t mid = (lo + hi) >> 1;
        if (__ldg(&meta[mid].word_off) <= t0) lo = mid; else hi = mid;
    }
    return lo;
}

// what one lane needs to enumerate the updates of the (up to) 32 k-mers that start in its packed word
struct LaneWord {
    uint64_t hi_w, lo_w, col;
    uint32_t i0_hi;       // 32 when the word is not the sequence's first: kmer_value2 then skips the leading-N term
    uint32_t cnt, firstN, bit;
};

__device__ __forceinline__ LaneWord load_lane_word(const HashParams& hp, const uint64_t* __restrict__ packed,
                                                   const SeqMeta* __restrict__ meta, const uint32_t* __restrict__ bin_ids,
                                                   uint32_t n_seqs, uint64_t s0, uint64_t t, uint64_t total_words)
{
    LaneWord w; w.hi_w = w.lo_w = w.col = 0; w.i0_hi = 0; w.cnt = 0; w.firstN = 0; w.bit = 0;
    if (t >= total_words) return w;
    const uint64_t s = seq_of_word(meta, n_seqs, s0, t);
    const uint64_t word_off = __ldg(&meta[s].word_off);
    const uint32_t len = __ldg(&meta[s].len), flags = __ldg(&meta[s].flags);
    if (len < hp.k) return w;                      // addFastaFile skips records shorter than k (:142)
    const uint64_t nk = (uint64_t)len - hp.k + 1;
    const uint64_t i0 = (t - word_off) * 32;
    if (i0 >= nk) return w;
    w.cnt = (uint32_t)((nk - i0) < 32 ? (nk - i0) : 32);
    w.hi_w = __ldg(packed + t); w.lo_w = __ldg(packed + t + 1);
    w.firstN = flags & 1u;
    const uint32_t bin = __ldg(bin_ids + s);
    if (bin >= hp.bins) {                          // device-resident bin ids the host could not check: flag, add nothing
        if (hp.err) atomicOr(hp.err, DYIBF_DEVERR_BIN_RANGE);
        w.cnt = 0;
        return w;
    }
    w.bit = bin & 63u; w.col = bin >> 6;
    w.i0_hi = i0 >= 32 ? 32u : 0u;
    return w;
}

// ---- ranking 32 simultaneous records into 32 buckets with five votes -------------------------------------------
// lanex[i] = all-ones where bit i of the lane number is 0: (ballot_i ^ lanex[i]) = lanes whose key bit i equals MY LANE's bit i.
struct LaneBits { uint32_t x[5]; };
__device__ __forceinline__ LaneBits lane_bits(uint32_t lane)
{
    LaneBits b;
    #pragma unroll
    for (int i = 0; i < 5; ++i) b.x[i] = ((lane >> i) & 1u) ? 0u : FULL;
    return b;
}

// mine: the valid lanes whose key equals my lane number (I count them: lane b owns bucket b's counter)
__device__ __forceinline__ uint32_t match_mine(uint32_t key, bool valid, const LaneBits& lb)
{
    uint32_t mine = __ballot_sync(FULL, valid);
    #pragma unroll
    for (int i = 0; i < 5; ++i) mine &= __ballot_sync(FULL, (key >> i) & 1u) ^ lb.x[i];
    return mine;
}

// load through the L2 (a line another lane of this warp has just written)
__device__ __forceinline__ uint32_t load_cg(const uint32_t* p) { return __ldcg(p); }
__device__ __forceinline__ uint64_t load_cg(const uint64_t* p) { return (uint64_t)__ldcg(reinterpret_cast<const unsigned long long*>(p)); }

// ---- a warp splitting a stream of records 32 ways ----------------------------------------------------------------
// State of one warp: a staging line of 32 records per bucket in shared memory, and in lane b's registers the state of
// bucket b: how many records are staged (fill < 32 between pushes), the next line of the bucket this warp may write, and
// how many lines are left of the chunk of CHUNK_LINES lines it reserved from the bucket's cursor.
//   push   : ranks the warp's 32 simultaneous records (five votes), drops those that fit into their bucket's staging
//            line, writes every line that filled up to its bucket as ONE coalesced 128-byte store (no partially written
//            sectors: a 4-byte store per record made the L2 evict nearly every sector half full, a DRAM read-modify-write
//            each), then drops the rest into the emptied lines.  A bucket's chunk is reserved with one atomic when its
//            first line is ready; only that flush (1 in CHUNK_LINES) waits for the atomic's round trip.
//   finish : at the end of the kernel the < 32 records still staged per bucket go straight into the filter, and the
//            unused lines of the last chunk are filled with copies of a record of the same bucket — OR is idempotent,
//            so a duplicate update is harmless and the consumer can read [0, cursor) without holes.
// A bucket whose cursor passed its capacity is full: its lines are OR-ed straight into the filter, 32 atomics per line.
static const uint32_t CHUNK_LINES = 8;                 // lines (of 32 records) per reservation; a power of two
static const uint32_t LINE_NONE   = 0xffffffffu;
static const uint32_t WARP_STAGE  = 1024 + 32;        // records of shared memory per warp: 32 lines of 32 + the 32 fill counts

template <typename Rec, bool MATCH = false>
struct Splitter {
    Rec* stage;                   // this warp's [32 buckets][32 records]
    uint32_t* sfill;              // MATCH: this warp's 32 fill counts in shared memory (instead of lane b's `fill`)
    Rec* out;                     // bucket 0; bucket b starts at out + b * cap
    unsigned long long* cursors;  // bucket b's fill count in records (a multiple of 32 * CHUNK_LINES)
    uint64_t* filter;
    uint64_t word0;               // first filter word of bucket 0's range
    uint32_t cap, shift;          // records per bucket; log2(words per bucket range): record = word-in-range << 6 | bit
    uint32_t lane, lt_mask;
    LaneBits lb;
    uint32_t fill, line;          // lane b: bucket b.  line = next line of the WHOLE out array (b * cap / 32 + line in bucket)

    __device__ __forceinline__ void init(Rec* stage_, Rec* out_, unsigned long long* cursors_, uint32_t cap_, uint64_t* filter_,
                                         uint32_t shift_, uint64_t word0_, uint32_t lane_)
    {
        stage = stage_; out = out_; cursors = cursors_; cap = cap_; filter = filter_; shift = shift_; word0 = word0_;
        lane = lane_; lt_mask = (1u << lane_) - 1u; lb = lane_bits(lane_);
        fill = 0; line = 0;
        sfill = reinterpret_cast<uint32_t*>(stage_ + 1024);          // the tail of this warp's WARP_STAGE records
        if (MATCH) { sfill[lane_] = 0; __syncwarp(); }
    }

    __device__ __forceinline__ void direct(uint32_t b, Rec rec) const
    {
        atomicOr(reinterpret_cast<unsigned long long*>(filter + word0 + ((uint64_t)b << shift) + (uint64_t)(rec >> 6)),
                 1ull << ((uint32_t)rec & 63u));
    }

    // write bucket b's staged line (all 32 lanes call this together)
    __device__ __forceinline__ void flush(uint32_t b)
    {
        // chunks are CHUNK_LINES lines long and aligned (cap / 32 is a multiple of CHUNK_LINES), so "the chunk is used up"
        // (or none was reserved yet) is line % CHUNK_LINES == 0; LINE_NONE (all ones) never looks like that
        uint32_t ln = __shfl_sync(FULL, line, b);
        if ((ln & (CHUNK_LINES - 1)) == 0) {
            if (lane == b) {
                const unsigned long long c = atomicAdd(&cursors[b], (unsigned long long)(32 * CHUNK_LINES));
                line = c >= cap ? LINE_NONE : b * (cap >> 5) + (uint32_t)(c >> 5);
            }
            ln = __shfl_sync(FULL, line, b);
        }
        const Rec v = stage[b * 32 + lane];
        if (ln != LINE_NONE) out[(uint64_t)ln * 32 + lane] = v;         // 32 buckets x cap / 32 lines fit 32 bits
        else direct(b, v);
        if (lane == b && ln != LINE_NONE) ++line;
    }

    __device__ __forceinline__ void push(bool valid, uint32_t key, Rec rec)
    {
        if (MATCH) {
            // one MATCH instead of five votes: my peers are the valid lanes with my key; the lowest of them keeps the
            // bucket's fill count (shared memory, one word per bucket)
            const uint32_t peers = __match_any_sync(FULL, valid ? key : 32u + lane);
            const uint32_t base = valid ? sfill[key] : 0u;
            const uint32_t cnt = __popc(peers), at = base + __popc(peers & lt_mask);          // at: 0 .. 62
            const bool leader = valid && (peers & lt_mask) == 0;
            __syncwarp();
            if (leader) sfill[key] = (base + cnt) & 31u;
            if (valid && at < 32) stage[key * 32 + at] = rec;
            __syncwarp();
            uint32_t m = __ballot_sync(FULL, leader && base + cnt >= 32);
            while (m) {
                const uint32_t b = __shfl_sync(FULL, key, __ffs(m) - 1);
                m &= m - 1;
                flush(b);
            }
            __syncwarp();
            if (valid && at >= 32) stage[key * 32 + at - 32] = rec;
            return;
        }
        // lane b learns which valid lanes hold key b (five votes); my peers are what the lane named by my key learned
        const uint32_t mine = match_mine(key, valid, lb);
        const uint32_t peers = __shfl_sync(FULL, mine, key);
        const uint32_t at = __shfl_sync(FULL, fill, key) + __popc(peers & lt_mask);      // 0 .. 62
        const uint32_t total = fill + __popc(mine);                                      // lane b: bucket b after this push
        if (valid && at < 32) stage[key * 32 + at] = rec;
        __syncwarp();
        uint32_t m = __ballot_sync(FULL, total >= 32);
        while (m) {
            const uint32_t b = __ffs(m) - 1;
            m &= m - 1;
            flush(b);
        }
        fill = total >= 32 ? total - 32 : total;
        __syncwarp();
        if (valid && at >= 32) stage[key * 32 + at - 32] = rec;
    }

    __device__ __forceinline__ void finish()
    {
        __syncwarp();
        #pragma unroll 1
        for (uint32_t b = 0; b < 32; ++b) {
            const uint32_t f = MATCH ? sfill[b] : __shfl_sync(FULL, fill, b);
            if (lane < f) direct(b, stage[b * 32 + lane]);
            const uint32_t ln = __shfl_sync(FULL, line, b);
            const uint32_t lf = ln == LINE_NONE ? 0u : (CHUNK_LINES - (ln & (CHUNK_LINES - 1))) & (CHUNK_LINES - 1);   // lines left of the chunk
            if (lf > 0) {
                // the chunk was reserved for a line that was then written, so line ln-1 exists and belongs to this warp
                const Rec pad = load_cg(out + (uint64_t)(ln - 1) * 32);
                for (uint32_t j = 0; j < lf; ++j) out[(uint64_t)(ln + j) * 32 + lane] = pad;
                if (lane == b) line += lf;
            }
        }
        fill = 0;
        if (MATCH) { __syncwarp(); sfill[lane] = 0; __syncwarp(); }
    }
};

// ---- level 1: hashed updates -> 32 buckets by key = word index >> shift -----------------------------------------
// out: 32 buckets of cap1 records (Rec = u32, or u64 when shift + 6 > 32); record = (word & (2^shift - 1)) << 6 | bit.
// Two-level plans call it with shift = R + 5 (super-buckets), single-level plans with shift = R (final buckets).
// A warp claims items of SPLIT1_ITEM / 32 groups of 32 packed words; the loads of the next group (sequence table,
// packed words, bin) are issued before the current group's 96 x 32 updates are pushed.
template <int H, typename Rec, bool MATCH>
__global__ void __launch_bounds__(256) split1_kernel(const HashParams hp, uint64_t* __restrict__ filter,
        const uint64_t* __restrict__ packed, const SeqMeta* __restrict__ meta, const uint32_t* __restrict__ bin_ids,
        uint32_t n_seqs, uint64_t word_begin, uint64_t total_words, uint32_t shift, uint32_t cap1,
        Rec* __restrict__ out, unsigned long long* __restrict__ cursors1, unsigned long long* __restrict__ work_counter)
{
    // packed words [word_begin, total_words) of the batch; word_begin is a multiple of SPLIT1_ITEM
    extern __shared__ __align__(16) unsigned char split_smem[];
    const uint32_t lane = threadIdx.x & 31;
    Splitter<Rec, MATCH> sp;
    sp.init(reinterpret_cast<Rec*>(split_smem) + (threadIdx.x >> 5) * WARP_STAGE, out, cursors1, cap1, filter, shift, 0, lane);
    const uint32_t h = H ? H : hp.h;
    const uint64_t low_mask = (1ull << shift) - 1;
    const uint64_t n_items = (total_words - word_begin + SPLIT1_ITEM - 1) / SPLIT1_ITEM;
    const uint32_t G = SPLIT1_ITEM / 32;
    uint64_t hint = 0;
    for (;;) {
        unsigned long long item = 0;
        if (lane == 0) item = atomicAdd(work_counter, 1ull);
        item = __shfl_sync(FULL, item, 0);
        if (item >= n_items) break;
        const uint64_t item0 = word_begin + item * SPLIT1_ITEM;
        hint = seq_of_first_word(meta, n_seqs, hint, item0);
        LaneWord next = load_lane_word(hp, packed, meta, bin_ids, n_seqs, hint, item0 + lane, total_words);
        #pragma unroll 1
        for (uint32_t g = 0; g < G; ++g) {
            const LaneWord w = next;
            if (g + 1 < G) {
                const uint64_t t0 = item0 + (uint64_t)(g + 1) * 32;
                if (t0 < total_words) hint = seq_of_first_word(meta, n_seqs, hint, t0);
                next = load_lane_word(hp, packed, meta, bin_ids, n_seqs, hint, t0 + lane, total_words);
            }
            const uint32_t max_cnt = __reduce_max_sync(FULL, w.cnt);
            #pragma unroll 1
            for (uint32_t x = 0; x < max_cnt; ++x) {
                const bool valid = x < w.cnt;
                const uint64_t kv = kmer_value2(w.hi_w, w.lo_w, x | w.i0_hi, hp.k, w.firstN);
                #pragma unroll
                for (int jh = 0; jh < (H ? H : DYIBF_MAX_HASH); ++jh) {
                    if (jh < (int)h) {
                        const uint64_t widx = block_index(kv, jh, hp) * hp.W + w.col;
                        sp.push(valid, (uint32_t)(widx >> shift) & 31u, (Rec)(((widx & low_mask) << 6) | w.bit));
                    }
                }
            }
        }
    }
    sp.finish();
}

// ---- level 2: one super-bucket's records -> its 32 regions' final buckets ---------------------------------------
// in: n_super buckets of cap1 records (word-in-super-bucket << 6 | bit); out: n_regions buckets of cap u32 records
// (word-in-region << 6 | bit); region = super * 32 + ((record >> (6 + R)) & 31).
// A warp serves ONE super-bucket for the whole kernel (warp w -> super-bucket w mod n_super, so that its staging lines
// stay valid) and claims items of SPLIT2_ITEM consecutive records from that super-bucket's own work counter; every lane
// fetches 8 records with 16-byte loads before any of them is pushed (1 KiB in flight per warp).
template <typename Rec> struct RecBatch {
    static const int PER_VEC = 16 / sizeof(Rec);       // records per 16-byte load
    static const int N_VEC = 8 / PER_VEC;              // loads per lane per batch
    Rec r[8];
};

template <typename Rec>
__device__ __forceinline__ void load_batch(RecBatch<Rec>& b, const Rec* __restrict__ src, uint64_t ib, uint64_t end, uint32_t lane,
                                           uint32_t* valid_mask)
{
    // record e of vector v of this lane is record ib + (v*32 + lane)*PER_VEC + e of the bucket.  A vector that starts
    // before `end` is loaded whole: buckets are 128-byte aligned and `end` is a multiple of 32 records (whole lines).
    uint32_t vm = 0;
    #pragma unroll
    for (int v = 0; v < RecBatch<Rec>::N_VEC; ++v) {
        const uint64_t first = ib + (uint64_t)(v * 32 + lane) * RecBatch<Rec>::PER_VEC;
        uint4 x = make_uint4(0, 0, 0, 0);
        if (first < end) x = __ldcs(reinterpret_cast<const uint4*>(src + first));
        if constexpr (sizeof(Rec) == 4) {
            b.r[v * 4 + 0] = x.x; b.r[v * 4 + 1] = x.y; b.r[v * 4 + 2] = x.z; b.r[v * 4 + 3] = x.w;
        } else {
            b.r[v * 2 + 0] = ((uint64_t)x.y << 32) | x.x; b.r[v * 2 + 1] = ((uint64_t)x.w << 32) | x.z;
        }
        // `end` is a multiple of 32 records, a vector never straddles it
        if (first < end) vm |= ((1u << RecBatch<Rec>::PER_VEC) - 1u) << (v * RecBatch<Rec>::PER_VEC);
    }
    *valid_mask = vm;
}

template <typename Rec, bool MATCH>
__global__ void __launch_bounds__(256) split2_kernel(uint64_t* __restrict__ filter, const Rec* __restrict__ in,
        const unsigned long long* __restrict__ cursors1, uint64_t cap1, uint32_t n_super, uint32_t R, uint32_t cap,
        uint32_t* __restrict__ out, unsigned long long* __restrict__ cursors, unsigned long long* __restrict__ work_counters)
{
    extern __shared__ __align__(16) unsigned char split_smem[];
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t sb = (uint32_t)((blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) % n_super);
    Splitter<uint32_t, MATCH> sp;
    sp.init(reinterpret_cast<uint32_t*>(split_smem) + (threadIdx.x >> 5) * WARP_STAGE, out + (uint64_t)sb * 32 * cap, cursors + sb * 32,
            cap, filter, R, (uint64_t)sb << (R + 5), lane);
    unsigned long long n = cursors1[sb];
    if (n > cap1) n = cap1;
    const uint64_t n_items = (n + SPLIT2_ITEM - 1) / SPLIT2_ITEM;
    const Rec* src = in + (uint64_t)sb * cap1;
    const uint32_t rec_mask = (uint32_t)((1ull << (R + 6)) - 1);
    for (;;) {
        unsigned long long item = 0;
        if (lane == 0) item = atomicAdd(&work_counters[sb], 1ull);
        item = __shfl_sync(FULL, item, 0);
        if (item >= n_items) break;
        const uint64_t begin = item * SPLIT2_ITEM;
        const uint64_t end = begin + SPLIT2_ITEM < n ? begin + SPLIT2_ITEM : n;
        #pragma unroll 1
        for (uint64_t ib = begin; ib < end; ib += 256) {
            RecBatch<Rec> b; uint32_t vm;
            load_batch<Rec>(b, src, ib, end, lane, &vm);
            #pragma unroll
            for (int e = 0; e < 8; ++e)
                sp.push((vm >> e) & 1u, (uint32_t)(b.r[e] >> (R + 6)) & 31u, (uint32_t)b.r[e] & rec_mask);
        }
    }
    sp.finish();
}

// ---- apply: OR one region's records into the region ---------------------------------------------------------------
// Block j of region r first asks the L2 for its share of the region's lines (sequential: DRAM streams them in, and the
// ORs that follow — from this block and from the region's other blocks — find their sectors resident), then ORs its
// chunk of the region's records in.
__global__ void __launch_bounds__(256) bucket_apply_kernel(uint64_t* __restrict__ filter, uint64_t data_words,
        const uint32_t* __restrict__ records, const unsigned long long* __restrict__ cursors, uint32_t R, uint64_t cap,
        uint32_t blocks_per_region)
{
    const uint32_t r = blockIdx.x / blocks_per_region, j = blockIdx.x - r * blocks_per_region;
    unsigned long long cnt = cursors[r];
    if (cnt > cap) cnt = cap;
    const uint64_t begin = (uint64_t)j * APPLY_CHUNK;
    if (begin >= cnt) return;
    const uint64_t end = begin + APPLY_CHUNK < cnt ? begin + APPLY_CHUNK : cnt;
    const uint64_t region_first = (uint64_t)r << R;
    unsigned long long* region = reinterpret_cast<unsigned long long*>(filter + region_first);
    {
        const uint64_t n_active = (cnt + APPLY_CHUNK - 1) / APPLY_CHUNK;
        uint64_t region_words = 1ull << R;
        if (region_first + region_words > data_words) region_words = data_words - region_first;
        const uint64_t n_lines = (region_words + 15) / 16;
        for (uint64_t line = (uint64_t)j * 256 + threadIdx.x; line < n_lines; line += n_active * 256)
            asm volatile("prefetch.global.L2 [%0];" :: "l"(region + line * 16));
    }
    const uint32_t* rec = records + (uint64_t)r * cap;
    // a record (64-bit word << 6 | bit) is also (32-bit word << 5 | bit) of the same little-endian table: 32-bit REDs
    unsigned int* region32 = reinterpret_cast<unsigned int*>(region);
    // four records per load (buckets are 128-byte aligned), streaming: the records are read once
    const uint64_t end4 = begin + ((end - begin) & ~3ull);
    for (uint64_t i = begin + 4ull * threadIdx.x; i < end4; i += 4 * 256) {
        const uint4 v = __ldcs(reinterpret_cast<const uint4*>(rec + i));
        atomicOr(region32 + (v.x >> 5), 1u << (v.x & 31u));
        atomicOr(region32 + (v.y >> 5), 1u << (v.y & 31u));
        atomicOr(region32 + (v.z >> 5), 1u << (v.z & 31u));
        atomicOr(region32 + (v.w >> 5), 1u << (v.w & 31u));
    }
    for (uint64_t i = end4 + threadIdx.x; i < end; i += 256) {
        const uint32_t v = __ldcs(rec + i);
        atomicOr(region32 + (v >> 5), 1u << (v & 31u));
    }
}

// ---- geometry and launchers --------------------------------------------------------------------------------------
// DYIBF_SPLIT_MATCH=1 selects the MATCH.ANY ranking of the Splitter instead of the five votes (measured slower on B200:
// split2 67.5 ms vs 55.1 ms per 12.3 G records — kept as a switch for the next architecture); read once
static bool split_use_match()
{
    static const bool v = [] { const char* e = getenv("DYIBF_SPLIT_MATCH"); return e && e[0] == '1'; }();
    return v;
}

bool bucket_plan(uint64_t data_words, uint64_t budget_bytes, BucketPlan* p)
{
    uint32_t R = 21;                                              // 2^21 words = 16 MiB regions
    while (((data_words + (1ull << R) - 1) >> R) > MAX_REGIONS) ++R;
    if (R > 26) return false;                                     // word-in-region must fit 26 bits (+ 6 bits of bit index)
    const uint64_t n_regions = (data_words + (1ull << R) - 1) >> R;
    if (n_regions < 8) return false;                              // fits the L2 anyway: direct atomics
    p->R = R; p->n_regions = (uint32_t)n_regions;
    p->n_super = n_regions > 32 ? (uint32_t)((n_regions + 31) / 32) : 1;
    p->rec1_bytes = p->n_super > 1 ? (R + 5 + 6 > 32 ? 8 : 4) : 0;
    // final buckets: n_regions x cap x 4 bytes; level-1 buckets: n_super x (8 x cap) x rec1_bytes
    uint64_t cap = budget_bytes / (4 * n_regions + 8ull * (p->n_super > 1 ? p->n_super : 0) * p->rec1_bytes);
    if (cap > MAX_BUCKET_CAP) cap = MAX_BUCKET_CAP;
    cap -= cap % (32 * CHUNK_LINES);                              // whole reservation chunks of 128-byte lines
    if (cap < 65536) return false;
    p->cap = cap; p->cap1 = p->n_super > 1 ? 8 * cap : 0;
    p->bytes_per_record_x4 = 16 + p->rec1_bytes;
    p->bytes_final = n_regions * cap * 4;
    p->bytes_level1 = (uint64_t)p->n_super * p->cap1 * p->rec1_bytes;
    return true;
}

template <typename Rec>
static cudaError_t split1_launch(const HashParams& hp, uint64_t* filter, const uint64_t* packed, const SeqMeta* meta,
                                 const uint32_t* bin_ids, uint32_t n_seqs, uint64_t word_begin, uint64_t total_words,
                                 uint32_t shift, uint32_t cap1, Rec* out, unsigned long long* cursors1,
                                 unsigned long long* work_counter, unsigned blocks, cudaStream_t st)
{
    dim3 grid(blocks), block(256);
    const size_t smem = 8 * WARP_STAGE * sizeof(Rec);
    const bool match = split_use_match();
#define DYIBF_S1(HH) do { \
        if (smem > 48 * 1024) { \
            cudaError_t ea = cudaFuncSetAttribute(split1_kernel<HH, Rec, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
            if (ea == cudaSuccess) ea = cudaFuncSetAttribute(split1_kernel<HH, Rec, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
            if (ea != cudaSuccess) return ea; \
        } \
        if (match) split1_kernel<HH, Rec, true><<<grid, block, smem, st>>>(hp, filter, packed, meta, bin_ids, n_seqs, word_begin, total_words, shift, cap1, out, cursors1, work_counter); \
        else       split1_kernel<HH, Rec, false><<<grid, block, smem, st>>>(hp, filter, packed, meta, bin_ids, n_seqs, word_begin, total_words, shift, cap1, out, cursors1, work_counter); \
    } while (0)
    switch (hp.h) {
        case 2: DYIBF_S1(2); break;
        case 3: DYIBF_S1(3); break;
        case 4: DYIBF_S1(4); break;
        case 5: DYIBF_S1(5); break;
        default: DYIBF_S1(0); break;
    }
#undef DYIBF_S1
    return cudaGetLastError();
}

// cursor array: [n_regions final fill counts][32 level-1 fill counts][2 level-1 work counters, 2 pad][32 level-2 work counters]
cudaError_t launch_bucket_split(const HashParams& hp, uint64_t* filter, const uint64_t* packed, const SeqMeta* meta,
                                const uint32_t* bin_ids, uint32_t n_seqs, uint64_t word_begin, uint64_t total_words,
                                const BucketPlan& p, void* records, unsigned long long* cursors, uint32_t counter_slot,
                                cudaStream_t st)
{
    if (n_seqs == 0 || total_words <= word_begin) return cudaSuccess;
    if (word_begin % BUILD_SCATTER_ALIGN || counter_slot > 1) return cudaErrorInvalidValue;
    unsigned long long* work_counter = cursors + p.n_regions + 32 + counter_slot;
    cudaError_t e = cudaMemsetAsync(work_counter, 0, sizeof(unsigned long long), st);
    if (e != cudaSuccess) return e;
    const uint64_t items = (total_words - word_begin + SPLIT1_ITEM - 1) / SPLIT1_ITEM;
    uint64_t blocks = (items + 7) / 8;
    if (blocks > sm_count_current() * 6) blocks = sm_count_current() * 6;                  // what fits (32 KiB of staging lines per block); items are claimed
    char* base = static_cast<char*>(records);
    if (p.n_super <= 1)                                            // single level: straight into the final buckets
        return split1_launch<uint32_t>(hp, filter, packed, meta, bin_ids, n_seqs, word_begin, total_words, p.R, (uint32_t)p.cap,
                                       reinterpret_cast<uint32_t*>(base), cursors, work_counter, (unsigned)blocks, st);
    unsigned long long* cursors1 = cursors + p.n_regions;
    if (p.rec1_bytes == 8)
        return split1_launch<uint64_t>(hp, filter, packed, meta, bin_ids, n_seqs, word_begin, total_words, p.R + 5, (uint32_t)p.cap1,
                                       reinterpret_cast<uint64_t*>(base + p.bytes_final), cursors1, work_counter, (unsigned)blocks, st);
    return split1_launch<uint32_t>(hp, filter, packed, meta, bin_ids, n_seqs, word_begin, total_words, p.R + 5, (uint32_t)p.cap1,
                                   reinterpret_cast<uint32_t*>(base + p.bytes_final), cursors1, work_counter, (unsigned)blocks, st);
}

cudaError_t launch_bucket_split2(uint64_t* filter, const BucketPlan& p, void* records, unsigned long long* cursors,
                                 uint64_t max_fill1, cudaStream_t st)
{
    if (p.n_super <= 1 || max_fill1 == 0) return cudaSuccess;
    char* base = static_cast<char*>(records);
    const uint64_t fill1 = max_fill1 < p.cap1 ? max_fill1 : p.cap1;
    unsigned long long* work_counters = cursors + p.n_regions + 32 + 4;
    cudaError_t e = cudaMemsetAsync(work_counters, 0, 32 * sizeof(unsigned long long), st);
    if (e != cudaSuccess) return e;
    uint64_t blocks = (((fill1 + SPLIT2_ITEM - 1) / SPLIT2_ITEM) * p.n_super + 7) / 8;
    if (blocks > sm_count_current() * 6) blocks = sm_count_current() * 6;
    if (blocks < 4) blocks = 4;                                    // 32 warps: every super-bucket has at least one
    const size_t smem = 8 * WARP_STAGE * sizeof(uint32_t);
    const bool match = split_use_match();
#define DYIBF_S2(REC, M) split2_kernel<REC, M><<<(unsigned)blocks, 256, smem, st>>>(filter, reinterpret_cast<const REC*>(base + p.bytes_final), \
            cursors + p.n_regions, p.cap1, p.n_super, p.R, (uint32_t)p.cap, reinterpret_cast<uint32_t*>(base), cursors, work_counters)
    if (p.rec1_bytes == 8) { if (match) DYIBF_S2(uint64_t, true); else DYIBF_S2(uint64_t, false); }
    else                   { if (match) DYIBF_S2(uint32_t, true); else DYIBF_S2(uint32_t, false); }
#undef DYIBF_S2
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    return cudaMemsetAsync(cursors + p.n_regions, 0, 32 * sizeof(unsigned long long), st);
}

cudaError_t launch_bucket_apply(uint64_t* filter, uint64_t data_words, const BucketPlan& p, void* records,
                                unsigned long long* cursors, uint64_t max_fill, cudaStream_t st)
{
    // max_fill: upper bound of the records appended since the last apply, known to the host; bounds the grid
    if (max_fill == 0) return cudaSuccess;
    char* base = static_cast<char*>(records);
    cudaError_t e;
    // padding lines (copies of real records) count like records: every bucket may hold up to its capacity
    const uint64_t fill = max_fill + (uint64_t)32 * CHUNK_LINES * 148 * 48 < p.cap ? max_fill + (uint64_t)32 * CHUNK_LINES * 148 * 48 : p.cap;
    const uint32_t bpr = (uint32_t)((fill + APPLY_CHUNK - 1) / APPLY_CHUNK);
    bucket_apply_kernel<<<p.n_regions * bpr, 256, 0, st>>>(filter, data_words, reinterpret_cast<const uint32_t*>(base), cursors,
                                                           p.R, p.cap, bpr);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    return cudaMemsetAsync(cursors, 0, (size_t)p.n_regions * sizeof(unsigned long long), st);
}

} // namespace dyibf
