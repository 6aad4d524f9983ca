// dyibf_internal.h — shared between the translation units of libdyibf.so (not installed).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/dyibf.h"

namespace dyibf {

// SMs of the CURRENT device (148 on a B200), looked up once per device: launchers size their grids in multiples of it
inline unsigned long long sm_count_current()
{
    static int cached[64];                       // zero = not looked up yet; a benign race writes the same value twice
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    int n = cached[dev];
    if (!n) {
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) { cudaGetLastError(); n = 148; }
        cached[dev] = n;
    }
    return (unsigned long long)n;
}

// Everything a kernel needs to turn a k-mer value into h block indices
// (reference: _init d_bloom_filter.h:245-254, getHashValue :286-291).
struct HashParams {
    uint64_t pre[DYIBF_MAX_HASH];   // _preCalcValues[i] = i ^ (k * seed)
    uint64_t P;                     // _noOfHashPos
    uint64_t magic;                 // floor(2^64 / P) when P is not a power of two
    uint32_t pow2;                  // 1: block = v & (P-1)
    uint32_t h;
    uint32_t k;
    uint32_t W;                     // 64-bit words per block
    uint32_t bins;
    // kmer_direct filter (SeqAnKDXFilter, src/d_kdx_filter.h): the one "hash" is the k-mer value itself; k-mers whose value
    // is not a block of the table go to block `sink` (query: an all-zero block behind the data, build: a scratch block)
    uint32_t direct;
    uint64_t sink;
    // sticky device-side error word of the handle (DYIBF_DEVERR_* bits, include/dyibf.h): a kernel that meets input the
    // host could not validate (device-resident bin ids, a sequence longer than the caller's bound) ORs a bit in and
    // skips the offending work instead of writing out of bounds
    uint32_t* err;
};

// One packed sequence: 2 bits per base, first base in the two top bits of word 0 (so that a k-mer is a
// funnel shift of two neighbouring words), N stored as A (the reference hashes N as A, alphabet_residue.h:680-683).
struct SeqMeta {
    uint64_t word_off;   // first u64 of this sequence in the packed buffer (ceil(len/32)+1 words are valid)
    uint32_t len;        // bases
    uint32_t flags;      // bit 0: first base is N (leading-N term of SeqAn's hashInit, shape_base.h:489)
};

// up to 4 byte ranges of the caller's sequence set that make up one device batch
struct PackSource {
    const uint8_t*  bytes[4];     // device pointer such that bytes[g][offsets[g][i]] is base 0 of sequence i of range g
    const uint64_t* offsets[4];   // device, cnt+1 entries each (original, un-rebased values)
    uint64_t        pack_base[4]; // byte position of range g inside the batch (monotone; defines word_off)
    uint64_t        first_off[4]; // offsets[g][0]
    uint32_t        n_ranges;
    uint32_t        cnt;          // sequences per range
};

struct Profiler;

// ---- launchers (each in its own .cu) ---------------------------------------------------------
// pack: bytes -> packed words + meta for n_in = n_ranges*cnt sequences, and (derive_rc) their reverse
// complements as sequences n_in .. 2*n_in-1.
cudaError_t launch_pack(const PackSource& src, int encoding, int derive_rc, uint64_t* packed, SeqMeta* meta,
                        uint64_t total_words, uint64_t rc_word_base, cudaStream_t st);
// words needed for a batch whose ranges total `bytes` bytes and n_in sequences (per strand)
static inline uint64_t packed_words_for(uint64_t bytes, uint64_t n_in) { return bytes / 32 + 2 * n_in + 2; }

// flags of launch_query
static const uint32_t QUERY_EARLY_EXIT = 1u;   // stop a sequence once no further k-mer can change its bin set (rows only)

// lookup_stat (optional, device): += number of (sequence, column tile, k-mer) lookups actually performed
cudaError_t launch_query(const HashParams& hp, const uint64_t* filter, const uint64_t* packed, const SeqMeta* meta,
                         uint32_t n_rows, uint32_t group, uint32_t threshold, uint32_t max_kmers,
                         uint64_t* out_rows, uint16_t* out_counts, uint32_t* work_counter, uint32_t flags,
                         unsigned long long* lookup_stat, int sm_count, cudaStream_t st);

// build: every k-mer of the n_seqs packed forward sequences ORs h bits (bin = bin_ids[s]) into the filter
cudaError_t launch_build(const HashParams& hp, uint64_t* filter, const uint64_t* packed, const SeqMeta* meta,
                         const uint32_t* bin_ids, uint32_t n_seqs, uint64_t total_words, cudaStream_t st);

// bucketed build (build_bucket.cu): a radix split appends 32-bit update records to per-region buckets, apply ORs them in.
struct BucketPlan {
    uint32_t R;            // log2(words per region)
    uint32_t n_regions;    // final buckets (<= 1024)
    uint32_t n_super;      // level-1 buckets of a two-level split (regions / 32, rounded up); 1: single level
    uint32_t rec1_bytes;   // size of a level-1 record: 4, or 8 when R + 5 + 6 bits do not fit 32; 0: single level
    uint64_t cap;          // records per final bucket
    uint64_t cap1;         // records per level-1 bucket (8 * cap: a quarter of its 32 regions' buckets — level 2 empties
                           // the level-1 buffer several times per sweep over the table); 0: single level
    uint64_t bytes_per_record_x4;   // buffer bytes per record of final capacity, times 4
    uint64_t bytes_final;  // the record buffer holds the final buckets first, then the level-1 buckets
    uint64_t bytes_level1;
};
bool        bucket_plan(uint64_t data_words, uint64_t budget_bytes, BucketPlan* plan);
// cursors: n_regions final fill counts, 32 level-1 fill counts, 2 level-1 work counters (+2 pad), 32 level-2 work counters
static inline uint64_t bucket_cursor_slots(const BucketPlan& p) { return (uint64_t)p.n_regions + 32 + 4 + 32; }
static const uint64_t BUILD_SCATTER_ALIGN = 1024;   // word_begin of a split launch must be a multiple of this
// level 1 (or the only level) over packed words [word_begin, total_words); launches that may overlap on different
// streams use different counter slots (0 / 1)
cudaError_t launch_bucket_split(const HashParams& hp, uint64_t* filter, const uint64_t* packed, const SeqMeta* meta,
                                const uint32_t* bin_ids, uint32_t n_seqs, uint64_t word_begin, uint64_t total_words,
                                const BucketPlan& plan, void* records, unsigned long long* cursors, uint32_t counter_slot,
                                cudaStream_t st);
// level 2 (two-level plans): super-buckets -> final buckets, resets the level-1 cursors; max_fill1: upper bound of the
// records in the level-1 buffer
cudaError_t launch_bucket_split2(uint64_t* filter, const BucketPlan& plan, void* records, unsigned long long* cursors,
                                 uint64_t max_fill1, cudaStream_t st);
// apply + cursor reset; max_fill: upper bound of the records appended since the last apply
cudaError_t launch_bucket_apply(uint64_t* filter, uint64_t data_words, const BucketPlan& plan, void* records,
                                unsigned long long* cursors, uint64_t max_fill, cudaStream_t st);

// bit-sliced build (build_sliced.cu): per-bin Bloom bit vectors that live in the L2, then interleaved into the table.
struct SlicedPlan {
    uint64_t pw;               // 32-bit words per bin vector (ceil(P / 32), rounded up to 32 = whole 128-byte lines)
    uint64_t n_tiles;          // tiles of 1024 blocks (one line of every bin's vector): pw / 32
    uint32_t cols_per_group;   // word columns (64 bins each) per group: 4, fewer for narrow tables / small budgets
    uint32_t group_bins;       // 64 * cols_per_group
    uint32_t n_groups;         // passes that cover all W columns
    uint64_t shadow_bytes;     // group_bins * pw * 4: one shadow buffer
    uint32_t n_buffers;        // 2 when there are several groups (scatter of group g+1 overlaps the interleave of group g), else 1
    uint32_t scatter_blocks;   // grid of the scatter kernel (8 warps x 1024 bases per block in flight): sized so that the
                               // bins in flight (window / average run of one bin) keep their vectors in the L2
};
static const uint32_t SLICED_MAX_BLOCKS = 148 * 8, SLICED_MIN_BLOCKS = 74;
static const uint64_t SLICED_BLOCK_BASES = 8 * 1024;                 // bases the 8 warps of one scatter block span
bool        sliced_plan(const HashParams& hp, uint64_t budget_bytes, SlicedPlan* plan);
// mask[g] = 1 for every group with a sequence; *switches = number of bin changes along the batch (device buffers)
cudaError_t launch_sliced_presence(const uint32_t* bin_ids, uint32_t n_seqs, uint32_t bins, const SlicedPlan& plan, uint32_t* mask,
                                   unsigned long long* switches, cudaStream_t st);
cudaError_t launch_sliced_zero(void* shadow, const SlicedPlan& plan, cudaStream_t st);
cudaError_t launch_sliced_scatter(const HashParams& hp, void* shadow, const SlicedPlan& plan, uint32_t group, const uint64_t* packed,
                                  const SeqMeta* meta, const uint32_t* bin_ids, uint32_t n_seqs, uint64_t total_words,
                                  unsigned long long* work_counter, cudaStream_t st);
cudaError_t launch_sliced_interleave(const HashParams& hp, uint64_t* filter, const void* shadow, const SlicedPlan& plan, uint32_t group,
                                     cudaStream_t st);

// clear: for every hash block, words cols[i] &= ~masks[i]
cudaError_t launch_clear(const HashParams& hp, uint64_t* filter, const uint32_t* cols, const uint64_t* masks, uint32_t n_cols,
                         cudaStream_t st);

// compaction: rows [n_rows x W] -> CSR.  workspace: counts matrix u64[(bins * n_chunks) + 1]
static const uint32_t COMPACT_CHUNK_ROWS = 2048;
size_t      compact_workspace_bytes(uint32_t bins, uint32_t n_rows);
cudaError_t launch_compact_count(const uint64_t* rows, uint32_t n_rows, uint32_t bins, uint64_t* ws, cudaStream_t st);
cudaError_t launch_compact_scan(uint64_t* ws, uint32_t bins, uint32_t n_rows, uint64_t* bin_offsets, cudaStream_t st);
// id_base: added to the row numbers written (a device that filters rows [id_base, id_base + n_rows) of a batch)
cudaError_t launch_compact_fill(const uint64_t* rows, uint32_t n_rows, uint32_t bins, const uint64_t* ws,
                                uint32_t* read_ids, uint64_t capacity, uint32_t id_base, cudaStream_t st);

// exclusive scan of u64 a[0..n) in place, total in a[n]; `a` must hold scan_workspace_elems(n) elements
size_t      scan_workspace_elems(uint64_t n);
cudaError_t launch_scan_u64(uint64_t* a, uint64_t n, cudaStream_t st);

// gather (loadFilteredReads, d_mapper.h:393-443): the read set one bin's mapper receives.
//   fwd sequences of the resident batch: bytes[offsets[i] - first_off ...], n_fwd of them (SE: reads; PE: m1 then m2)
//   ids[m]: the bin's read (pair) ids.  Output sequence j of m*group: section j/m = fwd | (mate fwd) | rc...
struct GatherArgs {
    const uint8_t*  bytes;        // device, base such that bytes[offsets[i]] is base 0 of forward sequence i
    const uint64_t* offsets;      // device, n_fwd + 1
    const uint32_t* ids;          // device, m
    uint32_t        m, group, n_fwd, encoding;
};
// pass 1: lens[j] (u64, to be scanned) and out_ids[j]
cudaError_t launch_gather_plan(const GatherArgs& a, uint64_t* lens, uint32_t* out_ids, cudaStream_t st);
// pass 2: copy / reverse-complement into out_bytes at limits[j]
cudaError_t launch_gather_copy(const GatherArgs& a, const uint64_t* limits, uint8_t* out_bytes, cudaStream_t st);

// ingest (ingest_kernel.cu): FASTA / FASTQ text -> sequence characters + limits, SeqAn's readRecords on the device
size_t      ingest_workspace_bytes(uint64_t n_bytes);
cudaError_t launch_ingest_count(const uint8_t* text, uint64_t n, int format, int iupac, uint64_t* ws, cudaStream_t st);
// totals: records, sequence characters, quality characters, first '>' offset, effective end, first error word
// (offset << 8 | code; all ones = none); synchronises the stream
cudaError_t ingest_read_totals(const uint64_t* ws, uint64_t n, uint64_t totals[6], cudaStream_t st);
cudaError_t launch_ingest_emit(const uint8_t* text, uint64_t n, int format, int iupac, uint64_t* ws, uint8_t* out_seqs,
                               uint64_t* out_limits, uint64_t limit_base, uint64_t n_records, uint64_t n_bases, cudaStream_t st,
                               uint64_t* out_rec_pos = nullptr);    // optional (device): text offset of every record's start character
cudaError_t launch_max_len(const uint64_t* limits, uint64_t n, unsigned long long* out, cudaStream_t st);

} // namespace dyibf
