/*
 * dyibf.h — C ABI of libdyibf.so: the B200 (sm_100a) implementation of DREAM-Yara's interleaved
 * Bloom filter (IBF) read->bin prefilter.
 *
 * This is the drop-in boundary.  The reference (temehi/dream_yara) has no FFI for this path; its
 * boundary is the C++ class seqan::SeqAnBloomFilter in src/d_bloom_filter.h:41-331, used by
 * src/d_mapper.h:308-387 (clasifyLoadedReads), src/d_build_filter.cpp:217-272 and
 * src/d_update_filter.cpp:188-252.  Every entry point below names the reference member it replaces.
 * include/d_bloom_filter.h in this repo re-creates the class on top of these calls.
 *
 * Rules: plain pointers and sizes only; no CPU fallback anywhere (a missing GPU is an error);
 * one handle = one filter (or one bin-shard of a filter) resident in the HBM of one GPU; a dyibf_group is one filter
 * over several GPUs of one process (replicated and read-sharded, or bin-sharded).
 * Calls on one handle must be serialised by the caller.  All functions return DYIBF_OK (0) or a
 * negative error code; dyibf_last_error() gives the message of the calling thread's last failure.
 *
 * Notation (same as the reference members): B bins, W = ceil(B/64) 64-bit words per hash block
 * (_binIntWidth), nbits = _noOfBits (includes the 256 metadata bits), P = (nbits-256)/(64 W)
 * hash blocks (_noOfHashPos), h hash functions, k k-mer size.
 */
#ifndef DYIBF_H
#define DYIBF_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DYIBF_OK                 0
#define DYIBF_ERR_ARG           -1   /* bad argument                                              */
#define DYIBF_ERR_IO            -2   /* file could not be opened / read / written                 */
#define DYIBF_ERR_CUDA          -3   /* CUDA runtime error (incl. "no device")                    */
#define DYIBF_ERR_FORMAT        -4   /* not an sdsl bit_vector file / bad trailer                 */
#define DYIBF_ERR_SIZE_MISMATCH -5   /* file bit count != expected (d_bloom_filter.h:71-76)       */
#define DYIBF_ERR_CAPACITY      -6   /* caller's output buffer too small                          */
#define DYIBF_ERR_NOMEM         -7
#define DYIBF_ERR_BUSY          -8   /* every asynchronous ticket of the handle is in flight          */

/* sticky device-side error bits (dyibf_device_errors): input that only the kernels could see was invalid */
#define DYIBF_DEVERR_BIN_RANGE     1u  /* a device-resident bin id >= bins: the sequence was skipped             */
#define DYIBF_DEVERR_SEQ_TOO_LONG  2u  /* a sequence had more k-mers than the caller's max_seq_len allows: it selected nothing */

#define DYIBF_MAX_HASH           8

/* how a sequence byte maps to a Dna5 code (A,C,G,T,N = 0..4) */
#define DYIBF_ENC_DNA5   0   /* seqan::Dna5 ordValue; anything > 4 is N                              */
#define DYIBF_ENC_DNA5Q  1   /* seqan::Dna5Q byte (reads in the mapper, store_seqs.h:57): >=252 is N */
#define DYIBF_ENC_ASCII  2   /* FASTA characters through Iupac->Dna5: ACGTU (any case), rest N       */

/* where the caller's buffers live */
#define DYIBF_MEM_HOST   0
#define DYIBF_MEM_DEVICE 1

typedef struct dyibf_filter dyibf_filter;
typedef struct dyibf_ticket dyibf_ticket;   /* one asynchronous classify in flight (dyibf_classify_begin)         */
typedef struct dyibf_group  dyibf_group;    /* one filter over several GPUs of one process (dyibf_group_create)   */

typedef struct dyibf_params {
    uint32_t bins;        /* _noOfBins      */
    uint32_t n_hash;      /* _noOfHashFunc  */
    uint32_t kmer;        /* _kmerSize      */
    uint32_t words_per_block; /* _binIntWidth (W) */
    uint64_t nbits;       /* _noOfBits      */
    uint64_t n_blocks;    /* _noOfHashPos (P) */
    uint64_t n_words;     /* ceil(nbits/64): the whole sdsl word array incl. trailer */
    int      device;      /* CUDA device ordinal */
} dyibf_params;

/* accumulated device time of the kernels of one handle (CUDA events; enable with dyibf_profile_enable) */
typedef struct dyibf_profile {
    double   pack_ms, query_ms, compact_ms, build_ms, clear_ms;
    uint64_t pack_launches, query_launches, compact_launches, build_launches, clear_launches;
    uint64_t query_lookups;   /* k-mer lookups (h block fetches each) the query kernel performed while profiling was on */
    double   build_apply_ms;  /* the part of build_ms / build_launches spent in the bucketed build's apply sweeps */
    uint64_t build_apply_launches;
    double   build_split2_ms; /* ... and in the second level of its radix split */
    uint64_t build_split2_launches;
} dyibf_profile;

const char* dyibf_last_error(void);
int  dyibf_device_count(int* count);

/* ---- construction / destruction ------------------------------------------------------------ */

/* SeqAnBloomFilter(n_bins, n_hash_func, kmer_size, vec_size), d_bloom_filter.h:48-57: zeroed filter.
 * n_hash == 0 makes a kmer_direct filter — SeqAnKDXFilter(n_bins, kmer_size, vec_size), src/d_kdx_filter.h:49-56
 * (`-ft kmer_direct`): block = the k-mer value itself, no hashing; its trailer is [bins, 0, k, 0] (:234-239), so
 * dyibf_load_file recognises such a file by the zero.  Every other entry point works on both kinds.  A k-mer whose
 * value is not a block of the table ((nbits-256)/(64 W) blocks) adds nothing and matches nothing (the reference
 * indexes out of bounds there: leading-N k-mers, or a table smaller than 4^k blocks). */
int  dyibf_create(uint32_t bins, uint32_t n_hash, uint32_t kmer, uint64_t nbits, int device,
                  dyibf_filter** out);
/* same, but the P*W data words live in caller-owned device memory (e.g. a buffer that was the target
 * of an NCCL broadcast); the handle neither zeroes nor frees it */
int  dyibf_attach(uint32_t bins, uint32_t n_hash, uint32_t kmer, uint64_t nbits, int device,
                  void* device_words, dyibf_filter** out);
/* SeqAnBloomFilter(fileName), d_bloom_filter.h:79-89: parameters come from the file's trailer */
int  dyibf_load_file(const char* path, int device, dyibf_filter** out);
/* SeqAnBloomFilter(fileName, n_bins, n_hash_func, kmer_size, vec_size), d_bloom_filter.h:59-77 */
int  dyibf_load_file_expect(const char* path, uint32_t bins, uint32_t n_hash, uint32_t kmer,
                            uint64_t nbits, int device, dyibf_filter** out);
/* readFilterMetadata, d_mapper.cpp:532-548 (reads exactly the 32-byte trailer + the header) */
int  dyibf_peek_file(const char* path, uint32_t* bins, uint32_t* n_hash, uint32_t* kmer, uint64_t* nbits);
void dyibf_destroy(dyibf_filter* f);

int  dyibf_get_params(const dyibf_filter* f, dyibf_params* out);       /* getNumberOfBins/getKmerSize :233-241 */
double dyibf_size_mb(const dyibf_filter* f);                          /* size_mb :156-159 */
void* dyibf_device_words(dyibf_filter* f);                            /* device pointer to the P*W data words */

/* raw word transfer: words [first, first+n) of the sdsl array (data words only; n_words-4.. is the trailer,
 * which the handle synthesises).  host <-> HBM through pinned staging. */
int  dyibf_upload_words(dyibf_filter* f, const uint64_t* host_words, uint64_t first, uint64_t n);
int  dyibf_download_words(dyibf_filter* f, uint64_t* host_words, uint64_t first, uint64_t n);
/* save(fileName), d_bloom_filter.h:150-154 -> sdsl::store_to_file: byte-identical file */
int  dyibf_save_file(dyibf_filter* f, const char* path);

/* The loader / writer behind dyibf_load_file* / dyibf_save_file (and their group forms) streams the file through pinned
 * host buffers with several file threads (pread / pwrite) and one thread issuing the asynchronous copies; this is the stage the mapper prints as
 * "Filter loading time" (src/d_mapper.h:913-921); dyibf_upload_words / dyibf_download_words of whole blocks use the same
 * pipeline.  dyibf_set_io_threads: 1..64, 0 = default (half of the host's cores, between 4 and 16), process-wide.
 * Measured on B200 boxes with 16 host cores, 16 GiB filter in the page cache: 7.3 / 26–33 / 43–51 / 44–45 GB/s with 1 / 4 / 8 / 16
 * threads (pinned H2D of the box: 52 GB/s).
 * dyibf_last_io_stats: the calling thread's last load or save. */
typedef struct dyibf_io_stats {
    uint64_t bytes;               /* data-word bytes moved between the file and HBM                         */
    double   seconds;             /* wall time of the transfer                                              */
    double   file_seconds;        /* longest time any one thread spent inside pread / pwrite                */
    double   copy_wait_seconds;   /* longest time any one thread waited for its copies                      */
    int      threads, devices;
    double   submit_seconds;      /* time the submitting thread spent inside the CUDA copy calls            */
    double   starve_seconds;      /* time it waited for the file threads (load: data to copy; save: free buffers) */
    double   pool_seconds;        /* pinned staging allocation (first transfer of the process only)          */
    uint64_t chunk_bytes;         /* bytes per pinned chunk                                                  */
} dyibf_io_stats;
int  dyibf_set_io_threads(int n);
int  dyibf_set_io_chunk_mib(int mib);   /* 1..1024, 0 = default (DYIBF_IO_CHUNK_MIB from the environment, else 4) */
int  dyibf_last_io_stats(dyibf_io_stats* out);

/* ---- build --------------------------------------------------------------------------------- */

/* addKmers(text, binNo), d_bloom_filter.h:91-94,293-315, for n_seqs sequences at once (atomic-OR kernel).
 * seqs: concatenated sequence bytes; offsets[n_seqs+1]; bin_ids[n_seqs].  Sequences shorter than k are
 * skipped (:142).  mem says where seqs/offsets/bin_ids live.  Host mode validates bin_ids (DYIBF_ERR_ARG); device mode
 * cannot: a device-resident bin id >= bins adds nothing and sets DYIBF_DEVERR_BIN_RANGE (dyibf_device_errors). */
int  dyibf_add_sequences(dyibf_filter* f, const uint8_t* seqs, const uint64_t* offsets, uint32_t n_seqs,
                         const uint32_t* bin_ids, int encoding, int mem, void* stream);
/* addFastaFile(fileName, binNo), d_bloom_filter.h:127-147, for n files at once and from their TEXT (uncompressed FASTA or FASTQ,
 * host memory; formats[i] DYIBF_FMT_*, NULL = all FASTA): text i holds the contigs of bin bins[i].  The texts go up once, SeqAn's
 * readRecord loop runs on the device (the rules of dyibf_parse_records with DYIBF_ALPHA_IUPAC: every record a sequence, a FASTA
 * record ends at any '>', a character outside the IUPAC alphabet is an error), records shorter than k are skipped (:142), and all
 * records are added with ONE build call, so a batch of whole files takes the sliced build.  On DYIBF_ERR_FORMAT nothing of
 * the batch has been added and *bad_index (may be NULL) is the offending text (0xffffffff when it cannot be told).
 * Measured: tools/dyibf_build_filter reads 1024 FASTA files of 4 Mbp through this call. */
int  dyibf_add_fasta_texts(dyibf_filter* f, const char* const* texts, const uint64_t* sizes, const uint32_t* bins, const int* formats,
                           uint32_t n, uint32_t* bad_index);
/* clearBins(bins2clear, threads), d_bloom_filter.h:99-123 */
int  dyibf_clear_bins(dyibf_filter* f, const uint32_t* bins, uint32_t n);

/* ---- query --------------------------------------------------------------------------------- */

/* DisOptions::getThreshold, d_mapper.h:83-92 with errorRate = (float)(e_percent/100.0) (d_mapper.cpp:340-342) */
uint16_t dyibf_threshold(int e_percent, uint32_t kmer, uint16_t read_len);
uint16_t dyibf_threshold_rate(float error_rate, uint32_t kmer, uint16_t read_len);

/* The filtering loop of clasifyLoadedReads (d_mapper.h:338-359) = whichBins (d_bloom_filter.h:161-224)
 * on every strand/mate of every read, OR-ed per read.
 *   seqs/offsets : n_seqs sequences.  If derive_rc != 0 the set is first extended to 2*n_seqs by appending
 *                  the reverse complement of every sequence in order (appendReverseComplement,
 *                  store_seqs.h:271-290; N stays N) — on the device, so only forward bytes cross PCIe.
 *   group        : sequences per output row: 1 (plain whichBins per sequence), 2 (single-end: [fwd | rc]),
 *                  4 (paired-end: [m1 | m2 | m1 rc | m2 rc]).  n_rows = n_total / group and row r is the OR
 *                  over g < group of whichBins(seq[r + g*n_rows], threshold).
 *   out_rows     : n_rows x W u64, bit b of word w <=> bin 64w+b selected.
 *   out_counts   : optional (NULL to skip) n_total x bins u16: the per-bin k-mer counts (`counts`, :165).
 *   max_seq_len  : longest sequence, 0 = unknown (host mode computes it; device mode then uses the
 *                  16-bit counters).  Sequences with more than 65535 k-mers are rejected (u16 in the reference).
 *                  A non-zero bound is trusted: a sequence with more than 255 k-mers under a bound that promised at most
 *                  255 selects nothing and sets DYIBF_DEVERR_SEQ_TOO_LONG (dyibf_device_errors).
 */
int  dyibf_query_batch(dyibf_filter* f, const uint8_t* seqs, const uint64_t* offsets, uint32_t n_seqs,
                       int encoding, int group, int derive_rc, uint32_t threshold, uint32_t max_seq_len,
                       uint64_t* out_rows, uint16_t* out_counts, int mem, void* stream);

/* origReadIdMap as `--threads 1` fills it (d_mapper.h:360-368): per bin the ascending ids of the rows that
 * selected it, as CSR.  bin_offsets[bins+1] (u64), read_ids[capacity] (u32).  *total receives the number of
 * (row,bin) pairs; DYIBF_ERR_CAPACITY if it exceeds capacity (bin_offsets is still complete). */
int  dyibf_compact(dyibf_filter* f, const uint64_t* rows, uint32_t n_rows, uint64_t* bin_offsets,
                   uint32_t* read_ids, uint64_t capacity, uint64_t* total, int mem, void* stream);

/* dyibf_compact for a caller that must not allocate device memory.  No counterpart in the reference (its lists are
 * std::vector push_backs on the host heap).
 * Why it exists: dyibf_compact keeps its per-bin counters in a workspace that it cudaMallocs on first use and re-allocates
 * when a larger batch arrives.  A host program that has given the card away before the mapper stage runs — a 64 GiB filter
 * shard plus another component's memory pool, a process under a fixed MPS memory limit — cannot count on that cudaMalloc to
 * succeed, and a service thread feeding 100 000-read batches cannot afford the device-wide stall an allocation causes.  This
 * entry point needs no handle and calls no cudaMalloc / cudaFree (nor anything else that allocates): rows, bin_offsets and
 * read_ids are DEVICE pointers from the caller's own arena on `device`, and the counters live in one array of
 * DYIBF_COMPACT_STATIC_BYTES (8 MiB) inside the library's image — every GPU that loads libdyibf.so carries those 8 MiB until
 * the process exits, whether or not this function is ever called.
 *   bins, rows (n_rows x ceil(bins/64) u64), bin_offsets[bins+1], read_ids[capacity], *total (host, may be NULL), stream:
 *   as dyibf_compact with mem = DYIBF_MEM_DEVICE; the lists are identical to its lists.  DYIBF_ERR_CAPACITY likewise.
 * Limit: the counters of the batch must fit the array, else DYIBF_ERR_NOMEM and nothing is launched.  Needed are
 * 8 * (c + ceil(c / 4096) + 2) bytes with c = bins * ceil(n_rows / r), r = n_rows * ceil(bins / 1024) / 768 rounded up to a
 * multiple of 32 and clamped to [32, 2048].  That admits about 2 000 000 rows at 1024 bins and 260 000 rows at 8192 bins;
 * cut larger batches into pieces (ids of a piece start at 0: add the piece's first row).
 * Sharing: there is ONE array per GPU and every call on that GPU uses it.  The call clears what it reads, so earlier
 * contents never matter — but the last kernel (the one that writes read_ids) still reads the array after the call has
 * returned, so two calls whose work can overlap on the GPU (different streams, or two host threads) corrupt each other.
 * Keep all calls for one GPU on one stream, or wait for the stream before the next call from elsewhere.
 * First call on each GPU: looks up where the array lives on that GPU (cudaGetSymbolAddress, which makes the driver load the
 * library's device code there) and remembers it.  Make that first call at start-up, before any stream capture begins. */
#define DYIBF_COMPACT_STATIC_BYTES (8ull << 20)
int  dyibf_compact_static(int device, uint32_t bins, const uint64_t* rows, uint32_t n_rows, uint64_t* bin_offsets,
                          uint32_t* read_ids, uint64_t capacity, uint64_t* total, void* stream);

/* clasifyLoadedReads (d_mapper.h:308-375) in one call, host buffers: query + compaction with the rows kept in
 * HBM, so only the per-bin lists come back.  Arguments as dyibf_query_batch / dyibf_compact. */
int  dyibf_classify(dyibf_filter* f, const uint8_t* seqs, const uint64_t* offsets, uint32_t n_seqs,
                    int encoding, int group, int derive_rc, uint32_t threshold,
                    uint64_t* bin_offsets, uint32_t* read_ids, uint64_t capacity, uint64_t* total);

/* The per-bin lists of the last dyibf_classify again (they stay in HBM): for a caller whose first buffer was too
 * small (DYIBF_ERR_CAPACITY) — no second pass over the reads. */
int  dyibf_classify_fetch(dyibf_filter* f, uint64_t* bin_offsets, uint32_t* read_ids, uint64_t capacity, uint64_t* total);

/* Asynchronous form of dyibf_classify for the reference's real batching: clasifyLoadedReads is called once per 100 000
 * reads (src/mapper.h:109) while a prefetch thread already loads the next batch (src/file_prefetched.h:137-205).
 * dyibf_classify_begin enqueues upload, pack, query and compaction of one batch on a stream and buffers of its own and
 * returns a ticket at once; up to eight tickets per handle may be in flight (DYIBF_ERR_BUSY beyond; four is where the
 * throughput levels off), so batch i+1 uploads
 * and starts while batch i is in the query kernel.  seqs / offsets must stay valid until dyibf_classify_end, which waits
 * for the ticket and copies the lists out exactly as dyibf_classify returns them.  On DYIBF_ERR_CAPACITY (bin_offsets and
 * *total are filled) the ticket stays in flight: call dyibf_classify_end again with a larger buffer.  After a successful
 * end the ticket's batch is the handle's resident batch (dyibf_gather_reads, dyibf_classify_fetch).  dyibf_classify_poll:
 * *ready = 1 once dyibf_classify_end would not block.  dyibf_classify_cancel abandons a ticket. */
int  dyibf_classify_begin(dyibf_filter* f, const uint8_t* seqs, const uint64_t* offsets, uint32_t n_seqs,
                          int encoding, int group, int derive_rc, uint32_t threshold, dyibf_ticket** ticket);
int  dyibf_classify_poll(dyibf_ticket* ticket, int* ready);
int  dyibf_classify_end(dyibf_ticket* ticket, uint64_t* bin_offsets, uint32_t* read_ids, uint64_t capacity, uint64_t* total);
void dyibf_classify_cancel(dyibf_ticket* ticket);

/* sortedBins, d_mapper.h:843-854: bin numbers by descending list size (std::sort, same comparator). bin_offsets[bins+1]
 * as dyibf_classify / dyibf_compact return them. */
int  dyibf_sorted_bins(const uint64_t* bin_offsets, uint32_t bins, uint32_t* out_bins);

/* loadFilteredReads, d_mapper.h:393-443, for one bin after dyibf_classify(derive_rc != 0) — the read batch and the
 * lists are still resident in HBM, nothing is uploaded again.  With m = the bin's list length the output holds
 * m*group sequences in the reference's order: the m forward reads (first mates), [paired-end: the m second mates,]
 * then the reverse complements of all of those (Dna5Q: quality kept, N stays N), as bytes + limits (n+1 entries, a
 * ConcatDirect StringSet's layout), and out_ids = origReadIdMap[bin] as loadFilteredReads leaves it:
 * id, [id+numPairs,] then each of those + numReads.  *n_seqs_out / *n_bytes_out always receive the needed sizes;
 * DYIBF_ERR_CAPACITY if either buffer is too small. */
int  dyibf_gather_reads(dyibf_filter* f, uint32_t bin, uint8_t* out_seqs, uint64_t capacity_bytes, uint64_t* out_limits,
                        uint32_t* out_ids, uint64_t capacity_seqs, uint64_t* n_seqs_out, uint64_t* n_bytes_out);

/* ---- ingest -------------------------------------------------------------------------------- */

#define DYIBF_FMT_FASTA   0
#define DYIBF_FMT_FASTQ   1
#define DYIBF_ALPHA_DNA5  0   /* reads: SeqAn target Dna5Q, valid characters ACGTN (either case)                  */
#define DYIBF_ALPHA_IUPAC 1   /* contigs of addFastaFile: target IupacString, valid the 16 IUPAC letters           */

/* readRecords(me.names, me.seqs, fileIn) of loadReads (src/mapper.h:451, src/store_seqs.h:175-188) and the readRecord
 * loop of addFastaFile (d_bloom_filter.h:127-147), i.e. SeqAn's FASTA / FASTQ tokenizer (seq_io/fasta_fastq.h:268-345),
 * on the device: `text` (n_bytes of an uncompressed FASTA or FASTQ file, or of a chunk that holds whole records) ->
 * the sequence characters of all records with whitespace removed (out_seqs, case kept: feed them to dyibf_query_batch /
 * dyibf_add_sequences with DYIBF_ENC_ASCII) and out_limits[n_seqs + 1] = limit_base + the start of every record
 * (a ConcatDirect StringSet's limits).  Ids and qualities are not extracted: the filter does not read them.
 * *n_seqs / *n_bases always receive the sizes; out_seqs == NULL and out_limits == NULL is a sizing call.
 * FASTQ in the four-line layout goes through the streaming tokeniser (405 GB/s of text); anything else SeqAn's sequential
 * reader accepts — sequences / qualities wrapped over several lines, blank lines between records — through a speculative
 * one (every '@' after white space is parsed as if it opened a record, pointer jumping picks the real ones; tens of GB/s).
 * Errors: DYIBF_ERR_FORMAT where SeqAn throws (a character outside the alphabet, no record in a non-empty text, a
 * FASTQ quality string whose length differs from its sequence's) — and for the one layout SeqAn accepts and this parser does
 * not: a FASTQ record that starts right behind the previous record's qualities with no white space in between;
 * DYIBF_ERR_CAPACITY when an output buffer is too small.
 * mem says where text, out_seqs and out_limits live (all three on the same side). */
int  dyibf_parse_records(int device, const char* text, uint64_t n_bytes, int format, int alphabet,
                         uint8_t* out_seqs, uint64_t capacity_bytes, uint64_t* out_limits, uint64_t capacity_seqs,
                         uint64_t limit_base, uint64_t* n_seqs, uint64_t* n_bases, int mem, void* stream);

/* loadReads (src/mapper.h:444-469) + clasifyLoadedReads (d_mapper.h:308-375) from the TEXT of the reads file(s), in one
 * call: the text (host memory; mates_text NULL for single-end) is copied to the device once, tokenised there
 * (dyibf_parse_records), the threshold is taken from the batch's average read length and e_percent as the reference does
 * (d_mapper.h:313-315), and only the per-bin lists come back (as dyibf_classify).  The parsed batch stays resident for
 * dyibf_gather_reads (ASCII sequences).  *n_reads receives the number of reads (pairs), *threshold the threshold used
 * (0: the caller applies the reference's "everything everywhere" rule, d_mapper.h:320-333). */
int  dyibf_classify_text(dyibf_filter* f, const char* reads_text, uint64_t reads_bytes, const char* mates_text, uint64_t mates_bytes,
                         int format, int e_percent, uint64_t* bin_offsets, uint32_t* read_ids, uint64_t capacity, uint64_t* total,
                         uint32_t* n_reads, uint32_t* threshold);
/* The same as a ticket (see dyibf_classify_begin): upload + tokenising of this batch's text overlap the query of the batches
 * begun before it — the double-buffered reader in front of the filter (src/file_prefetched.h:137-205).  The call returns once
 * the records are counted (it waits for the text upload and the tokeniser, not for the query); the text may be reused then.
 * The lists come back through dyibf_classify_end. */
int  dyibf_classify_text_begin(dyibf_filter* f, const char* reads_text, uint64_t reads_bytes, const char* mates_text, uint64_t mates_bytes,
                               int format, int e_percent, uint32_t* n_reads, uint32_t* threshold, dyibf_ticket** ticket);

/* ---- several GPUs, one process ----------------------------------------------------------------
 *
 * The reference's callers are single processes: spawnDisMapper constructs ONE filter (src/d_mapper.h:911-922),
 * clasifyLoadedReads splits ONE read batch over its worker threads (:338-375), build_filter splits the bins over its
 * threads (src/d_build_filter.cpp:227-260).  A group is the same with GPUs as the workers (dyibf_comm_init
 * = dyibf_group_create / dyibf_group_load_file, dyibf_broadcast_filter, dyibf_shard_bins; §8(e)):
 *   DYIBF_GROUP_READ_SHARDED  every device holds a replica (loaded once, fanned out over NVLink peer copies); device r
 *                             filters rows [r n/G, (r+1) n/G) (dyibf_shard_reads) of the caller's one host batch; per-bin lists
 *                             concatenate in device order = ascending ids.
 *   DYIBF_GROUP_BIN_SHARDED   for filters larger than one GPU's HBM: device r owns the word columns [r W/G, (r+1) W/G) of
 *                             every hash block = bins dyibf_shard_bins(r) as a self-contained filter with the same number of
 *                             blocks; every device sees every read; lists of the bin ranges concatenate.  Needs G | W.
 * Neither needs a reduction (the threshold is per bin and per read).  Results are identical to one device's.
 * dyibf_group_add_sequences routes every sequence to the device owning its bin's word column (no two devices OR into the
 * same word) and, for replicas, then exchanges the columns between the devices' HBM over NVLink so that all replicas are
 * identical again.  dyibf_group_member gives the per-device handle (options, profiling, device pointers).
 * dyibf_group_classify_begin hands whole batches to successive replicas (tickets end with dyibf_classify_end). */
#define DYIBF_GROUP_READ_SHARDED 0
#define DYIBF_GROUP_BIN_SHARDED  1
int  dyibf_shard_bins(uint32_t bins, int rank, int nranks, uint32_t* first_bin, uint32_t* n_bins);
int  dyibf_shard_reads(uint32_t n_reads, int rank, int nranks, uint32_t* first, uint32_t* last);
int  dyibf_group_create(uint32_t bins, uint32_t n_hash, uint32_t kmer, uint64_t nbits, const int* devices, int n_devices,
                        int mode, dyibf_group** out);
int  dyibf_group_load_file(const char* path, const int* devices, int n_devices, int mode, dyibf_group** out);
/* SeqAnBloomFilter(fileName, n_bins, n_hash_func, kmer_size, vec_size), d_bloom_filter.h:59-77, over several GPUs: like
 * dyibf_load_file_expect the members take bins / n_hash / kmer from the CALLER and only the file's bit count is checked
 * (DYIBF_ERR_SIZE_MISMATCH, the reference's message); the trailer's own values are not consulted */
int  dyibf_group_load_file_expect(const char* path, uint32_t bins, uint32_t n_hash, uint32_t kmer, uint64_t nbits,
                                  const int* devices, int n_devices, int mode, dyibf_group** out);
void dyibf_group_destroy(dyibf_group* g);
int  dyibf_group_get_params(const dyibf_group* g, dyibf_params* whole, int* n_devices, int* mode);
int  dyibf_group_member(dyibf_group* g, int index, dyibf_filter** member, uint32_t* first_bin, uint32_t* n_bins);
int  dyibf_group_set_option(dyibf_group* g, int option, int64_t value);
int  dyibf_group_save_file(dyibf_group* g, const char* path);
/* replicate member src_member's words to every other replica (after a build through that member's own handle) */
int  dyibf_broadcast_filter(dyibf_group* g, int src_member);
int  dyibf_group_classify(dyibf_group* g, const uint8_t* seqs, const uint64_t* offsets, uint32_t n_seqs, int encoding,
                          int group, int derive_rc, uint32_t threshold, uint64_t* bin_offsets, uint32_t* read_ids,
                          uint64_t capacity, uint64_t* total);
int  dyibf_group_classify_fetch(dyibf_group* g, uint64_t* bin_offsets, uint32_t* read_ids, uint64_t capacity, uint64_t* total);
int  dyibf_group_classify_begin(dyibf_group* g, const uint8_t* seqs, const uint64_t* offsets, uint32_t n_seqs, int encoding,
                                int group, int derive_rc, uint32_t threshold, dyibf_ticket** ticket);
int  dyibf_group_add_sequences(dyibf_group* g, const uint8_t* seqs, const uint64_t* offsets, uint32_t n_seqs,
                               const uint32_t* bin_ids, int encoding);
int  dyibf_group_clear_bins(dyibf_group* g, const uint32_t* bins, uint32_t n);

/* ---- options ------------------------------------------------------------------------------- */

/* DYIBF_OPT_EARLY_EXIT (default 1).  whichBins (d_bloom_filter.h:178-217) looks every k-mer of a sequence up before it
 * thresholds the counts (:219-223).  With this option on, and when no per-bin counts are requested, the query kernel
 * stops a sequence as soon as its bin set is decided: every bin has either reached the threshold, was already selected
 * by an earlier strand/mate of the same row (the selection only ever ORs), or cannot reach the threshold with the
 * k-mers left (count + remaining < threshold).  The rows are bit-identical to the full pass (checked against the oracle at
 * every threshold in tests/test_gpu_parity.py); only memory traffic drops (150 bp, k=18, e=3: 59 % of the lookups, 1.66 x
 * the reads/s).  On by default since round 2 — the mapper never asks for counts; 0 restores the reference's every-k-mer
 * pass, which is what bench.py's headline value / e2e / roofline measure (fixed algorithmic bytes per read). */
#define DYIBF_OPT_EARLY_EXIT 1
/* DYIBF_OPT_BUILD_BUCKET_MIB (default -1 = auto: what the call's updates need, up to 64 GiB, at most half of the free HBM; 0 = off).  _insertKmer's
 * single-bit ORs land at uniformly random words of the table.  For tables larger than the L2 and calls with at least 32 M
 * updates, dyibf_add_sequences first appends the updates as 32-bit records to per-region buckets in a scratch buffer of
 * this size and then ORs them in region by region, so that DRAM sees streaming instead of random sector traffic.  The
 * filter is bit-identical either way (OR is commutative and idempotent). */
#define DYIBF_OPT_BUILD_BUCKET_MIB 2
/* DYIBF_OPT_BUILD_MODE (default 0 = auto).  How dyibf_add_sequences turns _insertKmer's random single-bit ORs into memory traffic:
 *   1 direct    one atomic OR per (k-mer, hash function) straight into the table (what small calls and L2-sized tables use);
 *   2 bucketed  the radix-split + apply build described at DYIBF_OPT_BUILD_BUCKET_MIB;
 *   3 sliced    per-bin Bloom bit vectors (P bits = 16 MiB at 16 GiB / 1024 bins) that stay in the L2 while a bin's genome is
 *               hashed, then interleaved into the table with bit-matrix transposes — one L2 atomic per update and two streaming
 *               passes; needs the batch clustered by bin (build_filter adds one FASTA file per bin) and a scratch of all the
 *               bins' vectors (as large as the table: 16 GiB at cfg2) or, when that does not fit half of the free HBM, of
 *               256 / 128 / 64 bins' vectors with one pass per column group.
 * Auto picks sliced for bin-clustered batches of at least 256 M updates into tables larger than the L2, else bucketed (at
 * least 32 M updates) / direct.
 * The filter is bit-identical in every mode. */
#define DYIBF_OPT_BUILD_MODE 5
#define DYIBF_BUILD_AUTO     0
#define DYIBF_BUILD_DIRECT   1
#define DYIBF_BUILD_BUCKETED 2
#define DYIBF_BUILD_SLICED   3
/* read-only (dyibf_get_option): the DYIBF_BUILD_* path the last dyibf_add_sequences of the handle took */
#define DYIBF_OPT_LAST_BUILD_MODE 6
/* DYIBF_OPT_BUILD_CHUNK_MIB (default 0 = auto): dyibf_add_sequences(host) stages whole sequences in chunks of about this
 * many MiB of bases, alternating between two streams so that the copy of one chunk overlaps the kernels of the other.  Auto:
 * 64 MiB under the sliced build, 256 MiB under the bucketed / direct ones.  Pageable caller memory is staged through pinned
 * pieces by helper threads (the driver's own pageable path moves ~10 GB/s and holds the calling thread). */
#define DYIBF_OPT_BUILD_CHUNK_MIB 3
/* DYIBF_OPT_QUERY_CHUNK_MIB (default 0 = auto): dyibf_query_batch(host) / dyibf_classify stage the batch in chunks of rows,
 * alternating between two streams so that the upload of one chunk overlaps the kernels of the other.  Auto: chunks of about
 * 48 MiB of staged data, and a batch that would fit one chunk (the reference's 100 000-read batches, src/mapper.h:109) is
 * still cut into six (at least 8192 rows each) so that its own upload and query overlap.  A value > 0 fixes the chunk
 * size in MiB and switches the six-way cut off. */
#define DYIBF_OPT_QUERY_CHUNK_MIB 4
int  dyibf_set_option(dyibf_filter* f, int option, int64_t value);
int  dyibf_get_option(const dyibf_filter* f, int option, int64_t* value);

/* ---- instrumentation ----------------------------------------------------------------------- */
int  dyibf_profile_enable(dyibf_filter* f, int on);
int  dyibf_profile_get(dyibf_filter* f, dyibf_profile* out, int reset);
/* kernels this handle has launched since it was created (always counted, no profiling needed) */
uint64_t dyibf_launch_count(const dyibf_filter* f);
/* the sticky DYIBF_DEVERR_* bits of the handle; waits for the device first */
int  dyibf_device_errors(dyibf_filter* f, uint32_t* flags, int reset);

/* pinned host memory helpers (callers that want full PCIe rate allocate their buffers here) */
int  dyibf_host_alloc(void** ptr, size_t bytes);
void dyibf_host_free(void* ptr);

#ifdef __cplusplus
}
#endif
#endif /* DYIBF_H */
