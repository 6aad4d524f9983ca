/*
 * oracle/ibf_oracle.c — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C CPU restatement of DREAM-Yara's interleaved-Bloom-filter (IBF) read->bin prefilter,
 * written from the algorithm in /root/reference (temehi/dream_yara @ e1bf6b5); every function
 * cites the reference file:line it follows.  It is the checker the CUDA path is compared with.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load it; the product (dream_yara_b200/, libdyibf.so) never does and has no CPU fallback.
 *
 * PARITY IS PINNED: tests/test_oracle_golden.py checks this file against
 *   - the reference's own gold md5 (tests/gold/64-viral-binned-genomes.filter.md5),
 *   - golden vectors produced in the build container by the UNMODIFIED reference compiled from
 *     /root/reference (oracle/ref_shim.cpp, oracle/ref_mapper_shim.cpp -> oracle/_ref/*.so),
 *     committed under tests/golden/ with the generating script tests/golden/make_golden.py,
 *   - and, where oracle/_ref/libref_ibf.so is present, against that library directly.
 *
 * Notation: B bins, W = ceil(B/64) words per hash block, S = 64 W bits per block,
 * N = nbits (incl. 256 metadata bits), P = (N-256)/S hash positions, h hash functions, k k-mer size.
 */
#include <stdint.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>
#include <math.h>

#define IBF_SEED   0x90b45d39fb6da1faULL   /* d_bloom_filter.h:329  _seedValue  */
#define IBF_SHIFT  27                      /* d_bloom_filter.h:328  _shiftValue */
#define IBF_META_BITS 256                  /* d_misc_options.h:40   filterMetadataSize */
#define IBF_MAX_HASH 16

typedef struct {
    uint32_t bins;      /* _noOfBins      */
    uint32_t h;         /* _noOfHashFunc  */
    uint32_t k;         /* _kmerSize      */
    uint32_t W;         /* _binIntWidth   */
    uint64_t S;         /* _blockBitSize  */
    uint64_t nbits;     /* _noOfBits      */
    uint64_t P;         /* _noOfHashPos   */
    uint64_t pre[IBF_MAX_HASH]; /* _preCalcValues */
} ibf_params;

/* _init, d_bloom_filter.h:245-254 */
int ibf_params_init(ibf_params *p, uint32_t bins, uint32_t h, uint32_t k, uint64_t nbits)
{
    if (!bins || !h || h > IBF_MAX_HASH || !k || k > 32 || nbits < IBF_META_BITS + 64) return -1;
    p->bins = bins; p->h = h; p->k = k; p->nbits = nbits;
    p->W = (uint32_t)ceilf((float)bins / 64.0f);          /* std::ceil((float)_noOfBins / INT_WIDTH) */
    p->S = (uint64_t)p->W * 64u;
    p->P = (nbits - IBF_META_BITS) / p->S;
    if (!p->P) return -1;
    for (uint32_t i = 0; i < h; ++i)
        p->pre[i] = (uint64_t)i ^ ((uint64_t)k * IBF_SEED); /* i ^ (_kmerSize * _seedValue) */
    return 0;
}

uint64_t ibf_nwords(uint64_t nbits) { return (nbits + 63) >> 6; }   /* sdsl capacity()>>6 */

/* char -> Iupac -> Dna5: A,C,G,T/U (either case) -> 0..3, everything else -> N=4
 * (seqan basic/alphabet_residue_tabs.h:270-290 via alphabet_residue.h:728-731) */
uint8_t ibf_dna5_from_char(char c)
{
    switch (c) {
        case 'A': case 'a': return 0;
        case 'C': case 'c': return 1;
        case 'G': case 'g': return 2;
        case 'T': case 't': case 'U': case 'u': return 3;
        default: return 4;
    }
}

void ibf_encode_ascii(const char *text, uint64_t len, uint8_t *codes)
{
    for (uint64_t i = 0; i < len; ++i) codes[i] = ibf_dna5_from_char(text[i]);
}

/* Dna5Q byte -> Dna5 (alphabet_residue.h:1220-1244): >= 252 is N, else low two bits */
uint8_t ibf_dna5_from_dna5q(uint8_t v) { return v >= 252 ? 4 : (v & 3); }

/* reverse complement on Dna5 codes; N stays N (modifier_functors.h:170; store_seqs.h:271-290) */
void ibf_revcomp(const uint8_t *codes, uint64_t len, uint8_t *out)
{
    for (uint64_t i = 0; i < len; ++i) {
        uint8_t c = codes[len - 1 - i];
        out[i] = c >= 4 ? 4 : (uint8_t)(3 - c);
    }
}

/* SeqAn Shape<Dna,SimpleShape> rolling hash as whichBins/_addKmers drive it
 * (index/shape_base.h:403-411 resize, :479-494 hashInit, :699-714 hashNext;
 *  d_bloom_filter.h:168-176, :306-314).  The first character enters RAW (N = 4), all later
 * ones through the Dna5->Dna cast (&3, alphabet_residue.h:680-683).  Returns #hashes. */
uint64_t ibf_kmer_hashes(const uint8_t *codes, uint64_t len, uint32_t k, uint64_t *out)
{
    if (len < k) return 0;
    uint64_t leftFactor = 1;
    for (uint32_t i = 1; i < k; ++i) leftFactor *= 4;          /* _intPow(4, span-1), wraps never: k<=32 */
    uint64_t hv = codes[0];                                    /* ordValue(*it): raw Dna5 value */
    for (uint32_t i = 2; i < k; ++i) hv = hv * 4 + (codes[i - 1] & 3u);
    uint64_t left = 0;
    uint64_t n = len - k + 1;
    for (uint64_t i = 0; i < n; ++i) {
        hv = (hv - left * leftFactor) * 4 + (codes[i + k - 1] & 3u);
        left = codes[i] & 3u;                                  /* leftChar is a Dna: masked */
        out[i] = hv;
    }
    return n;
}

/* getHashValue, d_bloom_filter.h:286-291, applied to pre[i]*kmerHash (:183,:297): bit offset of the block */
static inline uint64_t ibf_block_bit(const ibf_params *p, uint32_t i, uint64_t kmerHash)
{
    uint64_t v = p->pre[i] * kmerHash;
    v ^= v >> IBF_SHIFT;
    v %= p->P;
    v *= p->S;
    return v;
}

/* exported for tests: the h block indices (v / S) of one k-mer hash */
void ibf_positions(const ibf_params *p, uint64_t kmerHash, uint64_t *block_idx)
{
    for (uint32_t i = 0; i < p->h; ++i) block_idx[i] = ibf_block_bit(p, i, kmerHash) / p->S;
}

/* addKmers/_addKmers/_insertKmer, d_bloom_filter.h:91-94,293-315 (caller skips len<k, :142) */
void ibf_add_kmers(uint64_t *words, const ibf_params *p, const uint8_t *codes, uint64_t len, uint32_t bin)
{
    if (len < p->k) return;
    uint64_t n = len - p->k + 1;
    uint64_t *hs = (uint64_t *)malloc(n * sizeof(uint64_t));
    ibf_kmer_hashes(codes, len, p->k, hs);
    for (uint64_t j = 0; j < n; ++j)
        for (uint32_t i = 0; i < p->h; ++i) {
            uint64_t bit = ibf_block_bit(p, i, hs[j]) + bin;
            words[bit >> 6] |= 1ULL << (bit & 63);             /* int_vector.hpp:721-728 */
        }
    free(hs);
}

/* per-bin k-mer counts of whichBins (the local `counts`, d_bloom_filter.h:165,178-217).
 * counts[bins] is ADDED to (caller zeroes). len<k selects nothing (reference: UB, :163). */
void ibf_count_bins(const uint64_t *words, const ibf_params *p, const uint8_t *codes, uint64_t len, uint16_t *counts)
{
    if (len < p->k) return;
    uint64_t n = len - p->k + 1;
    uint64_t *hs = (uint64_t *)malloc(n * sizeof(uint64_t));
    ibf_kmer_hashes(codes, len, p->k, hs);
    for (uint64_t j = 0; j < n; ++j) {
        uint64_t base[IBF_MAX_HASH];
        for (uint32_t i = 0; i < p->h; ++i) base[i] = ibf_block_bit(p, i, hs[j]) >> 6;
        for (uint32_t w = 0; w < p->W; ++w) {
            uint64_t t = words[base[0] + w];                   /* get_int(vecIndices[0], 64) */
            for (uint32_t i = 1; i < p->h; ++i) t &= words[base[i] + w];
            while (t) {                                        /* :196-211 = plain set-bit iteration */
                uint32_t b = (uint32_t)__builtin_ctzll(t);
                t &= t - 1;
                uint32_t bin = 64u * w + b;
                if (bin < p->bins) ++counts[bin];              /* pad bits >= bins: never set by build; ignored */
            }
        }
    }
    free(hs);
}

/* whichBins(selected, text, threshold), d_bloom_filter.h:161-224: OR-accumulates, never clears */
void ibf_which_bins(const uint64_t *words, const ibf_params *p, const uint8_t *codes, uint64_t len,
                    uint16_t threshold, uint8_t *selected)
{
    uint16_t *counts = (uint16_t *)calloc(p->bins, sizeof(uint16_t));
    ibf_count_bins(words, p, codes, len, counts);
    if (len >= p->k)
        for (uint32_t b = 0; b < p->bins; ++b)
            if (counts[b] >= threshold) selected[b] = 1;
    free(counts);
}

/* DisOptions::getThreshold, d_mapper.h:83-92; errorRate is a float set to e/100.0 (d_mapper.cpp:340-342,
 * mapper.h:68) and the product errorRate*readLen is truncated to uint16_t */
uint16_t ibf_threshold(int e_percent, uint32_t k, uint16_t readLen)
{
    float errorRate = (float)(e_percent / 100.0);
    uint16_t maxError = (uint16_t)(errorRate * readLen);
    if (k * (1u + maxError) > readLen) return 0;
    return (uint16_t)(readLen - k * (1u + maxError) + 1);
}

/* the filtering stage, clasifyLoadedReads d_mapper.h:308-375, over flat buffers.
 * seqs: n_seqs Dna5-code sequences (offsets[n_seqs+1]) in the layout loadReads +
 * appendReverseComplement produce: SE [fwd | rc], PE [m1 | m2 | m1 rc | m2 rc].
 * numReads = n_seqs/2 (SE) or n_seqs/4 (PE). out_rows: [numReads x W] u64 bitmask
 * (bit b of word w = bin 64w+b), zeroed here. */
void ibf_classify(const uint64_t *words, const ibf_params *p, const uint8_t *seqs, const uint64_t *offsets,
                  uint32_t n_seqs, int paired, uint16_t threshold, uint64_t *out_rows)
{
    uint32_t group = paired ? 4 : 2;
    uint32_t numReads = n_seqs / group;
    #pragma omp parallel
    {
        uint8_t *sel = (uint8_t *)malloc(p->bins);
        #pragma omp for schedule(dynamic, 64)
        for (uint32_t r = 0; r < numReads; ++r) {
            memset(sel, 0, p->bins);
            for (uint32_t g = 0; g < group; ++g) {
                uint32_t s = r + g * numReads;
                ibf_which_bins(words, p, seqs + offsets[s], offsets[s + 1] - offsets[s], threshold, sel);
            }
            uint64_t *row = out_rows + (uint64_t)r * p->W;
            memset(row, 0, p->W * sizeof(uint64_t));
            for (uint32_t b = 0; b < p->bins; ++b)
                if (sel[b]) row[b >> 6] |= 1ULL << (b & 63);
        }
        free(sel);
    }
}

/* origReadIdMap as `--threads 1` fills it (d_mapper.h:360-368): per bin, ascending read ids.
 * CSR: bin_offsets[bins+1], read_ids[total]. Returns total; pass read_ids=NULL to size. */
uint64_t ibf_compact(const uint64_t *rows, uint32_t numReads, uint32_t bins, uint32_t *bin_offsets, uint32_t *read_ids)
{
    uint32_t W = (bins + 63) / 64;
    uint64_t total = 0;
    for (uint32_t b = 0; b < bins; ++b) {
        bin_offsets[b] = (uint32_t)total;
        for (uint32_t r = 0; r < numReads; ++r)
            if ((rows[(uint64_t)r * W + (b >> 6)] >> (b & 63)) & 1) {
                if (read_ids) read_ids[total] = r;
                ++total;
            }
    }
    bin_offsets[bins] = (uint32_t)total;
    return total;
}

/* reverseComplement of a Dna5Q string as appendReverseComplement applies it to the reads (store_seqs.h:271-290):
 * FunctorComplement<Dna5Q> (modifier_functors.h:276-284) complements the base and re-assigns the quality;
 * N (byte 252, no quality: alphabet_residue.h:287,357-365) stays 252. */
void ibf_revcomp_dna5q(const uint8_t *bytes, uint64_t len, uint8_t *out)
{
    for (uint64_t i = 0; i < len; ++i) {
        uint8_t b = bytes[len - 1 - i];
        out[i] = b >= 252 ? 252 : (uint8_t)((b & 0xFC) | (3 - (b & 3)));
    }
}

/* loadFilteredReads, d_mapper.h:393-443 (filterType != NONE), over flat buffers.
 * seqs/offsets: the main mapper's FULL read set (n_seqs sequences incl. reverse complements; any byte encoding —
 * sequences are copied verbatim).  ids[m] = origReadIdMap[bin] as the filter left it.  Output: the child read set
 * (out_bytes, out_limits[n_out+1]) and the extended id list out_ids[n_out]; n_out = 2m (SE) or 4m (PE).
 * Returns n_out. */
uint64_t ibf_load_filtered_reads(const uint8_t *seqs, const uint64_t *offsets, uint32_t n_seqs, int paired,
                                 const uint32_t *ids, uint32_t m, uint8_t *out_bytes, uint64_t *out_limits, uint32_t *out_ids)
{
    uint32_t numReads = n_seqs / 2;            /* getReadsCount, bits_reads.h:57-62 */
    uint32_t numPairs = n_seqs / 4;            /* getPairsCount, bits_reads.h:64-69 */
    uint64_t n_out = 0, nb = 0;
    uint32_t numFiltered = m;
    out_limits[0] = 0;
    for (uint32_t i = 0; i < m; ++i) out_ids[i] = ids[i];
    #define IBF_APPEND(src_) do { uint32_t s__ = (src_); uint64_t l__ = offsets[s__ + 1] - offsets[s__]; \
        memcpy(out_bytes + nb, seqs + offsets[s__], l__); nb += l__; out_limits[++n_out] = nb; } while (0)
    for (uint32_t i = 0; i < numFiltered; ++i) IBF_APPEND(out_ids[i]);                       /* :413-417 forward reads */
    if (paired) {
        for (uint32_t i = 0; i < numFiltered; ++i) {                                          /* :423-428 mates */
            IBF_APPEND(out_ids[i] + numPairs);
            out_ids[numFiltered + i] = out_ids[i] + numPairs;
        }
        numFiltered *= 2;                                                                     /* :429 */
    }
    for (uint32_t i = 0; i < numFiltered; ++i) {                                              /* :433-438 reverse reads */
        IBF_APPEND(out_ids[i] + numReads);
        out_ids[numFiltered + i] = out_ids[i] + numReads;
    }
    #undef IBF_APPEND
    return n_out;
}

/* clearBins, d_bloom_filter.h:99-123: clear bit p*S+bin for every hash block */
void ibf_clear_bins(uint64_t *words, const ibf_params *p, const uint32_t *bins, uint32_t n)
{
    for (uint64_t hb = 0; hb < p->P; ++hb)
        for (uint32_t j = 0; j < n; ++j) {
            uint64_t bit = hb * p->S + bins[j];
            words[bit >> 6] &= ~(1ULL << (bit & 63));
        }
}

/* unaligned 64-bit get/set on the bit vector (sdsl bits::read_int / write_int, bits.hpp:767-778) */
static uint64_t get_int64(const uint64_t *words, uint64_t bit)
{
    uint64_t w = bit >> 6, o = bit & 63;
    return o ? (words[w] >> o) | (words[w + 1] << (64 - o)) : words[w];
}
static void set_int64(uint64_t *words, uint64_t bit, uint64_t x)
{
    uint64_t w = bit >> 6, o = bit & 63;
    if (!o) { words[w] = x; return; }
    words[w] = (words[w] & ((1ULL << o) - 1)) | (x << o);
    words[w + 1] = (words[w + 1] & ~((1ULL << o) - 1)) | (x >> (64 - o));
}

/* _setMetadata, d_bloom_filter.h:267-277: trailer [bins, h, k] at bits N-256, N-192, N-128 */
void ibf_set_metadata(uint64_t *words, const ibf_params *p)
{
    uint64_t m = p->nbits - IBF_META_BITS;
    set_int64(words, m, p->bins);
    set_int64(words, m + 64, p->h);
    set_int64(words, m + 128, p->k);
}

/* _getMetadata, d_bloom_filter.h:255-265 (u32/u8/u8 truncation as the members have) */
void ibf_get_metadata(const uint64_t *words, uint64_t nbits, uint32_t *bins, uint32_t *h, uint32_t *k)
{
    uint64_t m = nbits - IBF_META_BITS;
    *bins = (uint32_t)get_int64(words, m);
    *h = (uint8_t)get_int64(words, m + 64);
    *k = (uint8_t)get_int64(words, m + 128);
}

/* save, d_bloom_filter.h:150-154 -> sdsl::store_to_file (io.hpp:857-873) -> serialize
 * (int_vector.hpp:1463-1472): 8-byte header (width 1 << 56 | nbits) + ceil(nbits/64) LE words */
int ibf_save(uint64_t *words, const ibf_params *p, const char *path)
{
    ibf_set_metadata(words, p);
    FILE *f = fopen(path, "wb");
    if (!f) return -1;
    uint64_t header = (1ULL << 56) | p->nbits;
    uint64_t nw = ibf_nwords(p->nbits);
    int ok = fwrite(&header, 8, 1, f) == 1 && fwrite(words, 8, nw, f) == nw;
    ok = (fclose(f) == 0) && ok;
    return ok ? 0 : -1;
}

/* header peek (int_vector.hpp:515-531): returns nbits or 0 */
uint64_t ibf_file_nbits(const char *path)
{
    FILE *f = fopen(path, "rb");
    if (!f) return 0;
    uint64_t header = 0;
    size_t got = fread(&header, 8, 1, f);
    fclose(f);
    if (got != 1 || (header >> 56) != 1) return 0;
    return header & ((1ULL << 56) - 1);
}

/* ctor-from-file, d_bloom_filter.h:79-89: caller allocates ibf_nwords(ibf_file_nbits()) words */
int ibf_load(const char *path, uint64_t *words, uint64_t nwords, ibf_params *p)
{
    FILE *f = fopen(path, "rb");
    if (!f) return -1;
    uint64_t header = 0;
    if (fread(&header, 8, 1, f) != 1) { fclose(f); return -1; }
    uint64_t nbits = header & ((1ULL << 56) - 1);
    if (ibf_nwords(nbits) != nwords || fread(words, 8, nwords, f) != nwords) { fclose(f); return -1; }
    fclose(f);
    uint32_t bins, h, k;
    ibf_get_metadata(words, nbits, &bins, &h, &k);
    return ibf_params_init(p, bins, h, k, nbits);
}

/* ---- reading sequence records (the step before the filter) ----------------------------------------------------
 * readRecords(me.names, me.seqs, fileIn) (src/store_seqs.h:175-188, called by loadReads src/mapper.h:451) and the
 * readRecord loop of addFastaFile (src/d_bloom_filter.h:127-147) over SeqAn's tokenizer:
 * seqan seq_io/fasta_fastq.h:268-282 (Fasta), :302-345 (Fastq, qualities inside seq), sequence_file.h:340-343 (loop
 * until atEnd), stream/tokenization.h:51-58 (IsInAlphabet), :145-149 (IsWhitespace = ' ' '\t' '\n' '\r'). */
static int ibf_is_ws(unsigned char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r'; }
static int ibf_is_nl(unsigned char c) { return c == '\n' || c == '\r'; }

/* IsInAlphabet<TValue>: convert<char>(TValue(c)) == toUpperValue(c).  Dna5Q (reads): ACGTN either case; Iupac (contigs
 * of addFastaFile): the sixteen letters "UTAWCYMHGKRDSBVN" (alphabet_residue_tabs.h:58-62) either case. */
int ibf_in_alphabet(unsigned char c, int iupac)
{
    unsigned char u = (c >= 'a' && c <= 'z') ? (unsigned char)(c - 32) : c;
    if (!iupac) return u == 'A' || u == 'C' || u == 'G' || u == 'T' || u == 'N';
    return u != 0 && strchr("UTAWCYMHGKRDSBVN", u) != NULL;
}

/* format 0 = Fasta, 1 = Fastq.  out_seqs receives the sequence characters as they stand in the text (ignored
 * whitespace removed, case kept), limits[r] / limits[r+1] bracket record r.  Returns the number of records, -1 on a
 * ParseError / UnexpectedEnd of the reference, -2 when an output capacity is exceeded. */
int64_t ibf_parse_records(const char *text, uint64_t n, int format, int iupac, uint8_t *out_seqs, uint64_t cap_bytes,
                          uint64_t *limits, uint64_t cap_seqs, uint64_t *n_bases)
{
    const unsigned char *t = (const unsigned char *)text;
    const unsigned char begin = format ? '@' : '>';
    uint64_t i = 0, nb = 0; int64_t nr = 0;
    limits[0] = 0;
    while (i < n) {                                            /* for (; !atEnd(file); ) readRecord(...) */
        while (i < n && t[i] != begin) ++i;                    /* skipUntil(iter, TFastaBegin / TFastqBegin) */
        if (i >= n) return -1;                                 /* skipOne at the end: UnexpectedEnd */
        ++i;                                                   /* skipOne */
        while (i < n && !ibf_is_nl(t[i])) ++i;                 /* readLine(meta): to the newline ... */
        if (i < n) { if (t[i] == '\r' && i + 1 < n && t[i + 1] == '\n') ++i; ++i; }   /* ... and over "\r\n" / "\n" / "\r" */
        if ((uint64_t)nr >= cap_seqs) return -2;
        const unsigned char stop = format ? '+' : '>';
        uint64_t len = 0;
        for (; i < n && t[i] != stop; ++i) {                   /* readUntil(seq, iter, stop, IgnoreOrAssert) */
            if (ibf_is_ws(t[i])) continue;
            if (!ibf_in_alphabet(t[i], iupac)) return -1;      /* AssertFunctor<IsInAlphabet>: ParseError */
            if (nb >= cap_bytes) return -2;
            out_seqs[nb++] = t[i]; ++len;
        }
        if (format) {
            if (i >= n) return -1;                             /* skipOne(iter, '+') at the end: UnexpectedEnd */
            ++i;
            while (i < n && !ibf_is_nl(t[i])) ++i;             /* skipLine */
            if (i < n) { if (t[i] == '\r' && i + 1 < n && t[i + 1] == '\n') ++i; ++i; }
            uint64_t left = len;                               /* CountDownFunctor<Not<IsWhitespace>>(length(seq)) */
            for (; i < n && left; ++i) if (!ibf_is_ws(t[i])) --left;
            while (i < n && ibf_is_ws(t[i])) ++i;              /* skipUntil(NotFunctor<IsWhitespace>) */
            if (i < n && t[i] != '@') return -1;               /* "quality string is expected to be of the same length" */
        }
        limits[++nr] = nb;
    }
    *n_bases = nb;
    return nr;
}

/* ---- kmer_direct filter: SeqAnKDXFilter, src/d_kdx_filter.h ------------------------------------------------------
 * Same bit layout and trailer as the IBF (trailer [bins, 0, k, 0]: _setMetadata :234-239 writes no hash count), but the
 * block of a k-mer is its SeqAn hash value itself.  _init (:212-217) sets _noOfHashPos = (N-256)/_binIntWidth — 64 times
 * the number of blocks the vector really has — and `kmerHash >= _noOfHashPos` (:171, :288) is the only guard, so for
 * hashes in [(N-256)/S, (N-256)/W) the reference reads / writes beyond the vector (undefined behaviour; reachable with
 * leading-N sequences or a vector smaller than 4^k blocks).  Here those k-mers are skipped; inside the table the
 * restatement is exact.  Use ibf_params_init-like fields: W, S as for the IBF; P = true number of blocks. */
int kdx_params_init(ibf_params *p, uint32_t bins, uint32_t k, uint64_t nbits)
{
    if (!bins || !k || k > 32 || nbits < IBF_META_BITS + 64) return -1;
    memset(p, 0, sizeof *p);
    p->bins = bins; p->h = 0; p->k = k; p->nbits = nbits;
    p->W = (uint32_t)ceilf((float)bins / 64.0f);
    p->S = (uint64_t)p->W * 64u;
    p->P = (nbits - IBF_META_BITS) / p->S;
    return p->P ? 0 : -1;
}

/* _addKmers, d_kdx_filter.h:279-292 */
void kdx_add_kmers(uint64_t *words, const ibf_params *p, const uint8_t *codes, uint64_t len, uint32_t bin)
{
    if (len < p->k) return;
    uint64_t n = len - p->k + 1;
    uint64_t *hashes = (uint64_t *)malloc(n * sizeof(uint64_t));
    ibf_kmer_hashes(codes, len, p->k, hashes);
    for (uint64_t i = 0; i < n; ++i) {
        if (hashes[i] >= p->P) continue;                       /* reference: >= (N-256)/W, then out of bounds */
        uint64_t bit = p->S * hashes[i] + bin;                 /* _filterVector[_blockBitSize * kmerHash + binNo] = 1 */
        words[bit >> 6] |= 1ULL << (bit & 63);
    }
    free(hashes);
}

/* whichBins, d_kdx_filter.h:152-207: counts[bin] = number of k-mers of the text whose block has the bin's bit set */
void kdx_count_bins(const uint64_t *words, const ibf_params *p, const uint8_t *codes, uint64_t len, uint16_t *counts)
{
    memset(counts, 0, p->bins * sizeof(uint16_t));
    if (len < p->k) return;
    uint64_t n = len - p->k + 1;
    uint64_t *hashes = (uint64_t *)malloc(n * sizeof(uint64_t));
    ibf_kmer_hashes(codes, len, p->k, hashes);
    for (uint64_t i = 0; i < n; ++i) {
        if (hashes[i] >= p->P) continue;
        const uint64_t *blk = words + hashes[i] * p->W;        /* get_int(kmerHash * S + 64 batchNo, 64) */
        for (uint32_t w = 0; w < p->W; ++w) {
            uint64_t t = blk[w];
            while (t) {                                        /* :181-196: every set bit, LSB first */
                uint32_t b = 64 * w + (uint32_t)__builtin_ctzll(t);
                t &= t - 1;
                if (b < p->bins) ++counts[b];                  /* bits >= bins are never set by the build */
            }
        }
    }
    free(hashes);
}

/* the thresholding tail of whichBins (:200-205): OR-accumulates into `selected` */
void kdx_which_bins(const uint64_t *words, const ibf_params *p, const uint8_t *codes, uint64_t len, uint16_t threshold, uint8_t *selected)
{
    uint16_t *counts = (uint16_t *)malloc(p->bins * sizeof(uint16_t));
    kdx_count_bins(words, p, codes, len, counts);
    if (len >= p->k)
        for (uint32_t b = 0; b < p->bins; ++b) if (counts[b] >= threshold) selected[b] = 1;
    free(counts);
}
