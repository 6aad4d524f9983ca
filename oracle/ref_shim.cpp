// oracle/ref_shim.cpp — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Thin extern "C" wrapper that compiles the UNMODIFIED reference filter class
// (seqan::SeqAnBloomFilter, /root/reference/src/d_bloom_filter.h:41-331) from where it lies
// under /root/reference (never copied into this repo) into oracle/_ref/libref_ibf.so.
// It is the authority the C restatement (oracle/ibf_oracle.c) and the CUDA path are pinned to,
// and the `cpu_baseline.kind == "reference"` arm of bench.py.  Only tests/, smoke() and the
// cpu_baseline / --impl reference legs of bench.py may load it.
//
// Build: see oracle/Makefile (target _ref/libref_ibf.so). Needs /root/reference; the built .so
// travels to the GPU box with the gpurun snapshot.
//
// glibc >= 2.38 defines a C23 macro INT_WIDTH in <limits.h>; the reference declares
// `static const uint8_t INT_WIDTH` (src/d_misc_options.h:41) — undefine the macro first.
#include <climits>
#include <limits.h>
#undef INT_WIDTH

#include <string>
#include <vector>
#include <mutex>
#include <condition_variable>
#include <future>
#include <thread>
#include <cstring>
#include <dirent.h>
#include <sys/stat.h>

#include <seqan/index.h>
#include <seqan/seq_io.h>
#include <seqan/arg_parse.h>

#include "store_seqs.h"
#include "misc_timer.h"
#include "misc_types.h"
#include "bits_matches.h"
#include "misc_options.h"
#include "d_misc_options.h"
// the raw word array of the filter is a private member; the bench needs to hand the reference the
// very same filter words the GPU holds, and tests need to read them back without a 16 GiB file.
#define private public
#include "d_bloom_filter.h"
#include "d_kdx_filter.h"
#undef private

using namespace seqan;
typedef SeqAnBloomFilter<> TFilter;

static Dna5String codes_to_dna5(const uint8_t* codes, uint64_t len)
{
    Dna5String s;
    resize(s, len);
    for (uint64_t i = 0; i < len; ++i) s[i] = Dna5(codes[i] > 4 ? 4 : codes[i]);
    return s;
}

extern "C" {

void* ref_new(uint32_t bins, uint32_t h, uint32_t k, uint64_t nbits)
{
    return new TFilter(bins, (uint8_t)h, (uint8_t)k, nbits);
}

void* ref_load(const char* path) { return new TFilter(path); }

void ref_free(void* f) { delete (TFilter*)f; }

void ref_params(void* f, uint32_t* bins, uint32_t* h, uint32_t* k, uint64_t* nbits,
                uint32_t* W, uint64_t* P)
{
    TFilter* x = (TFilter*)f;
    *bins = x->_noOfBins; *h = x->_noOfHashFunc; *k = x->_kmerSize; *nbits = x->_noOfBits;
    *W = x->_binIntWidth; *P = x->_noOfHashPos;
}

// raw 64-bit words of the sdsl bit_vector (ceil(nbits/64) of them)
uint64_t* ref_words(void* f) { return (uint64_t*)((TFilter*)f)->_filterVector.data(); }
uint64_t ref_nwords(void* f) { return (((TFilter*)f)->_filterVector.bit_size() + 63) >> 6; }

// addKmers (d_bloom_filter.h:91) on a sequence of Dna5 codes 0..4
void ref_add_kmers(void* f, const uint8_t* codes, uint64_t len, uint32_t bin)
{
    Dna5String s = codes_to_dna5(codes, len);
    ((TFilter*)f)->addKmers(s, bin);
}

// addFastaFile (d_bloom_filter.h:127): the reference's own FASTA reader and Iupac->Dna5 conversion
int ref_add_fasta(void* f, const char* path, uint32_t bin)
{
    try { ((TFilter*)f)->addFastaFile(CharString(path), bin); }
    catch (...) { return -1; }
    return 0;
}

void ref_clear_bins(void* f, const uint32_t* bins, uint32_t n, uint32_t threads)
{
    std::vector<uint32_t> v(bins, bins + n);
    ((TFilter*)f)->clearBins(v, threads);
}

int ref_save(void* f, const char* path) { return ((TFilter*)f)->save(path) ? 0 : -1; }

// whichBins (d_bloom_filter.h:161): OR-accumulates into selected[bins] (bytes 0/1)
void ref_which_bins(void* f, const uint8_t* codes, uint64_t len, uint16_t threshold, uint8_t* selected)
{
    TFilter* x = (TFilter*)f;
    Dna5String s = codes_to_dna5(codes, len);
    std::vector<bool> sel(x->getNumberOfBins(), false);
    for (uint32_t b = 0; b < x->getNumberOfBins(); ++b) sel[b] = selected[b] != 0;
    x->whichBins(sel, s, threshold);
    for (uint32_t b = 0; b < x->getNumberOfBins(); ++b) selected[b] = sel[b];
}

// SeqAn rolling hash exactly as whichBins drives it (d_bloom_filter.h:168-176)
uint64_t ref_kmer_hashes(const uint8_t* codes, uint64_t len, uint32_t k, uint64_t* out)
{
    if (len < k) return 0;
    Dna5String text = codes_to_dna5(codes, len);
    Shape<Dna, SimpleShape> shape;
    resize(shape, k);
    hashInit(shape, begin(text));
    auto it = begin(text);
    uint64_t n = len - k + 1;
    for (uint64_t i = 0; i < n; ++i) { out[i] = hashNext(shape, it); ++it; }
    return n;
}

// reverse complement through SeqAn (store_seqs.h:271-290 uses the same functor on Dna5Q reads)
void ref_revcomp(const uint8_t* codes, uint64_t len, uint8_t* out)
{
    Dna5String s = codes_to_dna5(codes, len);
    reverseComplement(s);
    for (uint64_t i = 0; i < len; ++i) out[i] = ordValue(s[i]);
}

// char -> Iupac -> Dna5 as addFastaFile converts records (IupacString -> TString const &)
void ref_ascii_to_dna5(const char* text, uint64_t len, uint8_t* out)
{
    IupacString s;
    resize(s, len);
    for (uint64_t i = 0; i < len; ++i) s[i] = Iupac(text[i]);
    Dna5String d = s;
    for (uint64_t i = 0; i < len; ++i) out[i] = ordValue(d[i]);
}

// The filtering stage of the mapper (src/d_mapper.h:338-375), re-driven over flat buffers:
// `seqs` = n_seqs sequences of Dna5 codes (offsets[n_seqs+1]) laid out as loadReads() +
// appendReverseComplement() leave them: SE [fwd | rc], PE [m1 | m2 | m1 rc | m2 rc];
// numReads = n_seqs/2 (SE) or n_seqs/4 (PE). Contiguous read ranges over std::async tasks,
// one shared const filter, whichBins on every strand/mate of a read OR-ed into one vector<bool>.
// Output: bitmask rows [numReads x W] u64 (bit b of word w = bin 64w+b).
void ref_classify(void* f, const uint8_t* seqs, const uint64_t* offsets, uint32_t n_seqs, int paired,
                  uint16_t threshold, uint32_t threads, uint64_t* out_rows)
{
    TFilter const& filter = *(TFilter*)f;
    uint32_t bins = ((TFilter*)f)->getNumberOfBins();
    uint32_t W = (bins + 63) / 64;
    uint32_t numReads = paired ? n_seqs / 4 : n_seqs / 2;
    uint32_t numThr = threads ? threads : 1;
    uint32_t batchSize = numReads / numThr;
    if (batchSize * numThr < numReads) ++batchSize;
    std::vector<std::future<void>> tasks;
    for (uint32_t taskNo = 0; taskNo < numThr; ++taskNo)
    {
        tasks.emplace_back(std::async(std::launch::async, [=, &filter] {
            for (uint32_t readID = taskNo * batchSize; readID < numReads && readID < (taskNo + 1) * batchSize; ++readID)
            {
                std::vector<bool> selectedBins(bins, false);
                int n = paired ? 4 : 2;
                for (int g = 0; g < n; ++g)
                {
                    uint32_t s = readID + g * numReads;
                    Dna5String text = codes_to_dna5(seqs + offsets[s], offsets[s + 1] - offsets[s]);
                    filter.whichBins(selectedBins, text, threshold);
                }
                uint64_t* row = out_rows + (uint64_t)readID * W;
                for (uint32_t w = 0; w < W; ++w) row[w] = 0;
                for (uint32_t b = 0; b < bins; ++b)
                    if (selectedBins[b]) row[b >> 6] |= 1ULL << (b & 63);
            }
        }));
    }
    for (auto&& t : tasks) t.get();
}

// readRecords(me.names, me.seqs, fileIn) exactly as the mapper's loadReads drives it (src/store_seqs.h:175-188 with
// SeqConfig: Dna5Q, ConcatDirect; src/mapper.h:451), or — iupac != 0 — the readRecord loop of addFastaFile
// (src/d_bloom_filter.h:127-147: IupacString records, implicitly converted to Dna5 for addKmers; records shorter
// than k are NOT dropped here).  Output: Dna5 codes of all records + limits.  Returns the number of records, -1 when
// SeqAn throws (ParseError / UnexpectedEnd), -2 when a capacity is exceeded, -3 when the file cannot be opened.
int64_t ref_read_records(const char* path, int iupac, uint8_t* codes, uint64_t cap_bytes, uint64_t* limits, uint64_t cap_seqs,
                         uint64_t* n_bases)
{
    SeqFileIn in;
    if (!open(in, path)) return -3;
    uint64_t nb = 0; int64_t nr = 0;
    limits[0] = 0;
    try {
        if (iupac) {
            CharString id; IupacString seq;
            while (!atEnd(in)) {
                readRecord(id, seq, in);
                Dna5String d = seq;
                if ((uint64_t)nr >= cap_seqs || nb + length(d) > cap_bytes) return -2;
                for (uint64_t i = 0; i < length(d); ++i) codes[nb++] = ordValue(d[i]);
                limits[++nr] = nb;
            }
        } else {
            SeqStore<void, SeqConfig<void> > store;
            readRecords(store, in);
            nr = length(store.seqs);
            if ((uint64_t)nr > cap_seqs || lengthSum(store.seqs) > cap_bytes) return -2;
            for (int64_t r = 0; r < nr; ++r) {
                for (uint64_t i = 0; i < length(store.seqs[r]); ++i) codes[nb++] = ordValue(Dna5(store.seqs[r][i]));
                limits[r + 1] = nb;
            }
        }
    } catch (seqan::Exception const&) {
        return -1;
    }
    *n_bases = nb;
    return nr;
}

// ---- kmer_direct filter: the UNMODIFIED seqan::SeqAnKDXFilter (src/d_kdx_filter.h:41-311) ------------------------------
typedef SeqAnKDXFilter<> TKdx;
void* ref_kdx_new(uint32_t bins, uint32_t k, uint64_t nbits) { return new TKdx(bins, (uint8_t)k, nbits); }
void* ref_kdx_load(const char* path) { return new TKdx(path); }
void ref_kdx_free(void* f) { delete (TKdx*)f; }
uint64_t* ref_kdx_words(void* f) { return (uint64_t*)((TKdx*)f)->_filterVector.data(); }
uint64_t ref_kdx_nwords(void* f) { return (((TKdx*)f)->_filterVector.bit_size() + 63) >> 6; }
void ref_kdx_params(void* f, uint32_t* bins, uint32_t* k, uint64_t* nbits, uint64_t* hashpos)
{
    TKdx* x = (TKdx*)f;
    *bins = x->_noOfBins; *k = x->_kmerSize; *nbits = x->_noOfBits; *hashpos = x->_noOfHashPos;
}
void ref_kdx_add_kmers(void* f, const uint8_t* codes, uint64_t len, uint32_t bin)
{
    Dna5String s = codes_to_dna5(codes, len);
    ((TKdx*)f)->addKmers(s, bin);
}
int ref_kdx_add_fasta(void* f, const char* path, uint32_t bin)
{
    try { ((TKdx*)f)->addFastaFile(CharString(path), bin); } catch (...) { return -1; }
    return 0;
}
void ref_kdx_clear_bins(void* f, const uint32_t* bins, uint32_t n, uint32_t threads)
{
    std::vector<uint32_t> v(bins, bins + n);
    ((TKdx*)f)->clearBins(v, threads);
}
int ref_kdx_save(void* f, const char* path) { return ((TKdx*)f)->save(path) ? 0 : -1; }
void ref_kdx_which_bins(void* f, const uint8_t* codes, uint64_t len, uint16_t threshold, uint8_t* selected)
{
    TKdx* x = (TKdx*)f;
    uint32_t bins = x->getNumberOfBins();
    std::vector<bool> sel(bins);
    for (uint32_t b = 0; b < bins; ++b) sel[b] = selected[b] != 0;
    Dna5String s = codes_to_dna5(codes, len);
    x->whichBins(sel, s, threshold);
    for (uint32_t b = 0; b < bins; ++b) selected[b] = sel[b] ? 1 : 0;
}

} // extern "C"
