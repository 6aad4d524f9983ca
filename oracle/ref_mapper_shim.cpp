// oracle/ref_mapper_shim.cpp — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Compiles the reference's OWN filtering stage — DisOptions::getThreshold
// (/root/reference/src/d_mapper.h:83-92) and clasifyLoadedReads (d_mapper.h:308-387) — from the
// unmodified headers under /root/reference into oracle/_ref/libref_mapper.so, so the per-batch
// threshold and the SE/PE read-set layout are pinned by the reference itself rather than by a
// restatement.  Only the two templates used here are instantiated (the full dream_yara_mapper TU
// takes ~16 min; this takes about a minute).  Used in THIS container only, to generate the golden
// vectors under tests/golden/ (tests/golden/make_golden.py).
#include <climits>
#include <limits.h>
#undef INT_WIDTH        // glibc C23 macro vs src/d_misc_options.h:41

struct Options;          // forward, as d_mapper.cpp:41 has it

#include <vector>
#include <string>
#include <random>
#include <math.h>
#include <fstream>
#include <iostream>
#include <mutex>
#include <condition_variable>
#include <future>
#include <thread>
#include <dirent.h>
#include <sys/stat.h>

#include <seqan/basic.h>
#include <seqan/sequence.h>
#include <seqan/index.h>
#include <seqan/parallel.h>

#include "basic_alphabet.h"
#include "file_pair.h"
#include "file_prefetched.h"
#include "store_seqs.h"
#include "misc_timer.h"
#include "misc_tags.h"
#include "misc_types.h"
#include "index_fm.h"
#include "bits_reads.h"
#include "bits_hits.h"
#include "bits_context.h"
#include "bits_matches.h"
#include "bits_seeds.h"
#include "bits_bucket.h"
#include "find_verifier.h"
#include "find_extender.h"
#include "misc_options.h"
#include "d_misc_options.h"
#include "d_kdx_filter.h"
#include "d_bloom_filter.h"
#include "mapper_collector.h"
#include "mapper_classifier.h"
#include "mapper_ranker.h"
#include "mapper_filter.h"
#include "mapper_extender.h"
#include "mapper_verifier.h"
#include "mapper_aligner.h"
#include "mapper_writer.h"
#include "mapper.h"
#include "d_mapper.h"

using namespace seqan;

template <typename TSequencing>
static void run_classify(const char* filterFile, const uint8_t* codes, const uint64_t* offsets,
                         uint32_t n_seqs, int e_percent, uint32_t threads,
                         uint16_t* thr_out, uint32_t* bin_offsets, uint32_t* read_ids, uint64_t cap)
{
    typedef ReadMapperConfig<Parallel, TSequencing, EditDistance, uint8_t, uint32_t, uint32_t> TMainConfig;
    DisOptions disOptions;
    disOptions.errorRate = e_percent / 100.0;           // d_mapper.cpp:340-342
    disOptions.threadsCount = threads;
    disOptions.verbose = 0;
    SeqAnBloomFilter<> filter(filterFile);              // d_mapper.h:913
    disOptions.kmerSize = filter.getKmerSize();         // d_mapper.h:915
    disOptions.numberOfBins = filter.getNumberOfBins(); // d_mapper.cpp:545
    Mapper<void, TMainConfig> mainMapper(disOptions);
    // what readRecords() leaves in reads.seqs: SE all reads; PE all first mates then all second mates
    for (uint32_t s = 0; s < n_seqs; ++s)
    {
        String<Dna5Q> r;
        resize(r, offsets[s + 1] - offsets[s]);
        for (uint64_t i = 0; i < length(r); ++i) r[i] = Dna5(codes[offsets[s] + i]);
        appendValue(mainMapper.reads.seqs, r);
    }
    appendReverseComplement(mainMapper.reads);          // mapper.h:457
    clasifyLoadedReads(mainMapper, filter, disOptions); // d_mapper.h:308
    uint16_t avgReadLen = lengthSum(mainMapper.reads.seqs) / (getReadsCount(mainMapper.reads.seqs) * 2);
    *thr_out = disOptions.getThreshold(avgReadLen);
    uint64_t pos = 0;
    for (uint32_t b = 0; b < disOptions.numberOfBins; ++b)
    {
        bin_offsets[b] = pos;
        for (uint32_t id : disOptions.origReadIdMap[b])
            if (pos < cap) read_ids[pos++] = id;
    }
    bin_offsets[disOptions.numberOfBins] = pos;
}

// loadFilteredReads (d_mapper.h:393-443) + sortedBins (:843-854) after clasifyLoadedReads, for the requested bins.
// The reads carry qualities so that the Dna5Q bytes (and their reverse complements) are exercised.
template <typename TSequencing>
static void run_gather(const char* filterFile, const uint8_t* codes, const uint8_t* quals, const uint64_t* offsets,
                       uint32_t n_seqs, int e_percent, const uint32_t* bins, uint32_t n_req,
                       uint8_t* fwd_bytes, uint8_t* out_bytes, uint64_t cap_bytes, uint64_t* out_limits, uint32_t* out_ids,
                       uint64_t cap_seqs, uint64_t* seq_start, uint32_t* sorted_bins)
{
    typedef ReadMapperConfig<Parallel, TSequencing, EditDistance, uint8_t, uint32_t, uint32_t> TMainConfig;
    DisOptions disOptions;
    disOptions.errorRate = e_percent / 100.0;
    disOptions.threadsCount = 1;
    disOptions.verbose = 0;
    disOptions.filterType = BLOOM;
    SeqAnBloomFilter<> filter(filterFile);
    disOptions.kmerSize = filter.getKmerSize();
    disOptions.numberOfBins = filter.getNumberOfBins();
    Mapper<void, TMainConfig> mainMapper(disOptions);
    for (uint32_t s = 0; s < n_seqs; ++s)
    {
        String<Dna5Q> r;
        resize(r, offsets[s + 1] - offsets[s]);
        for (uint64_t i = 0; i < length(r); ++i)
        {
            r[i] = Dna5(codes[offsets[s] + i]);
            assignQualityValue(r[i], (int)quals[offsets[s] + i]);
        }
        appendValue(mainMapper.reads.seqs, r);
    }
    for (uint64_t i = 0; i < offsets[n_seqs]; ++i)         // the forward bytes exactly as the ConcatDirect set holds them
        fwd_bytes[i] = concat(mainMapper.reads.seqs)[i].value;
    appendReverseComplement(mainMapper.reads);
    clasifyLoadedReads(mainMapper, filter, disOptions);
    std::vector<uint32_t> sb = sortedBins(disOptions);
    for (uint32_t b = 0; b < disOptions.numberOfBins; ++b) sorted_bins[b] = sb[b];
    uint64_t n_out = 0, n_bytes = 0;
    out_limits[0] = 0;
    for (uint32_t q = 0; q < n_req; ++q)
    {
        seq_start[q] = n_out;
        disOptions.currentBinNo = bins[q];
        Mapper<void, TMainConfig> me(disOptions);
        loadFilteredReads(me, mainMapper, disOptions);
        std::vector<uint32_t> const & ids = disOptions.origReadIdMap[bins[q]];
        for (uint32_t i = 0; i < length(me.reads.seqs); ++i)
        {
            if (n_out >= cap_seqs) return;
            for (uint64_t x = 0; x < length(me.reads.seqs[i]); ++x)
                if (n_bytes < cap_bytes) out_bytes[n_bytes++] = me.reads.seqs[i][x].value;
            out_ids[n_out] = ids[i];
            out_limits[++n_out] = n_bytes;
        }
    }
    seq_start[n_req] = n_out;
}

extern "C" {

void refm_gather(const char* filterFile, const uint8_t* codes, const uint8_t* quals, const uint64_t* offsets, uint32_t n_seqs,
                 int paired, int e_percent, const uint32_t* bins, uint32_t n_req, uint8_t* fwd_bytes,
                 uint8_t* out_bytes, uint64_t cap_bytes, uint64_t* out_limits, uint32_t* out_ids, uint64_t cap_seqs,
                 uint64_t* seq_start, uint32_t* sorted_bins)
{
    if (paired) run_gather<PairedEnd>(filterFile, codes, quals, offsets, n_seqs, e_percent, bins, n_req, fwd_bytes, out_bytes, cap_bytes, out_limits, out_ids, cap_seqs, seq_start, sorted_bins);
    else        run_gather<SingleEnd>(filterFile, codes, quals, offsets, n_seqs, e_percent, bins, n_req, fwd_bytes, out_bytes, cap_bytes, out_limits, out_ids, cap_seqs, seq_start, sorted_bins);
}

// DisOptions::getThreshold with errorRate set the way the CLI parser sets it
uint16_t refm_threshold(int e_percent, uint32_t kmerSize, uint16_t readLen)
{
    DisOptions o;
    o.errorRate = e_percent / 100.0;
    o.kmerSize = kmerSize;
    return o.getThreshold(readLen);
}

// clasifyLoadedReads on n_seqs forward sequences (SE: the reads; PE: all m1 then all m2),
// Dna5 codes. Output CSR: bin_offsets[bins+1], read_ids (in origReadIdMap order).
void refm_classify(const char* filterFile, const uint8_t* codes, const uint64_t* offsets, uint32_t n_seqs,
                   int paired, int e_percent, uint32_t threads,
                   uint16_t* thr_out, uint32_t* bin_offsets, uint32_t* read_ids, uint64_t cap)
{
    if (paired) run_classify<PairedEnd>(filterFile, codes, offsets, n_seqs, e_percent, threads, thr_out, bin_offsets, read_ids, cap);
    else        run_classify<SingleEnd>(filterFile, codes, offsets, n_seqs, e_percent, threads, thr_out, bin_offsets, read_ids, cap);
}

}
