// ==========================================================================
//                     d_kdx_filter.h  (B200 drop-in)
// ==========================================================================
// Drop-in replacement for DREAM-Yara's src/d_kdx_filter.h (temehi/dream_yara, lines 41-311): the class
// seqan::SeqAnKDXFilter<TString> — the `-ft kmer_direct` filter: same bit layout, file format and members as the
// interleaved Bloom filter, but the block of a k-mer is its value (no hashing) — with the same public members and
// argument meaning.  On the C ABI (include/dyibf.h) it is a filter with ZERO hash functions, so the class below is
// the B200 SeqAnBloomFilter with the two reference constructors.  Included after d_bloom_filter.h, as the three
// translation units of the reference do (src/d_build_filter.cpp:60-61, src/d_update_filter.cpp:60-61,
// src/d_mapper.cpp:91-92 include it just before; here the order is the other way round, see the #include below).
//
// Where the reference indexes out of bounds the B200 class is defined instead: a k-mer whose value is not a block of
// the table (the reference only guards with _noOfHashPos = (N-256)/_binIntWidth, 64 x the blocks there are, :215) —
// sequences that start with N, tables smaller than 4^k blocks — adds nothing and matches nothing; clearBins covers the
// real blocks (the reference's walks 64 x as far and crashes, :71-97).
#ifndef APP_YARA_DIS_KDX_FILTER_B200_H_
#define APP_YARA_DIS_KDX_FILTER_B200_H_

#include "d_bloom_filter.h"

namespace seqan
{
    template<typename TString = Dna5String>
    class SeqAnKDXFilter : public SeqAnBloomFilter<TString>
    {
        typedef SeqAnBloomFilter<TString> TBase;
    public:
        // d_kdx_filter.h:49-56
        SeqAnKDXFilter(uint32_t n_bins, uint8_t kmer_size, uint64_t vec_size, int device = 0) :
            TBase(n_bins, (uint8_t)0, kmer_size, vec_size, device) {}

        // d_kdx_filter.h:58-67 — load failure: message on std::cerr, exit(1)
        explicit SeqAnKDXFilter(const char *fileName, int device = 0) : TBase(fileName, device) {}

        SeqAnKDXFilter(SeqAnKDXFilter &&) = default;
        // whichBins(text, threshold) -> vector<bool>, d_kdx_filter.h:199-204, takes a uint32_t threshold in the reference;
        // the inherited uint16_t overload accepts the same calls
    };
}

#endif  // APP_YARA_DIS_KDX_FILTER_B200_H_
