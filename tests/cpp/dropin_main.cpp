// tests/cpp/dropin_main.cpp — exercises include/d_bloom_filter.h and include/d_mapper_filter.h (the C++ drop-in
// for DREAM-Yara's SeqAnBloomFilter and clasifyLoadedReads) the way the reference's mains use them, with real
// SeqAn types.  Compiled in the build container against the SeqAn vendored under /root/reference/include (needed
// at compile time only); the binary travels to the GPU box, where tests/test_cpp_dropin.py runs it and compares
// its output with the oracle.
//
//   dropin_test build    <bins> <h> <k> <nbits> <dir with 0.fa..> <out.filter>
//   dropin_test query    <filter> <reads.fa> <threshold> <out.txt>      per-read whichBins(fwd) | whichBins(rc)
//   dropin_test classify <filter> <reads.fa> <e percent> <paired 0/1> <out.txt>
//   dropin_test gather   <filter> <reads.fa> <e percent> <paired 0/1> <out.txt>   classify, then sortedBins and
//                        loadFilteredReads for every bin in that order (reads get synthetic qualities)
//   dropin_test kdx      <bins> <k> <dir with 0.fa..> <out.filter> <reads.fa> <threshold> <out.txt>
//                        SeqAnKDXFilter (-ft kmer_direct): build as build_filter does, save, reload, per-read whichBins
#include <climits>
#include <limits.h>
#undef INT_WIDTH

#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include <seqan/sequence.h>
#include <seqan/seq_io.h>
#include <seqan/index.h>
#include <seqan/modifier.h>

#include "d_bloom_filter.h"
#include "d_kdx_filter.h"
#include "d_mapper_filter.h"

using namespace seqan;

struct MiniTimer { double t = 0; };
inline void start(MiniTimer & t) { t.t = sysTime(); }
inline void stop(MiniTimer & t) { t.t = sysTime() - t.t; }
inline double getValue(MiniTimer & t) { return t.t; }

// the fields of DisOptions the filtering stage touches (src/d_mapper.h:49-94)
struct MiniDisOptions
{
    float       errorRate = 0.05f;
    uint32_t    kmerSize = 20;
    uint32_t    numberOfBins = 64;
    int         verbose = 0;
    double      filterReads = 0;
    uint32_t    currentBinNo = 0;
    std::vector<std::vector<uint32_t> > origReadIdMap;
    uint16_t getThreshold(uint16_t readLen) { return dyibf_threshold_rate(errorRate, kmerSize, readLen); }
};

int main(int argc, char ** argv)
{
    if (argc < 2) return 2;
    std::string mode = argv[1];
    try
    {
        if (mode == "build" && argc == 8)
        {
            uint32_t bins = std::stoul(argv[2]);
            SeqAnBloomFilter<> filter(bins, (uint8_t)std::stoul(argv[3]), (uint8_t)std::stoul(argv[4]), std::stoull(argv[5]));
            for (uint32_t b = 0; b < bins; ++b)
            {
                CharString path = std::string(argv[6]) + "/" + std::to_string(b) + ".fa";
                filter.addFastaFile(path, b);
            }
            std::cerr << "size_mb " << filter.size_mb() << std::endl;
            return filter.save(argv[7]) ? 0 : 1;
        }
        if (mode == "kdx" && argc == 9)
        {
            uint32_t bins = std::stoul(argv[2]);
            uint8_t k = (uint8_t)std::stoul(argv[3]);
            {
                uint64_t vec_size = (1u << (2 * k));                 // src/d_build_filter.cpp:307-310
                vec_size *= bins;
                vec_size += 256;
                SeqAnKDXFilter<> filter(bins, k, vec_size);
                for (uint32_t b = 0; b < bins; ++b)
                {
                    CharString path = std::string(argv[4]) + "/" + std::to_string(b) + ".fa";
                    filter.addFastaFile(path, b);
                }
                if (!filter.save(argv[5])) return 1;
            }
            SeqAnKDXFilter<> filter(argv[5]);                        // src/d_mapper.h:926
            if (filter.getNumberOfBins() != bins || filter.getKmerSize() != k) return 3;
            uint16_t thr = (uint16_t)std::stoul(argv[7]);
            SeqFileIn in(argv[6]);
            std::ofstream out(argv[8]);
            CharString id; String<Dna5Q> seq;
            while (!atEnd(in))
            {
                readRecord(id, seq, in);
                String<Dna5Q> rc = seq; reverseComplement(rc);
                std::vector<bool> sel(filter.getNumberOfBins(), false);
                filter.whichBins(sel, seq, thr);
                filter.whichBins(sel, rc, thr);
                for (uint32_t b = 0; b < sel.size(); ++b) if (sel[b]) out << b << ' ';
                out << '\n';
            }
            return 0;
        }
        if (mode == "query" && argc == 6)
        {
            // through the move constructor: the moved-to object must still know its devices and sharding (DYIBF_DEVICES /
            // DYIBF_GROUP_MODE=bin: whichBins has to ask every bin shard, not member 0 only)
            SeqAnBloomFilter<> loaded(argv[2]);
            SeqAnBloomFilter<> filter(std::move(loaded));
            std::cerr << "devices " << filter.deviceCount() << std::endl;
            uint16_t thr = (uint16_t)std::stoul(argv[4]);
            SeqFileIn in(argv[3]);
            std::ofstream out(argv[5]);
            CharString id; String<Dna5Q> seq;
            while (!atEnd(in))
            {
                readRecord(id, seq, in);
                String<Dna5Q> rc = seq; reverseComplement(rc);
                std::vector<bool> sel(filter.getNumberOfBins(), false);
                filter.whichBins(sel, seq, thr);           // implicit Dna5Q -> Dna5String temporary, as d_mapper.h:351
                filter.whichBins(sel, rc, thr);
                for (uint32_t b = 0; b < sel.size(); ++b) if (sel[b]) out << b << ' ';
                out << '\n';
            }
            return 0;
        }
        if (mode == "classify" && argc == 7)
        {
            SeqAnBloomFilter<> filter(argv[2]);
            bool paired = std::stoi(argv[5]) != 0;
            StringSet<String<Dna5Q>, Owner<ConcatDirect<> > > seqs;
            SeqFileIn in(argv[3]);
            CharString id; String<Dna5Q> seq;
            while (!atEnd(in)) { readRecord(id, seq, in); appendValue(seqs, seq); }
            uint32_t n = length(seqs);
            for (uint32_t i = 0; i < n; ++i) { String<Dna5Q> rc = seqs[i]; reverseComplement(rc); appendValue(seqs, rc); }
            MiniDisOptions opt;
            opt.errorRate = std::stoi(argv[4]) / 100.0;        // d_mapper.cpp:340-342
            opt.kmerSize = filter.getKmerSize(); opt.numberOfBins = filter.getNumberOfBins(); opt.verbose = 2;
            MiniTimer timer;
            clasifyLoadedReadsB200(seqs, filter, opt, paired, timer);
            std::ofstream out(argv[6]);
            for (uint32_t b = 0; b < opt.numberOfBins; ++b)
            {
                for (uint32_t r : opt.origReadIdMap[b]) out << r << ' ';
                out << '\n';
            }
            return 0;
        }
        if (mode == "gather" && argc == 7)
        {
            SeqAnBloomFilter<> filter(argv[2]);
            bool paired = std::stoi(argv[5]) != 0;
            StringSet<String<Dna5Q>, Owner<ConcatDirect<> > > seqs;
            SeqFileIn in(argv[3]);
            CharString id; String<Dna5Q> seq;
            uint64_t pos = 0;
            while (!atEnd(in))
            {
                readRecord(id, seq, in);
                for (uint64_t i = 0; i < length(seq); ++i, ++pos) assignQualityValue(seq[i], (int)((pos * 7 + 11) % 63));
                appendValue(seqs, seq);
            }
            uint32_t n = length(seqs);
            for (uint32_t i = 0; i < n; ++i) { String<Dna5Q> rc = seqs[i]; reverseComplement(rc); appendValue(seqs, rc); }
            MiniDisOptions opt;
            opt.errorRate = std::stoi(argv[4]) / 100.0;
            opt.kmerSize = filter.getKmerSize(); opt.numberOfBins = filter.getNumberOfBins();
            MiniTimer timer;
            clasifyLoadedReadsB200(seqs, filter, opt, paired, timer);
            std::ofstream out(argv[6]);
            for (uint32_t b : sortedBinsB200(opt))
            {
                opt.currentBinNo = b;
                StringSet<String<Dna5Q>, Owner<ConcatDirect<> > > child;
                loadFilteredReadsB200(child, filter, opt);
                out << "bin " << b << ' ' << length(child) << '\n';
                for (uint32_t r : opt.origReadIdMap[b]) out << r << ' ';
                out << '\n';
                for (uint32_t i = 0; i < length(child); ++i)
                {
                    for (uint64_t x = 0; x < length(child[i]); ++x) out << (unsigned)child[i][x].value << ' ';
                    out << '\n';
                }
            }
            return 0;
        }
    }
    catch (std::exception const & e) { std::cerr << "error: " << e.what() << std::endl; return 3; }
    std::cerr << "usage: see source" << std::endl;
    return 2;
}
