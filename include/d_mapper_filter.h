// ==========================================================================
//                     d_mapper_filter.h  (B200 filtering stage)
// ==========================================================================
// Replacement body for clasifyLoadedReads (temehi/dream_yara src/d_mapper.h:308-387): the read batch the main
// mapper holds is filtered in ONE call — H2D of the forward reads, reverse complements and the IBF lookup on
// the GPU, compaction into per-bin ascending read-id lists on the GPU, D2H of the lists — instead of
// threadsCount std::async tasks calling whichBins four times per read and pushing ids under a global mutex.
//
// Same observable behaviour as the reference with --threads 1:
//   * one threshold per batch from the AVERAGE read length (:313-315) via disOptions.getThreshold (:83-92);
//   * threshold == 0: the warning, every id pushed to every bin, and then — as the reference has no early
//     return (:320-333) — the normal pass still runs, so ids appear twice;
//   * origReadIdMap[bin] ascending; single-end ids are read ids, paired-end ids are pair ids (:335-336);
//   * disOptions.filterReads accumulates the stage's wall time; verbose > 1 prints "bin <b>\t<count>".
//
// Integration: in src/d_mapper.h replace the body of clasifyLoadedReads by
//     clasifyLoadedReadsB200(mainMapper.reads.seqs, filter, disOptions,
//                            IsSameType<typename TMainConfig::TSequencing, PairedEnd>::VALUE, mainMapper.timer);
#ifndef APP_YARA_DIS_MAPPER_FILTER_B200_H_
#define APP_YARA_DIS_MAPPER_FILTER_B200_H_

#include <cstdint>
#include <cstring>
#include <iostream>
#include <vector>

namespace seqan
{

// TReadSeqs: StringSet<String<Dna5Q>, Owner<ConcatDirect<> > > as loadReads + appendReverseComplement leave it
// (src/store_seqs.h:54-61,271-290): [fwd | rc] or [m1 | m2 | m1 rc | m2 rc].  Only the forward half is sent to
// the device; the reverse complements are re-derived there (identical by construction: N stays N).
template <typename TReadSeqs, typename TFilter, typename TDisOptions, typename TTimer>
inline void clasifyLoadedReadsB200(TReadSeqs & readSeqs, TFilter const & filter, TDisOptions & disOptions,
                                   bool pairedEnd, TTimer & timer)
{
    start(timer);

    uint32_t numReads = length(readSeqs) / 2;                                   // getReadsCount, bits_reads.h:57-62
    uint16_t avgReadLen = numReads ? lengthSum(readSeqs) / (numReads * 2) : 0;
    uint16_t threshold = disOptions.getThreshold(avgReadLen);

    disOptions.origReadIdMap.clear();
    disOptions.origReadIdMap.resize(disOptions.numberOfBins);

    std::vector<uint32_t> everyRead;
    if (threshold == 0)
    {
        std::cerr <<"[WARNING!] 0 k-mer is required to filter a read!\n";
        std::cerr <<"All reads will pass filteration and be mapped everywhere.\n ";
        std::cerr <<"This will be extremly slow.\n ";
        std::cerr <<"Choose an approprate error rate based on kmer size and read length\n";
        everyRead.resize(numReads);
        for (uint32_t readID = 0; readID < numReads; ++readID)
            everyRead[readID] = readID;
    }

    // forward half of the set: SE the n reads, PE the 2p mates
    uint32_t numFwd = length(readSeqs) / 2;
    static_assert(sizeof(typename Value<typename Value<TReadSeqs>::Type>::Type) == 1, "one byte per base expected");
    const uint8_t * bytes = numFwd ? reinterpret_cast<const uint8_t *>(&*begin(concat(readSeqs), Standard())) : nullptr;
    std::vector<uint64_t> limits(numFwd + 1);
    for (uint32_t i = 0; i <= numFwd; ++i)
        limits[i] = stringSetLimits(readSeqs)[i];

    std::vector<std::vector<uint32_t> > lists;
    if (numFwd)
        filter.classify(lists, bytes, limits.data(), numFwd, DYIBF_ENC_DNA5Q, pairedEnd ? 4 : 2, true, threshold);
    else
        lists.resize(disOptions.numberOfBins);

    for (uint32_t binNo = 0; binNo < disOptions.numberOfBins; ++binNo)
    {
        std::vector<uint32_t> & dst = disOptions.origReadIdMap[binNo];
        dst = everyRead;                                                        // empty unless threshold == 0
        dst.insert(dst.end(), lists[binNo].begin(), lists[binNo].end());
    }

    stop(timer);
    disOptions.filterReads += getValue(timer);

    if (disOptions.verbose > 1)
    {
        for (uint32_t binNo = 0; binNo < disOptions.numberOfBins; ++binNo)
            std::cerr << "bin " << binNo << "\t" << disOptions.origReadIdMap[binNo].size() << std::endl;
    }
}

// sortedBins (src/d_mapper.h:843-854): same std::sort call, same comparator.
template <typename TDisOptions>
inline std::vector<uint32_t> sortedBinsB200(TDisOptions const & disOptions)
{
    std::vector<uint64_t> offs(disOptions.numberOfBins + 1, 0);
    for (uint32_t b = 0; b < disOptions.numberOfBins; ++b)
        offs[b + 1] = offs[b] + disOptions.origReadIdMap[b].size();
    std::vector<uint32_t> out(disOptions.numberOfBins);
    dyibf_sorted_bins(offs.data(), disOptions.numberOfBins, out.data());
    return out;
}

// Replacement body for loadFilteredReads (src/d_mapper.h:393-443), filterType != NONE: the child mapper's read set
// [fwd | (mates) | reverse complements] is assembled on the GPU from the batch clasifyLoadedReadsB200 left resident
// and arrives as one bytes + limits pair; origReadIdMap[currentBinNo] is extended exactly as the reference does.
// Not valid for threshold == 0 batches (the reference's double-pushed lists): fall back to the reference body there.
//     loadFilteredReadsB200(me.reads.seqs, filter, disOptions);
template <typename TReadSeqs, typename TFilter, typename TDisOptions>
inline void loadFilteredReadsB200(TReadSeqs & childSeqs, TFilter const & filter, TDisOptions & disOptions)
{
    std::vector<uint8_t> bytes;
    std::vector<uint64_t> limits;
    std::vector<uint32_t> ids;
    filter.gatherReads(disOptions.currentBinNo, bytes, limits, ids);
    typedef typename Value<typename Value<TReadSeqs>::Type>::Type TAlphabet;
    static_assert(sizeof(TAlphabet) == 1, "one byte per base expected");
    clear(childSeqs);
    resize(concat(childSeqs), bytes.size(), Exact());
    if (!bytes.empty())
        std::memcpy(reinterpret_cast<uint8_t *>(&*begin(concat(childSeqs), Standard())), bytes.data(), bytes.size());
    resize(stringSetLimits(childSeqs), limits.size(), Exact());
    for (size_t i = 0; i < limits.size(); ++i)
        stringSetLimits(childSeqs)[i] = limits[i];
    disOptions.origReadIdMap[disOptions.currentBinNo] = ids;
}

}

#endif  // APP_YARA_DIS_MAPPER_FILTER_B200_H_
