// tools/fasta_batch.h — file reading and batching for the standalone builder/updater.
//
// SeqAnBloomFilter::addFastaFile (temehi/dream_yara src/d_bloom_filter.h:127-147) reads a file record by record on the host and
// hashes every record of at least k characters into the file's bin.  Here the host only reads (and inflates) the files: their
// TEXT goes to the device in batches of whole files (dyibf_add_fasta_texts), where SeqAn's reader rules and the Iupac -> Dna5
// conversion (ACGTU any case -> 0..3, everything else -> N) are applied and the records are added in one build call.  Reader
// threads own contiguous bin ranges (the reference's std::async tasks, src/d_build_filter.cpp:227-260); whoever holds a full
// batch feeds the GPU.
#ifndef DYIBF_TOOLS_FASTA_BATCH_H_
#define DYIBF_TOOLS_FASTA_BATCH_H_

#include <zlib.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "dyibf.h"

namespace dyibf_tools {

// the extensions SeqAn's SeqFileIn accepts for FASTA/FASTQ, plain and gzip (seq_io/sequence_file.h, stream/formatted_file.h)
inline const std::vector<std::string> & sequenceExtensions()
{
    static const std::vector<std::string> ext = {
        ".fa", ".fasta", ".fna", ".ffn", ".fas", ".faa", ".frn", ".fq", ".fastq",
        ".fa.gz", ".fasta.gz", ".fna.gz", ".ffn.gz", ".fas.gz", ".faa.gz", ".frn.gz", ".fq.gz", ".fastq.gz"};
    return ext;
}

inline bool fileReadable(const std::string & path)
{
    FILE * fp = fopen(path.c_str(), "rb");
    if (!fp) return false;
    fclose(fp);
    return true;
}

// commonExtension, src/d_misc_options.h:196-223: the first extension for which <dir>/<bin><ext> exists for all bins
inline std::string commonExtension(const std::string & dir, uint32_t numberOfBins)
{
    for (const std::string & ext : sequenceExtensions())
    {
        uint32_t count = 0;
        for (; count < numberOfBins; ++count)
            if (!fileReadable(dir + std::to_string(count) + ext))
                break;
        if (count == numberOfBins)
            return ext;
    }
    fprintf(stderr, "The given directory:\n\t%s\ndoes not contain the fasta files of all the bins!\n"
                    "All files should have identical valid extension!\n", dir.c_str());
    exit(1);
}

// the text of whole files on its way to the device: dyibf_add_fasta_texts tokenises it there with SeqAn's rules
// (addFastaFile, src/d_bloom_filter.h:127-147: every record of a file is a sequence of the file's bin, a FASTA record ends at
// any '>', characters outside the IUPAC alphabet abort, records shorter than k are skipped) and adds all of it in one build call
struct TextBatch
{
    std::vector<std::vector<char>> texts;
    std::vector<uint32_t> binIds;
    std::vector<int> formats;
    uint64_t bytes = 0;
    void clear() { texts.clear(); binIds.clear(); formats.clear(); bytes = 0; }
    bool empty() const { return texts.empty(); }
};

inline bool endsWith(const std::string & s, const char * suf)
{
    const std::string t = suf;
    return s.size() >= t.size() && s.compare(s.size() - t.size(), t.size(), t) == 0;
}

// SeqFileIn picks the format from the extension (".gz" aside)
inline int formatOfPath(std::string path)
{
    if (endsWith(path, ".gz")) path.resize(path.size() - 3);
    return endsWith(path, ".fq") || endsWith(path, ".fastq") ? 1 /* DYIBF_FMT_FASTQ */ : 0 /* DYIBF_FMT_FASTA */;
}

// the (inflated) bytes of one file become one text of the batch
inline void appendFileText(TextBatch & out, const std::string & path, uint32_t binNo)
{
    gzFile gz = gzopen(path.c_str(), "rb");                          // plain files pass through
    if (!gz)
        throw std::runtime_error("Unable to open contigs File: " + path);
    gzbuffer(gz, 1 << 20);
    // a plain file's size is known (one allocation, no regrowth); a gzip file's text is a few times its size
    std::vector<char> text;
    {
        FILE * fp = fopen(path.c_str(), "rb");
        long long sz = 0;
        if (fp) { if (fseeko(fp, 0, SEEK_END) == 0) sz = (long long)ftello(fp); fclose(fp); }
        const bool gzip = endsWith(path, ".gz");
        text.resize((size_t)std::max<long long>(gzip ? 4 * sz : sz + 1, 1 << 16));
    }
    size_t used = 0;
    int n;
    while ((n = gzread(gz, text.data() + used, (unsigned)std::min<size_t>(text.size() - used, 1u << 30))) > 0)
    {
        used += (size_t)n;
        if (used == text.size()) text.resize(text.size() * 2);
    }
    const bool bad = n < 0;
    gzclose(gz);
    if (bad)
        throw std::runtime_error("Error while reading contigs File: " + path);
    text.resize(used);
    out.bytes += used;
    out.texts.push_back(std::move(text));
    out.binIds.push_back(binNo);
    out.formats.push_back(formatOfPath(path));
}

// one build call for the batch; throws what the library reports (a ParseError names the file)
inline void addTextBatch(dyibf_filter * filter, TextBatch & batch, const std::vector<std::string> * names = nullptr)
{
    if (batch.empty()) return;
    std::vector<const char *> ptrs; std::vector<uint64_t> sizes;
    for (auto const & t : batch.texts) { ptrs.push_back(t.data()); sizes.push_back(t.size()); }
    uint32_t bad = 0xffffffffu;
    if (dyibf_add_fasta_texts(filter, ptrs.data(), sizes.data(), batch.binIds.data(), batch.formats.data(), (uint32_t)ptrs.size(), &bad) != DYIBF_OK)
    {
        std::string msg = dyibf_last_error();
        if (bad != 0xffffffffu) msg = "ParseError in the contigs of bin " + std::to_string(batch.binIds[bad]) +
                                      (names && bad < names->size() ? " (" + (*names)[bad] + ")" : std::string()) + ": " + msg;
        throw std::runtime_error(msg);
    }
    batch.clear();
}

} // namespace dyibf_tools

#endif  // DYIBF_TOOLS_FASTA_BATCH_H_
