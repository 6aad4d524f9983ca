// tools/fasta_batch.h — host-side FASTA/FASTQ(.gz) reading and batching for the standalone builder/updater.
//
// Mirrors what SeqAnBloomFilter::addFastaFile (temehi/dream_yara src/d_bloom_filter.h:127-147) does on the host:
// every record of the file is one sequence (no k-mers across records), records shorter than k are skipped (:142),
// and the characters go through Iupac -> Dna5 (ACGTU any case -> 0..3, everything else -> N; on the device,
// DYIBF_ENC_ASCII).  Reader threads parse files concurrently (the reference's std::async bin ranges,
// src/d_build_filter.cpp:227-260, become host readers); one thread feeds the GPU with batches of whole files.
#ifndef DYIBF_TOOLS_FASTA_BATCH_H_
#define DYIBF_TOOLS_FASTA_BATCH_H_

#include <zlib.h>

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

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

// one file's records, appended to a batch: concat bytes (ASCII), limits, bin ids
struct Batch
{
    std::vector<uint8_t>  concat;
    std::vector<uint64_t> limits{0};
    std::vector<uint32_t> binIds;
    void clear() { concat.clear(); limits.assign(1, 0); binIds.clear(); }
    bool empty() const { return binIds.empty(); }
};

// Reads FASTA or FASTQ (decided by the first '>' or '@'), plain or gzip, with the rules of SeqAn's readRecord into an
// IupacString (seq_io/fasta_fastq.h:268-345), which is what addFastaFile uses: a FASTA record ends at ANY '>' (not only at a
// line start); white space inside the sequence is dropped; a character outside the IUPAC alphabet is a ParseError (the
// reference tool aborts: here a std::runtime_error, the tools print it and exit 1); FASTQ qualities are counted without
// white space until there are as many as bases, and the next record must follow.  Records shorter than k are skipped.
// (The device tokeniser dyibf_parse_records(DYIBF_ALPHA_IUPAC) applies the same rules; tests/test_tools.py feeds both the
// same malformed files.)
inline bool iupacChar(char c)
{
    switch (c >= 'a' && c <= 'z' ? c - 32 : c)
    {
        case 'A': case 'C': case 'G': case 'T': case 'U': case 'M': case 'R': case 'W': case 'S': case 'Y': case 'K':
        case 'V': case 'H': case 'D': case 'B': case 'N': return true;
        default: return false;
    }
}

inline void appendFastaRecords(Batch & out, const std::string & path, uint32_t binNo, uint32_t k)
{
    gzFile gz = gzopen(path.c_str(), "rb");
    if (!gz)
        throw std::runtime_error("Unable to open contigs File: " + path);
    gzbuffer(gz, 1 << 20);
    std::vector<char> buf(1 << 20);
    enum { START, HEADER, SEQ, FQ_SEQ, FQ_PLUS, FQ_QUAL } st = START;
    bool fastq = false, afterFastqRecord = false;
    uint64_t recStart = out.concat.size();
    uint64_t qualLeft = 0;
    auto finish = [&]() {
        if (out.concat.size() - recStart >= k) { out.limits.push_back(out.concat.size()); out.binIds.push_back(binNo); }
        else out.concat.resize(recStart);
        recStart = out.concat.size();
    };
    auto blank = [](char c) { return c == '\n' || c == '\r' || c == ' ' || c == '\t' || c == '\v' || c == '\f'; };
    auto base = [&](char c) {
        if (!iupacChar(c)) { gzclose(gz); throw std::runtime_error("ParseError in " + path + ": character '" + std::string(1, c) + "' is not in the IUPAC alphabet"); }
        out.concat.push_back((uint8_t)c);
    };
    int n;
    while ((n = gzread(gz, buf.data(), (unsigned)buf.size())) > 0)
    {
        for (int i = 0; i < n; ++i)
        {
            const char c = buf[i];
            switch (st)
            {
                case START:
                    if (c == '>' && !afterFastqRecord) { fastq = false; st = HEADER; }
                    else if (c == '@') { fastq = true; st = HEADER; }
                    else if (afterFastqRecord && !blank(c))
                    { gzclose(gz); throw std::runtime_error("ParseError in " + path + ": Fastq quality string is expected to be of the same length as the sequence"); }
                    break;
                case HEADER:
                    if (c == '\n') { st = fastq ? FQ_SEQ : SEQ; recStart = out.concat.size(); }
                    break;
                case SEQ:
                    if (c == '>') { finish(); st = HEADER; }
                    else if (!blank(c)) base(c);
                    break;
                case FQ_SEQ:
                    if (c == '+') { st = FQ_PLUS; qualLeft = out.concat.size() - recStart; }
                    else if (!blank(c)) base(c);
                    break;
                case FQ_PLUS:
                    if (c == '\n') { st = FQ_QUAL; if (qualLeft == 0) { finish(); st = START; afterFastqRecord = true; } }
                    break;
                case FQ_QUAL:
                    if (!blank(c) && --qualLeft == 0) { finish(); st = START; afterFastqRecord = true; }
                    break;
            }
        }
    }
    const bool bad = n < 0;
    gzclose(gz);
    if (bad)
        throw std::runtime_error("Error while reading contigs File: " + path);
    if (st == SEQ) finish();
    else if (st == FQ_SEQ || st == FQ_QUAL || st == FQ_PLUS || (st == HEADER && fastq))
        throw std::runtime_error("UnexpectedEnd in " + path + ": the file ends inside a record");
}

} // namespace dyibf_tools

#endif  // DYIBF_TOOLS_FASTA_BATCH_H_
