// tools/dyibf_filter_reads.cpp — the filtering stage of dream_yara_mapper as a standalone program on top of the C ABI
// (include/dyibf.h): what runDisMapper does between loading the filter and handing read sets to the per-bin mappers
// (temehi/dream_yara src/d_mapper.h:900-959, :308-387, :843-854), and nothing of the mapping itself.
//
//   dyibf_filter_reads -fi <FILTER> [-e 5] [-rb 100000] [-v|-vv] [-o lists.txt] [--device 0 | --devices 0,1,..] [--inflight 3]
//                      <READS> [<MATES>]
//
// The batches are filtered asynchronously, as the reference's prefetching reader (src/file_prefetched.h:137-205) would feed
// them: up to --inflight batches are in flight (dyibf_classify_begin / dyibf_classify_end), so the upload of batch i+1 and
// the list download of batch i-1 overlap the query of batch i; results are consumed in batch order.  With --devices the
// filter is replicated over several GPUs of this one process (dyibf_group_load_file) and successive batches go to
// successive GPUs.
//
// Per batch of -rb reads (pairs), as the mapper loads them (src/mapper.h:109, d_mapper.h:870-896): the records are
// tokenised ON THE DEVICE (dyibf_parse_records = SeqAn's readRecords; the host only reads / inflates the file), the
// threshold comes from the batch's average read length (d_mapper.h:313-315), dyibf_classify fills origReadIdMap.
// -vv prints "bin <b>\t<size>" per batch (:380-386); -v the timers and "Avg reads per bin" (:947-958); -o writes, per
// batch, one line per bin in sortedBins order (:843-854): "<batch> <bin> <n> id id ...".
#include <zlib.h>

#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include "dyibf.h"

namespace {

double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

[[noreturn]] void usage(const char * msg)
{
    if (msg) std::cerr << "dyibf_filter_reads: " << msg << "\n";
    std::cerr << "usage: dyibf_filter_reads -fi FILTER [OPTIONS] <READS FILE> [<MATES FILE>]\n"
                 "  -fi, --bloom-filter FILE   *.filter (bloom or kmer_direct; the kind is in the file)\n"
                 "  -e,  --error-rate INT      percent, 0..10 (default 5)\n"
                 "  -rb, --reads-batch INT     reads (pairs) per batch, 1000..5000000 (default 100000)\n"
                 "  -o,  --output-file FILE    per-bin read id lists\n"
                 "  -v / -vv                   statistics / per-bin list sizes\n"
                 "       --device INT          CUDA device (default 0)\n"
                 "       --devices LIST        several GPUs, e.g. 0,1,2,3 or all: one replica each, batches round-robin\n"
                 "       --inflight INT        batches in flight, 1..16 (default 3; 1 = synchronous calls)\n";
    exit(msg ? 1 : 0);
}

bool endsWith(const std::string & s, const char * suf)
{
    const std::string t = suf;
    return s.size() >= t.size() && s.compare(s.size() - t.size(), t.size(), t) == 0;
}

// SeqFileIn picks the format from the extension; the content decides when it says nothing
int formatOf(std::string path, const std::vector<char> & data)
{
    if (endsWith(path, ".gz")) path.resize(path.size() - 3);
    if (endsWith(path, ".fq") || endsWith(path, ".fastq")) return DYIBF_FMT_FASTQ;
    if (endsWith(path, ".fa") || endsWith(path, ".fasta") || endsWith(path, ".fna")) return DYIBF_FMT_FASTA;
    return !data.empty() && data[0] == '@' ? DYIBF_FMT_FASTQ : DYIBF_FMT_FASTA;
}

void readAll(const std::string & path, std::vector<char> & out)
{
    gzFile gz = gzopen(path.c_str(), "rb");                       // plain files pass through
    if (!gz) { std::cerr << "Unable to open reads file: " << path << std::endl; exit(1); }
    gzbuffer(gz, 1 << 20);
    out.clear();
    std::vector<char> buf(8 << 20);
    int n;
    while ((n = gzread(gz, buf.data(), (unsigned)buf.size())) > 0) out.insert(out.end(), buf.begin(), buf.begin() + n);
    gzclose(gz);
    if (n < 0) { std::cerr << "Error while reading: " << path << std::endl; exit(1); }
}

struct Records { std::vector<uint8_t> seqs; std::vector<uint64_t> limits; };

// the batch handed to dyibf_classify lives in pinned host memory (its upload then runs at PCIe speed), grown on demand
struct PinnedBytes
{
    uint8_t * p = nullptr; size_t cap = 0, n = 0;
    ~PinnedBytes() { dyibf_host_free(p); }
    void clear() { n = 0; }
    void append(const uint8_t * src, size_t len)
    {
        if (n + len > cap)
        {
            size_t want = std::max<size_t>(2 * cap, n + len + 4096);
            void * q = nullptr;
            if (dyibf_host_alloc(&q, want) != DYIBF_OK) { std::cerr << dyibf_last_error() << std::endl; exit(1); }
            if (n) memcpy(q, p, n);
            dyibf_host_free(p);
            p = (uint8_t *)q; cap = want;
        }
        memcpy(p + n, src, len);
        n += len;
    }
};

void parse(int device, const std::string & path, Records & r)
{
    std::vector<char> text;
    readAll(path, text);
    uint64_t n = 0, nb = 0;
    const int fmt = formatOf(path, text);
    int rc = dyibf_parse_records(device, text.data(), text.size(), fmt, DYIBF_ALPHA_DNA5, nullptr, 0, nullptr, 0, 0, &n, &nb, DYIBF_MEM_HOST, nullptr);
    if (rc == DYIBF_OK)
    {
        r.seqs.resize(nb + 1); r.limits.resize(n + 1);
        rc = dyibf_parse_records(device, text.data(), text.size(), fmt, DYIBF_ALPHA_DNA5, r.seqs.data(), nb, r.limits.data(), n, 0, &n, &nb,
                                 DYIBF_MEM_HOST, nullptr);
    }
    if (rc != DYIBF_OK) { std::cerr << path << ": " << dyibf_last_error() << std::endl; exit(1); }
}

} // namespace

int main(int argc, char ** argv)
{
    std::string filterFile, outFile;
    std::vector<std::string> inputs;
    int e = 5, verbose = 0, device = 0, inflight = 3;
    std::vector<int> devices;
    uint64_t readsBatch = 100000;
    for (int i = 1; i < argc; ++i)
    {
        const std::string a = argv[i];
        auto value = [&](const char * name) -> std::string { if (i + 1 >= argc) usage((std::string("missing value for ") + name).c_str()); return argv[++i]; };
        if (a == "-h" || a == "--help") usage(nullptr);
        else if (a == "-v" || a == "--verbose") verbose = std::max(verbose, 1);
        else if (a == "-vv" || a == "--very-verbose") verbose = 2;
        else if (a == "-fi" || a == "--bloom-filter") filterFile = value("-fi");
        else if (a == "-o" || a == "--output-file") outFile = value("-o");
        else if (a == "-e" || a == "--error-rate") { e = atoi(value("-e").c_str()); if (e < 0 || e > 10) usage("-e out of range 0..10"); }
        else if (a == "-rb" || a == "--reads-batch") { readsBatch = strtoull(value("-rb").c_str(), nullptr, 10); if (readsBatch < 1000 || readsBatch > 5000000) usage("-rb out of range"); }
        else if (a == "--device") device = atoi(value("--device").c_str());
        else if (a == "--inflight") { inflight = atoi(value("--inflight").c_str()); if (inflight < 1 || inflight > 16) usage("--inflight out of range 1..16"); }
        else if (a == "--devices")
        {
            const std::string v = value("--devices");
            if (v == "all") { int n = 0; dyibf_device_count(&n); for (int d = 0; d < n; ++d) devices.push_back(d); }
            else for (size_t at = 0; at <= v.size();) { size_t c = v.find(',', at); if (c == std::string::npos) c = v.size(); if (c > at) devices.push_back(atoi(v.substr(at, c - at).c_str())); at = c + 1; }
            if (devices.empty()) usage("--devices needs a list");
        }
        else if (!a.empty() && a[0] == '-') usage(("unknown option " + a).c_str());
        else inputs.push_back(a);
    }
    if (filterFile.empty() || inputs.empty() || inputs.size() > 2) usage("need -fi FILTER and one or two reads files");
    const bool paired = inputs.size() == 2;
    inflight = std::min<int>(inflight, 4 * (int)std::max<size_t>(devices.size(), 1));   // four tickets per GPU (dyibf.h)

    const double t0 = now();
    dyibf_filter * filter = nullptr;
    dyibf_group * grp = nullptr;
    if (devices.size() > 1) device = devices[0];
    const int lrc = devices.size() > 1 ? dyibf_group_load_file(filterFile.c_str(), devices.data(), (int)devices.size(), DYIBF_GROUP_READ_SHARDED, &grp)
                                       : dyibf_load_file(filterFile.c_str(), devices.empty() ? device : devices[0], &filter);
    if (lrc != DYIBF_OK)
    {
        std::cerr << dyibf_last_error() << std::endl;                     // d_bloom_filter.h:81-85: message, exit(1)
        return 1;
    }
    dyibf_params prm;
    if (grp) dyibf_group_get_params(grp, &prm, nullptr, nullptr); else dyibf_get_params(filter, &prm);
    const double loadFilter = now() - t0;

    const double t1 = now();
    Records m1, m2;
    parse(device, inputs[0], m1);
    if (paired)
    {
        parse(device, inputs[1], m2);
        if (m2.limits.size() != m1.limits.size()) { std::cerr << "The two reads files hold different numbers of records." << std::endl; return 1; }
    }
    const uint64_t nReads = m1.limits.size() - 1;                        // reads, or pairs
    const double loadReads = now() - t1;

    std::ofstream out;
    if (!outFile.empty()) { out.open(outFile); if (!out) { std::cerr << "Unable to open " << outFile << std::endl; return 1; } }
    double filterReads = 0;
    uint64_t filteredReads = 0, batchNo = 0;
    std::vector<uint64_t> binOffsets(prm.bins + 1); std::vector<uint32_t> ids, order(prm.bins);
    // a ring of batches in flight: each owns its pinned read bytes and limits until its ticket has ended
    struct InFlight { PinnedBytes seqs; std::vector<uint64_t> limits; dyibf_ticket * ticket = nullptr; uint64_t n = 0, batchNo = 0; uint16_t threshold = 0; };
    std::vector<InFlight> ring((size_t)inflight);
    auto finish = [&](InFlight & fl) -> bool
    {
        uint64_t total = 0;
        ids.resize(std::max<uint64_t>(ids.size(), 4 * fl.n));
        int rc = dyibf_classify_end(fl.ticket, binOffsets.data(), ids.data(), ids.size(), &total);
        if (rc == DYIBF_ERR_CAPACITY)                                    // the ticket stays in flight: again, with room
        {
            ids.resize(total);
            rc = dyibf_classify_end(fl.ticket, binOffsets.data(), ids.data(), ids.size(), &total);
        }
        fl.ticket = nullptr;
        if (rc != DYIBF_OK) { std::cerr << dyibf_last_error() << std::endl; return false; }
        // threshold 0: the reference pushes every read id (of the SE numbering) to every bin first and then still runs the loop (:320-333)
        const uint64_t pre = fl.threshold == 0 ? (paired ? 2 * fl.n : fl.n) : 0;
        if (verbose > 1)
            for (uint32_t b = 0; b < prm.bins; ++b)
                std::cerr << "bin " << b << "\t" << pre + binOffsets[b + 1] - binOffsets[b] << std::endl;
        for (uint32_t b = 0; b < prm.bins; ++b) filteredReads += (pre + binOffsets[b + 1] - binOffsets[b]) * (paired ? 2 : 1);   // :442
        if (out.is_open())
        {
            dyibf_sorted_bins(binOffsets.data(), prm.bins, order.data());
            for (uint32_t b : order)
            {
                out << fl.batchNo << ' ' << b << ' ' << pre + binOffsets[b + 1] - binOffsets[b];
                for (uint64_t i = 0; i < pre; ++i) out << ' ' << i;
                for (uint64_t i = binOffsets[b]; i < binOffsets[b + 1]; ++i) out << ' ' << ids[i];
                out << '\n';
            }
        }
        return true;
    };
    const double tFilter = now();
    for (uint64_t first = 0; first < nReads; first += readsBatch, ++batchNo)
    {
        InFlight & fl = ring[batchNo % ring.size()];
        if (fl.ticket && !finish(fl)) return 1;                          // the oldest batch leaves the ring, in batch order
        const uint64_t n = std::min(readsBatch, nReads - first);
        // the batch's forward read set: [reads] or [mates 1 | mates 2] (src/file_pair.h:82-88)
        fl.seqs.clear(); fl.limits.assign(1, 0); fl.n = n; fl.batchNo = batchNo;
        for (const Records * r : {(const Records *)&m1, paired ? (const Records *)&m2 : (const Records *)nullptr})
        {
            if (!r) continue;
            const uint64_t b0 = r->limits[first], b1 = r->limits[first + n];
            const uint64_t base = fl.seqs.n;
            fl.seqs.append(r->seqs.data() + b0, b1 - b0);
            for (uint64_t i = 1; i <= n; ++i) fl.limits.push_back(base + r->limits[first + i] - b0);
        }
        // clasifyLoadedReads, d_mapper.h:308-375
        const uint64_t nSeqs = fl.limits.size() - 1;                     // forward sequences; the set the reference sees is twice that
        const uint16_t avgReadLen = (uint16_t)(nSeqs ? fl.limits.back() / nSeqs : 0);   // lengthSum / (2 numReads), both doubled by the rc half
        fl.threshold = dyibf_threshold(e, prm.kmer, avgReadLen);
        if (fl.threshold == 0)
            std::cerr << "[WARNING!] 0 k-mer is required to filter a read!\nAll reads will pass filteration and be mapped everywhere.\n "
                         "This will be extremly slow.\n Choose an approprate error rate based on kmer size and read length\n";
        if (fl.seqs.n == 0) { const uint8_t zero = 0; fl.seqs.append(&zero, 1); }
        const int rc = grp ? dyibf_group_classify_begin(grp, fl.seqs.p, fl.limits.data(), (uint32_t)nSeqs, DYIBF_ENC_ASCII, paired ? 4 : 2, 1, fl.threshold, &fl.ticket)
                           : dyibf_classify_begin(filter, fl.seqs.p, fl.limits.data(), (uint32_t)nSeqs, DYIBF_ENC_ASCII, paired ? 4 : 2, 1, fl.threshold, &fl.ticket);
        if (rc != DYIBF_OK) { std::cerr << dyibf_last_error() << std::endl; return 1; }
    }
    for (uint64_t i = 0; i < ring.size(); ++i)                           // drain, oldest first
    {
        InFlight & fl = ring[(batchNo + i) % ring.size()];
        if (fl.ticket && !finish(fl)) return 1;
    }
    filterReads = now() - tFilter;
    if (verbose > 0)
    {
        const double total = (now() - t0) / 100.0;
        std::cerr << "\nFilter loading time:\t\t" << loadFilter << " sec\t\t" << loadFilter / total << " %" << std::endl;
        std::cerr << "Reads loading time:\t\t" << loadReads << " sec\t\t" << loadReads / total << " %" << std::endl;
        std::cerr << "Reads filtering time:\t\t" << filterReads << " sec\t\t" << filterReads / total << " %" << std::endl;
        std::cerr << "Reads count:\t\t\t" << nReads * (paired ? 2 : 1) << std::endl;
        std::cerr << "Avg reads per bin:\t\t" << (double)filteredReads / prm.bins << std::endl;
    }
    dyibf_destroy(filter);
    dyibf_group_destroy(grp);
    return 0;
}
