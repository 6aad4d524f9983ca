// tools/dyibf_filter_reads.cpp — the filtering stage of dream_yara_mapper as a standalone program on top of the C ABI
// (include/dyibf.h): what runDisMapper does between loading the filter and handing read sets to the per-bin mappers
// (temehi/dream_yara src/d_mapper.h:900-959, :308-387, :843-854), and nothing of the mapping itself.
//
//   dyibf_filter_reads -fi <FILTER> [-e 5] [-rb 100000] [-v|-vv] [-o lists.txt] [--device 0 | --devices 0,1,..] [--inflight 3]
//                      <READS> [<MATES>]
//
// The reads file(s) are STREAMED: a reader thread inflates / reads them piece by piece and cuts the text into batches of
// -rb records (four-line FASTQ: every fourth line; FASTA: every '>' at a line start), exactly the batches the mapper's
// loadReads forms (src/mapper.h:109, d_mapper.h:870-896); the main thread hands each batch's TEXT to the GPU as a ticket
// (dyibf_classify_text_begin): it is uploaded, tokenised on the device (SeqAn's readRecords, dyibf_parse_records), filtered
// and compacted there, and only the per-bin lists come back (dyibf_classify_end).  Up to --inflight batches are in flight, so
// file reading, the upload + tokenising of batch i+1, the query of batch i and the list download of batch i-1 all overlap —
// the reference's prefetching reader (src/file_prefetched.h:137-205) with the GPU as the consumer.  Text the cutter cannot
// split by lines alone (FASTQ wrapped over several lines, '>' inside a FASTA line, junk before the first record) is handed
// over whole from that record on: it is tokenised in one piece on the device and batched from the parsed records.
// With --devices the filter is replicated over several GPUs of this one process (dyibf_group_load_file) and successive
// batches go to successive GPUs.
//
// Per batch: the threshold comes from the batch's average read length (d_mapper.h:313-315), the lists are origReadIdMap.
// -vv prints "bin <b>\t<size>" per batch (:380-386); -v the timers and "Avg reads per bin" (:947-958); -o writes, per
// batch, one line per bin in sortedBins order (:843-854): "<batch> <bin> <n> id id ...".
#include <zlib.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <unistd.h>
#include <immintrin.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <mutex>
#include <string>
#include <thread>
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
                 "       --inflight INT        batches in flight, 1..16 (default 4; 1 = synchronous calls)\n";
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

// batch text lives in pinned host memory (its upload then runs at PCIe speed), grown on demand
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

// the text of one batch from one file: a view into the mapped file, or an own pinned copy (inflated input)
struct Piece
{
    PinnedBytes own;
    const uint8_t * p = nullptr; size_t n = 0;
    void clear() { own.clear(); p = nullptr; n = 0; }
    void view(const uint8_t * src, size_t len) { p = src; n = len; }
    void copy(const uint8_t * src, size_t len) { own.append(src, len); p = own.p; n = own.n; }
};

// the AVX2 form of the block loop below (128 bytes per step), used when the CPU has it
__attribute__((target("avx2")))
size_t skipLinesAvx2(const char * p, size_t n, uint64_t want, uint64_t * cnt_io)
{
    uint64_t cnt = *cnt_io;
    size_t i = 0;
    const __m256i nl = _mm256_set1_epi8('\n');
    while (i + 128 <= n && cnt + 128 < want)
    {
        const uint32_t m0 = (uint32_t)_mm256_movemask_epi8(_mm256_cmpeq_epi8(_mm256_loadu_si256((const __m256i *)(p + i)), nl));
        const uint32_t m1 = (uint32_t)_mm256_movemask_epi8(_mm256_cmpeq_epi8(_mm256_loadu_si256((const __m256i *)(p + i + 32)), nl));
        const uint32_t m2 = (uint32_t)_mm256_movemask_epi8(_mm256_cmpeq_epi8(_mm256_loadu_si256((const __m256i *)(p + i + 64)), nl));
        const uint32_t m3 = (uint32_t)_mm256_movemask_epi8(_mm256_cmpeq_epi8(_mm256_loadu_si256((const __m256i *)(p + i + 96)), nl));
        cnt += (uint64_t)__builtin_popcountll((uint64_t)m0 | ((uint64_t)m1 << 32)) + (uint64_t)__builtin_popcountll((uint64_t)m2 | ((uint64_t)m3 << 32));
        i += 128;
    }
    *cnt_io = cnt;
    return i;
}

// position behind the want-th '\n' of [p, p + n), or n (and *found < want) when there are fewer
size_t skipLines(const char * p, size_t n, uint64_t want, uint64_t * found)
{
    uint64_t cnt = 0;
    size_t i = 0;
    static const bool avx2 = __builtin_cpu_supports("avx2");
    if (avx2) i = skipLinesAvx2(p, n, want, &cnt);
    const __m128i nl = _mm_set1_epi8('\n');
    // whole 64-byte blocks while the remaining lines cannot end inside one
    while (i + 64 <= n && cnt + 64 < want)
    {
        const unsigned m0 = (unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(_mm_loadu_si128((const __m128i *)(p + i)), nl));
        const unsigned m1 = (unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(_mm_loadu_si128((const __m128i *)(p + i + 16)), nl));
        const unsigned m2 = (unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(_mm_loadu_si128((const __m128i *)(p + i + 32)), nl));
        const unsigned m3 = (unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(_mm_loadu_si128((const __m128i *)(p + i + 48)), nl));
        cnt += (uint64_t)__builtin_popcountll((uint64_t)m0 | ((uint64_t)m1 << 16) | ((uint64_t)m2 << 32) | ((uint64_t)m3 << 48));
        i += 64;
    }
    for (; i < n && cnt < want; ++i) cnt += p[i] == '\n';
    *found = cnt;
    return i;
}

// One input file cut into batches of whole records.  A plain file is mapped and the batches are views into the mapping (no
// copy in this program: the library stages pageable memory through pinned pieces with its helper threads); a gzip file is
// inflated piece by piece into a buffer and the batches are pinned copies.
struct Source
{
    gzFile gz = nullptr;
    std::string path;
    std::vector<char> buf;            // inflated text (gzip input)
    const char * map = nullptr; size_t mapSize = 0;
    size_t begin = 0, end = 0;        // unread text: [begin, end) of data()
    bool eof = false, started = false;
    int fmt = -1;
    std::vector<std::thread> prefault;
    std::atomic<bool> stopPrefault{false};

    const char * data() const { return map ? map : buf.data(); }
    void open(const std::string & p)
    {
        path = p;
        unsigned char magic[2] = {0, 0};
        FILE * fp = fopen(p.c_str(), "rb");
        if (!fp) { std::cerr << "Unable to open reads file: " << p << std::endl; exit(1); }
        const size_t got = fread(magic, 1, 2, fp);
        fseeko(fp, 0, SEEK_END);
        const off_t size = ftello(fp);
        fclose(fp);
        if (!(got == 2 && magic[0] == 0x1f && magic[1] == 0x8b) && size > 0)
        {
            const int fd = ::open(p.c_str(), O_RDONLY);
            void * m = fd >= 0 ? mmap(nullptr, (size_t)size, PROT_READ, MAP_PRIVATE, fd, 0) : MAP_FAILED;
            if (fd >= 0) close(fd);
            if (m != MAP_FAILED)
            {
                map = (const char *)m; mapSize = end = (size_t)size; eof = true;
                madvise(m, mapSize, MADV_SEQUENTIAL);
                // page-cache pages are mapped on first touch (~0.5 us per 4 KiB page: 0.4 s for a 3 GB file, as long as the
                // whole scan): four helper threads populate the page tables ahead of the cutter, stripe by stripe
                for (int t = 0; t < 4; ++t)
                    prefault.emplace_back([this, t]
                    {
                        const size_t stripe = 32u << 20;
                        for (size_t off = (size_t)t * stripe; off < mapSize && !stopPrefault; off += 4 * stripe)
                        {
                            const size_t len = std::min(stripe, mapSize - off);
#ifdef MADV_POPULATE_READ
                            if (madvise((void *)(map + off), len, MADV_POPULATE_READ) == 0) continue;
#endif
                            volatile char sink = 0;
                            for (size_t i = 0; i < len; i += 4096) sink = sink + map[off + i];
                        }
                    });
            }
        }
        if (!map)
        {
            gz = gzopen(p.c_str(), "rb");                          // plain files that could not be mapped pass through
            if (!gz) { std::cerr << "Unable to open reads file: " << p << std::endl; exit(1); }
            gzbuffer(gz, 1 << 20);
            buf.resize(32 << 20);
            more();
        }
        fmt = formatOf(p, std::vector<char>(data(), data() + std::min<size_t>(end, 1)));
    }
    ~Source()
    {
        stopPrefault = true;
        for (auto & t : prefault) t.join();
        if (gz) gzclose(gz);
        if (map) munmap((void *)map, mapSize);
    }
    // more text behind `end` (positions stay valid: the buffer only grows here); false at the end of the file
    bool more()
    {
        if (eof) return false;
        if (end == buf.size()) buf.resize(buf.size() * 2);
        const int n = gzread(gz, buf.data() + end, (unsigned)std::min<size_t>(buf.size() - end, 1u << 30));
        if (n < 0) { std::cerr << "Error while reading: " << path << std::endl; exit(1); }
        if (n == 0) { eof = true; return false; }
        end += (size_t)n;
        return true;
    }
    bool have(size_t p) { while (p >= end) if (!more()) return false; return true; }
    static bool blank(char c) { return c == '\n' || c == '\r' || c == ' ' || c == '\t'; }
    // position behind the next newline at or after p; `end` (with *open_line set) when the file ends inside the line
    size_t lineEnd(size_t p, bool * open_line)
    {
        *open_line = false;
        for (;;)
        {
            if (!have(p)) { *open_line = true; return end; }
            const char * nl = (const char *)memchr(data() + p, '\n', end - p);
            if (nl) return (size_t)(nl - data()) + 1;
            p = end;
        }
    }
    // The text of the next (up to) `want` records -> out; returns the records cut.  *irregular: the text that follows is not
    // something lines alone can split (nothing of it was consumed).
    uint64_t cut(uint64_t want, Piece & out, bool * irregular)
    {
        *irregular = false;
        if (!map && begin > 0) { memmove(buf.data(), buf.data() + begin, end - begin); end -= begin; begin = 0; }
        size_t pos = begin;                                        // records [begin, pos) are complete
        uint64_t n = 0;
        if (!started)
        {
            if (!have(pos)) return 0;                              // empty file
            if (data()[pos] != (fmt == DYIBF_FMT_FASTQ ? '@' : '>')) { *irregular = true; return 0; }
            if (fmt == DYIBF_FMT_FASTQ)
            {
                // Is this four-line FASTQ at all?  The head of the file is checked line by line ('@' / sequence / '+' / qualities
                // of equal length, 256 records); a wrapped or blank-line-separated file goes over whole (the device reader's
                // general FASTQ parser handles it).  Later batches are cut by counting lines only; the record count the
                // tokeniser reports must then equal the number cut, so a layout change in mid-file cannot go unnoticed.
                size_t p = pos;
                for (int r = 0; r < 256 && have(p); ++r)
                {
                    size_t len[4]; bool open_line = false, ok = true;
                    const size_t rec0 = p;
                    for (int line = 0; line < 4 && ok; ++line)
                    {
                        ok = have(p) && (line != 0 || data()[p] == '@') && (line != 2 || data()[p] == '+');
                        if (!ok) break;
                        const size_t q = lineEnd(p, &open_line);
                        size_t e = q;
                        while (e > p && (data()[e - 1] == '\n' || data()[e - 1] == '\r')) --e;
                        len[line] = e - p;
                        if (open_line && line != 3) ok = false;
                        p = q;
                    }
                    if (ok && len[1] != len[3]) ok = false;
                    if (!ok)
                    {
                        // white space only behind the last record is fine
                        size_t w = rec0; while (have(w) && blank(data()[w])) ++w;
                        if (!have(w) && r > 0) break;
                        *irregular = true; return 0;
                    }
                }
            }
            started = true;
        }
        while (n < want && have(pos))
        {
            if (blank(data()[pos]))
            {
                // white space after a record: fine when nothing but white space follows (SeqAn skips it, fasta_fastq.h:339)
                size_t p = pos;
                while (have(p) && blank(data()[p])) ++p;
                if (!have(p)) { pos = end; break; }
                *irregular = true;
                break;
            }
            if (fmt == DYIBF_FMT_FASTQ)
            {
                // Four lines per record: the batch ends behind its 4 * n-th newline.  Only the first line is looked at here; the
                // device tokeniser checks the layout of every record ('@' / '+' lines, lengths) and refuses what is not
                // four-line FASTQ, so a cut at the wrong place cannot go unnoticed.
                if (data()[pos] != '@') { *irregular = true; break; }
                uint64_t found = 0;
                const uint64_t lines = 4 * (want - n);
                size_t p = pos;
                for (;;)
                {
                    uint64_t f = 0;
                    p += skipLines(data() + p, end - p, lines - found, &f);
                    found += f;
                    if (found == lines || !more()) break;
                }
                if (found < lines)
                {
                    // the file ends here: white space behind the last record does not count (SeqAn skips it), and the last
                    // line may lack its newline
                    size_t e = end;
                    while (e > pos && blank(data()[e - 1])) --e;
                    skipLines(data() + pos, e - pos, ~0ull, &found);
                    found += e > pos ? 1 : 0;
                    p = end;
                }
                const uint64_t recs = found / 4;
                if (found % 4 != 0 || recs == 0) { *irregular = true; break; }     // the file ends inside a record: the tokeniser reports it
                n += recs; pos = p;
            }
            else
            {
                // '>' at a line start opens a record; it runs to the next such '>'.  SeqAn ends a record at ANY '>': one inside
                // a line makes the layout irregular
                size_t p = pos;
                bool ok = data()[p] == '>';
                for (++p; ok;)
                {
                    if (!have(p)) { p = end; break; }
                    const char * gt = (const char *)memchr(data() + p, '>', end - p);
                    if (!gt) { p = end; continue; }
                    p = (size_t)(gt - data());
                    ok = data()[p - 1] == '\n';
                    break;
                }
                if (!ok) { *irregular = true; break; }
                pos = p; ++n;
            }
        }
        if (map) out.view((const uint8_t *)data() + begin, pos - begin);
        else out.copy((const uint8_t *)data() + begin, pos - begin);
        begin = pos;
        return n;
    }
    // everything that is left
    void rest(std::vector<char> & out)
    {
        while (more()) {}
        out.assign(data() + begin, data() + end);
        begin = end;
    }
};

struct Records { std::vector<uint8_t> seqs; std::vector<uint64_t> limits; };

void parseText(int device, const std::string & path, int fmt, const std::vector<char> & text, Records & r)
{
    uint64_t n = 0, nb = 0;
    int rc = dyibf_parse_records(device, text.data(), text.size(), fmt, DYIBF_ALPHA_DNA5, nullptr, 0, nullptr, 0, 0, &n, &nb, DYIBF_MEM_HOST, nullptr);
    if (rc == DYIBF_OK)
    {
        r.seqs.resize(nb + 1); r.limits.resize(n + 1);
        rc = dyibf_parse_records(device, text.data(), text.size(), fmt, DYIBF_ALPHA_DNA5, r.seqs.data(), nb, r.limits.data(), n, 0, &n, &nb,
                                 DYIBF_MEM_HOST, nullptr);
    }
    if (rc != DYIBF_OK) { std::cerr << path << ": " << dyibf_last_error() << std::endl; exit(1); }
}

// a batch of text on its way from the reader thread to the GPU
struct TextBatch
{
    Piece text[2];
    uint64_t n = 0;                   // records (pairs) in it
    bool rest = false, last = false;  // rest: restText holds everything from here on (irregular layout)
    std::vector<char> restText[2];
};

struct BatchQueue
{
    std::mutex m; std::condition_variable cv;
    std::vector<TextBatch *> ready, spare;
    TextBatch * take(std::vector<TextBatch *> & q) { std::unique_lock<std::mutex> l(m); cv.wait(l, [&] { return !q.empty(); }); TextBatch * b = q.front(); q.erase(q.begin()); return b; }
    void give(std::vector<TextBatch *> & q, TextBatch * b) { { std::lock_guard<std::mutex> l(m); q.push_back(b); } cv.notify_all(); }
};

} // namespace

int main(int argc, char ** argv)
{
    std::string filterFile, outFile;
    std::vector<std::string> inputs;
    int e = 5, verbose = 0, device = 0, inflight = 4;
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
    inflight = std::min<int>(inflight, 8 * (int)std::max<size_t>(devices.size(), 1));   // eight tickets per GPU (dyibf.h)

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

    // ---- the reader thread: files -> batches of text ----
    Source src[2];
    src[0].open(inputs[0]);
    if (paired) { src[1].open(inputs[1]); if (src[1].fmt != src[0].fmt) { std::cerr << "The two reads files differ in format." << std::endl; return 1; } }
    const int fmt = src[0].fmt;
    BatchQueue bq;
    std::vector<TextBatch> pool((size_t)inflight + 3);
    for (auto & b : pool) bq.spare.push_back(&b);
    double loadReads = 0;
    std::thread reader([&]
    {
        for (bool done = false; !done;)
        {
            TextBatch * b = bq.take(bq.spare);
            const double ta = now();
            b->text[0].clear(); b->text[1].clear(); b->rest = false; b->last = false;
            b->restText[0].clear(); b->restText[1].clear();
            bool irr0 = false, irr1 = false;
            const uint64_t n0 = src[0].cut(readsBatch, b->text[0], &irr0);
            const uint64_t n1 = paired ? src[1].cut(n0, b->text[1], &irr1) : n0;   // the mates: the same number of records
            b->n = n0;
            if (!irr0 && !irr1 && n1 == n0)
            {
                b->last = done = n0 < readsBatch;                    // a short batch: the end of the file
                loadReads += now() - ta;
                bq.give(bq.ready, b);
                continue;
            }
            // From here on lines alone cannot split the text (or the two files fell out of step): the rest of the file(s) goes
            // over in one piece.  Records cut in step just before go first as a normal batch; otherwise their text leads the rest.
            TextBatch * r = b;
            if (n1 == n0 && n0 > 0)
            {
                bq.give(bq.ready, b);
                r = bq.take(bq.spare);
                r->text[0].clear(); r->text[1].clear(); r->restText[0].clear(); r->restText[1].clear();
            }
            else
                for (int i = 0; i < 2; ++i) r->restText[i].assign((const char *)b->text[i].p, (const char *)b->text[i].p + b->text[i].n);
            std::vector<char> tail;
            for (int i = 0; i < (paired ? 2 : 1); ++i)
            {
                src[i].rest(tail);
                r->restText[i].insert(r->restText[i].end(), tail.begin(), tail.end());
            }
            r->n = 0; r->rest = true; r->last = done = true;
            loadReads += now() - ta;
            bq.give(bq.ready, r);
        }
    });

    std::ofstream out;
    if (!outFile.empty()) { out.open(outFile); if (!out) { std::cerr << "Unable to open " << outFile << std::endl; return 1; } }
    double filterReads = 0;
    uint64_t filteredReads = 0, batchNo = 0, nReads = 0;
    std::vector<uint64_t> binOffsets(prm.bins + 1); std::vector<uint32_t> ids, order(prm.bins);
    // a ring of batches in flight: each owns its ticket (and, on the parsed path, its pinned read bytes and limits) until it has ended
    struct InFlight { PinnedBytes seqs; std::vector<uint64_t> limits; dyibf_ticket * ticket = nullptr; uint64_t n = 0, batchNo = 0; uint32_t threshold = 0; TextBatch * text = nullptr; };
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
        if (fl.text) { bq.give(bq.spare, fl.text); fl.text = nullptr; }
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
    auto warnZero = [] {
        std::cerr << "[WARNING!] 0 k-mer is required to filter a read!\nAll reads will pass filteration and be mapped everywhere.\n "
                     "This will be extremly slow.\n Choose an approprate error rate based on kmer size and read length\n";
    };
    // the parsed path: batches cut from parsed records (irregular layouts)
    auto runParsed = [&](const Records & m1, const Records & m2) -> bool
    {
        const uint64_t nRec = m1.limits.size() - 1;
        for (uint64_t first = 0; first < nRec; first += readsBatch, ++batchNo)
        {
            InFlight & fl = ring[batchNo % ring.size()];
            if (fl.ticket && !finish(fl)) return false;                  // the oldest batch leaves the ring, in batch order
            const uint64_t n = std::min(readsBatch, nRec - first);
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
            const uint64_t nSeqs = fl.limits.size() - 1;                 // forward sequences; the set the reference sees is twice that
            const uint16_t avgReadLen = (uint16_t)(nSeqs ? fl.limits.back() / nSeqs : 0);   // lengthSum / (2 numReads), both doubled by the rc half
            fl.threshold = dyibf_threshold(e, prm.kmer, avgReadLen);
            if (fl.threshold == 0) warnZero();
            if (fl.seqs.n == 0) { const uint8_t zero = 0; fl.seqs.append(&zero, 1); }
            const int rc = grp ? dyibf_group_classify_begin(grp, fl.seqs.p, fl.limits.data(), (uint32_t)nSeqs, DYIBF_ENC_ASCII, paired ? 4 : 2, 1, fl.threshold, &fl.ticket)
                               : dyibf_classify_begin(filter, fl.seqs.p, fl.limits.data(), (uint32_t)nSeqs, DYIBF_ENC_ASCII, paired ? 4 : 2, 1, fl.threshold, &fl.ticket);
            if (rc != DYIBF_OK) { std::cerr << dyibf_last_error() << std::endl; return false; }
            nReads += n;
        }
        return true;
    };
    const double tFilter = now();
    bool ok = true;
    for (bool last = false; ok && !last;)
    {
        TextBatch * b = bq.take(bq.ready);
        last = b->last;
        if (b->rest)
        {
            Records m1, m2;
            const double ta = now();
            parseText(device, inputs[0], fmt, b->restText[0], m1);
            if (paired)
            {
                parseText(device, inputs[1], fmt, b->restText[1], m2);
                if (m2.limits.size() != m1.limits.size()) { std::cerr << "The two reads files hold different numbers of records." << std::endl; ok = false; }
            }
            loadReads += now() - ta;
            ok = ok && runParsed(m1, m2);
            bq.give(bq.spare, b);
            continue;
        }
        if (b->n == 0) { bq.give(bq.spare, b); continue; }
        InFlight & fl = ring[batchNo % ring.size()];
        if (fl.ticket && !finish(fl)) { ok = false; break; }
        fl.n = b->n; fl.batchNo = batchNo++; fl.text = b;
        uint32_t nr = 0;
        // several GPUs: successive batches to successive replicas, through the member handles
        dyibf_filter * target = filter;
        if (grp) dyibf_group_member(grp, (int)(fl.batchNo % devices.size()), &target, nullptr, nullptr);
        const int rc = dyibf_classify_text_begin(target, (const char *)b->text[0].p, b->text[0].n, paired ? (const char *)b->text[1].p : nullptr,
                                                 paired ? b->text[1].n : 0, fmt, e, &nr, &fl.threshold, &fl.ticket);
        if (rc != DYIBF_OK) { std::cerr << dyibf_last_error() << std::endl; ok = false; break; }
        if (nr != b->n) { std::cerr << "batch " << fl.batchNo << ": " << nr << " records tokenised, " << b->n << " cut" << std::endl; ok = false; break; }
        if (fl.threshold == 0) warnZero();
        nReads += b->n;
        bq.give(bq.spare, b); fl.text = nullptr;                         // the text is on the device: the reader may refill the buffer
    }
    for (uint64_t i = 0; ok && i < ring.size(); ++i)                     // drain, oldest first
    {
        InFlight & fl = ring[(batchNo + i) % ring.size()];
        if (fl.ticket && !finish(fl)) ok = false;
    }
    if (!ok)
    {
        // the reader thread may be blocked on the batch queue: leave without unwinding it (everything worth flushing is on std::cerr)
        std::cerr.flush();
        std::_Exit(1);
    }
    reader.join();
    filterReads = now() - tFilter;
    if (verbose > 0)
    {
        const double total = (now() - t0) / 100.0;
        std::cerr << "\nFilter loading time:\t\t" << loadFilter << " sec\t\t" << loadFilter / total << " %" << std::endl;
        std::cerr << "Reads loading time:\t\t" << loadReads << " sec\t\t" << loadReads / total << " %  (reader thread, overlapped)" << std::endl;
        std::cerr << "Reads filtering time:\t\t" << filterReads << " sec\t\t" << filterReads / total << " %" << std::endl;
        std::cerr << "Reads count:\t\t\t" << nReads * (paired ? 2 : 1) << std::endl;
        std::cerr << "Avg reads per bin:\t\t" << (double)filteredReads / prm.bins << std::endl;
    }
    dyibf_destroy(filter);
    dyibf_group_destroy(grp);
    return 0;
}
