// tools/dyibf_build_filter.cpp — standalone equivalent of dream_yara_build_filter (temehi/dream_yara
// src/d_build_filter.cpp) for `-ft bloom`, on top of the C ABI (include/dyibf.h): same positional argument and
// flags (-b -t -k -nh -bs -o -v), same defaults (:88-96), same checks, and an output file byte-identical to the
// reference's.  The per-bin std::async ranges of build_filter (:227-260) become host reader threads; the bits are set
// by the atomic-OR kernel in batches of whole files.
//
//   dyibf_build_filter [-b 64] [-t 1] [-k 20] [-nh 4] [-bs 1] [-o out.filter] [-v] [--device 0] <REFERENCE FILES DIR>
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <future>
#include <iostream>
#include <mutex>
#include <string>
#include <vector>

#include "dyibf.h"
#include "fasta_batch.h"

using namespace dyibf_tools;

namespace {

struct Options
{
    std::string contigsDir, filterFile;
    uint32_t kmerSize = 20, numberOfBins = 64, numberOfHashes = 4, threadsCount = 1;     // d_build_filter.cpp:88-96
    uint64_t bloomFilterSize = 8589934592ull + 256;                                       // 1 GiB + filterMetadataSize
    bool verbose = false;
    bool kmerDirect = false;
    int device = 0;
};

double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

[[noreturn]] void usage(const char * msg)
{
    if (msg) std::cerr << "dyibf_build_filter: " << msg << "\n";
    std::cerr << "usage: dyibf_build_filter [OPTIONS] <REFERENCE FILES DIR>\n"
                 "  -b, --number-of-bins INT   1..4194300 (default 64)\n"
                 "  -t, --threads INT          host reader threads (default 1)\n"
                 "  -ft, --filter-type STR     bloom | kmer_direct (default bloom; kmer_direct: 4^k blocks, -k <= 15, -nh / -bs ignored)\n"
                 "  -k, --kmer-size INT        14..32 (default 20)\n"
                 "  -nh, --num-hash INT        2..5 (default 4)\n"
                 "  -bs, --bloom-size INT      GB, power of two, 1..512 (default 1)\n"
                 "  -o, --output-file FILE     *.filter (default <dir>bloom.filter)\n"
                 "  -v, --verbose\n      --device INT           CUDA device (default 0)\n";
    exit(msg ? 1 : 0);
}

uint64_t intArg(int & i, int argc, char ** argv, uint64_t lo, uint64_t hi, const char * name)
{
    if (i + 1 >= argc) usage((std::string("missing value for ") + name).c_str());
    char * end = nullptr;
    const std::string txt = argv[++i];
    unsigned long long v = strtoull(txt.c_str(), &end, 10);     // accepts the test suite's "-nh ' 2'" like SeqAn's lexical cast
    if (end == txt.c_str() || *end != 0 || v < lo || v > hi)
        usage((std::string("value of ") + name + " out of range").c_str());
    return v;
}

Options parse(int argc, char ** argv)
{
    Options o;
    bool haveOut = false;
    for (int i = 1; i < argc; ++i)
    {
        const std::string a = argv[i];
        if (a == "-h" || a == "--help") usage(nullptr);
        else if (a == "-v" || a == "--verbose") o.verbose = true;
        else if (a == "-b" || a == "--number-of-bins") o.numberOfBins = (uint32_t)intArg(i, argc, argv, 1, 4194300, "-b");
        else if (a == "-t" || a == "--threads") o.threadsCount = (uint32_t)intArg(i, argc, argv, 1, 2048, "-t");
        else if (a == "-k" || a == "--kmer-size") o.kmerSize = (uint32_t)intArg(i, argc, argv, 14, 32, "-k");
        else if (a == "-nh" || a == "--num-hash") o.numberOfHashes = (uint32_t)intArg(i, argc, argv, 2, 5, "-nh");
        else if (a == "--device") o.device = (int)intArg(i, argc, argv, 0, 1023, "--device");
        else if (a == "-bs" || a == "--bloom-size")
        {
            uint64_t bs = intArg(i, argc, argv, 1, 512, "-bs");
            if (bs & (bs - 1))                                                            // d_build_filter.cpp:198-210
            {
                std::cerr << "[ERROR] --bloom-size (-bs) parameter should be a power of 2!" << std::endl;
                exit(1);
            }
            o.bloomFilterSize = bs * 8589934592ull + 256;
        }
        else if (a == "-ft" || a == "--filter-type")
        {
            if (i + 1 >= argc) usage("missing value for -ft");
            const std::string v = argv[++i];
            if (v == "kmer_direct") o.kmerDirect = true;
            else if (v != "bloom") usage("-ft must be bloom or kmer_direct");
        }
        else if (a == "-o" || a == "--output-file")
        {
            if (i + 1 >= argc) usage("missing value for -o");
            o.filterFile = argv[++i]; haveOut = true;
            if (o.filterFile.size() < 7 || o.filterFile.substr(o.filterFile.size() - 7) != ".filter")
                usage("the output file must have the extension .filter");
        }
        else if (!a.empty() && a[0] == '-') usage(("unknown option " + a).c_str());
        else if (o.contigsDir.empty()) o.contigsDir = a;
        else usage("too many arguments");
    }
    if (o.contigsDir.empty()) usage("missing <REFERENCE FILES DIR>");
    if (o.contigsDir.back() != '/') o.contigsDir += '/';                                  // appendTrailingSlash
    if (!haveOut) o.filterFile = o.contigsDir + "bloom.filter";                           // :186-190
    if (o.kmerDirect)                                                                     // d_build_filter.cpp:305-310
    {
        if (o.kmerSize > 15) usage("-ft kmer_direct needs -k <= 15 (the reference computes 1u << 2k)");
        o.bloomFilterSize = (uint64_t)(1u << (2 * o.kmerSize)) * o.numberOfBins + 256;
        o.numberOfHashes = 0;                                                             // the C ABI's mark of a kmer_direct filter
    }
    return o;
}

} // namespace

int main(int argc, char ** argv)
{
    Options options = parse(argc, argv);
    const double t0 = now();
    // The filter (CUDA start-up + 16 GiB of HBM: one to three seconds on a fresh process) is created beside the reader tasks: the
    // first files are being read meanwhile; whoever feeds the GPU first waits for it.
    dyibf_filter * filter = nullptr;
    double tCreated = t0;
    std::string createError;
    std::shared_future<bool> created = std::async(std::launch::async, [&] {
        const bool ok = dyibf_create(options.numberOfBins, options.numberOfHashes, options.kmerSize, options.bloomFilterSize,
                                     options.device, &filter) == DYIBF_OK;
        if (!ok) createError = dyibf_last_error();                    // the message is per thread
        tCreated = now();
        return ok;
    }).share();
    const std::string comExt = commonExtension(options.contigsDir, options.numberOfBins);

    // reader tasks own contiguous bin ranges like the reference's tasks; whoever holds a full batch feeds the GPU
    const uint32_t numThr = options.threadsCount;
    uint32_t batchSize = options.numberOfBins / numThr;
    if (batchSize * numThr < options.numberOfBins) ++batchSize;
    // Reader tasks own contiguous bin ranges; what they have read goes into ONE shared batch (whole files, in the order the
    // tasks hand them over: every file is a run of one bin), and whoever makes it reach ~1 GiB of text feeds the GPU with it
    // while the others keep reading: few large build calls instead of one per task.
    std::mutex gpuMutex, poolMutex, logMutex;
    const uint64_t flushBytes = 1ull << 30, handBytes = 64ull << 20;
    TextBatch pool;
    double deviceSeconds = 0; uint32_t deviceCalls = 0;               // under gpuMutex
    auto hand_over = [&](TextBatch & mine, bool last) {
        TextBatch full;
        {
            std::lock_guard<std::mutex> lock(poolMutex);
            for (size_t i = 0; i < mine.texts.size(); ++i)
            {
                pool.bytes += mine.texts[i].size();
                pool.texts.push_back(std::move(mine.texts[i])); pool.binIds.push_back(mine.binIds[i]); pool.formats.push_back(mine.formats[i]);
            }
            mine.clear();
            if (pool.bytes >= flushBytes || last) std::swap(full, pool);
        }
        if (full.empty()) return;
        std::lock_guard<std::mutex> lock(gpuMutex);
        if (!created.get()) throw std::runtime_error(createError);
        const double tg = now();
        addTextBatch(filter, full);
        deviceCalls += 1; deviceSeconds += now() - tg;
    };
    std::atomic<uint32_t> running(numThr);
    std::vector<std::future<void>> tasks;
    for (uint32_t taskNo = 0; taskNo < numThr; ++taskNo)
    {
        tasks.emplace_back(std::async(std::launch::async, [&, taskNo] {
            TextBatch batch;
            try
            {
                for (uint32_t binNo = taskNo * batchSize; binNo < options.numberOfBins && binNo < (taskNo + 1) * batchSize; ++binNo)
                {
                    const double tb = now();
                    appendFileText(batch, options.contigsDir + std::to_string(binNo) + comExt, binNo);
                    if (batch.bytes >= handBytes) hand_over(batch, false);
                    if (options.verbose)
                    {
                        std::lock_guard<std::mutex> lock(logMutex);
                        std::cerr << "[bin " << binNo << "] Done adding kmers!\t\t\t" << now() - tb << " sec" << std::endl;
                    }
                }
            }
            catch (...) { --running; throw; }
            hand_over(batch, --running == 0);                        // the last task to finish flushes what is left
        }));
    }
    try
    {
        for (auto & t : tasks) t.get();
        // "last" above is decided outside poolMutex: a task that was not last can still push its files after the last one has
        // flushed.  Whatever is in the pool once every task has returned goes to the GPU here.
        TextBatch none;
        hand_over(none, true);
        if (!created.get()) throw std::runtime_error(createError);
    }
    catch (std::exception const & e) { std::cerr << "dyibf_build_filter: " << e.what() << std::endl; return 1; }
    if (options.verbose)
    {
        std::cerr << "Device calls (text up, tokeniser, build):\t" << deviceSeconds << " sec in " << deviceCalls << " calls; filter created after "
                  << tCreated - t0 << " sec" << std::endl;
        std::cerr << "All bins are done adding kmers!\t\t" << now() - t0 << " sec" << std::endl;
    }

    const double ts = now();
    if (dyibf_save_file(filter, options.filterFile.c_str()) != DYIBF_OK)
    {
        std::cerr << "dyibf_build_filter: " << dyibf_last_error() << std::endl;
        return 1;
    }
    if (options.verbose)
        std::cerr << "Done saving filter (" << dyibf_size_mb(filter) << " MB)\t\t" << now() - ts << " sec" << std::endl;
    dyibf_destroy(filter);
    std::cerr << "\nFinshed in \t\t\t" << now() - t0 << " sec" << std::endl;
    return 0;
}
