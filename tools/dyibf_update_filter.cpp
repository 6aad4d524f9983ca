// tools/dyibf_update_filter.cpp — standalone equivalent of dream_yara_update_filter (temehi/dream_yara
// src/d_update_filter.cpp:188-252) for bloom filters, on top of the C ABI (include/dyibf.h): load the .filter into
// HBM, clearBins for the bins named by the FASTA files (file name = bin number, :276-285 of d_misc_options.h), add the
// new k-mers, save in place.
//
//   dyibf_update_filter [-t 1] [-v] [--device 0] <BLOOM-FILTER FILE> <4.fna> <7.fna> ...
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <map>
#include <string>
#include <vector>

#include "dyibf.h"
#include "fasta_batch.h"

using namespace dyibf_tools;

namespace {

double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// getBinNoFromFile, d_misc_options.h:276-285: the file name without directory and (last) extension must be a number
bool getBinNoFromFile(uint32_t & binNo, const std::string & path)
{
    const size_t slash = path.find_last_of('/');
    std::string name = slash == std::string::npos ? path : path.substr(slash + 1);
    if (name.size() > 3 && name.compare(name.size() - 3, 3, ".gz") == 0) name.resize(name.size() - 3);   // 4.fna.gz -> 4.fna
    const size_t dot = name.find_last_of('.');
    if (dot != std::string::npos) name.resize(dot);                                                     // trimExtension
    if (name.empty()) return false;
    char * p = nullptr;
    binNo = (uint32_t)std::strtol(name.c_str(), &p, 10);
    return *p == 0;
}

[[noreturn]] void usage(const char * msg)
{
    if (msg) std::cerr << "dyibf_update_filter: " << msg << "\n";
    std::cerr << "usage: dyibf_update_filter [-t INT] [-ft bloom] [-v] [--device INT] <BLOOM-FILTER FILE> <4.fna> <7.fna> ...\n";
    exit(msg ? 1 : 0);
}

} // namespace

int main(int argc, char ** argv)
{
    std::string filterFile;
    std::map<uint32_t, std::string> binContigs;
    bool verbose = false;
    int device = 0;
    for (int i = 1; i < argc; ++i)
    {
        const std::string a = argv[i];
        if (a == "-h" || a == "--help") usage(nullptr);
        else if (a == "-v" || a == "--verbose") verbose = true;
        else if (a == "-t" || a == "--threads") { if (++i >= argc) usage("missing value for -t"); }   // the GPU does the pass
        else if (a == "--device") { if (++i >= argc) usage("missing value for --device"); device = atoi(argv[i]); }
        else if (a == "-ft" || a == "--filter-type")
        {
            if (++i >= argc) usage("missing value for -ft");
            const std::string v = argv[i];                              // the kind is read from the file's trailer anyway
            if (v != "bloom" && v != "kmer_direct") usage("-ft must be bloom or kmer_direct");
        }
        else if (!a.empty() && a[0] == '-') usage(("unknown option " + a).c_str());
        else if (filterFile.empty()) filterFile = a;
        else
        {
            uint32_t binNo = 0;
            if (!getBinNoFromFile(binNo, a))
            {
                std::cerr << "File: " << a << "\ndoesn't have a valid name\n";
                exit(1);
            }
            binContigs[binNo] = a;
        }
    }
    if (filterFile.empty() || binContigs.empty()) usage("need a filter file and at least one FASTA file");

    const double t0 = now();
    dyibf_filter * filter = nullptr;
    if (dyibf_load_file(filterFile.c_str(), device, &filter) != DYIBF_OK)
    {
        std::cerr << dyibf_last_error() << std::endl;                  // "File ... could not be read." + exit(1), d_bloom_filter.h:81-85
        return 1;
    }
    dyibf_params prm;
    dyibf_get_params(filter, &prm);
    std::vector<uint32_t> bins2update;
    for (auto const & kv : binContigs)
    {
        if (kv.first >= prm.bins)                                      // d_update_filter.cpp:197-203
        {
            std::cerr << "The provided bloom filter has only " << prm.bins << " Bins.\nRetry after removing " << kv.second
                      << " from arguments!" << std::endl;
            return 1;
        }
        bins2update.push_back(kv.first);
    }
    if (dyibf_clear_bins(filter, bins2update.data(), (uint32_t)bins2update.size()) != DYIBF_OK)
    {
        std::cerr << "dyibf_update_filter: " << dyibf_last_error() << std::endl;
        return 1;
    }
    TextBatch batch;
    try
    {
        for (auto const & kv : binContigs)
        {
            const double tb = now();
            appendFileText(batch, kv.second, kv.first);
            if (verbose)
                std::cerr << "[bin " << kv.first << "] updated using " << kv.second << "!\t\t" << now() - tb << " sec" << std::endl;
            if (batch.bytes >= (1ull << 30) || kv.first == binContigs.rbegin()->first) addTextBatch(filter, batch);
        }
    }
    catch (std::exception const & e) { std::cerr << "dyibf_update_filter: " << e.what() << std::endl; return 1; }
    if (verbose) std::cerr << "All given bins are updated!\t\t" << now() - t0 << " sec" << std::endl;
    const double ts = now();
    if (dyibf_save_file(filter, filterFile.c_str()) != DYIBF_OK)
    {
        std::cerr << "dyibf_update_filter: " << dyibf_last_error() << std::endl;
        return 1;
    }
    if (verbose) std::cerr << "Done saving filter (" << dyibf_size_mb(filter) << " MB)\t\t" << now() - ts << " sec" << std::endl;
    dyibf_destroy(filter);
    std::cerr << "\nFinshed in \t\t\t" << now() - t0 << " sec" << std::endl;
    return 0;
}
