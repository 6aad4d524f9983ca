// ==========================================================================
//                     d_bloom_filter.h  (B200 drop-in)
// ==========================================================================
// Drop-in replacement for DREAM-Yara's src/d_bloom_filter.h (temehi/dream_yara, lines 41-331):
// the same class seqan::SeqAnBloomFilter<TString> with the same public members, argument meaning and error
// behaviour, but the filter lives in the HBM of a B200 and every member forwards to the C ABI of libdyibf.so
// (include/dyibf.h; hand-written sm_100a CUDA).  There is no CPU fallback.
//
// Use: put this directory before the reference's src/ on the include path (or overwrite the file), add
// `-ldyibf` to the link line.  The three translation units that include the header compile unchanged
// (src/d_build_filter.cpp:61, src/d_update_filter.cpp:61, src/d_mapper.cpp:92).
//
// Per-sequence whichBins()/addKmers() are batches of one — correct, but a PCIe round trip each.  The two call
// sites that matter should use the batched members added at the bottom (whichBinsBatch / classify /
// addSequences); include/d_mapper_filter.h holds the replacement body of clasifyLoadedReads, INTEGRATION.md the
// patches.
//
// Like the original, this header expects <seqan/sequence.h>, <seqan/seq_io.h> and d_misc_options.h
// (INT_WIDTH, filterMetadataSize) to have been included before it.
#ifndef APP_YARA_DIS_BLOOM_FILTER_B200_H_
#define APP_YARA_DIS_BLOOM_FILTER_B200_H_

#include <cstdint>
#include <cstdlib>
#include <iostream>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "dyibf.h"

// src/d_mapper.cpp:537 (readFilterMetadata) uses sdsl::int_vector<64> and gets it through the reference header's
// #include <sdsl/bit_vectors.hpp> (src/d_bloom_filter.h:35); keep that TU compiling unchanged where sdsl is on the path.
// This class itself does not use sdsl.
#if defined(__has_include)
#  if __has_include(<sdsl/bit_vectors.hpp>)
#    include <sdsl/bit_vectors.hpp>
#  endif
#endif
#include <valarray>
#include <algorithm>

namespace seqan
{
    template<typename TString = Dna5String>
    class SeqAnBloomFilter
    {
    public:
        typedef Shape<Dna, SimpleShape> TShape;     // kept for source compatibility; hashing happens on the GPU

        // Which GPUs: the reference's call sites construct the filter without saying (src/d_mapper.h:913, src/d_build_filter.cpp:298,
        // src/d_update_filter.cpp:276), so the unmodified programs pick the devices up from the environment:
        //   DYIBF_DEVICES = "all" | "0,1,2,3"   (default: device 0 only)
        //   DYIBF_GROUP_MODE = "read" (default: a replica per GPU, every read batch split over them) | "bin" (the GPUs own
        //                      contiguous bin ranges of a filter larger than one GPU's HBM)
        // or a caller passes them: SeqAnBloomFilter(fileName, devices, mode).  More than one device makes the object a
        // dyibf_group (include/dyibf.h); every member below works on either.

        // d_bloom_filter.h:48
        SeqAnBloomFilter(uint32_t n_bins, uint8_t n_hash_func, uint8_t kmer_size, uint64_t vec_size, int device = -1)
        {
            // spawnDisMapper's placeholder for `-ft none` (src/d_mapper.h:941: SeqAnBloomFilter<> filter(64, 3, 20, 1)) is
            // never queried: a vector too small for one hash block makes an empty object instead of an error
            if (vec_size < 256 + 64ull * ((n_bins + 63) / 64)) { _prm = dyibf_params(); _prm.bins = n_bins; _prm.n_hash = n_hash_func; _prm.kmer = kmer_size; return; }
            std::vector<int> devs = _devicesFor(device);
            if (devs.size() > 1) _check(dyibf_group_create(n_bins, n_hash_func, kmer_size, vec_size, devs.data(), (int)devs.size(), _envMode(), &_g));
            else                 _check(dyibf_create(n_bins, n_hash_func, kmer_size, vec_size, devs[0], &_h));
            _cacheParams();
        }

        // d_bloom_filter.h:59 — load failure / size mismatch: message on std::cerr, exit(1)
        SeqAnBloomFilter(const char *fileName, uint32_t n_bins, uint8_t n_hash_func, uint8_t kmer_size, uint64_t vec_size,
                         int device = -1)
        {
            std::vector<int> devs = _devicesFor(device);
            if (devs.size() > 1)    // the caller's parameters, not the trailer's, as on one device and in the reference
                _exitOnLoadFailure(dyibf_group_load_file_expect(fileName, n_bins, n_hash_func, kmer_size, vec_size, devs.data(), (int)devs.size(), _envMode(), &_g));
            else
                _exitOnLoadFailure(dyibf_load_file_expect(fileName, n_bins, n_hash_func, kmer_size, vec_size, devs[0], &_h));
            _cacheParams();
        }

        // d_bloom_filter.h:79 — parameters from the file's trailer
        explicit SeqAnBloomFilter(const char *fileName, int device = -1)
        {
            std::vector<int> devs = _devicesFor(device);
            if (devs.size() > 1) _exitOnLoadFailure(dyibf_group_load_file(fileName, devs.data(), (int)devs.size(), _envMode(), &_g));
            else                 _exitOnLoadFailure(dyibf_load_file(fileName, devs[0], &_h));
            _cacheParams();
        }

        // explicit device lists (mode: DYIBF_GROUP_READ_SHARDED / DYIBF_GROUP_BIN_SHARDED)
        SeqAnBloomFilter(const char *fileName, std::vector<int> const & devices, int mode)
        {
            _exitOnLoadFailure(dyibf_group_load_file(fileName, devices.data(), (int)devices.size(), mode, &_g));
            _cacheParams();
        }
        SeqAnBloomFilter(uint32_t n_bins, uint8_t n_hash_func, uint8_t kmer_size, uint64_t vec_size, std::vector<int> const & devices, int mode)
        {
            _check(dyibf_group_create(n_bins, n_hash_func, kmer_size, vec_size, devices.data(), (int)devices.size(), mode, &_g));
            _cacheParams();
        }

        ~SeqAnBloomFilter() { dyibf_destroy(_h); dyibf_group_destroy(_g); }

        // the object owns a device handle: movable, not copyable (the reference object was never copied either)
        SeqAnBloomFilter(SeqAnBloomFilter const &) = delete;
        SeqAnBloomFilter & operator=(SeqAnBloomFilter const &) = delete;
        SeqAnBloomFilter(SeqAnBloomFilter && o) : _h(o._h), _g(o._g), _nDevices(o._nDevices), _mode(o._mode), _prm(o._prm), _mtx(std::move(o._mtx))
        {
            o._h = nullptr; o._g = nullptr;
            o._mtx.reset(new std::mutex);               // the moved-from object stays destructible and lockable
        }

        // d_bloom_filter.h:91
        void addKmers(TString const & text, uint32_t const & binNo)
        {
            if (length(text) < _prm.kmer) return;               // reference: out-of-bounds hashing; here: nothing
            uint64_t offsets[2] = {0, (uint64_t)length(text)};
            std::lock_guard<std::mutex> lock(*_mtx);
            _add(_bytes(text), offsets, 1, &binNo, _encoding(text));
        }

        // d_bloom_filter.h:99 — threadsCount is accepted for source compatibility; the GPU does the pass
        void clearBins(std::vector<uint32_t> & bins2clear, uint32_t & /*threadsCount*/)
        {
            std::lock_guard<std::mutex> lock(*_mtx);
            if (_g) _check(dyibf_group_clear_bins(_g, bins2clear.data(), (uint32_t)bins2clear.size()));
            else    _check(dyibf_clear_bins(_need(), bins2clear.data(), (uint32_t)bins2clear.size()));
        }

        // d_bloom_filter.h:127 — same reader, same Iupac -> TString conversion, one kernel batch per file
        void addFastaFile(CharString const & fastaFile, uint32_t const & binNo)
        {
            CharString id;
            IupacString seq;
            SeqFileIn seqFileIn;
            if (!open(seqFileIn, toCString(fastaFile)))
            {
                std::string msg = "Unable to open contigs File: ";
                msg += toCString(fastaFile);
                throw std::runtime_error(msg);                   // reference throws a dangling const char* (:133-138)
            }
            std::vector<uint8_t> concat;
            std::vector<uint64_t> offsets(1, 0);
            while (!atEnd(seqFileIn))
            {
                readRecord(id, seq, seqFileIn);
                if (length(seq) < _prm.kmer)
                    continue;
                TString text = seq;                               // Iupac -> Dna5 exactly as the reference converts
                const uint8_t * p = _bytes(text);
                concat.insert(concat.end(), p, p + length(text));
                offsets.push_back(concat.size());
            }
            close(seqFileIn);
            if (offsets.size() > 1)
            {
                std::vector<uint32_t> bins(offsets.size() - 1, binNo);
                std::lock_guard<std::mutex> lock(*_mtx);     // the reference calls this from several threads (one per bin range)
                _add(concat.data(), offsets.data(), (uint32_t)bins.size(), bins.data(), _encodingOf((typename Value<TString>::Type *)nullptr));
            }
        }

        // d_bloom_filter.h:150
        bool save(const char *fileName)
        {
            std::lock_guard<std::mutex> lock(*_mtx);
            if (_g) return dyibf_group_save_file(_g, fileName) == DYIBF_OK;
            return _h && dyibf_save_file(_h, fileName) == DYIBF_OK;
        }

        // d_bloom_filter.h:156
        double size_mb()
        {
            // sdsl::size_in_mega_bytes = serialized bytes / 2^20: 8-byte header + words
            return (double)(8 + _prm.n_words * 8) / (1024.0 * 1024.0);
        }

        // d_bloom_filter.h:161 — OR-accumulates into `selected`, never clears
        inline void whichBins(std::vector<bool> & selected, TString const & text, uint16_t const & threshold) const
        {
            if (length(text) < _prm.kmer) return;               // reference: uint16 wrap + out-of-bounds reads
            uint64_t offsets[2] = {0, (uint64_t)length(text)};
            std::vector<uint64_t> row;
            std::lock_guard<std::mutex> lock(*_mtx);             // called from threadsCount std::async tasks (d_mapper.h:338-375)
            _rows(row, _bytes(text), offsets, 1, _encoding(text), 1, 0, threshold);
            for (uint32_t binNo = 0; binNo < _prm.bins; ++binNo)
                if ((row[binNo >> 6] >> (binNo & 63)) & 1)
                    selected[binNo] = true;
        }

        // d_bloom_filter.h:226
        std::vector<bool> whichBins(TString const & text, uint16_t const & threshold) const
        {
            std::vector<bool> selected(_prm.bins, false);
            whichBins(selected, text, threshold);
            return selected;
        }

        uint32_t getNumberOfBins() const { return _prm.bins; }          // :233
        uint8_t  getKmerSize() const { return (uint8_t)_prm.kmer; }     // :238

        // ------------------------------------------------------------------------------------------------
        // batched entries (new): what the patched clasifyLoadedReads / build_filter call
        // ------------------------------------------------------------------------------------------------

        // Many sequences at once, laid out as a ConcatDirect StringSet holds them: `concat` bytes + `limits`.
        // group/derive_rc as in dyibf_query_batch. rows: n_rows x getBinWidth() u64.
        void whichBinsBatch(std::vector<uint64_t> & rows, const uint8_t * concat, const uint64_t * limits, uint32_t n_seqs,
                            int encoding, int group, bool derive_rc, uint16_t threshold) const
        {
            std::lock_guard<std::mutex> lock(*_mtx);
            _rows(rows, concat, limits, n_seqs, encoding, group, derive_rc ? 1 : 0, threshold);
        }

        // Query + compaction: per bin the ascending ids of the reads (pairs) that select it.
        void classify(std::vector<std::vector<uint32_t> > & readIdsPerBin, const uint8_t * concat, const uint64_t * limits,
                      uint32_t n_seqs, int encoding, int group, bool derive_rc, uint16_t threshold) const
        {
            std::vector<uint64_t> binOffsets(_prm.bins + 1, 0);
            uint64_t n_rows = (derive_rc ? 2ull * n_seqs : n_seqs) / group;
            std::vector<uint32_t> ids(std::max<uint64_t>(4 * n_rows, 1024));
            uint64_t total = 0;
            std::lock_guard<std::mutex> lock(*_mtx);
            int rc = _g ? dyibf_group_classify(_g, concat, limits, n_seqs, encoding, group, derive_rc ? 1 : 0, threshold,
                                               binOffsets.data(), ids.data(), ids.size(), &total)
                        : dyibf_classify(_need(), concat, limits, n_seqs, encoding, group, derive_rc ? 1 : 0, threshold,
                                         binOffsets.data(), ids.data(), ids.size(), &total);
            if (rc == DYIBF_ERR_CAPACITY)                        // the lists are still in HBM: fetch, no second pass
            {
                ids.resize(total);
                rc = _g ? dyibf_group_classify_fetch(_g, binOffsets.data(), ids.data(), ids.size(), &total)
                        : dyibf_classify_fetch(_h, binOffsets.data(), ids.data(), ids.size(), &total);
            }
            _check(rc);
            readIdsPerBin.assign(_prm.bins, std::vector<uint32_t>());
            for (uint32_t b = 0; b < _prm.bins; ++b)
                readIdsPerBin[b].assign(ids.begin() + binOffsets[b], ids.begin() + binOffsets[b + 1]);
        }

        // Asynchronous classify for the reference's batching (one clasifyLoadedReads call per 100 000 reads, src/mapper.h:109,
        // while file_prefetched.h:137-205 already loads the next batch): classifyBegin returns at once, up to four batches per
        // GPU in flight; concat / limits must stay valid until classifyEnd, which returns what classify returns.
        dyibf_ticket * classifyBegin(const uint8_t * concat, const uint64_t * limits, uint32_t n_seqs, int encoding, int group,
                                     bool derive_rc, uint16_t threshold) const
        {
            dyibf_ticket * t = nullptr;
            std::lock_guard<std::mutex> lock(*_mtx);
            if (_g && _mode == DYIBF_GROUP_READ_SHARDED)
                _check(dyibf_group_classify_begin(_g, concat, limits, n_seqs, encoding, group, derive_rc ? 1 : 0, threshold, &t));
            else if (_g)
                throw std::runtime_error("dyibf: classifyBegin hands whole batches to replicas; a bin-sharded filter uses classify()");
            else
                _check(dyibf_classify_begin(_need(), concat, limits, n_seqs, encoding, group, derive_rc ? 1 : 0, threshold, &t));
            return t;
        }

        void classifyEnd(dyibf_ticket * ticket, std::vector<std::vector<uint32_t> > & readIdsPerBin, uint32_t n_rows) const
        {
            std::vector<uint64_t> binOffsets(_prm.bins + 1, 0);
            std::vector<uint32_t> ids(std::max<uint64_t>(4ull * n_rows, 1024));
            uint64_t total = 0;
            std::lock_guard<std::mutex> lock(*_mtx);
            int rc = dyibf_classify_end(ticket, binOffsets.data(), ids.data(), ids.size(), &total);
            if (rc == DYIBF_ERR_CAPACITY)                        // the ticket stays in flight: again, with room
            {
                ids.resize(total);
                rc = dyibf_classify_end(ticket, binOffsets.data(), ids.data(), ids.size(), &total);
            }
            _check(rc);
            readIdsPerBin.assign(_prm.bins, std::vector<uint32_t>());
            for (uint32_t b = 0; b < _prm.bins; ++b)
                readIdsPerBin[b].assign(ids.begin() + binOffsets[b], ids.begin() + binOffsets[b + 1]);
        }

        // loadFilteredReads (d_mapper.h:393-443) for one bin of the last classify(derive_rc = true): the read set the
        // bin's mapper receives (bytes + limits, a ConcatDirect layout) and origReadIdMap[bin] as the reference extends it.
        void gatherReads(uint32_t binNo, std::vector<uint8_t> & bytes, std::vector<uint64_t> & limits,
                         std::vector<uint32_t> & ids) const
        {
            uint64_t ns = 0, nb = 0;
            std::lock_guard<std::mutex> lock(*_mtx);
            if (_g) throw std::runtime_error("dyibf: gatherReads needs the batch on one GPU (the batch of a multi-GPU filter is split)");
            int rc = dyibf_gather_reads(_need(), binNo, nullptr, 0, nullptr, nullptr, 0, &ns, &nb);
            if (rc != DYIBF_OK && rc != DYIBF_ERR_CAPACITY) _check(rc);
            bytes.resize(nb); limits.assign(ns + 1, 0); ids.resize(ns);
            if (ns)
                _check(dyibf_gather_reads(_h, binNo, bytes.data(), nb, limits.data(), ids.data(), ns, &ns, &nb));
        }

        // sortedBins (d_mapper.h:843-854) from per-bin list sizes
        static std::vector<uint32_t> sortedBins(std::vector<std::vector<uint32_t> > const & readIdsPerBin)
        {
            std::vector<uint64_t> offs(readIdsPerBin.size() + 1, 0);
            for (size_t b = 0; b < readIdsPerBin.size(); ++b) offs[b + 1] = offs[b] + readIdsPerBin[b].size();
            std::vector<uint32_t> out(readIdsPerBin.size());
            _check(dyibf_sorted_bins(offs.data(), (uint32_t)readIdsPerBin.size(), out.data()));
            return out;
        }

        // DYIBF_OPT_EARLY_EXIT (see dyibf.h): identical bin sets from fewer lookups
        void setEarlyExit(bool on)
        {
            if (_g) _check(dyibf_group_set_option(_g, DYIBF_OPT_EARLY_EXIT, on ? 1 : 0));
            else    _check(dyibf_set_option(_need(), DYIBF_OPT_EARLY_EXIT, on ? 1 : 0));
        }

        void addSequences(const uint8_t * concat, const uint64_t * limits, uint32_t n_seqs, const uint32_t * binIds, int encoding)
        {
            std::lock_guard<std::mutex> lock(*_mtx);
            _add(concat, limits, n_seqs, binIds, encoding);
        }

        uint32_t getBinWidth() const { return _prm.words_per_block; }
        dyibf_filter * handle() const { return _h; }           // single-GPU objects
        dyibf_group * group() const { return _g; }             // multi-GPU objects
        int deviceCount() const { return _nDevices; }

    private:
        static void _check(int rc)
        {
            if (rc != DYIBF_OK)
                throw std::runtime_error(std::string("dyibf: ") + dyibf_last_error());
        }

        static void _exitOnLoadFailure(int rc)
        {
            if (rc == DYIBF_OK) return;
            std::cerr << dyibf_last_error() << std::endl;         // "File ... could not be read." / "Size mismatch: ..."
            exit(1);
        }

        void _cacheParams()
        {
            if (_g) _check(dyibf_group_get_params(_g, &_prm, &_nDevices, &_mode));
            else    _check(dyibf_get_params(_h, &_prm));
        }

        dyibf_filter * _need() const
        {
            if (!_h) throw std::runtime_error("dyibf: this filter object is empty (vector size too small for one hash block)");
            return _h;
        }

        // DYIBF_DEVICES / DYIBF_GROUP_MODE (see the constructors)
        static std::vector<int> _devicesFor(int device)
        {
            if (device >= 0) return std::vector<int>(1, device);
            std::vector<int> devs;
            const char * env = std::getenv("DYIBF_DEVICES");
            if (env && std::string(env) == "all")
            {
                int n = 0;
                _check(dyibf_device_count(&n));
                for (int i = 0; i < n; ++i) devs.push_back(i);
            }
            else if (env && *env)
            {
                std::string s(env);
                size_t at = 0;
                while (at <= s.size())
                {
                    size_t comma = s.find(',', at);
                    if (comma == std::string::npos) comma = s.size();
                    if (comma > at) devs.push_back(std::atoi(s.substr(at, comma - at).c_str()));
                    at = comma + 1;
                }
            }
            if (devs.empty()) devs.push_back(0);
            return devs;
        }
        static int _envMode()
        {
            const char * env = std::getenv("DYIBF_GROUP_MODE");
            return (env && std::string(env) == "bin") ? DYIBF_GROUP_BIN_SHARDED : DYIBF_GROUP_READ_SHARDED;
        }

        void _add(const uint8_t * concat, const uint64_t * limits, uint32_t n_seqs, const uint32_t * binIds, int encoding)
        {
            if (_g) _check(dyibf_group_add_sequences(_g, concat, limits, n_seqs, binIds, encoding));
            else    _check(dyibf_add_sequences(_need(), concat, limits, n_seqs, binIds, encoding, DYIBF_MEM_HOST, nullptr));
        }

        // selection rows of a batch; on a multi-GPU object through the member handles (a compatibility path: the fast
        // multi-GPU entry is classify): replicas -> member 0, bin shards -> every member's column slice side by side
        void _rows(std::vector<uint64_t> & rows, const uint8_t * concat, const uint64_t * limits, uint32_t n_seqs, int encoding,
                   int group, int derive_rc, uint16_t threshold) const
        {
            const uint64_t n_rows = (derive_rc ? 2ull * n_seqs : n_seqs) / group;
            const uint32_t W = _prm.words_per_block;
            rows.assign(n_rows * W, 0);
            if (!_g)
            {
                _check(dyibf_query_batch(_need(), concat, limits, n_seqs, encoding, group, derive_rc, threshold, 0, rows.data(), nullptr,
                                         DYIBF_MEM_HOST, nullptr));
                return;
            }
            const int members = _mode == DYIBF_GROUP_BIN_SHARDED ? _nDevices : 1;
            std::vector<uint64_t> part;
            for (int m = 0; m < members; ++m)
            {
                dyibf_filter * f = nullptr; uint32_t firstBin = 0, nBins = 0;
                _check(dyibf_group_member(_g, m, &f, &firstBin, &nBins));
                dyibf_params p; _check(dyibf_get_params(f, &p));
                part.assign(n_rows * p.words_per_block, 0);
                _check(dyibf_query_batch(f, concat, limits, n_seqs, encoding, group, derive_rc, threshold, 0, part.data(), nullptr,
                                         DYIBF_MEM_HOST, nullptr));
                for (uint64_t r = 0; r < n_rows; ++r)
                    for (uint32_t w = 0; w < p.words_per_block; ++w)
                        rows[r * W + firstBin / 64 + w] = part[r * p.words_per_block + w];
            }
        }

        // SeqAn alphabets are SimpleType<unsigned char, ...>: one byte per character, contiguous in Alloc strings
        static int _encodingOf(Dna5 *)  { return DYIBF_ENC_DNA5; }
        static int _encodingOf(Dna5Q *) { return DYIBF_ENC_DNA5Q; }
        static int _encodingOf(Dna *)   { return DYIBF_ENC_DNA5; }
        static int _encodingOf(char *)  { return DYIBF_ENC_ASCII; }
        template <typename T> static int _encoding(T const &) { return _encodingOf((typename Value<T>::Type *)nullptr); }
        template <typename T> static const uint8_t * _bytes(T const & s)
        {
            static_assert(sizeof(typename Value<T>::Type) == 1, "one byte per character expected");
            return reinterpret_cast<const uint8_t *>(&*begin(s, Standard()));
        }

        dyibf_filter *  _h = nullptr;          // one GPU ...
        dyibf_group *   _g = nullptr;          // ... or several (exactly one of the two is set; neither: the empty object)
        int             _nDevices = 1, _mode = DYIBF_GROUP_READ_SHARDED;
        dyibf_params    _prm;
        // a handle is stream-ordered: calls are serialised here, as the reference's callers use one object from many threads
        std::unique_ptr<std::mutex> _mtx{new std::mutex()};
    };
}

#endif  // APP_YARA_DIS_BLOOM_FILTER_B200_H_
