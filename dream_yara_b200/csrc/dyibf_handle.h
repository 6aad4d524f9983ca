// dyibf_handle.h — the handle behind the C ABI and the helpers the translation units of the host layer share
// (dyibf_api.cu: single-GPU entry points; dyibf_async.cu: ticketed classify; dyibf_group.cu: one process, several GPUs;
// dyibf_io.cu: the .filter loader / writer).  Not installed.
#pragma once
#include "dyibf_internal.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cstdarg>
#include <mutex>
#include <string>
#include <vector>
#include <algorithm>

namespace dyibf {

int fail(int code, const char* fmt, ...);                 // sets the calling thread's dyibf_last_error(), returns code
const std::string& last_error_string();
void set_last_error(const std::string& s);

#define CU(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) \
    return ::dyibf::fail(DYIBF_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); } while (0)

static const uint64_t SEED = 0x90b45d39fb6da1faULL;      // _seedValue, d_bloom_filter.h:329
static const uint64_t META_BITS = 256;                    // filterMetadataSize, d_misc_options.h:40
static const size_t   STAGE_BYTES = 32u << 20;            // pinned staging buffers for word transfers

struct DevBuf {
    void* p = nullptr; size_t cap = 0;
    cudaError_t reserve(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) { cudaError_t e = cudaFree(p); p = nullptr; cap = 0; if (e != cudaSuccess) return e; }
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    void swap(DevBuf& o) { std::swap(p, o.p); std::swap(cap, o.cap); }
};

struct Slot {                     // one stage of the host-mode pipeline
    cudaStream_t st = nullptr;
    DevBuf in_bytes, in_offsets, packed, meta, rows, counts, ctr;
};

enum Kind { K_PACK = 0, K_QUERY, K_COMPACT, K_BUILD, K_CLEAR, K_APPLY, K_SPLIT2, K_COPY /* trace only */, K_N };

struct Pending { cudaEvent_t a, b; int kind; cudaStream_t st; };

struct ResidentBatch { bool valid = false; uint32_t n_fwd = 0, group = 0, encoding = 0; uint64_t first_off = 0; };

// One asynchronous dyibf_classify_begin/_end in flight (dyibf_async.cu): its own stream and buffers
struct Ticket;

} // namespace dyibf

static const int DYIBF_MAX_TICKETS = 8;

struct dyibf_filter {
    dyibf_params prm;
    dyibf::HashParams   hp;
    int          sm_count = 148;
    uint64_t*    d_words = nullptr;       // P*W data words
    uint64_t     data_words = 0;
    bool         owns_words = true;
    dyibf::Slot  slot[2];
    dyibf::DevBuf misc, compact_ws, build_bins, rows_all, stats, dev_err;
    // the read batch and the per-bin lists of the last dyibf_classify stay resident for dyibf_gather_reads
    dyibf::DevBuf batch_bytes, batch_offsets, csr_off, csr_ids, gather_ws, gather_out, gather_ids;
    dyibf::ResidentBatch batch;
    std::vector<uint64_t> h_csr_off;
    uint32_t     query_flags = dyibf::QUERY_EARLY_EXIT;   // QUERY_* (dyibf_set_option); early exit is on by default
    int64_t      bucket_mib = -1;         // DYIBF_OPT_BUILD_BUCKET_MIB: -1 auto, 0 direct atomics only
    int64_t      query_chunk_mib = 0;     // DYIBF_OPT_QUERY_CHUNK_MIB: 0 = auto
    int64_t      build_chunk_mib = 0;     // DYIBF_OPT_BUILD_CHUNK_MIB: bases per host-mode build chunk (0 = auto)
    int64_t      build_mode = 0;          // DYIBF_OPT_BUILD_MODE
    int          last_build_mode = 0;     // what the last dyibf_add_sequences used (DYIBF_BUILD_*)
    dyibf::DevBuf bucket_records, bucket_cursors;
    cudaStream_t sliced_aux = nullptr;    // sliced build: interleave + zero of the filled shadow buffer run here, beside the next scatter
    cudaEvent_t  sliced_filled[2] = {nullptr, nullptr}, sliced_free[2] = {nullptr, nullptr};
    const void*  sliced_clean_ptr = nullptr;   // bucket_records.p when its first sliced_clean_bytes bytes are known to be zero (left so by the last sliced build)
    uint64_t     sliced_clean_bytes = 0;
    void*        stage[2] = {nullptr, nullptr};   // pinned
    cudaEvent_t  stage_ev[2] = {nullptr, nullptr};
    bool         profiling = false;
    cudaEvent_t  trace_base = nullptr;    // DYIBF_TRACE=<file>: per-scope start / end times relative to this event
    std::vector<dyibf::Pending> pending;
    std::vector<cudaEvent_t> ev_pool;
    double       ms[dyibf::K_N] = {0, 0, 0, 0, 0, 0, 0};
    uint64_t     launches[dyibf::K_N] = {0, 0, 0, 0, 0, 0, 0};
    uint64_t     kernel_launches = 0;     // every kernel this handle has launched (dyibf_launch_count)
    dyibf::Ticket* tickets[DYIBF_MAX_TICKETS] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
};

namespace dyibf {

struct DeviceGuard {
    int prev = -1; bool ok = true;
    explicit DeviceGuard(int dev) { if (cudaGetDevice(&prev) != cudaSuccess) prev = -1; ok = cudaSetDevice(dev) == cudaSuccess; }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// profiling + launch counting: bracket `n_kernels` kernel launches on a stream
struct Scope {
    dyibf_filter* f; cudaStream_t st; int kind; cudaEvent_t a = nullptr, b = nullptr;
    static cudaEvent_t get(dyibf_filter* f) {
        if (!f->ev_pool.empty()) { cudaEvent_t e = f->ev_pool.back(); f->ev_pool.pop_back(); return e; }
        cudaEvent_t e = nullptr; cudaEventCreate(&e); return e;
    }
    Scope(dyibf_filter* f_, cudaStream_t st_, int kind_, unsigned n_kernels = 1) : f(f_), st(st_), kind(kind_) {
        f->kernel_launches += n_kernels;
        if (f->profiling) { a = get(f); b = get(f); cudaEventRecord(a, st); }
    }
    ~Scope() { if (a) { cudaEventRecord(b, st); f->pending.push_back({a, b, kind, st}); } }
};

void drain_profile(dyibf_filter* f);

// ---- pieces of dyibf_api.cu the other translation units build on ------------------------------------------------
int init_params(uint32_t bins, uint32_t h, uint32_t k, uint64_t nbits, dyibf_params* prm, HashParams* hp);
int new_handle(uint32_t bins, uint32_t h, uint32_t k, uint64_t nbits, int device, void* external, bool zero, dyibf_filter** out);
int ensure_stage(dyibf_filter* f);
void synth_tail(const dyibf_filter* f, uint64_t first, uint64_t n, uint64_t* out);
int check_query_args(dyibf_filter* f, const uint8_t* seqs, const uint64_t* offsets, uint32_t n_seqs, int encoding,
                     int group, int derive_rc, uint32_t threshold);
int query_on_device(dyibf_filter* f, Slot& sl, const PackSource& src, uint64_t total_bytes, int encoding, uint32_t group,
                    int derive_rc, uint32_t threshold, uint32_t max_seq_len, uint64_t* d_rows, uint16_t* d_counts, cudaStream_t st);
// Host-mode query of rows [row0, row1) of the batch (the whole batch: 0, n_rows); rows land at d_rows_all / h_rows
// relative to row0.
int query_host(dyibf_filter* f, const uint8_t* seqs, const uint64_t* offsets, uint32_t n_seqs, int encoding, int group,
               int derive_rc, uint32_t threshold, uint32_t max_seq_len, uint64_t* h_rows, uint64_t* d_rows_all,
               uint16_t* h_counts, bool keep_batch, uint32_t row0, uint32_t row1);
int compact_on_device(dyibf_filter* f, const uint64_t* d_rows, uint32_t n_rows, uint32_t id_base, uint64_t* d_bin_offsets,
                      uint32_t* d_read_ids, uint64_t capacity, uint64_t* total, cudaStream_t st);
// query + compaction of rows [row0, row1) into the handle's resident CSR (csr_off / csr_ids / h_csr_off), ids = row numbers
// of the whole batch.  The lists stay in HBM; classify_fetch copies them out.
int classify_range(dyibf_filter* f, const uint8_t* seqs, const uint64_t* offsets, uint32_t n_seqs, int encoding, int group,
                   int derive_rc, uint32_t threshold, uint32_t row0, uint32_t row1);
int classify_fetch(dyibf_filter* f, uint64_t* bin_offsets, uint32_t* read_ids, uint64_t capacity, uint64_t* total);
void free_tickets(dyibf_filter* f);
void trim_build_scratch(dyibf_filter* f);

// dyibf_hostcopy.cu: pageable caller memory -> pinned staging -> device
void host_copy_streaming(void* dst, const void* src, size_t n);              // memcpy with non-temporal stores
void host_copy_parallel(void* dst, const void* src, size_t n, int threads);  // ... split over up to 8 threads
bool is_pageable(const void* p);
int staged_h2d(void* d_dst, const void* h_src, size_t n, cudaStream_t st);   // returns once the range has been staged

// bulk word transfer between one table and host memory through the threaded pinned pipeline of dyibf_io.cu
int io_copy_words(dyibf_filter* f, uint64_t* host_words, uint64_t first_block, uint64_t n_blocks, bool to_device);

// ingest_general.cu: FASTQ in SeqAn's full generality (wrapped sequences / qualities, blank lines between records) — the fall-back
// of dyibf_parse_records for texts the four-line tokeniser refuses.  count() leaves the record table on the device for emit().
class FastqGeneral {
public:
    int count(const uint8_t* text, uint64_t n, int iupac, cudaStream_t st, uint64_t* n_records, uint64_t* n_bases, uint64_t* err_word);
    int emit(uint8_t* out_seqs, uint64_t* out_limits, uint64_t limit_base, cudaStream_t st);
    ~FastqGeneral();
private:
    const uint8_t* text_ = nullptr; uint64_t n_ = 0, n_cand_ = 0;
    DevBuf globals_, tiles_, cand_, rec_, jump_, mark_, scan_a_, scan_b_;
};

int parse_records_impl(int device, const char* text, uint64_t n_bytes, int format, int alphabet, uint8_t* out_seqs,
                       uint64_t capacity_bytes, uint64_t* out_limits, uint64_t capacity_seqs, uint64_t limit_base,
                       uint64_t* n_seqs, uint64_t* n_bases, int mem, void* stream, uint64_t* out_rec_pos);

struct FileHeader { uint64_t nbits; uint32_t bins, h, k; uint64_t file_size; };
int read_file_header(FILE* fp, const char* path, FileHeader* fh);

} // namespace dyibf
