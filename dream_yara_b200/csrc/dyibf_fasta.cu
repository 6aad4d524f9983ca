// dyibf_fasta.cu — addFastaFile on the device, for many files at once: dyibf_add_fasta_texts.
//
// Reference code replaced: SeqAnBloomFilter::addFastaFile (src/d_bloom_filter.h:127-147) as build_filter drives it — one FASTA
// file per bin, read record by record with SeqAn's reader into an IupacString, every record of at least k characters hashed
// into the bin (src/d_build_filter.cpp:227-260, one std::async task per range of bins).  On the host that reader is a
// byte-wise state machine (a few hundred MB/s per thread) and becomes the bottleneck of a build once the hashing runs on the
// GPU at 50 G k-mers/s.  Here the files' TEXT goes to the device: the texts of a batch are laid end to end (a newline
// between them), tokenised in ONE pass of the streaming tokeniser (ingest_kernel.cu, DYIBF_ALPHA_IUPAC — same rules: a record
// ends at any '>', white space is dropped, a character outside the IUPAC alphabet is a ParseError), every record gets the bin
// of the file its '>' lies in, and the whole batch is added with one dyibf_add_sequences (so the sliced build sees it).
//
// Texts that cannot be laid end to end without changing their meaning — FASTQ, or FASTA with text before the first '>' (which
// SeqAn skips per file) — are tokenised one by one instead (still on the device) and added together.
#include "dyibf_handle.h"

using namespace dyibf;

namespace dyibf {

namespace {

// bin of every record: the file whose text range holds the record's start character
__global__ void __launch_bounds__(256) fasta_record_bins(const uint64_t* __restrict__ rec_pos, uint64_t n_records,
                                                         const uint64_t* __restrict__ file_begin, const uint32_t* __restrict__ file_bin,
                                                         uint32_t n_files, uint32_t* __restrict__ out_bins)
{
    const uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_records) return;
    const uint64_t p = rec_pos[r];
    uint32_t lo = 0, hi = n_files;                                   // file_begin[lo] <= p < file_begin[hi]
    while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (file_begin[mid] <= p) lo = mid; else hi = mid; }
    out_bins[r] = file_bin[lo];
}

struct Scratch {                                                     // device buffers of one call
    void* p[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    ~Scratch() { for (void* q : p) if (q) cudaFree(q); }
};

} // namespace

} // namespace dyibf

extern "C" {

int dyibf_add_fasta_texts(dyibf_filter* f, const char* const* texts, const uint64_t* sizes, const uint32_t* bins, const int* formats,
                          uint32_t n, uint32_t* bad_index)
{
    if (bad_index) *bad_index = 0xffffffffu;
    if (!f) return fail(DYIBF_ERR_ARG, "filter is NULL");
    if (n == 0) return DYIBF_OK;
    if (!texts || !sizes || !bins) return fail(DYIBF_ERR_ARG, "NULL argument");
    bool joinable = true;
    uint64_t total = 0;
    for (uint32_t i = 0; i < n; ++i) {
        if (bins[i] >= f->prm.bins) return fail(DYIBF_ERR_ARG, "bin %u out of range (%u bins)", bins[i], f->prm.bins);
        if (sizes[i] && !texts[i]) return fail(DYIBF_ERR_ARG, "text %u is NULL", i);
        const int fmt = formats ? formats[i] : DYIBF_FMT_FASTA;
        if (fmt != DYIBF_FMT_FASTA && fmt != DYIBF_FMT_FASTQ) return fail(DYIBF_ERR_ARG, "text %u: bad format", i);
        if (fmt != DYIBF_FMT_FASTA || (sizes[i] && texts[i][0] != '>')) joinable = false;
        total += sizes[i] + 1;
    }
    if (total == n) return DYIBF_OK;                                  // nothing but empty files: no records (SeqAn: atEnd at once)
    DeviceGuard guard(f->prm.device);
    cudaStream_t st = f->slot[0].st;
    Scratch tmp;
    auto upload = [&](void* dst, const void* src, uint64_t len) -> int {
        if (!len) return DYIBF_OK;
        if (len >= (1u << 20) && is_pageable(src)) return staged_h2d(dst, src, len, st);
        CU(cudaMemcpyAsync(dst, src, len, cudaMemcpyHostToDevice, st));
        return DYIBF_OK;
    };
    uint8_t* d_seqs = nullptr; uint64_t* d_limits = nullptr; uint32_t* d_bins = nullptr;
    uint64_t n_records = 0;
    if (joinable) {
        // ---- one text: file i at file_begin[i], a newline behind every file (an id line that the file's end cuts short must
        // not swallow the next file's '>')
        std::vector<uint64_t> file_begin(n + 1);
        uint64_t at = 0;
        for (uint32_t i = 0; i < n; ++i) { file_begin[i] = at; at += sizes[i] + 1; }
        file_begin[n] = at;
        CU(cudaMalloc(&tmp.p[0], total + 32));
        char* d_text = static_cast<char*>(tmp.p[0]);
        CU(cudaMemsetAsync(d_text, '\n', total + 32, st));
        for (uint32_t i = 0; i < n; ++i) { int rc = upload(d_text + file_begin[i], texts[i], sizes[i]); if (rc) return rc; }
        uint64_t ns = 0, nb = 0;
        int rc = parse_records_impl(f->prm.device, d_text, total, DYIBF_FMT_FASTA, DYIBF_ALPHA_IUPAC, nullptr, 0, nullptr, 0, 0, &ns, &nb,
                                    DYIBF_MEM_DEVICE, st, nullptr);
        if (rc == DYIBF_OK && ns) {
            CU(cudaMalloc(&tmp.p[1], nb + 32));
            CU(cudaMalloc(&tmp.p[2], (ns + 1) * 8));
            CU(cudaMalloc(&tmp.p[3], ns * 8));
            d_seqs = static_cast<uint8_t*>(tmp.p[1]); d_limits = static_cast<uint64_t*>(tmp.p[2]);
            uint64_t* d_pos = static_cast<uint64_t*>(tmp.p[3]);
            rc = parse_records_impl(f->prm.device, d_text, total, DYIBF_FMT_FASTA, DYIBF_ALPHA_IUPAC, d_seqs, nb + 16, d_limits, ns, 0, &ns, &nb,
                                    DYIBF_MEM_DEVICE, st, d_pos);
            if (rc == DYIBF_OK) {
                CU(cudaMalloc(&tmp.p[4], (uint64_t)(n + 1) * 8));
                CU(cudaMalloc(&tmp.p[5], (uint64_t)n * 4));
                CU(cudaMalloc(&tmp.p[6], ns * 4));
                d_bins = static_cast<uint32_t*>(tmp.p[6]);
                CU(cudaMemcpyAsync(tmp.p[4], file_begin.data(), (uint64_t)(n + 1) * 8, cudaMemcpyHostToDevice, st));
                CU(cudaMemcpyAsync(tmp.p[5], bins, (uint64_t)n * 4, cudaMemcpyHostToDevice, st));
                fasta_record_bins<<<(unsigned)((ns + 255) / 256), 256, 0, st>>>(d_pos, ns, static_cast<const uint64_t*>(tmp.p[4]),
                                                                              static_cast<const uint32_t*>(tmp.p[5]), n, d_bins);
                CU(cudaGetLastError());
                f->kernel_launches += 11;
            }
        }
        if (rc == DYIBF_ERR_FORMAT && bad_index) {
            // "... at byte <pos>" of the joined text -> the file
            const std::string& msg = last_error_string();
            const size_t k = msg.rfind("at byte ");
            if (k != std::string::npos) {
                const uint64_t pos = strtoull(msg.c_str() + k + 8, nullptr, 10);
                uint32_t i = 0;
                while (i + 1 < n && file_begin[i + 1] <= pos) ++i;
                *bad_index = i;
                std::string m2 = msg.substr(0, k) + "at byte " + std::to_string(pos - file_begin[i]) + " of text " + std::to_string(i);
                set_last_error(m2);
            }
        }
        if (rc) return rc;
        n_records = ns;
    } else {
        // ---- one by one: sizes first, then every file's records behind the previous file's
        std::vector<uint64_t> ns(n, 0), nb(n, 0), t_off(n + 1, 0);
        for (uint32_t i = 0; i < n; ++i) t_off[i + 1] = t_off[i] + ((sizes[i] + 15) & ~15ull);
        CU(cudaMalloc(&tmp.p[0], t_off[n] + 32));
        char* d_text = static_cast<char*>(tmp.p[0]);
        uint64_t recs = 0, bases = 0;
        for (uint32_t i = 0; i < n; ++i) {
            int rc = upload(d_text + t_off[i], texts[i], sizes[i]);
            if (rc == DYIBF_OK)
                rc = parse_records_impl(f->prm.device, d_text + t_off[i], sizes[i], formats ? formats[i] : DYIBF_FMT_FASTA, DYIBF_ALPHA_IUPAC,
                                        nullptr, 0, nullptr, 0, 0, &ns[i], &nb[i], DYIBF_MEM_DEVICE, st, nullptr);
            if (rc) { if (bad_index) *bad_index = i; return rc; }
            recs += ns[i]; bases += nb[i];
        }
        if (recs) {
            CU(cudaMalloc(&tmp.p[1], bases + 32));
            CU(cudaMalloc(&tmp.p[2], (recs + 1) * 8));
            CU(cudaMalloc(&tmp.p[6], recs * 4));
            d_seqs = static_cast<uint8_t*>(tmp.p[1]); d_limits = static_cast<uint64_t*>(tmp.p[2]); d_bins = static_cast<uint32_t*>(tmp.p[6]);
            std::vector<uint32_t> h_bins; h_bins.reserve(recs);
            uint64_t r0 = 0, b0 = 0;
            for (uint32_t i = 0; i < n; ++i) {
                uint64_t a = 0, b = 0;
                int rc = parse_records_impl(f->prm.device, d_text + t_off[i], sizes[i], formats ? formats[i] : DYIBF_FMT_FASTA, DYIBF_ALPHA_IUPAC,
                                            d_seqs + b0, nb[i] + 16, d_limits + r0, ns[i], b0, &a, &b, DYIBF_MEM_DEVICE, st, nullptr);
                if (rc) { if (bad_index) *bad_index = i; return rc; }
                h_bins.insert(h_bins.end(), (size_t)ns[i], bins[i]);
                r0 += ns[i]; b0 += nb[i];
            }
            CU(cudaMemcpyAsync(d_bins, h_bins.data(), recs * 4, cudaMemcpyHostToDevice, st));
            CU(cudaStreamSynchronize(st));                           // h_bins goes out of scope
        }
        n_records = recs;
    }
    if (n_records == 0) return DYIBF_OK;
    if (n_records > 0xffffffffull) return fail(DYIBF_ERR_ARG, "too many records in one batch");
    // the records (ASCII, IUPAC -> Dna5 on the way into the packer: everything but ACGTU is N, alphabet_residue_tabs.h:276-290) go in;
    // records shorter than k are skipped there (d_bloom_filter.h:142)
    int rc = dyibf_add_sequences(f, d_seqs, d_limits, (uint32_t)n_records, d_bins, DYIBF_ENC_ASCII, DYIBF_MEM_DEVICE, st);
    if (rc) return rc;
    CU(cudaStreamSynchronize(st));                                   // the scratch is freed on return
    return DYIBF_OK;
}

} // extern "C"
