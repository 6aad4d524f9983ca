// dyibf_api.cu — the extern "C" layer of libdyibf.so (include/dyibf.h): handle, HBM residency of the filter,
// sdsl-compatible .filter I/O, and the host drivers that feed the kernels.  No CPU implementation of the
// filter lives here: every call either runs the sm_100a kernels or fails.
//
// On-disk format (sdsl::int_vector<1>::serialize, include/sdsl-lite/include/sdsl/int_vector.hpp:1463-1472,
// write_header :534-544; trailer d_bloom_filter.h:255-277):
//   u64 (1<<56 | nbits) ; ceil(nbits/64) little-endian u64 words ; the last four words are [bins, h, k, 0].
// In HBM only the P*W data words are kept; the (all-zero) gap before the trailer and the trailer itself are
// synthesised when the filter is downloaded or saved.
#include "dyibf_handle.h"

#include <sys/syscall.h>
#include <unistd.h>

#include <cmath>

using namespace dyibf;

namespace dyibf {

static thread_local std::string g_err;

int fail(int code, const char* fmt, ...)
{
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    g_err = buf;
    return code;
}
const std::string& last_error_string() { return g_err; }
void set_last_error(const std::string& s) { g_err = s; }

int init_params(uint32_t bins, uint32_t h, uint32_t k, uint64_t nbits, dyibf_params* prm, HashParams* hp)
{
    // _init, d_bloom_filter.h:245-254
    if (bins == 0) return fail(DYIBF_ERR_ARG, "number of bins must be > 0");
    // h == 0: a kmer_direct filter (SeqAnKDXFilter::_init, d_kdx_filter.h:212-217; its trailer has no hash count :234-239)
    if (h > DYIBF_MAX_HASH) return fail(DYIBF_ERR_ARG, "number of hash functions %u not in 0..%d", h, DYIBF_MAX_HASH);
    if (k == 0 || k > 32) return fail(DYIBF_ERR_ARG, "k-mer size %u not in 1..32", k);
    uint32_t W = (uint32_t)std::ceil((float)bins / 64.0f);           // std::ceil((float)_noOfBins / INT_WIDTH)
    if (W > 65535) return fail(DYIBF_ERR_ARG, "%u bins need %u words per block; _binIntWidth is 16 bit", bins, W);
    if ((uint64_t)W * 64 < bins) return fail(DYIBF_ERR_ARG, "bins %u not representable in float ceil", bins);
    uint64_t S = (uint64_t)W * 64;
    if (nbits < META_BITS + S) return fail(DYIBF_ERR_ARG, "vector of %llu bits too small for one block of %llu bits",
                                           (unsigned long long)nbits, (unsigned long long)S);
    uint64_t P = (nbits - META_BITS) / S;
    prm->bins = bins; prm->n_hash = h; prm->kmer = k; prm->words_per_block = W;
    prm->nbits = nbits; prm->n_blocks = P; prm->n_words = (nbits + 63) >> 6;
    memset(hp, 0, sizeof *hp);
    for (uint32_t i = 0; i < h; ++i) hp->pre[i] = (uint64_t)i ^ ((uint64_t)k * SEED);
    hp->P = P; hp->pow2 = (P & (P - 1)) == 0;
    hp->magic = hp->pow2 ? 0 : (uint64_t)((((unsigned __int128)1) << 64) / P);
    hp->h = h ? h : 1; hp->k = k; hp->W = W; hp->bins = bins;
    hp->direct = h == 0; hp->sink = P;
    return DYIBF_OK;
}

int new_handle(uint32_t bins, uint32_t h, uint32_t k, uint64_t nbits, int device, void* external, bool zero,
               dyibf_filter** out)
{
    if (!out) return fail(DYIBF_ERR_ARG, "out is NULL");
    *out = nullptr;
    dyibf_params prm; HashParams hp;
    int rc = init_params(bins, h, k, nbits, &prm, &hp);
    if (rc) return rc;
    int ndev = 0;
    CU(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail(DYIBF_ERR_CUDA, "CUDA device %d not present (%d devices)", device, ndev);
    DeviceGuard guard(device);
    if (!guard.ok) return fail(DYIBF_ERR_CUDA, "cudaSetDevice(%d) failed", device);
    dyibf_filter* f = new dyibf_filter();
    f->prm = prm; f->prm.device = device; f->hp = hp;
    f->data_words = prm.n_blocks * prm.words_per_block;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) f->sm_count = prop.multiProcessorCount;
    // kmer_direct: two extra blocks behind the data — block P stays zero (queries of k-mers outside the table), block P+1
    // takes the builds of such k-mers; neither is ever saved
    const uint64_t extra_words = hp.direct ? 2ull * prm.words_per_block : 0;
    if (external && hp.direct) { delete f; return fail(DYIBF_ERR_ARG, "dyibf_attach: kmer_direct filters own their memory"); }
    if (external) { f->d_words = (uint64_t*)external; f->owns_words = false; }
    else {
        cudaError_t e = cudaMalloc((void**)&f->d_words, (f->data_words + extra_words) * 8);
        if (e != cudaSuccess) { delete f; return fail(e == cudaErrorMemoryAllocation ? DYIBF_ERR_NOMEM : DYIBF_ERR_CUDA,
                                                      "cudaMalloc of %llu filter bytes: %s", (unsigned long long)(f->data_words * 8), cudaGetErrorString(e)); }
        if (e == cudaSuccess && extra_words) e = cudaMemset(f->d_words + f->data_words, 0, extra_words * 8);
        if (e != cudaSuccess) { cudaFree(f->d_words); delete f; return fail(DYIBF_ERR_CUDA, "cudaMemset: %s", cudaGetErrorString(e)); }
        if (zero) { e = cudaMemset(f->d_words, 0, f->data_words * 8); if (e != cudaSuccess) { cudaFree(f->d_words); delete f; return fail(DYIBF_ERR_CUDA, "cudaMemset: %s", cudaGetErrorString(e)); } }
    }
    // sticky device-side error word (DYIBF_DEVERR_*): kernels OR into it, dyibf_device_errors reads it
    {
        cudaError_t e = f->dev_err.reserve(8);
        if (e == cudaSuccess) e = cudaMemset(f->dev_err.p, 0, 8);
        if (e != cudaSuccess) { if (f->owns_words) cudaFree(f->d_words); delete f; return fail(DYIBF_ERR_CUDA, "device error word: %s", cudaGetErrorString(e)); }
        f->hp.err = static_cast<uint32_t*>(f->dev_err.p);
    }
    // The fills above run on the legacy default stream and are asynchronous to the host; every later operation of this
    // handle runs on non-blocking streams that are NOT ordered behind it.  Wait here, once, so that no upload, build or
    // caller-stream kernel can reach words the zero fill has not written yet.
    {
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { if (f->owns_words) cudaFree(f->d_words); f->dev_err.release(); delete f; return fail(DYIBF_ERR_CUDA, "zero fill: %s", cudaGetErrorString(e)); }
    }
    for (int i = 0; i < 2; ++i) {
        if (cudaStreamCreateWithFlags(&f->slot[i].st, cudaStreamNonBlocking) != cudaSuccess ||
            cudaEventCreateWithFlags(&f->stage_ev[i], cudaEventDisableTiming) != cudaSuccess) {
            dyibf_destroy(f); return fail(DYIBF_ERR_CUDA, "stream/event creation failed");
        }
    }
    *out = f;
    return DYIBF_OK;
}

int ensure_stage(dyibf_filter* f)
{
    for (int i = 0; i < 2; ++i)
        if (!f->stage[i]) CU(cudaMallocHost(&f->stage[i], STAGE_BYTES));
    return DYIBF_OK;
}

void drain_profile(dyibf_filter* f)
{
    // debugging aid: DYIBF_TRACE=<file> appends one line per profiled scope — kind, stream, start and end in ms since
    // profiling was switched on — i.e. a timeline of the handle's kernels across its streams
    static const char* trace_path = getenv("DYIBF_TRACE");
    FILE* tr = (trace_path && f->trace_base && !f->pending.empty()) ? fopen(trace_path, "a") : nullptr;
    for (auto& p : f->pending) {
        float ms = 0.f;
        if (cudaEventSynchronize(p.b) == cudaSuccess && cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) {
            f->ms[p.kind] += ms; f->launches[p.kind] += 1;
            float t0 = 0.f;
            if (tr && cudaEventElapsedTime(&t0, f->trace_base, p.a) == cudaSuccess)
                fprintf(tr, "%d %p %.4f %.4f\n", p.kind, (void*)p.st, t0, t0 + ms);
        }
        f->ev_pool.push_back(p.a); f->ev_pool.push_back(p.b);
    }
    f->pending.clear();
    if (tr) fclose(tr);
}

// trailer + gap words for download/save: word index `w` of the sdsl array, w >= data_words
void synth_tail(const dyibf_filter* f, uint64_t first, uint64_t n, uint64_t* out)
{
    memset(out, 0, n * 8);
    const uint64_t m = f->prm.nbits - META_BITS;
    const uint64_t vals[3] = {f->prm.bins, f->prm.n_hash, f->prm.kmer};
    for (int i = 0; i < 3; ++i) {                                     // set_int at bit m + 64 i (_setMetadata :267-277)
        uint64_t bit = m + 64 * i, w = bit >> 6, o = bit & 63;
        if (w >= first && w < first + n) out[w - first] |= vals[i] << o;
        if (o && w + 1 >= first && w + 1 < first + n) out[w + 1 - first] |= vals[i] >> (64 - o);
    }
}

} // namespace dyibf

extern "C" {

const char* dyibf_last_error(void) { return g_err.c_str(); }

int dyibf_device_count(int* count)
{
    if (!count) return fail(DYIBF_ERR_ARG, "count is NULL");
    CU(cudaGetDeviceCount(count));
    return DYIBF_OK;
}

int dyibf_create(uint32_t bins, uint32_t n_hash, uint32_t kmer, uint64_t nbits, int device, dyibf_filter** out)
{
    return new_handle(bins, n_hash, kmer, nbits, device, nullptr, true, out);
}

int dyibf_attach(uint32_t bins, uint32_t n_hash, uint32_t kmer, uint64_t nbits, int device, void* device_words,
                 dyibf_filter** out)
{
    if (!device_words) return fail(DYIBF_ERR_ARG, "device_words is NULL");
    if (((uintptr_t)device_words & 15) != 0) return fail(DYIBF_ERR_ARG, "device_words must be 16-byte aligned");
    return new_handle(bins, n_hash, kmer, nbits, device, device_words, false, out);
}

void dyibf_destroy(dyibf_filter* f)
{
    if (!f) return;
    DeviceGuard guard(f->prm.device);
    cudaDeviceSynchronize();
    drain_profile(f);
    for (auto e : f->ev_pool) cudaEventDestroy(e);
    if (f->trace_base) cudaEventDestroy(f->trace_base);
    if (f->sliced_aux) { cudaStreamSynchronize(f->sliced_aux); cudaStreamDestroy(f->sliced_aux); }
    for (int i = 0; i < 2; ++i) { if (f->sliced_filled[i]) cudaEventDestroy(f->sliced_filled[i]); if (f->sliced_free[i]) cudaEventDestroy(f->sliced_free[i]); }
    for (int i = 0; i < 2; ++i) {
        Slot& s = f->slot[i];
        s.in_bytes.release(); s.in_offsets.release(); s.packed.release(); s.meta.release(); s.rows.release(); s.counts.release(); s.ctr.release();
        if (s.st) cudaStreamDestroy(s.st);
        if (f->stage[i]) cudaFreeHost(f->stage[i]);
        if (f->stage_ev[i]) cudaEventDestroy(f->stage_ev[i]);
    }
    free_tickets(f);
    f->misc.release(); f->compact_ws.release(); f->build_bins.release(); f->rows_all.release(); f->stats.release(); f->dev_err.release();
    f->batch_bytes.release(); f->batch_offsets.release(); f->csr_off.release(); f->csr_ids.release();
    f->gather_ws.release(); f->gather_out.release(); f->gather_ids.release();
    f->bucket_records.release(); f->bucket_cursors.release();
    if (f->owns_words && f->d_words) cudaFree(f->d_words);
    delete f;
}

int dyibf_get_params(const dyibf_filter* f, dyibf_params* out)
{
    if (!f || !out) return fail(DYIBF_ERR_ARG, "NULL argument");
    *out = f->prm;
    return DYIBF_OK;
}

double dyibf_size_mb(const dyibf_filter* f)
{
    // sdsl::size_in_mega_bytes = serialized bytes / 2^20: 8-byte header + words
    return f ? (double)(8 + f->prm.n_words * 8) / (1024.0 * 1024.0) : 0.0;
}

void* dyibf_device_words(dyibf_filter* f) { return f ? f->d_words : nullptr; }

int dyibf_peek_file(const char* path, uint32_t* bins, uint32_t* n_hash, uint32_t* kmer, uint64_t* nbits)
{
    if (!path) return fail(DYIBF_ERR_ARG, "path is NULL");
    FILE* fp = fopen(path, "rb");
    if (!fp) return fail(DYIBF_ERR_IO, "File \"%s\" could not be read.", path);
    FileHeader fh; int rc = read_file_header(fp, path, &fh);
    fclose(fp);
    if (rc) return rc;
    if (bins) *bins = fh.bins; if (n_hash) *n_hash = fh.h; if (kmer) *kmer = fh.k; if (nbits) *nbits = fh.nbits;
    return DYIBF_OK;
}

int dyibf_upload_words(dyibf_filter* f, const uint64_t* host_words, uint64_t first, uint64_t n)
{
    if (!f || (!host_words && n)) return fail(DYIBF_ERR_ARG, "NULL argument");
    if (first + n > f->prm.n_words) return fail(DYIBF_ERR_ARG, "word range [%llu,+%llu) beyond %llu words",
                                                (unsigned long long)first, (unsigned long long)n, (unsigned long long)f->prm.n_words);
    DeviceGuard guard(f->prm.device);
    if (first >= f->data_words) return DYIBF_OK;                      // gap / trailer: synthesised
    n = std::min<uint64_t>(n, f->data_words - first);
    const uint32_t Wb = f->prm.words_per_block;
    if (first % Wb == 0 && n >= (uint64_t)Wb * 1024) {                // bulk: whole blocks through the threaded pinned pipeline
        const uint64_t nb = n / Wb;
        int brc = io_copy_words(f, const_cast<uint64_t*>(host_words), first / Wb, nb, true);
        if (brc) return brc;
        host_words += nb * Wb; first += nb * Wb; n -= nb * Wb;
        if (n == 0) return DYIBF_OK;
    }
    int rc = ensure_stage(f); if (rc) return rc;
    uint64_t left = n * 8, done = 0; int cur = 0;
    while (left) {
        size_t c = (size_t)std::min<uint64_t>(left, STAGE_BYTES);
        CU(cudaEventSynchronize(f->stage_ev[cur]));
        memcpy(f->stage[cur], (const char*)host_words + done, c);
        CU(cudaMemcpyAsync((char*)(f->d_words + first) + done, f->stage[cur], c, cudaMemcpyHostToDevice, f->slot[0].st));
        CU(cudaEventRecord(f->stage_ev[cur], f->slot[0].st));
        done += c; left -= c; cur ^= 1;
    }
    CU(cudaStreamSynchronize(f->slot[0].st));
    return DYIBF_OK;
}

int dyibf_download_words(dyibf_filter* f, uint64_t* host_words, uint64_t first, uint64_t n)
{
    if (!f || (!host_words && n)) return fail(DYIBF_ERR_ARG, "NULL argument");
    if (first + n > f->prm.n_words) return fail(DYIBF_ERR_ARG, "word range [%llu,+%llu) beyond %llu words",
                                                (unsigned long long)first, (unsigned long long)n, (unsigned long long)f->prm.n_words);
    DeviceGuard guard(f->prm.device);
    uint64_t nd = first < f->data_words ? std::min<uint64_t>(n, f->data_words - first) : 0;
    CU(cudaDeviceSynchronize());                                      // builds still in flight on other streams
    const uint32_t Wb = f->prm.words_per_block;
    uint64_t bulk = 0;
    if (nd && first % Wb == 0 && nd >= (uint64_t)Wb * 1024) {         // bulk: whole blocks through the threaded pinned pipeline
        const uint64_t nb = nd / Wb;
        int brc = io_copy_words(f, host_words, first / Wb, nb, false);
        if (brc) return brc;
        bulk = nb * Wb;
    }
    if (nd > bulk) {
        int rc = ensure_stage(f); if (rc) return rc;
        const uint64_t first0 = first; (void)first0;
        host_words += bulk; first += bulk;
        uint64_t left = (nd - bulk) * 8, done = 0;
        // two pinned buffers: the next D2H is in flight while the previous buffer is copied to the caller
        size_t c[2] = {0, 0}; uint64_t at[2] = {0, 0}; int cur = 0;
        auto issue = [&](int b) -> cudaError_t {
            c[b] = (size_t)std::min<uint64_t>(left, STAGE_BYTES); at[b] = done;
            cudaError_t e = cudaMemcpyAsync(f->stage[b], (const char*)(f->d_words + first) + done, c[b],
                                            cudaMemcpyDeviceToHost, f->slot[0].st);
            if (e == cudaSuccess) e = cudaEventRecord(f->stage_ev[b], f->slot[0].st);
            done += c[b]; left -= c[b];
            return e;
        };
        CU(issue(cur));
        while (c[cur]) {
            int nx = cur ^ 1; c[nx] = 0;
            if (left) CU(issue(nx));
            CU(cudaEventSynchronize(f->stage_ev[cur]));
            memcpy((char*)host_words + at[cur], f->stage[cur], c[cur]);
            cur = nx;
        }
    }
    if (n > nd) synth_tail(f, first + nd, n - nd, host_words + nd);
    return DYIBF_OK;
}

} // extern "C"

// ============================================================================================================
// query / compaction / build drivers
// ============================================================================================================
namespace dyibf {

// a build leaves its record buffer (up to tens of GiB) allocated for the next dyibf_add_sequences; a handle that goes on
// to answer queries or is saved gives it back here
void trim_build_scratch(dyibf_filter* f)
{
    if (f->bucket_records.cap > (1ull << 30)) { cudaDeviceSynchronize(); f->bucket_records.release(); f->sliced_clean_ptr = nullptr; }
}

// pack + query of one device-resident batch on stream st, using the workspace of slot sl
int query_on_device(dyibf_filter* f, Slot& sl, const PackSource& src, uint64_t total_bytes, int encoding, uint32_t group,
                    int derive_rc, uint32_t threshold, uint32_t max_seq_len, uint64_t* d_rows, uint16_t* d_counts,
                    cudaStream_t st)
{
    const uint64_t n_in = (uint64_t)src.n_ranges * src.cnt;
    const uint64_t n_total = derive_rc ? 2 * n_in : n_in;
    if (n_total == 0) return DYIBF_OK;
    trim_build_scratch(f);
    if (n_total > 0xffffffffull) return fail(DYIBF_ERR_ARG, "too many sequences in one batch");
    const uint32_t n_rows = (uint32_t)(n_total / group);
    const uint64_t words_one = packed_words_for(total_bytes, n_in);
    CU(sl.packed.reserve((derive_rc ? 2 : 1) * words_one * 8));
    CU(sl.meta.reserve(n_total * sizeof(SeqMeta)));
    CU(sl.ctr.reserve(256));
    {
        Scope sc(f, st, K_PACK);
        CU(launch_pack(src, encoding, derive_rc, (uint64_t*)sl.packed.p, (SeqMeta*)sl.meta.p, words_one, words_one, st));
    }
    uint32_t max_kmers = 0;                                     // 0 = unknown -> 16-bit counters
    if (max_seq_len) max_kmers = max_seq_len >= f->hp.k ? max_seq_len - f->hp.k + 1 : 1;
    {
        Scope sc(f, st, K_QUERY);
        CU(launch_query(f->hp, f->d_words, (const uint64_t*)sl.packed.p, (const SeqMeta*)sl.meta.p, n_rows, group,
                        threshold, max_kmers, d_rows, d_counts, (uint32_t*)sl.ctr.p, f->query_flags,
                        f->profiling ? (unsigned long long*)f->stats.p : nullptr, f->sm_count, st));
    }
    return DYIBF_OK;
}

int check_query_args(dyibf_filter* f, const uint8_t* seqs, const uint64_t* offsets, uint32_t n_seqs, int encoding,
                     int group, int derive_rc, uint32_t threshold)
{
    if (!f) return fail(DYIBF_ERR_ARG, "filter is NULL");
    if (n_seqs && (!offsets || !seqs)) return fail(DYIBF_ERR_ARG, "NULL sequence buffers");
    if (encoding < DYIBF_ENC_DNA5 || encoding > DYIBF_ENC_ASCII) return fail(DYIBF_ERR_ARG, "unknown encoding %d", encoding);
    if (group != 1 && group != 2 && group != 4) return fail(DYIBF_ERR_ARG, "group must be 1, 2 or 4 (got %d)", group);
    if (derive_rc && group == 1) return fail(DYIBF_ERR_ARG, "derive_rc needs group 2 (single-end) or 4 (paired-end)");
    const uint64_t n_total = derive_rc ? 2ull * n_seqs : n_seqs;
    if (n_total % group) return fail(DYIBF_ERR_ARG, "%llu sequences do not divide into groups of %d", (unsigned long long)n_total, group);
    if (threshold > 65535) return fail(DYIBF_ERR_ARG, "threshold %u exceeds uint16", threshold);
    return DYIBF_OK;
}

// Host-mode query: the rows of [0, n_rows) are produced chunk by chunk through two stream slots
// (H2D of chunk c+1 overlaps pack+query of chunk c overlaps D2H of chunk c-1).
// Destination: h_rows (host, per-chunk D2H) and/or d_rows_all (device, whole batch).
int query_host(dyibf_filter* f, const uint8_t* seqs, const uint64_t* offsets, uint32_t n_seqs, int encoding, int group,
               int derive_rc, uint32_t threshold, uint32_t max_seq_len, uint64_t* h_rows, uint64_t* d_rows_all,
               uint16_t* h_counts, bool keep_batch, uint32_t row0, uint32_t row1)
{
    const uint32_t group_in = derive_rc ? group / 2 : group;
    const uint64_t n_total = derive_rc ? 2ull * n_seqs : n_seqs;
    const uint32_t n_rows = (uint32_t)(n_total / group);
    if (row1 > n_rows) row1 = n_rows;
    if (n_rows == 0 || row0 >= row1) return DYIBF_OK;
    const uint32_t n_mine = row1 - row0;                               // rows this call produces
    const uint32_t W = f->prm.words_per_block, k = f->hp.k;
    // Without a caller-supplied bound the longest sequence is found chunk by chunk below, just before the chunk's copies are
    // issued: the scan of one chunk's limits then runs while the device works on the chunk before (a scan of all limits up
    // front costs ~8 ms per 10 M reads with the device idle).
    const bool scan_len = max_seq_len == 0;
    if (offsets[n_seqs] < offsets[0]) return fail(DYIBF_ERR_ARG, "offsets not monotone");
    if (!scan_len && max_seq_len >= k && max_seq_len - k + 1 > 65535)
        return fail(DYIBF_ERR_ARG, "a sequence of %u bases has more than 65535 k-mers (uint16 counts in the reference)", max_seq_len);

    const uint64_t total_bytes = offsets[n_seqs] - offsets[0];
    const uint64_t per_row = total_bytes / n_rows + 8ull * group_in + 8ull * W + 1;
    uint64_t chunk_rows = ((uint64_t)(f->query_chunk_mib > 0 ? f->query_chunk_mib : 48) << 20) / per_row;
    if (f->query_chunk_mib <= 0 && n_mine <= chunk_rows)             // small batch: overlap its own upload with its own query
        chunk_rows = std::max<uint64_t>(((uint64_t)n_mine + 5) / 6, 8192);   // measured on 100 000-read calls: 2/4/6/8/16 cuts
                                                                             // -> 35.1/39.4/41.9/38.6/35.6 M reads/s (one chunk 28.5)
    if (h_counts) chunk_rows = std::min<uint64_t>(chunk_rows, (256ull << 20) / ((uint64_t)group * f->prm.bins * 2 + 1));
    chunk_rows = std::max<uint64_t>(chunk_rows, 1024);
    chunk_rows = std::min<uint64_t>(chunk_rows, n_mine);

    if (keep_batch) {                     // the whole forward set stays in HBM, laid out like the caller's buffer
        f->batch.valid = false;
        CU(f->batch_bytes.reserve(total_bytes + 16));
        CU(f->batch_offsets.reserve((uint64_t)(n_seqs + 1) * 8));
    }
    // pageable caller memory goes through pinned pieces filled by helper threads (dyibf_hostcopy.cu); the driver's own
    // pageable path moves ~10 GB/s and drains the stream first
    const bool big = offsets[n_seqs] - offsets[0] >= (1u << 20);
    const bool pageable = big && is_pageable(seqs + offsets[0]), pageable_off = big && is_pageable(offsets);
    int rc = DYIBF_OK; uint32_t c = 0;
    for (uint64_t a = row0; a < row1 && !rc; a += chunk_rows, ++c) {
        const uint32_t cnt = (uint32_t)std::min<uint64_t>(chunk_rows, row1 - a);
        Slot& sl = f->slot[c & 1];
        PackSource src; memset(&src, 0, sizeof src);
        src.n_ranges = group_in; src.cnt = cnt;
        uint32_t chunk_max = max_seq_len;
        bool inside = true;                                             // every copy below stays inside [offsets[0], offsets[n_seqs]]
        for (uint32_t g = 0; g < group_in; ++g) {
            const uint64_t s0 = (uint64_t)g * n_rows + a;
            inside = inside && offsets[0] <= offsets[s0] && offsets[s0] <= offsets[s0 + cnt] && offsets[s0 + cnt] <= offsets[n_seqs];
        }
        if (!inside) { rc = fail(DYIBF_ERR_ARG, "offsets not monotone"); break; }
        if (scan_len) {
            uint64_t mx = 0;                                            // branch-free (vectorises): a decreasing pair wraps to > 2^32
            for (uint32_t g = 0; g < group_in; ++g) {
                const uint64_t* o = offsets + (uint64_t)g * n_rows + a;
                for (uint32_t s = 0; s < cnt; ++s) mx = std::max(mx, o[s + 1] - o[s]);
            }
            if (mx > 0xffffffffull) { rc = fail(DYIBF_ERR_ARG, "offsets not monotone, or a sequence longer than 2^32"); break; }
            chunk_max = (uint32_t)std::max<uint64_t>(mx, 1);
            if (chunk_max >= k && chunk_max - k + 1 > 65535) {
                rc = fail(DYIBF_ERR_ARG, "a sequence of %u bases has more than 65535 k-mers (uint16 counts in the reference)", chunk_max);
                break;
            }
        }
        uint64_t bytes = 0;
        for (uint32_t g = 0; g < group_in; ++g) {
            const uint64_t s0 = (uint64_t)g * n_rows + a;
            src.first_off[g] = offsets[s0]; src.pack_base[g] = bytes;
            bytes += offsets[s0 + cnt] - offsets[s0];
        }
        if (!keep_batch) {
            CU(sl.in_bytes.reserve(bytes + 16));
            CU(sl.in_offsets.reserve((uint64_t)group_in * (cnt + 1) * 8));
        }
        CU(sl.rows.reserve((uint64_t)cnt * W * 8));
        uint16_t* d_counts = nullptr;
        if (h_counts) { CU(sl.counts.reserve((uint64_t)group * cnt * f->prm.bins * 2)); d_counts = (uint16_t*)sl.counts.p; }
        for (uint32_t g = 0; g < group_in; ++g) {
            const uint64_t s0 = (uint64_t)g * n_rows + a;
            const uint64_t nb = offsets[s0 + cnt] - offsets[s0];
            uint8_t* dst = keep_batch ? (uint8_t*)f->batch_bytes.p + (offsets[s0] - offsets[0])
                                      : (uint8_t*)sl.in_bytes.p + src.pack_base[g];
            if (nb && pageable) { rc = staged_h2d(dst, seqs + offsets[s0], nb, sl.st); if (rc) break; }
            else if (nb) CU(cudaMemcpyAsync(dst, seqs + offsets[s0], nb, cudaMemcpyHostToDevice, sl.st));
            uint64_t* doff = keep_batch ? (uint64_t*)f->batch_offsets.p + s0
                                        : (uint64_t*)sl.in_offsets.p + (uint64_t)g * (cnt + 1);
            if (pageable_off) { rc = staged_h2d(doff, offsets + s0, (uint64_t)(cnt + 1) * 8, sl.st); if (rc) break; }
            else CU(cudaMemcpyAsync(doff, offsets + s0, (uint64_t)(cnt + 1) * 8, cudaMemcpyHostToDevice, sl.st));
            src.bytes[g] = dst - src.first_off[g];
            src.offsets[g] = doff;
        }
        if (rc) break;
        uint64_t* d_rows = d_rows_all ? d_rows_all + (a - row0) * W : (uint64_t*)sl.rows.p;
        rc = query_on_device(f, sl, src, bytes, encoding, (uint32_t)group, derive_rc, threshold, chunk_max, d_rows, d_counts, sl.st);
        if (rc) break;
        if (h_rows) CU(cudaMemcpyAsync(h_rows + (a - row0) * W, d_rows, (uint64_t)cnt * W * 8, cudaMemcpyDeviceToHost, sl.st));
        if (h_counts)
            for (uint32_t g = 0; g < (uint32_t)group; ++g)
                CU(cudaMemcpyAsync(h_counts + ((uint64_t)g * n_mine + (a - row0)) * f->prm.bins, d_counts + (uint64_t)g * cnt * f->prm.bins,
                                   (uint64_t)cnt * f->prm.bins * 2, cudaMemcpyDeviceToHost, sl.st));
    }
    cudaError_t e0 = cudaStreamSynchronize(f->slot[0].st), e1 = cudaStreamSynchronize(f->slot[1].st);
    if (rc) return rc;
    if (e0 != cudaSuccess || e1 != cudaSuccess)
        return fail(DYIBF_ERR_CUDA, "query pipeline: %s", cudaGetErrorString(e0 != cudaSuccess ? e0 : e1));
    if (keep_batch) {
        f->batch.valid = true; f->batch.n_fwd = n_seqs; f->batch.group = (uint32_t)group; f->batch.encoding = (uint32_t)encoding;
        f->batch.first_off = offsets[0];
    }
    return DYIBF_OK;
}

// compaction of device rows; bin_offsets/read_ids are device pointers; *total on the host (synchronises st)
int compact_on_device(dyibf_filter* f, const uint64_t* d_rows, uint32_t n_rows, uint32_t id_base, uint64_t* d_bin_offsets,
                      uint32_t* d_read_ids, uint64_t capacity, uint64_t* total, cudaStream_t st)
{
    const uint32_t bins = f->prm.bins;
    CU(f->compact_ws.reserve(compact_workspace_bytes(bins, n_rows)));
    uint64_t* ws = (uint64_t*)f->compact_ws.p;
    CU(cudaMemsetAsync(ws, 0, compact_workspace_bytes(bins, n_rows), st));
    uint64_t tot = 0;
    {
        Scope sc(f, st, K_COMPACT, 5);                               // count walk, three scan kernels, offsets gather
        CU(launch_compact_count(d_rows, n_rows, bins, ws, st));
        CU(launch_compact_scan(ws, bins, n_rows, d_bin_offsets, st));
    }
    CU(cudaMemcpyAsync(&tot, d_bin_offsets + bins, 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (total) *total = tot;
    if (tot > capacity) return fail(DYIBF_ERR_CAPACITY, "%llu (read,bin) pairs exceed the capacity of %llu",
                                    (unsigned long long)tot, (unsigned long long)capacity);
    {
        Scope sc(f, st, K_COMPACT);
        CU(launch_compact_fill(d_rows, n_rows, bins, ws, d_read_ids, capacity, id_base, st));
    }
    return DYIBF_OK;
}

} // namespace dyibf

extern "C" {

uint16_t dyibf_threshold_rate(float error_rate, uint32_t kmer, uint16_t read_len)
{
    // DisOptions::getThreshold, d_mapper.h:83-92 — float product truncated to uint16, unsigned arithmetic after
    uint16_t maxError = (uint16_t)(error_rate * read_len);
    if (kmer * (1u + maxError) > read_len) return 0;
    return (uint16_t)(read_len - kmer * (1u + maxError) + 1);
}

uint16_t dyibf_threshold(int e_percent, uint32_t kmer, uint16_t read_len)
{
    return dyibf_threshold_rate((float)(e_percent / 100.0), kmer, read_len);   // d_mapper.cpp:340-342
}

int dyibf_query_batch(dyibf_filter* f, const uint8_t* seqs, const uint64_t* offsets, uint32_t n_seqs, int encoding,
                      int group, int derive_rc, uint32_t threshold, uint32_t max_seq_len, uint64_t* out_rows,
                      uint16_t* out_counts, int mem, void* stream)
{
    int rc = check_query_args(f, seqs, offsets, n_seqs, encoding, group, derive_rc, threshold);
    if (rc) return rc;
    if (n_seqs && !out_rows) return fail(DYIBF_ERR_ARG, "out_rows is NULL");
    if (n_seqs == 0) return DYIBF_OK;
    DeviceGuard guard(f->prm.device);
    if (mem == DYIBF_MEM_HOST)
        return query_host(f, seqs, offsets, n_seqs, encoding, group, derive_rc, threshold, max_seq_len, out_rows, nullptr, out_counts, false, 0, 0xffffffffu);
    if (mem != DYIBF_MEM_DEVICE) return fail(DYIBF_ERR_ARG, "unknown memory space %d", mem);
    // device mode: one range, whole batch, caller's stream; total bytes are not known on the host without a
    // read-back, so the two offsets that bound the batch are fetched (16 bytes, synchronises the stream)
    cudaStream_t st = (cudaStream_t)stream;
    uint64_t first = 0, last = 0;
    CU(cudaMemcpyAsync(&first, offsets, 8, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(&last, offsets + n_seqs, 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (last < first) return fail(DYIBF_ERR_ARG, "offsets not monotone");
    if (max_seq_len >= f->hp.k && max_seq_len - f->hp.k + 1 > 65535)
        return fail(DYIBF_ERR_ARG, "a sequence of %u bases has more than 65535 k-mers (uint16 counts in the reference)", max_seq_len);
    PackSource src; memset(&src, 0, sizeof src);
    src.n_ranges = 1; src.cnt = n_seqs; src.bytes[0] = seqs; src.offsets[0] = offsets; src.first_off[0] = first; src.pack_base[0] = 0;
    // bytes[0][offsets[i]] must be base 0 of sequence i: seqs is indexed by the offsets as given
    return query_on_device(f, f->slot[0], src, last - first, encoding, (uint32_t)group, derive_rc, threshold, max_seq_len,
                           out_rows, out_counts, st);
}

int dyibf_compact(dyibf_filter* f, const uint64_t* rows, uint32_t n_rows, uint64_t* bin_offsets, uint32_t* read_ids,
                  uint64_t capacity, uint64_t* total, int mem, void* stream)
{
    if (!f || !bin_offsets || (n_rows && !rows) || (capacity && !read_ids)) return fail(DYIBF_ERR_ARG, "NULL argument");
    DeviceGuard guard(f->prm.device);
    const uint32_t bins = f->prm.bins, W = f->prm.words_per_block;
    if (mem == DYIBF_MEM_DEVICE)
        return compact_on_device(f, rows, n_rows, 0, bin_offsets, read_ids, capacity, total, (cudaStream_t)stream);
    if (mem != DYIBF_MEM_HOST) return fail(DYIBF_ERR_ARG, "unknown memory space %d", mem);
    Slot& sl = f->slot[0];
    CU(sl.rows.reserve((uint64_t)n_rows * W * 8 + 8));
    CU(f->misc.reserve((uint64_t)(bins + 1) * 8));
    if (n_rows) CU(cudaMemcpyAsync(sl.rows.p, rows, (uint64_t)n_rows * W * 8, cudaMemcpyHostToDevice, sl.st));
    // pass 1 + scan give the total; the id buffer is sized from it
    uint64_t tot = 0;
    int rc = compact_on_device(f, (const uint64_t*)sl.rows.p, n_rows, 0, (uint64_t*)f->misc.p, nullptr, 0, &tot, sl.st);
    if (rc != DYIBF_OK && rc != DYIBF_ERR_CAPACITY) return rc;
    CU(cudaMemcpyAsync(bin_offsets, f->misc.p, (uint64_t)(bins + 1) * 8, cudaMemcpyDeviceToHost, sl.st));
    CU(cudaStreamSynchronize(sl.st));
    if (total) *total = tot;
    if (tot > capacity) return fail(DYIBF_ERR_CAPACITY, "%llu (read,bin) pairs exceed the capacity of %llu",
                                    (unsigned long long)tot, (unsigned long long)capacity);
    if (tot == 0) return DYIBF_OK;
    CU(sl.counts.reserve(tot * 4));
    {
        Scope sc(f, sl.st, K_COMPACT);
        CU(launch_compact_fill((const uint64_t*)sl.rows.p, n_rows, bins, (const uint64_t*)f->compact_ws.p, (uint32_t*)sl.counts.p, tot, 0, sl.st));
    }
    CU(cudaMemcpyAsync(read_ids, sl.counts.p, tot * 4, cudaMemcpyDeviceToHost, sl.st));
    CU(cudaStreamSynchronize(sl.st));
    return DYIBF_OK;
}

} // extern "C"

namespace dyibf {

// copy the per-bin lists of the last classify to the host
int classify_fetch(dyibf_filter* f, uint64_t* bin_offsets, uint32_t* read_ids, uint64_t capacity, uint64_t* total)
{
    const uint32_t bins = f->prm.bins;
    if (f->h_csr_off.size() != (size_t)bins + 1) return fail(DYIBF_ERR_ARG, "no dyibf_classify result is resident");
    memcpy(bin_offsets, f->h_csr_off.data(), (size_t)(bins + 1) * 8);
    const uint64_t tot = f->h_csr_off[bins];
    if (total) *total = tot;
    if (tot > capacity) return fail(DYIBF_ERR_CAPACITY, "%llu (read,bin) pairs exceed the capacity of %llu",
                                    (unsigned long long)tot, (unsigned long long)capacity);
    if (tot == 0) return DYIBF_OK;
    cudaStream_t st = f->slot[0].st;
    CU(cudaMemcpyAsync(read_ids, f->csr_ids.p, tot * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return DYIBF_OK;
}

// rows [row0, row1) of the batch: query with the rows kept in HBM, compaction into the handle's resident CSR with the
// row numbers of the WHOLE batch as ids (a device of a read-sharded group filters a contiguous row range)
int classify_range(dyibf_filter* f, const uint8_t* seqs, const uint64_t* offsets, uint32_t n_seqs, int encoding, int group,
                   int derive_rc, uint32_t threshold, uint32_t row0, uint32_t row1)
{
    const uint32_t bins = f->prm.bins, W = f->prm.words_per_block;
    const uint64_t n_total = derive_rc ? 2ull * n_seqs : n_seqs;
    const uint32_t n_rows = (uint32_t)(n_total / group);
    if (row1 > n_rows) row1 = n_rows;
    if (row0 > row1) row0 = row1;
    const uint32_t n_mine = row1 - row0;
    f->h_csr_off.clear(); f->batch.valid = false;
    // rows of the whole range stay in HBM; only the CSR lists cross PCIe.  With derive_rc (and the whole batch on this
    // device) the forward read set stays resident as well, so that dyibf_gather_reads needs no second upload.
    CU(f->rows_all.reserve((uint64_t)n_mine * W * 8 + 8));
    uint64_t* d_rows = (uint64_t*)f->rows_all.p;
    if (n_mine) {
        int rc = query_host(f, seqs, offsets, n_seqs, encoding, group, derive_rc, threshold, 0, nullptr, d_rows, nullptr,
                            derive_rc != 0 && n_mine == n_rows, row0, row1);
        if (rc) return rc;
    }
    Slot& sl = f->slot[0];
    CU(f->csr_off.reserve((uint64_t)(bins + 1) * 8));
    uint64_t tot = 0;
    int rc = compact_on_device(f, d_rows, n_mine, row0, (uint64_t*)f->csr_off.p, nullptr, 0, &tot, sl.st);
    if (rc != DYIBF_OK && rc != DYIBF_ERR_CAPACITY) return rc;
    f->h_csr_off.resize((size_t)bins + 1);
    CU(cudaMemcpyAsync(f->h_csr_off.data(), f->csr_off.p, (uint64_t)(bins + 1) * 8, cudaMemcpyDeviceToHost, sl.st));
    if (tot) {
        CU(f->csr_ids.reserve(tot * 4));
        Scope sc(f, sl.st, K_COMPACT);
        CU(launch_compact_fill(d_rows, n_mine, bins, (const uint64_t*)f->compact_ws.p, (uint32_t*)f->csr_ids.p, tot, row0, sl.st));
    }
    CU(cudaStreamSynchronize(sl.st));
    return DYIBF_OK;
}

} // namespace dyibf

extern "C" {

int dyibf_classify(dyibf_filter* f, const uint8_t* seqs, const uint64_t* offsets, uint32_t n_seqs, int encoding,
                   int group, int derive_rc, uint32_t threshold, uint64_t* bin_offsets, uint32_t* read_ids,
                   uint64_t capacity, uint64_t* total)
{
    int rc = check_query_args(f, seqs, offsets, n_seqs, encoding, group, derive_rc, threshold);
    if (rc) return rc;
    if (!bin_offsets || (capacity && !read_ids)) return fail(DYIBF_ERR_ARG, "NULL argument");
    DeviceGuard guard(f->prm.device);
    rc = classify_range(f, seqs, offsets, n_seqs, encoding, group, derive_rc, threshold, 0, 0xffffffffu);
    if (rc) return rc;
    return classify_fetch(f, bin_offsets, read_ids, capacity, total);
}

int dyibf_classify_fetch(dyibf_filter* f, uint64_t* bin_offsets, uint32_t* read_ids, uint64_t capacity, uint64_t* total)
{
    if (!f || !bin_offsets || (capacity && !read_ids)) return fail(DYIBF_ERR_ARG, "NULL argument");
    DeviceGuard guard(f->prm.device);
    return classify_fetch(f, bin_offsets, read_ids, capacity, total);
}

int dyibf_sorted_bins(const uint64_t* bin_offsets, uint32_t bins, uint32_t* out_bins)
{
    // sortedBins, d_mapper.h:843-854: iota, then std::sort by descending list size — the same (unstable) std::sort
    // call with the same comparator, so ties come out as the reference's do when built with the same C++ library
    if (!bin_offsets || !out_bins) return fail(DYIBF_ERR_ARG, "NULL argument");
    std::vector<uint32_t> idx(bins);
    for (uint32_t i = 0; i < bins; ++i) idx[i] = i;
    std::sort(idx.begin(), idx.end(), [bin_offsets](size_t i1, size_t i2) {
        return bin_offsets[i1 + 1] - bin_offsets[i1] > bin_offsets[i2 + 1] - bin_offsets[i2]; });
    memcpy(out_bins, idx.data(), (size_t)bins * 4);
    return DYIBF_OK;
}

int dyibf_gather_reads(dyibf_filter* f, uint32_t bin, uint8_t* out_seqs, uint64_t capacity_bytes, uint64_t* out_limits,
                       uint32_t* out_ids, uint64_t capacity_seqs, uint64_t* n_seqs_out, uint64_t* n_bytes_out)
{
    if (!f) return fail(DYIBF_ERR_ARG, "filter is NULL");
    if (!f->batch.valid || f->h_csr_off.size() != (size_t)f->prm.bins + 1)
        return fail(DYIBF_ERR_ARG, "no read batch is resident: call dyibf_classify with derive_rc != 0 first");
    if (bin >= f->prm.bins) return fail(DYIBF_ERR_ARG, "bin %u out of range (%u bins)", bin, f->prm.bins);
    DeviceGuard guard(f->prm.device);
    const uint64_t m = f->h_csr_off[bin + 1] - f->h_csr_off[bin];
    const uint64_t n_out = m * f->batch.group;
    if (n_out > 0xffffffffull) return fail(DYIBF_ERR_ARG, "too many sequences");
    if (n_seqs_out) *n_seqs_out = n_out;
    if (n_bytes_out) *n_bytes_out = 0;
    if (n_out == 0) { if (out_limits) out_limits[0] = 0; return DYIBF_OK; }
    cudaStream_t st = f->slot[0].st;
    GatherArgs a;
    a.bytes = (const uint8_t*)f->batch_bytes.p - f->batch.first_off;
    a.offsets = (const uint64_t*)f->batch_offsets.p;
    a.ids = (const uint32_t*)f->csr_ids.p + f->h_csr_off[bin];
    a.m = (uint32_t)m; a.group = f->batch.group; a.n_fwd = f->batch.n_fwd; a.encoding = f->batch.encoding;
    CU(f->gather_ws.reserve(scan_workspace_elems(n_out) * 8));
    CU(f->gather_ids.reserve(n_out * 4));
    uint64_t* lens = (uint64_t*)f->gather_ws.p;
    CU(launch_gather_plan(a, lens, (uint32_t*)f->gather_ids.p, st));
    CU(launch_scan_u64(lens, n_out, st));
    uint64_t n_bytes = 0;
    CU(cudaMemcpyAsync(&n_bytes, lens + n_out, 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (n_bytes_out) *n_bytes_out = n_bytes;
    if (n_out > capacity_seqs || n_bytes > capacity_bytes)
        return fail(DYIBF_ERR_CAPACITY, "bin %u needs %llu sequences / %llu bytes", bin, (unsigned long long)n_out, (unsigned long long)n_bytes);
    if (!out_limits || !out_ids || (n_bytes && !out_seqs)) return fail(DYIBF_ERR_ARG, "NULL output");
    CU(f->gather_out.reserve(n_bytes + 16));
    CU(launch_gather_copy(a, lens, (uint8_t*)f->gather_out.p, st));
    if (n_bytes) CU(cudaMemcpyAsync(out_seqs, f->gather_out.p, n_bytes, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(out_limits, lens, (n_out + 1) * 8, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(out_ids, f->gather_ids.p, n_out * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return DYIBF_OK;
}

} // extern "C"

namespace dyibf {
/* ---- ingest: SeqAn's readRecords on the device (ingest_kernel.cu) ---------------------------------------- */
// out_rec_pos (device memory, capacity_seqs entries, MEM_DEVICE only, may be NULL): the text offset of every record's start
// character — what dyibf_add_fasta_texts needs to tell which file of a concatenation a record came from
int parse_records_impl(int device, const char* text, uint64_t n_bytes, int format, int alphabet, uint8_t* out_seqs,
                       uint64_t capacity_bytes, uint64_t* out_limits, uint64_t capacity_seqs, uint64_t limit_base,
                       uint64_t* n_seqs, uint64_t* n_bases, int mem, void* stream, uint64_t* out_rec_pos)
{
    if (!n_seqs || !n_bases) return fail(DYIBF_ERR_ARG, "dyibf_parse_records: NULL size output");
    if ((format != DYIBF_FMT_FASTA && format != DYIBF_FMT_FASTQ) || (alphabet != DYIBF_ALPHA_DNA5 && alphabet != DYIBF_ALPHA_IUPAC) ||
        (mem != DYIBF_MEM_HOST && mem != DYIBF_MEM_DEVICE) || (n_bytes && !text))
        return fail(DYIBF_ERR_ARG, "dyibf_parse_records: bad format / alphabet / mem / text");
    *n_seqs = 0; *n_bases = 0;
    const bool sizing = !out_seqs && !out_limits;
    DeviceGuard guard(device);
    if (!guard.ok) return fail(DYIBF_ERR_CUDA, "cudaSetDevice(%d) failed", device);
    cudaStream_t st = (cudaStream_t)stream;
    struct Tmp {                                              // scratch of this call
        void* p[4] = {nullptr, nullptr, nullptr, nullptr};
        ~Tmp() { for (void* q : p) if (q) cudaFree(q); }
    } tmp;
    if (n_bytes == 0) {                                       // readRecords on an empty file: no records
        if (sizing) return DYIBF_OK;
        if (!out_limits) return fail(DYIBF_ERR_ARG, "dyibf_parse_records: NULL out_limits");
        if (mem == DYIBF_MEM_HOST) { out_limits[0] = limit_base; return DYIBF_OK; }
        CU(cudaMemcpyAsync(out_limits, &limit_base, 8, cudaMemcpyHostToDevice, st));
        CU(cudaStreamSynchronize(st));
        return DYIBF_OK;
    }
    const uint8_t* d_text = reinterpret_cast<const uint8_t*>(text);
    if (mem == DYIBF_MEM_HOST) {
        CU(cudaMalloc(&tmp.p[0], n_bytes));
        CU(cudaMemcpyAsync(tmp.p[0], text, n_bytes, cudaMemcpyHostToDevice, st));
        d_text = static_cast<const uint8_t*>(tmp.p[0]);
    }
    // the tile tables are kept per device between calls (a cudaMalloc / cudaFree pair costs as much as a pass over the text)
    static std::mutex ws_mutex;
    static DevBuf ws_cache[64];
    std::lock_guard<std::mutex> ws_lock(ws_mutex);
    DevBuf& wsb = ws_cache[device & 63];
    CU(wsb.reserve(ingest_workspace_bytes(n_bytes)));
    uint64_t* ws = static_cast<uint64_t*>(wsb.p);
    const int iupac = alphabet == DYIBF_ALPHA_IUPAC;
    CU(launch_ingest_count(d_text, n_bytes, format, iupac, ws, st));
    uint64_t tot[6];
    CU(ingest_read_totals(ws, n_bytes, tot, st));
    static const char* const what[] = {"", "character outside the sequence alphabet", "not a FASTQ record",
                                       "quality string and sequence differ in length", "truncated FASTQ record",
                                       "a FASTQ record that starts right behind the previous qualities, without white space (not supported by the device reader)"};
    auto format_error = [&](uint64_t word) {
        const uint64_t pos = word >> 8; const unsigned code = (unsigned)(word & 0xff);
        return fail(DYIBF_ERR_FORMAT, "dyibf_parse_records: %s at byte %llu", code < 6 ? what[code] : "parse error", (unsigned long long)pos);
    };
    // FASTQ the four-line tokeniser refuses (wrapped lines, blank lines between records, or a real error): SeqAn's sequential
    // rules through the speculative parser of ingest_general.cu, which either yields the records or the reference's error
    auto general_fastq = [&]() -> int {
        FastqGeneral fq;
        uint64_t ns = 0, nb = 0, errw = ~0ull;
        int rc = fq.count(d_text, n_bytes, iupac, st, &ns, &nb, &errw);
        if (rc) return rc;
        if (errw != ~0ull) return format_error(errw);
        *n_seqs = ns; *n_bases = nb;
        if (sizing) return DYIBF_OK;
        if (!out_limits || (nb && !out_seqs)) return fail(DYIBF_ERR_ARG, "dyibf_parse_records: NULL output");
        if (ns > capacity_seqs || nb > capacity_bytes)
            return fail(DYIBF_ERR_CAPACITY, "dyibf_parse_records: %llu records / %llu sequence characters do not fit the output buffers",
                        (unsigned long long)ns, (unsigned long long)nb);
        uint8_t* d_seqs = out_seqs; uint64_t* d_limits = out_limits;
        if (mem == DYIBF_MEM_HOST) {
            CU(cudaMalloc(&tmp.p[2], nb + 16));
            CU(cudaMalloc(&tmp.p[3], (ns + 1) * 8));
            d_seqs = static_cast<uint8_t*>(tmp.p[2]); d_limits = static_cast<uint64_t*>(tmp.p[3]);
        }
        rc = fq.emit(d_seqs, d_limits, limit_base, st);
        if (rc) return rc;
        if (mem == DYIBF_MEM_HOST) {
            if (nb) CU(cudaMemcpyAsync(out_seqs, d_seqs, nb, cudaMemcpyDeviceToHost, st));
            CU(cudaMemcpyAsync(out_limits, d_limits, (ns + 1) * 8, cudaMemcpyDeviceToHost, st));
        }
        CU(cudaStreamSynchronize(st));                                // the record table is freed when fq goes out of scope
        return DYIBF_OK;
    };
    auto needs_general = [&](uint64_t word) {
        const unsigned code = (unsigned)(word & 0xff);
        return format == DYIBF_FMT_FASTQ && word != ~0ull && (code == 2 || code == 3 || code == 4);
    };
    if (format == DYIBF_FMT_FASTA && tot[3] == ~0ull)
        return fail(DYIBF_ERR_FORMAT, "dyibf_parse_records: no '>' in a non-empty FASTA text (SeqAn: UnexpectedEnd)");
    if (format == DYIBF_FMT_FASTQ && tot[4] == 0)
        return fail(DYIBF_ERR_FORMAT, "dyibf_parse_records: no '@' in a non-empty FASTQ text (SeqAn: UnexpectedEnd)");
    if (needs_general(tot[5])) return general_fastq();
    if (tot[5] != ~0ull) return format_error(tot[5]);
    *n_seqs = tot[0]; *n_bases = tot[1];
    // a sizing call runs the emit pass into scratch as well: that pass is the one that validates (alphabet, FASTQ layout / lengths)
    if (!sizing && (!out_limits || (tot[1] && !out_seqs))) return fail(DYIBF_ERR_ARG, "dyibf_parse_records: NULL output");
    if (!sizing && (tot[0] > capacity_seqs || tot[1] > capacity_bytes))
        return fail(DYIBF_ERR_CAPACITY, "dyibf_parse_records: %llu records / %llu sequence characters do not fit the output buffers",
                    (unsigned long long)tot[0], (unsigned long long)tot[1]);
    uint8_t* d_seqs = out_seqs; uint64_t* d_limits = out_limits;
    if (mem == DYIBF_MEM_HOST || sizing) {
        CU(cudaMalloc(&tmp.p[2], tot[1] + 16));
        CU(cudaMalloc(&tmp.p[3], (tot[0] + 1) * 8));
        d_seqs = static_cast<uint8_t*>(tmp.p[2]); d_limits = static_cast<uint64_t*>(tmp.p[3]);
    }
    CU(launch_ingest_emit(d_text, n_bytes, format, iupac, ws, d_seqs, d_limits, limit_base, tot[0], tot[1], st,
                          mem == DYIBF_MEM_DEVICE && !sizing ? out_rec_pos : nullptr));
    CU(ingest_read_totals(ws, n_bytes, tot, st));
    if (needs_general(tot[5])) {                                     // a length mismatch only the last pass sees
        *n_seqs = 0; *n_bases = 0;
        if (tmp.p[2]) { cudaFree(tmp.p[2]); tmp.p[2] = nullptr; }
        if (tmp.p[3]) { cudaFree(tmp.p[3]); tmp.p[3] = nullptr; }
        return general_fastq();
    }
    if (tot[5] != ~0ull) { *n_seqs = 0; *n_bases = 0; return format_error(tot[5]); }
    if (mem == DYIBF_MEM_HOST && !sizing) {
        if (tot[1]) CU(cudaMemcpyAsync(out_seqs, d_seqs, tot[1], cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(out_limits, d_limits, (tot[0] + 1) * 8, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
    }
    return DYIBF_OK;
}
} // namespace dyibf

extern "C" {

int dyibf_parse_records(int device, const char* text, uint64_t n_bytes, int format, int alphabet, uint8_t* out_seqs,
                        uint64_t capacity_bytes, uint64_t* out_limits, uint64_t capacity_seqs, uint64_t limit_base,
                        uint64_t* n_seqs, uint64_t* n_bases, int mem, void* stream)
{
    return parse_records_impl(device, text, n_bytes, format, alphabet, out_seqs, capacity_bytes, out_limits, capacity_seqs, limit_base,
                              n_seqs, n_bases, mem, stream, nullptr);
}

} // extern "C"

// ============================================================================================================
// build / clear / instrumentation
// ============================================================================================================
namespace {

// One dyibf_add_sequences call's use of the bucketed build (build_bucket.cu).  Update records accumulate in HBM across
// the call's chunks: level 1 of the split fills the (smaller) super-bucket buffer, level 2 empties it into the final
// per-region buckets, and the final buckets are OR-ed into the table (one sweep over the table) when they are nearly
// full and at the end of the call.
// the build's view of the hash parameters: k-mers outside a kmer_direct table OR into the scratch block
HashParams build_params(const dyibf_filter* f) { HashParams hp = f->hp; if (hp.direct) hp.sink = hp.P + 1; return hp; }

struct BucketRun {
    bool active = false;
    BucketPlan plan;
    uint64_t pending1 = 0, limit1 = 0;             // level-1 buffer: upper bound of the records in it / what it may hold
    uint64_t pending = 0, limit = 0;               // final buckets (pending + pending1 <= limit at all times)
};

int buckets_begin(dyibf_filter* f, uint64_t expected_records, BucketRun* br)
{
    br->active = false;
    if (f->bucket_mib == 0 || f->hp.direct) return DYIBF_OK;         // kmer_direct tables are built with the direct kernel
    if (f->data_words * 8 <= (256ull << 20)) return DYIBF_OK;       // table (nearly) fits the L2: direct atomics are L2 hits
    if (expected_records < (32ull << 20)) return DYIBF_OK;          // too few updates per region sector to pay for a sweep
    uint64_t budget = f->bucket_mib > 0 ? (uint64_t)f->bucket_mib << 20 : 64ull << 30;
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) return DYIBF_OK;
    free_b += f->bucket_records.cap;
    if (f->bucket_mib < 0 && budget > free_b / 2) budget = free_b / 2;
    if (!bucket_plan(f->data_words, budget, &br->plan)) return DYIBF_OK;
    // the whole call in one sweep if the budget allows (slack: see the limits below)
    const uint64_t per_rec = br->plan.bytes_per_record_x4;          // bytes per record of final capacity, times 4
    const uint64_t want = expected_records * per_rec / 4 + expected_records * per_rec / 32 + (256ull << 20);
    if (budget > want && !bucket_plan(f->data_words, want, &br->plan)) return DYIBF_OK;
    f->sliced_clean_ptr = nullptr;                                   // the buffer is about to hold records
    if (f->bucket_records.reserve(br->plan.bytes_final + br->plan.bytes_level1) != cudaSuccess) {
        cudaGetLastError();                                          // not enough memory: direct atomics
        return DYIBF_OK;
    }
    CU(f->bucket_cursors.reserve(bucket_cursor_slots(br->plan) * 8));
    // zeroed on a stream of the handle and waited for: the split kernels run on non-blocking streams (and, in device
    // mode, on the caller's), none of which is ordered behind the legacy default stream a plain cudaMemset uses
    CU(cudaMemsetAsync(f->bucket_cursors.p, 0, bucket_cursor_slots(br->plan) * 8, f->slot[0].st));
    CU(cudaStreamSynchronize(f->slot[0].st));
    // 10 % slack per bucket: the hash is uniform; every writing warp may leave a chunk of padding lines per bucket
    br->limit = (uint64_t)((double)br->plan.n_regions * (double)br->plan.cap / 1.10);
    br->limit1 = br->plan.n_super > 1 ? (uint64_t)((double)br->plan.n_super * (double)br->plan.cap1 / 1.10) : br->limit;
    br->pending = br->pending1 = 0;
    br->active = true;
    return DYIBF_OK;
}

// level 2: the super-bucket buffer -> the final buckets (two-level plans; stream-ordered on st)
int buckets_flush_level1(dyibf_filter* f, BucketRun* br, cudaStream_t st)
{
    if (!br->active || br->pending1 == 0) return DYIBF_OK;
    if (br->plan.n_super > 1) {
        Scope sc(f, st, K_SPLIT2);
        CU(launch_bucket_split2(f->d_words, br->plan, f->bucket_records.p, (unsigned long long*)f->bucket_cursors.p, br->pending1, st));
    }
    br->pending += br->pending1;
    br->pending1 = 0;
    return DYIBF_OK;
}

int buckets_apply(dyibf_filter* f, BucketRun* br, cudaStream_t st)
{
    if (!br->active) return DYIBF_OK;
    int rc = buckets_flush_level1(f, br, st);
    if (rc) return rc;
    if (br->pending == 0) return DYIBF_OK;
    Scope sc(f, st, K_APPLY);
    CU(launch_bucket_apply(f->d_words, f->data_words, br->plan, f->bucket_records.p, (unsigned long long*)f->bucket_cursors.p,
                           br->pending, st));
    br->pending = 0;
    return DYIBF_OK;
}

// Split the updates of packed words [0, words) of one pack buffer into the buckets, in word ranges that fit the buffers;
// makes room first (level 2, apply) where needed.  sync_all: the call uses both stream slots (host mode), so a buffer is
// emptied only after everything in flight on EITHER stream, and later splits on either stream wait for the emptying —
// ordered with events between the two streams (join_streams / fork_streams), never by blocking the host, so the host keeps
// enqueueing chunks and the uploads keep running.  [round 1 blocked the host here: 17.6 instead of 22 G k-mers/s]
// in flight is waited for before a buffer is emptied; otherwise the work is stream-ordered on st.
// st waits for everything enqueued so far on the other slot's stream
int join_streams(dyibf_filter* f, cudaStream_t st)
{
    cudaStream_t ot = f->slot[0].st == st ? f->slot[1].st : f->slot[0].st;
    CU(cudaEventRecord(f->stage_ev[0], ot));
    CU(cudaStreamWaitEvent(st, f->stage_ev[0], 0));
    return DYIBF_OK;
}
// the other slot's stream waits for everything enqueued so far on st
int fork_streams(dyibf_filter* f, cudaStream_t st)
{
    cudaStream_t ot = f->slot[0].st == st ? f->slot[1].st : f->slot[0].st;
    CU(cudaEventRecord(f->stage_ev[1], st));
    CU(cudaStreamWaitEvent(ot, f->stage_ev[1], 0));
    return DYIBF_OK;
}

int buckets_scatter(dyibf_filter* f, BucketRun* br, const uint64_t* packed, const SeqMeta* meta, const uint32_t* bin_ids,
                    uint32_t n_seqs, uint64_t words, uint32_t counter_slot, bool sync_all, cudaStream_t st)
{
    const uint64_t per_word = 32ull * f->hp.h;                      // upper bound of the records one packed word yields
    const bool two = br->plan.n_super > 1;
    uint64_t max_words = (two ? br->limit1 : br->limit) / per_word;
    max_words -= max_words % BUILD_SCATTER_ALIGN;
    if (max_words < BUILD_SCATTER_ALIGN) max_words = BUILD_SCATTER_ALIGN;
    for (uint64_t w0 = 0; w0 < words;) {
        uint64_t room = ((two ? br->limit1 - br->pending1 : br->limit - br->pending) / per_word);
        room -= room % BUILD_SCATTER_ALIGN;
        const uint64_t left = words - w0;
        if (room < left && room < max_words) {
            // not enough room for a full-sized range: empty the level-1 buffer, and the final buckets if they could not
            // take what level 1 may still hold plus a full range
            if (sync_all) { int jrc = join_streams(f, st); if (jrc) return jrc; }
            int rc = buckets_flush_level1(f, br, st);
            if (rc) return rc;
            const uint64_t next = std::min(left, max_words) * per_word;
            if (br->pending && br->pending + next > br->limit) { rc = buckets_apply(f, br, st); if (rc) return rc; }
            if (sync_all) { int frc = fork_streams(f, st); if (frc) return frc; }
            continue;
        }
        const uint64_t n = std::min(left, std::min(room, max_words));
        const uint64_t add = n * per_word;
        if (two && br->pending + br->pending1 + add > br->limit) {   // the final buckets must be able to take level 1's content
            if (sync_all) { int jrc = join_streams(f, st); if (jrc) return jrc; }
            int rc = buckets_apply(f, br, st);
            if (rc) return rc;
            if (sync_all) { int frc = fork_streams(f, st); if (frc) return frc; }
        }
        { Scope sc(f, st, K_BUILD);
          CU(launch_bucket_split(f->hp, f->d_words, packed, meta, bin_ids, n_seqs, w0, w0 + n, br->plan, f->bucket_records.p,
                                 (unsigned long long*)f->bucket_cursors.p, counter_slot, st)); }
        if (two) br->pending1 += add; else br->pending += add;
        w0 += n;
    }
    return DYIBF_OK;
}

void buckets_end(dyibf_filter* f, BucketRun* br)
{
    // The record buffer stays allocated for the next dyibf_add_sequences call: a cudaMalloc / cudaFree pair of tens of GiB
    // costs more than the build of a 4 Gbp batch.  trim_build_scratch (first query after a build, dyibf_save_file) frees it.
    (void)f;
    br->active = false;
}

// ---- the bit-sliced build (build_sliced.cu): per-bin bit vectors in the L2, then interleaved --------------------------
// One shadow buffer when the plan has one group (the usual case: all W columns); with several groups, two buffers and an
// auxiliary stream: when the batch moves on to the next column group, the filled buffer is handed to the auxiliary stream
// (interleave into the table, then zero it again) while the scatter of the next group fills the other one.  The buffers are
// left zeroed, so the next call on the handle starts without a memset.
struct SlicedRun {
    bool active = false;
    SlicedPlan plan;
    int cur = -1;                                   // group whose updates the current buffer holds (-1: none)
    int buf = 0;                                    // shadow buffer being filled
    bool zeroed[2] = {false, false};                // the buffer has been zeroed in this call (by the caller's stream or the aux stream)
    bool busy[2] = {false, false};                  // the aux stream still works on the buffer: wait for free_ev before filling it
};

// may this call use the sliced build at all (mode, table size, call size) — and with which plan
bool sliced_candidate(dyibf_filter* f, uint64_t expected_records, SlicedPlan* plan)
{
    const int64_t mode = f->build_mode;
    if (mode == DYIBF_BUILD_DIRECT || mode == DYIBF_BUILD_BUCKETED || f->hp.direct) return false;
    if (mode == DYIBF_BUILD_AUTO) {
        if (f->data_words * 8 <= (256ull << 20)) return false;     // table (nearly) fits the L2: direct atomics are L2 hits
        // too few updates to pay for the zero / interleave passes (a group's passes move ~4 x its share of the table: the
        // break-even against direct atomics at 21 G updates/s is ~70 M updates per group; calls between 32 M and 256 M updates
        // go to the bucketed build)
        if (expected_records < (256ull << 20)) return false;
    }
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) return false;
    free_b += f->bucket_records.cap;
    const uint64_t budget = std::min<uint64_t>(free_b / 2, 17ull << 30);
    return sliced_plan(f->hp, budget, plan);
}

int sliced_resources(dyibf_filter* f)
{
    if (f->sliced_aux) return DYIBF_OK;
    CU(cudaStreamCreateWithFlags(&f->sliced_aux, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
        CU(cudaEventCreateWithFlags(&f->sliced_filled[i], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&f->sliced_free[i], cudaEventDisableTiming));
    }
    return DYIBF_OK;
}

// bin changes along the batch decide whether the vectors of the bins in flight fit the L2
int sliced_begin(dyibf_filter* f, const SlicedPlan& plan, uint64_t total_bases, uint64_t switches, SlicedRun* sr)
{
    *sr = SlicedRun(); sr->plan = plan;
    // the scatter kernel's window (bases its resident warps span) is cut to about 1.5 average runs of one bin, so that two
    // or three bins' vectors are being filled at any time; the smallest useful grid still spans SLICED_MIN_BLOCKS * 8192 bases
    const uint64_t avg_run = std::max<uint64_t>(total_bases / (switches + 1), 1);
    const uint64_t want_blocks = (avg_run + avg_run / 2) / SLICED_BLOCK_BASES;
    sr->plan.scatter_blocks = (uint32_t)std::min<uint64_t>(std::max<uint64_t>(want_blocks, SLICED_MIN_BLOCKS), SLICED_MAX_BLOCKS);
    if (f->build_mode == DYIBF_BUILD_AUTO) {
        const uint64_t bins_in_flight = sr->plan.scatter_blocks * SLICED_BLOCK_BASES / avg_run + 2;
        if (bins_in_flight * plan.pw * 4 > (80ull << 20)) return DYIBF_OK;   // not clustered by bin: bucketed / direct
    }
    if (f->bucket_records.reserve(plan.shadow_bytes * plan.n_buffers) != cudaSuccess) {
        cudaGetLastError();                                          // not enough memory: the other builds
        return DYIBF_OK;
    }
    CU(f->bucket_cursors.reserve(256));
    int rc = sliced_resources(f);
    if (rc) return rc;
    // the last sliced build left its buffers zeroed (the auxiliary stream zeroes a buffer after interleaving it): no memset then
    if (f->sliced_clean_ptr == f->bucket_records.p && f->sliced_clean_bytes >= plan.shadow_bytes * plan.n_buffers)
        sr->zeroed[0] = sr->zeroed[1] = true;
    f->sliced_clean_ptr = nullptr;                                   // dirty from here on, until sliced_finish_group says otherwise
    sr->active = true;
    return DYIBF_OK;
}

inline void* sliced_buffer(dyibf_filter* f, const SlicedRun* sr, int i) { return static_cast<char*>(f->bucket_records.p) + (size_t)i * sr->plan.shadow_bytes; }

// the filled buffer goes to the auxiliary stream: interleave into the table, zero again.  `st` has the group's scatters.
int sliced_hand_over(dyibf_filter* f, SlicedRun* sr, cudaStream_t st)
{
    const int b = sr->buf;
    CU(cudaEventRecord(f->sliced_filled[b], st));
    CU(cudaStreamWaitEvent(f->sliced_aux, f->sliced_filled[b], 0));
    {
        Scope sc(f, f->sliced_aux, K_APPLY);
        CU(launch_sliced_interleave(f->hp, f->d_words, sliced_buffer(f, sr, b), sr->plan, (uint32_t)sr->cur, f->sliced_aux));
    }
    CU(launch_sliced_zero(sliced_buffer(f, sr, b), sr->plan, f->sliced_aux));
    CU(cudaEventRecord(f->sliced_free[b], f->sliced_aux));
    sr->busy[b] = true; sr->zeroed[b] = true;
    sr->cur = -1;
    return DYIBF_OK;
}

// end of the call: the last group goes into the table and `st` waits for everything the auxiliary stream still does
int sliced_finish_group(dyibf_filter* f, SlicedRun* sr, cudaStream_t st)
{
    if (!sr->active) return DYIBF_OK;
    if (sr->cur >= 0) { int rc = sliced_hand_over(f, sr, st); if (rc) return rc; }
    for (int b = 0; b < 2; ++b)
        if (sr->busy[b]) { CU(cudaStreamWaitEvent(st, f->sliced_free[b], 0)); sr->busy[b] = false; }
    bool all_clean = true;
    for (uint32_t b = 0; b < sr->plan.n_buffers; ++b) all_clean = all_clean && sr->zeroed[b];
    if (all_clean) { f->sliced_clean_ptr = f->bucket_records.p; f->sliced_clean_bytes = sr->plan.shadow_bytes * sr->plan.n_buffers; }
    return DYIBF_OK;
}

// the updates of group g's sequences of one packed batch -> the shadow; switches to the other buffer first where the group
// changes.  both_streams: the call alternates between the two slot streams (host mode) — the switch is ordered against the
// other stream with events; scatters of different chunks into one buffer may overlap (atomics).
int sliced_scatter_group(dyibf_filter* f, SlicedRun* sr, uint32_t g, const uint64_t* packed, const SeqMeta* meta,
                         const uint32_t* bin_ids, uint32_t n_seqs, uint64_t words, bool both_streams, cudaStream_t st)
{
    if (sr->cur != (int)g) {
        if (both_streams) { int rc = join_streams(f, st); if (rc) return rc; }
        if (sr->cur >= 0) {
            int rc = sliced_hand_over(f, sr, st);
            if (rc) return rc;
            sr->buf = (sr->buf + 1) % (int)sr->plan.n_buffers;
        }
        const int b = sr->buf;
        if (sr->busy[b]) { CU(cudaStreamWaitEvent(st, f->sliced_free[b], 0)); sr->busy[b] = false; }   // interleaved and zeroed again
        if (!sr->zeroed[b]) { CU(launch_sliced_zero(sliced_buffer(f, sr, b), sr->plan, st)); sr->zeroed[b] = true; }
        if (both_streams) { int rc = fork_streams(f, st); if (rc) return rc; }
        sr->cur = (int)g;
    }
    Scope sc(f, st, K_BUILD);
    // one work counter per slot stream: scatters of successive chunks may overlap
    unsigned long long* ctr = static_cast<unsigned long long*>(f->bucket_cursors.p) + (st == f->slot[1].st ? 8 : 0);
    CU(launch_sliced_scatter(f->hp, sliced_buffer(f, sr, sr->buf), sr->plan, g, packed, meta, bin_ids, n_seqs, words, ctr, st));
    return DYIBF_OK;
}

} // namespace

extern "C" {

int dyibf_add_sequences(dyibf_filter* f, const uint8_t* seqs, const uint64_t* offsets, uint32_t n_seqs,
                        const uint32_t* bin_ids, int encoding, int mem, void* stream)
{
    if (!f) return fail(DYIBF_ERR_ARG, "filter is NULL");
    if (n_seqs == 0) return DYIBF_OK;
    if (!seqs || !offsets || !bin_ids) return fail(DYIBF_ERR_ARG, "NULL argument");
    if (encoding < DYIBF_ENC_DNA5 || encoding > DYIBF_ENC_ASCII) return fail(DYIBF_ERR_ARG, "unknown encoding %d", encoding);
    DeviceGuard guard(f->prm.device);
    if (mem == DYIBF_MEM_DEVICE) {
        cudaStream_t st = (cudaStream_t)stream;
        uint64_t first = 0, last = 0;
        CU(cudaMemcpyAsync(&first, offsets, 8, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(&last, offsets + n_seqs, 8, cudaMemcpyDeviceToHost, st));
        // sliced build: which column groups occur and how clustered the bins are — asked of the device-resident bin ids
        // here, answered by the synchronisation that fetches the offsets anyway
        SlicedPlan splan; SlicedRun sr;
        std::vector<uint32_t> group_mask; unsigned long long switches = 0;
        // the call's size is not known before the sync: plan for a large one, decide after
        const bool s_cand = sliced_candidate(f, ~0ull, &splan) && splan.n_groups <= 4096;
        if (s_cand) {
            CU(f->build_bins.reserve((size_t)splan.n_groups * 4 + 16));
            uint32_t* d_mask = static_cast<uint32_t*>(f->build_bins.p);
            unsigned long long* d_sw = reinterpret_cast<unsigned long long*>(d_mask + ((splan.n_groups + 1) & ~1u));
            CU(launch_sliced_presence(bin_ids, n_seqs, f->prm.bins, splan, d_mask, d_sw, st));
            f->kernel_launches += 1;
            group_mask.resize(splan.n_groups);
            CU(cudaMemcpyAsync(group_mask.data(), d_mask, (size_t)splan.n_groups * 4, cudaMemcpyDeviceToHost, st));
            CU(cudaMemcpyAsync(&switches, d_sw, 8, cudaMemcpyDeviceToHost, st));
        }
        CU(cudaStreamSynchronize(st));
        if (last < first) return fail(DYIBF_ERR_ARG, "offsets not monotone");
        if (s_cand && (f->build_mode != DYIBF_BUILD_AUTO || (last - first) * f->hp.h >= (256ull << 20))) {
            int src_ = sliced_begin(f, splan, last - first, switches, &sr);
            if (src_) return src_;
        }
        Slot& sl = f->slot[0];
        PackSource src; memset(&src, 0, sizeof src);
        src.n_ranges = 1; src.cnt = n_seqs; src.bytes[0] = seqs; src.offsets[0] = offsets; src.first_off[0] = first;
        const uint64_t words = packed_words_for(last - first, n_seqs);
        CU(sl.packed.reserve(words * 8));
        CU(sl.meta.reserve((uint64_t)n_seqs * sizeof(SeqMeta)));
        { Scope sc(f, st, K_PACK);
          CU(launch_pack(src, encoding, 0, (uint64_t*)sl.packed.p, (SeqMeta*)sl.meta.p, words, words, st)); }
        if (sr.active) {
            f->last_build_mode = DYIBF_BUILD_SLICED;
            for (uint32_t g = 0; g < splan.n_groups; ++g) {
                if (!group_mask[g]) continue;
                int rc = sliced_scatter_group(f, &sr, g, (const uint64_t*)sl.packed.p, (const SeqMeta*)sl.meta.p, bin_ids, n_seqs, words, false, st);
                if (rc) return rc;
            }
            return sliced_finish_group(f, &sr, st);                  // the shadow stays allocated: stream-ordered call
        }
        BucketRun br;
        int brc = f->build_mode == DYIBF_BUILD_DIRECT ? (br.active = false, DYIBF_OK) : buckets_begin(f, (last - first) * f->hp.h, &br);
        if (brc) return brc;
        f->last_build_mode = br.active ? DYIBF_BUILD_BUCKETED : DYIBF_BUILD_DIRECT;
        if (!br.active) {
            Scope sc(f, st, K_BUILD);
            CU(launch_build(build_params(f), f->d_words, (const uint64_t*)sl.packed.p, (const SeqMeta*)sl.meta.p, bin_ids, n_seqs, words, st));
            return DYIBF_OK;
        }
        // stream-ordered: split, level 2, apply, next word range
        brc = buckets_scatter(f, &br, (const uint64_t*)sl.packed.p, (const SeqMeta*)sl.meta.p, bin_ids, n_seqs, words, 0, false, st);
        if (brc) return brc;
        // the record buffer stays allocated: the kernels above may still be running (stream-ordered call)
        return buckets_apply(f, &br, st);
    }
    if (mem != DYIBF_MEM_HOST) return fail(DYIBF_ERR_ARG, "unknown memory space %d", mem);
    for (uint32_t s = 0; s < n_seqs; ++s) {
        if (offsets[s + 1] < offsets[s]) return fail(DYIBF_ERR_ARG, "offsets not monotone at %u", s);
        if (offsets[s + 1] - offsets[s] > 0xffffffffull) return fail(DYIBF_ERR_ARG, "sequence %u longer than 2^32", s);
        if (bin_ids[s] >= f->prm.bins) return fail(DYIBF_ERR_ARG, "bin %u of sequence %u out of range (%u bins)", bin_ids[s], s, f->prm.bins);
    }
    // chunks of whole sequences, ~256 MB of bases each by default (8192 scatter work items: enough warps for every SM),
    // alternating between the two stream slots
    const uint64_t expected = (offsets[n_seqs] - offsets[0]) * f->hp.h;
    SlicedPlan splan; SlicedRun sr;
    if (sliced_candidate(f, expected, &splan)) {
        uint64_t switches = 0;
        for (uint32_t s = 1; s < n_seqs; ++s) switches += bin_ids[s] != bin_ids[s - 1];
        int src_ = sliced_begin(f, splan, offsets[n_seqs] - offsets[0], switches, &sr);
        if (src_) return src_;
    }
    BucketRun br;
    int brc = (sr.active || f->build_mode == DYIBF_BUILD_DIRECT) ? (br.active = false, DYIBF_OK) : buckets_begin(f, expected, &br);
    if (brc) return brc;
    // auto chunk size: 64 MiB under the sliced build (its scatter never holds the host, so small chunks only shorten the
    // pipeline's ramp), 256 MiB under the others (a bucket flush holds the host; larger chunks amortise it)
    const uint64_t target = (uint64_t)(f->build_chunk_mib ? f->build_chunk_mib : sr.active ? 64 : 256) << 20;
    f->last_build_mode = sr.active ? DYIBF_BUILD_SLICED : br.active ? DYIBF_BUILD_BUCKETED : DYIBF_BUILD_DIRECT;
    // The bases go up on a copy stream of their own, two chunks ahead of the kernels: the upload of chunk c+2 into slot c&1
    // starts as soon as pack(c) has consumed that slot's bytes, and keeps running while buckets_scatter blocks the host on
    // the two kernel streams to empty a full record buffer (level 2 / apply).  pack(c) waits for its own upload by event.
    struct Chunk { uint32_t a, cnt; uint64_t bytes, words; };
    std::vector<Chunk> chunks;
    uint64_t max_bytes = 0, max_words = 0; uint32_t max_cnt = 0;
    for (uint32_t a = 0; a < n_seqs;) {
        uint32_t b = a + 1;
        while (b < n_seqs && offsets[b + 1] - offsets[a] <= target) ++b;
        const Chunk ch{a, b - a, offsets[b] - offsets[a], packed_words_for(offsets[b] - offsets[a], b - a)};
        chunks.push_back(ch);
        max_bytes = std::max(max_bytes, ch.bytes); max_words = std::max(max_words, ch.words); max_cnt = std::max(max_cnt, ch.cnt);
        a = b;
    }
    for (int i = 0; i < 2 && i < (int)chunks.size(); ++i) {        // sized once: no reallocation under the pipeline
        Slot& sl = f->slot[i];
        CU(sl.in_bytes.reserve(max_bytes + 16));
        CU(sl.packed.reserve(max_words * 8));
        CU(sl.meta.reserve((uint64_t)max_cnt * sizeof(SeqMeta)));
    }
    struct Upload {                                                  // copy stream + per-slot events, released on every exit path
        cudaStream_t st = nullptr; cudaEvent_t copied[2] = {nullptr, nullptr}, packed[2] = {nullptr, nullptr};
        ~Upload() {
            if (st) { cudaStreamSynchronize(st); cudaStreamDestroy(st); }
            for (int i = 0; i < 2; ++i) { if (copied[i]) cudaEventDestroy(copied[i]); if (packed[i]) cudaEventDestroy(packed[i]); }
        }
    } up;
    CU(cudaStreamCreateWithFlags(&up.st, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
        CU(cudaEventCreateWithFlags(&up.copied[i], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&up.packed[i], cudaEventDisableTiming));
    }
    const bool pageable = offsets[n_seqs] - offsets[0] >= (8u << 20) && is_pageable(seqs + offsets[0]);
    auto upload = [&](uint32_t c) -> int {                           // bases + limits of chunk c into slot c&1, on the copy stream
        const Chunk& ch = chunks[c]; Slot& sl = f->slot[c & 1];
        if (ch.bytes && pageable) {                                  // through pinned pieces, copied by helper threads
            int rc = staged_h2d(sl.in_bytes.p, seqs + offsets[ch.a], ch.bytes, up.st);
            if (rc) return rc;
        } else if (ch.bytes) {
            CU(cudaMemcpyAsync(sl.in_bytes.p, seqs + offsets[ch.a], ch.bytes, cudaMemcpyHostToDevice, up.st));
        }
        CU(cudaEventRecord(up.copied[c & 1], up.st));
        return DYIBF_OK;
    };
    // the limits and bin ids of the whole call go up once, ahead of everything (a per-chunk copy from the caller's pageable
    // arrays would make the host wait for the chunk's stream to drain every time, serialising upload and kernels)
    const uint64_t off_bytes = ((uint64_t)(n_seqs + 1) * 8 + 15) & ~15ull;
    CU(f->build_bins.reserve(off_bytes + (uint64_t)n_seqs * 4 + 16));
    const uint64_t* d_offsets_all = static_cast<const uint64_t*>(f->build_bins.p);
    CU(cudaMemcpyAsync(f->build_bins.p, offsets, (uint64_t)(n_seqs + 1) * 8, cudaMemcpyHostToDevice, up.st));
    CU(cudaMemcpyAsync(static_cast<char*>(f->build_bins.p) + off_bytes, bin_ids, (uint64_t)n_seqs * 4, cudaMemcpyHostToDevice, up.st));
    cudaEvent_t bins_up = up.packed[0];                            // borrowed: recorded once here, before its per-slot use
    CU(cudaEventRecord(bins_up, up.st));
    CU(cudaStreamWaitEvent(f->slot[0].st, bins_up, 0));
    CU(cudaStreamWaitEvent(f->slot[1].st, bins_up, 0));
    const uint32_t* d_bins_all = reinterpret_cast<const uint32_t*>(static_cast<const char*>(f->build_bins.p) + off_bytes);
    int urc = upload(0);
    if (!urc && chunks.size() > 1) urc = upload(1);
    if (urc) return urc;
    for (uint32_t c = 0; c < chunks.size(); ++c) {
        const Chunk& cur = chunks[c];
        Slot& sl = f->slot[c & 1];
        const uint32_t* d_bins = d_bins_all + cur.a;
        CU(cudaStreamWaitEvent(sl.st, up.copied[c & 1], 0));
        PackSource src; memset(&src, 0, sizeof src);
        src.n_ranges = 1; src.cnt = cur.cnt; src.first_off[0] = offsets[cur.a];
        src.bytes[0] = (const uint8_t*)sl.in_bytes.p - offsets[cur.a]; src.offsets[0] = d_offsets_all + cur.a;
        { Scope sc(f, sl.st, K_PACK);
          CU(launch_pack(src, encoding, 0, (uint64_t*)sl.packed.p, (SeqMeta*)sl.meta.p, cur.words, cur.words, sl.st)); }
        const bool more = c + 2 < chunks.size();
        if (more) CU(cudaEventRecord(up.packed[c & 1], sl.st));      // the slot's bytes are free once this pack has run
        // Bucketed: buckets_scatter may hold the host to empty a full record buffer, so the next upload is queued first.
        // Otherwise the chunk's kernels are queued first: staging a pageable range holds the host instead.
        if (more && br.active) {
            CU(cudaStreamWaitEvent(up.st, up.packed[c & 1], 0));
            urc = upload(c + 2);
            if (urc) return urc;
        }
        if (sr.active) {
            // the column groups this chunk's sequences belong to, in order of first appearance (bin-sorted input: ascending,
            // every group once per call)
            std::vector<uint32_t> groups;
            for (uint32_t s = cur.a; s < cur.a + cur.cnt; ++s) {
                const uint32_t g = bin_ids[s] / splan.group_bins;
                if (groups.empty() || (groups.back() != g && std::find(groups.begin(), groups.end(), g) == groups.end())) groups.push_back(g);
            }
            if (!groups.empty() && sr.cur >= 0) {                    // keep the group that is already in the shadow first
                auto it = std::find(groups.begin(), groups.end(), (uint32_t)sr.cur);
                if (it != groups.end()) std::rotate(groups.begin(), it, it + 1);
            }
            for (uint32_t g : groups) {
                int rc = sliced_scatter_group(f, &sr, g, (const uint64_t*)sl.packed.p, (const SeqMeta*)sl.meta.p, d_bins,
                                              cur.cnt, cur.words, true, sl.st);
                if (rc) return rc;
            }
        } else if (br.active) {
            brc = buckets_scatter(f, &br, (const uint64_t*)sl.packed.p, (const SeqMeta*)sl.meta.p, d_bins,
                                  cur.cnt, cur.words, c & 1, true, sl.st);
            if (brc) return brc;
        } else {
            Scope sc(f, sl.st, K_BUILD);
            CU(launch_build(build_params(f), f->d_words, (const uint64_t*)sl.packed.p, (const SeqMeta*)sl.meta.p,
                            d_bins, cur.cnt, cur.words, sl.st));
        }
        if (more && !br.active) {
            CU(cudaStreamWaitEvent(up.st, up.packed[c & 1], 0));
            urc = upload(c + 2);
            if (urc) return urc;
        }
    }
    CU(cudaStreamSynchronize(f->slot[0].st));
    CU(cudaStreamSynchronize(f->slot[1].st));
    if (sr.active) {
        int rc = sliced_finish_group(f, &sr, f->slot[0].st);
        if (rc) return rc;
        CU(cudaStreamSynchronize(f->slot[0].st));
    }
    if (br.active) {
        brc = buckets_apply(f, &br, f->slot[0].st);
        if (brc) return brc;
        CU(cudaStreamSynchronize(f->slot[0].st));
        buckets_end(f, &br);
    }
    return DYIBF_OK;
}

int dyibf_clear_bins(dyibf_filter* f, const uint32_t* bins, uint32_t n)
{
    if (!f || (n && !bins)) return fail(DYIBF_ERR_ARG, "NULL argument");
    if (n == 0) return DYIBF_OK;
    std::vector<uint32_t> cols; std::vector<uint64_t> masks;
    {
        std::vector<uint32_t> sorted(bins, bins + n);
        std::sort(sorted.begin(), sorted.end());
        for (uint32_t b : sorted) {
            if (b >= f->prm.bins) return fail(DYIBF_ERR_ARG, "bin %u out of range (%u bins)", b, f->prm.bins);
            if (cols.empty() || cols.back() != (b >> 6)) { cols.push_back(b >> 6); masks.push_back(0); }
            masks.back() |= 1ull << (b & 63);
        }
    }
    DeviceGuard guard(f->prm.device);
    const uint32_t nc = (uint32_t)cols.size();
    CU(f->misc.reserve((uint64_t)nc * 16 + 16));
    uint64_t* d_masks = (uint64_t*)f->misc.p;
    uint32_t* d_cols = (uint32_t*)(d_masks + nc);
    cudaStream_t st = f->slot[0].st;
    CU(cudaMemcpyAsync(d_masks, masks.data(), (uint64_t)nc * 8, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(d_cols, cols.data(), (uint64_t)nc * 4, cudaMemcpyHostToDevice, st));
    { Scope sc(f, st, K_CLEAR);
      CU(launch_clear(f->hp, f->d_words, d_cols, d_masks, nc, st)); }
    CU(cudaStreamSynchronize(st));
    return DYIBF_OK;
}

int dyibf_set_option(dyibf_filter* f, int option, int64_t value)
{
    if (!f) return fail(DYIBF_ERR_ARG, "filter is NULL");
    switch (option) {
        case DYIBF_OPT_EARLY_EXIT:
            if (value) f->query_flags |= QUERY_EARLY_EXIT; else f->query_flags &= ~QUERY_EARLY_EXIT;
            return DYIBF_OK;
        case DYIBF_OPT_BUILD_BUCKET_MIB:
            if (value < -1) return fail(DYIBF_ERR_ARG, "DYIBF_OPT_BUILD_BUCKET_MIB must be -1 (auto), 0 (off) or a size in MiB");
            f->bucket_mib = value;
            return DYIBF_OK;
        case DYIBF_OPT_QUERY_CHUNK_MIB:
            if (value < 0 || value > 4096) return fail(DYIBF_ERR_ARG, "DYIBF_OPT_QUERY_CHUNK_MIB must be 0 (auto) or 1..4096");
            f->query_chunk_mib = value;
            return DYIBF_OK;
        case DYIBF_OPT_BUILD_CHUNK_MIB:
            if (value < 0 || value > 4096) return fail(DYIBF_ERR_ARG, "DYIBF_OPT_BUILD_CHUNK_MIB must be 0 (auto) or 1..4096");
            f->build_chunk_mib = value;
            return DYIBF_OK;
        case DYIBF_OPT_BUILD_MODE:
            if (value < DYIBF_BUILD_AUTO || value > DYIBF_BUILD_SLICED) return fail(DYIBF_ERR_ARG, "DYIBF_OPT_BUILD_MODE must be 0..3");
            f->build_mode = value;
            return DYIBF_OK;
        default:
            return fail(DYIBF_ERR_ARG, "unknown option %d", option);
    }
}

int dyibf_get_option(const dyibf_filter* f, int option, int64_t* value)
{
    if (!f || !value) return fail(DYIBF_ERR_ARG, "NULL argument");
    switch (option) {
        case DYIBF_OPT_EARLY_EXIT: *value = (f->query_flags & QUERY_EARLY_EXIT) ? 1 : 0; return DYIBF_OK;
        case DYIBF_OPT_BUILD_BUCKET_MIB: *value = f->bucket_mib; return DYIBF_OK;
        case DYIBF_OPT_BUILD_CHUNK_MIB: *value = f->build_chunk_mib; return DYIBF_OK;
        case DYIBF_OPT_QUERY_CHUNK_MIB: *value = f->query_chunk_mib; return DYIBF_OK;
        case DYIBF_OPT_BUILD_MODE: *value = f->build_mode; return DYIBF_OK;
        case DYIBF_OPT_LAST_BUILD_MODE: *value = f->last_build_mode; return DYIBF_OK;
        default: return fail(DYIBF_ERR_ARG, "unknown option %d", option);
    }
}

int dyibf_profile_enable(dyibf_filter* f, int on)
{
    if (!f) return fail(DYIBF_ERR_ARG, "filter is NULL");
    DeviceGuard guard(f->prm.device);
    if (on && !f->stats.p) {
        CU(f->stats.reserve(64));
        CU(cudaMemset(f->stats.p, 0, 64));
        CU(cudaDeviceSynchronize());
    }
    if (on && getenv("DYIBF_TRACE")) {                                // timeline origin (see drain_profile)
        if (!f->trace_base) CU(cudaEventCreate(&f->trace_base));
        CU(cudaDeviceSynchronize());
        CU(cudaEventRecord(f->trace_base, 0));
        CU(cudaEventSynchronize(f->trace_base));
    }
    f->profiling = on != 0;
    return DYIBF_OK;
}

int dyibf_profile_get(dyibf_filter* f, dyibf_profile* out, int reset)
{
    if (!f || !out) return fail(DYIBF_ERR_ARG, "NULL argument");
    DeviceGuard guard(f->prm.device);
    drain_profile(f);
    out->pack_ms = f->ms[K_PACK]; out->query_ms = f->ms[K_QUERY]; out->compact_ms = f->ms[K_COMPACT];
    out->build_ms = f->ms[K_BUILD] + f->ms[K_SPLIT2] + f->ms[K_APPLY]; out->clear_ms = f->ms[K_CLEAR];
    out->build_apply_ms = f->ms[K_APPLY]; out->build_apply_launches = f->launches[K_APPLY];
    out->build_split2_ms = f->ms[K_SPLIT2]; out->build_split2_launches = f->launches[K_SPLIT2];
    out->pack_launches = f->launches[K_PACK]; out->query_launches = f->launches[K_QUERY];
    out->compact_launches = f->launches[K_COMPACT]; out->build_launches = f->launches[K_BUILD] + f->launches[K_SPLIT2] + f->launches[K_APPLY];
    out->clear_launches = f->launches[K_CLEAR];
    out->query_lookups = 0;
    if (f->stats.p) {                                                  // all launches have drained above
        CU(cudaDeviceSynchronize());
        CU(cudaMemcpy(&out->query_lookups, f->stats.p, 8, cudaMemcpyDeviceToHost));
        if (reset) CU(cudaMemset(f->stats.p, 0, 64));
    }
    if (reset) for (int i = 0; i < K_N; ++i) { f->ms[i] = 0; f->launches[i] = 0; }
    return DYIBF_OK;
}

int dyibf_device_errors(dyibf_filter* f, uint32_t* flags, int reset)
{
    if (!f || !flags) return fail(DYIBF_ERR_ARG, "NULL argument");
    DeviceGuard guard(f->prm.device);
    CU(cudaDeviceSynchronize());                                       // everything that could still set a bit has run
    CU(cudaMemcpy(flags, f->dev_err.p, 4, cudaMemcpyDeviceToHost));
    if (reset && *flags) CU(cudaMemset(f->dev_err.p, 0, 4));
    if (reset) CU(cudaDeviceSynchronize());
    return DYIBF_OK;
}

uint64_t dyibf_launch_count(const dyibf_filter* f) { return f ? f->kernel_launches : 0; }

// Large pinned buffers can be spread over the host's NUMA nodes (DYIBF_HOST_ALLOC_INTERLEAVE=1): a buffer that several GPUs on
// different sockets read at once (dyibf_group_classify from one batch) is then served by all memory controllers instead of one.
// Off by default: a buffer read by one GPU is best on that GPU's own socket, which the first-touch default usually gives.
static bool interleave_policy(bool on)
{
#if defined(__linux__) && defined(SYS_set_mempolicy)
    unsigned long mask[16];                                          // the nodes that are online: "0-1", "0,2-3", ...
    FILE* fp = fopen("/sys/devices/system/node/online", "r");
    if (!fp) return false;
    int lo = 0, hi = 0; char sep = 0;
    memset(mask, 0, sizeof mask);
    bool any = false;
    while (fscanf(fp, "%d", &lo) == 1) {
        hi = lo;
        if (fscanf(fp, "%c", &sep) == 1 && sep == '-') { if (fscanf(fp, "%d", &hi) != 1) hi = lo; if (fscanf(fp, "%c", &sep) != 1) sep = 0; }
        for (int n = lo; n <= hi && n < 1024; ++n) { mask[n / (8 * sizeof(long))] |= 1ul << (n % (8 * sizeof(long))); any = true; }
        if (sep != ',') break;
    }
    fclose(fp);
    if (!any || hi == 0) return false;                               // one node: nothing to interleave
    const int MPOL_DEFAULT_ = 0, MPOL_INTERLEAVE_ = 3;
    return syscall(SYS_set_mempolicy, on ? MPOL_INTERLEAVE_ : MPOL_DEFAULT_, on ? mask : nullptr, on ? 1024ul : 0ul) == 0;
#else
    (void)on; return false;
#endif
}

int dyibf_host_alloc(void** ptr, size_t bytes)
{
    if (!ptr) return fail(DYIBF_ERR_ARG, "ptr is NULL");
    const char* env = getenv("DYIBF_HOST_ALLOC_INTERLEAVE");
    const bool spread = env && atoi(env) != 0 && bytes >= (64u << 20) && interleave_policy(true);
    cudaError_t e = cudaMallocHost(ptr, bytes ? bytes : 1);
    if (spread) interleave_policy(false);
    if (e != cudaSuccess) return fail(DYIBF_ERR_CUDA, "cudaMallocHost(%zu): %s", bytes, cudaGetErrorString(e));
    return DYIBF_OK;
}

void dyibf_host_free(void* ptr) { if (ptr) cudaFreeHost(ptr); }

} // extern "C"
