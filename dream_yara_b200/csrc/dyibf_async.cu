// dyibf_async.cu — ticketed (asynchronous) classify: dyibf_classify_begin / dyibf_classify_end.
//
// Why: the reference's mapper calls the filtering stage once per batch of 100 000 reads (src/mapper.h:109) while a
// prefetch thread already loads the next batch (src/file_prefetched.h:137-205).  A synchronous dyibf_classify of such a
// batch cannot hide its own upload, list download and the query kernel's ramp-up / drain behind anything.  A ticket owns
// a stream and a private set of buffers, so batch i+1 uploads and starts while batch i is still in the query kernel and
// batch i-1's lists travel back: the GPU sees a continuous stream of work at the reference's batch size.
//
// Scheduling (measured, B200, 100 000-read batches from pinned memory, tools/debug/async_probe.py and the DYIBF_TRACE timeline):
// every ticket has ONE stream of default priority and the query kernel keeps its persistent blocks.  2 / 3 / 4 tickets in
// flight: 49.7 / 52.2 / 53.4 M reads/s (one large batch: 56.4 M).  What is left is a drain bubble: the compaction of batch i waits
// until the query kernels of the batches queued behind it have released SM slots.  Tried and dropped: a high-priority
// stream for copies / pack / compaction with the query kernel on a low-priority stream and short-lived blocks (one work item
// per warp) so that the small kernels get slots at once — the timeline becomes gap-free, but three query kernels
// interleaved block by block run at 45.5 M reads/s in total (each alone: 56.7 M in either block form), i.e. co-resident
// query kernels cost more than the bubble; priorities with persistent blocks change nothing (52.3 M at 3 in flight).
//
// begin: H2D of the batch (straight from the caller's memory: truly asynchronous when it is pinned, staged by the driver
//        when it is pageable), pack, query, compaction count + scan + fill, D2H of the bin offsets — all enqueued, nothing
//        waited for.  The caller's seqs / offsets must stay valid until dyibf_classify_end.
// end:   waits for the ticket, copies the lists out (the only copy whose size is not known before), and makes the
//        ticket's batch the handle's resident batch (dyibf_gather_reads / dyibf_classify_fetch work on it).
#include "dyibf_handle.h"

using namespace dyibf;

namespace dyibf {

struct Ticket {
    dyibf_filter* f = nullptr;
    int           index = 0;
    bool          busy = false;
    Slot          sl;                         // stream + pack / query workspace
    DevBuf        bytes, offsets, rows, compact_ws, csr_off, csr_ids, text;
    uint64_t*     h_off = nullptr;            // pinned: bins + 1 bin offsets
    void*         h_in = nullptr; size_t h_in_cap = 0;   // pinned staging of a pageable caller's batch (bytes, then limits)
    cudaEvent_t   done = nullptr;
    uint32_t      n_rows = 0;
    uint64_t      ids_cap = 0;                // entries the fill kernel was allowed to write
    ResidentBatch batch;
};

void free_tickets(dyibf_filter* f)
{
    for (int i = 0; i < DYIBF_MAX_TICKETS; ++i) {
        Ticket* t = f->tickets[i];
        if (!t) continue;
        if (t->sl.st) { cudaStreamSynchronize(t->sl.st); cudaStreamDestroy(t->sl.st); }
        if (t->done) cudaEventDestroy(t->done);
        if (t->h_off) cudaFreeHost(t->h_off);
        if (t->h_in) cudaFreeHost(t->h_in);
        t->sl.in_bytes.release(); t->sl.in_offsets.release(); t->sl.packed.release(); t->sl.meta.release();
        t->sl.rows.release(); t->sl.counts.release(); t->sl.ctr.release();
        t->bytes.release(); t->offsets.release(); t->rows.release(); t->compact_ws.release(); t->csr_off.release(); t->csr_ids.release(); t->text.release();
        delete t;
        f->tickets[i] = nullptr;
    }
}

static int acquire_ticket(dyibf_filter* f, Ticket** out)
{
    for (int i = 0; i < DYIBF_MAX_TICKETS; ++i) {
        Ticket* t = f->tickets[i];
        if (t && t->busy) continue;
        if (!t) {
            t = new Ticket();
            t->f = f; t->index = i;
            if (cudaStreamCreateWithFlags(&t->sl.st, cudaStreamNonBlocking) != cudaSuccess ||
                cudaEventCreateWithFlags(&t->done, cudaEventDisableTiming) != cudaSuccess ||
                cudaMallocHost((void**)&t->h_off, ((size_t)f->prm.bins + 2) * 8) != cudaSuccess) {
                if (t->sl.st) cudaStreamDestroy(t->sl.st);
                if (t->done) cudaEventDestroy(t->done);
                delete t;
                return fail(DYIBF_ERR_CUDA, "dyibf_classify_begin: ticket resources could not be created");
            }
            f->tickets[i] = t;
        }
        *out = t;
        return DYIBF_OK;
    }
    return fail(DYIBF_ERR_BUSY, "dyibf_classify_begin: all %d tickets of this handle are in flight", DYIBF_MAX_TICKETS);
}

// A begin call that fails after its first asynchronous copy has been enqueued must not return while the DMA engine still
// reads the caller's buffers (or the ticket's pinned h_in, which the next begin on that ticket overwrites): unless
// released, the guard waits for the ticket's stream on the way out.
struct DrainOnError {
    cudaStream_t st; bool armed = true;
    explicit DrainOnError(cudaStream_t s) : st(s) {}
    ~DrainOnError() { if (armed) { cudaStreamSynchronize(st); cudaGetLastError(); } }
    void release() { armed = false; }
};

// compaction of the ticket's rows: count + scan, then the fill into a buffer sized for the usual case (<= 4 lists per row);
// a batch that needs more is re-filled in dyibf_classify_end
static int enqueue_compaction(Ticket* t, uint32_t n_rows)
{
    dyibf_filter* f = t->f;
    const uint32_t bins = f->prm.bins;
    cudaStream_t st = t->sl.st;
    const uint64_t* d_rows = (const uint64_t*)t->rows.p;
    const size_t ws_bytes = compact_workspace_bytes(bins, n_rows);
    CU(t->compact_ws.reserve(ws_bytes));
    t->ids_cap = std::max<uint64_t>(4ull * n_rows, 1024);
    CU(t->csr_ids.reserve(t->ids_cap * 4));
    t->ids_cap = t->csr_ids.cap / 4;
    uint64_t* ws = (uint64_t*)t->compact_ws.p;
    CU(cudaMemsetAsync(ws, 0, ws_bytes, st));
    Scope sc(f, st, K_COMPACT, 6);
    CU(launch_compact_count(d_rows, n_rows, bins, ws, st));
    CU(launch_compact_scan(ws, bins, n_rows, (uint64_t*)t->csr_off.p, st));
    CU(launch_compact_fill(d_rows, n_rows, bins, ws, (uint32_t*)t->csr_ids.p, t->ids_cap, 0, st));
    return DYIBF_OK;
}

// the bin offsets travel back; the ticket is in flight from here on
static int enqueue_finish(Ticket* t)
{
    cudaStream_t st = t->sl.st;
    CU(cudaMemcpyAsync(t->h_off, t->csr_off.p, (uint64_t)(t->f->prm.bins + 1) * 8, cudaMemcpyDeviceToHost, st));
    CU(cudaEventRecord(t->done, st));
    t->busy = true;
    return DYIBF_OK;
}

} // namespace dyibf

extern "C" {

int dyibf_classify_begin(dyibf_filter* f, const uint8_t* seqs, const uint64_t* offsets, uint32_t n_seqs, int encoding,
                         int group, int derive_rc, uint32_t threshold, dyibf_ticket** ticket)
{
    if (!ticket) return fail(DYIBF_ERR_ARG, "ticket is NULL");
    *ticket = nullptr;
    int rc = check_query_args(f, seqs, offsets, n_seqs, encoding, group, derive_rc, threshold);
    if (rc) return rc;
    DeviceGuard guard(f->prm.device);
    Ticket* t = nullptr;
    rc = acquire_ticket(f, &t);
    if (rc) return rc;
    const uint32_t bins = f->prm.bins, W = f->prm.words_per_block, k = f->hp.k;
    const uint32_t group_in = derive_rc ? group / 2 : group;
    const uint64_t n_total = derive_rc ? 2ull * n_seqs : n_seqs;
    const uint32_t n_rows = (uint32_t)(n_total / group);
    cudaStream_t st = t->sl.st;
    DrainOnError drain(st);
    t->n_rows = n_rows; t->batch.valid = false; t->ids_cap = 0;
    CU(t->csr_off.reserve((uint64_t)(bins + 1) * 8));
    if (n_rows == 0) {
        CU(cudaMemsetAsync(t->csr_off.p, 0, (uint64_t)(bins + 1) * 8, st));
    } else {
        // the longest sequence picks the counter width of the query kernel; one pass over the limits (~30 us per 100 000)
        uint64_t mx = 0;
        for (uint32_t s = 0; s < n_seqs; ++s) mx = std::max(mx, offsets[s + 1] - offsets[s]);
        if (offsets[n_seqs] < offsets[0] || mx > 0xffffffffull) return fail(DYIBF_ERR_ARG, "offsets not monotone, or a sequence longer than 2^32");
        const uint32_t max_len = (uint32_t)std::max<uint64_t>(mx, 1);
        if (max_len >= k && max_len - k + 1 > 65535)
            return fail(DYIBF_ERR_ARG, "a sequence of %u bases has more than 65535 k-mers (uint16 counts in the reference)", max_len);
        const uint64_t total_bytes = offsets[n_seqs] - offsets[0];
        CU(t->bytes.reserve(total_bytes + 16));
        CU(t->offsets.reserve((uint64_t)(n_seqs + 1) * 8));
        CU(t->rows.reserve((uint64_t)n_rows * W * 8 + 8));
        const uint8_t* src_bytes = seqs + offsets[0];
        const uint64_t* src_off = offsets;
        const size_t off_bytes = (size_t)(n_seqs + 1) * 8;
        // Pageable caller memory: cudaMemcpyAsync from it is staged by the driver on the calling thread at ~10 GB/s, which for
        // a 100 000-read batch (15 MB) is most of the batch's GPU time.  The batch is instead copied into a pinned buffer of
        // the ticket by a few helper threads (dyibf_hostcopy.cu) and uploaded from there asynchronously.
        if (total_bytes + off_bytes >= (1u << 20) && is_pageable(src_bytes)) {
            // the previous use of the staging buffer was waited for by that ticket's dyibf_classify_end
            const size_t b_pad = (total_bytes + 63) & ~size_t(63);
            if (b_pad + off_bytes > t->h_in_cap) {
                if (t->h_in) { cudaFreeHost(t->h_in); t->h_in = nullptr; t->h_in_cap = 0; }
                const size_t want = b_pad + off_bytes + (b_pad + off_bytes) / 4;
                CU(cudaMallocHost(&t->h_in, want));
                t->h_in_cap = want;
            }
            host_copy_parallel(t->h_in, src_bytes, total_bytes, 4);
            host_copy_streaming(static_cast<char*>(t->h_in) + b_pad, offsets, off_bytes);
            src_bytes = static_cast<const uint8_t*>(t->h_in);
            src_off = reinterpret_cast<const uint64_t*>(static_cast<char*>(t->h_in) + b_pad);
        }
        {
            Scope sc(f, st, K_COPY, 0);
            if (total_bytes) CU(cudaMemcpyAsync(t->bytes.p, src_bytes, total_bytes, cudaMemcpyHostToDevice, st));
            CU(cudaMemcpyAsync(t->offsets.p, src_off, off_bytes, cudaMemcpyHostToDevice, st));
        }
        // the batch as the pack kernel sees it: group_in ranges of n_rows sequences each, laid out as the caller's buffer
        PackSource src; memset(&src, 0, sizeof src);
        src.n_ranges = group_in; src.cnt = n_rows;
        for (uint32_t g = 0; g < group_in; ++g) {
            const uint64_t s0 = (uint64_t)g * n_rows;
            src.first_off[g] = offsets[s0]; src.pack_base[g] = offsets[s0] - offsets[0];
            src.bytes[g] = (const uint8_t*)t->bytes.p - offsets[0];
            src.offsets[g] = (const uint64_t*)t->offsets.p + s0;
        }
        uint64_t* d_rows = (uint64_t*)t->rows.p;
        rc = query_on_device(f, t->sl, src, total_bytes, encoding, (uint32_t)group, derive_rc, threshold, max_len, d_rows, nullptr, st);
        if (rc) return rc;
        rc = enqueue_compaction(t, n_rows);
        if (rc) return rc;
        if (derive_rc) {
            t->batch.valid = true; t->batch.n_fwd = n_seqs; t->batch.group = (uint32_t)group; t->batch.encoding = (uint32_t)encoding;
            t->batch.first_off = offsets[0];
        }
    }
    rc = enqueue_finish(t);
    if (rc) return rc;
    drain.release();
    *ticket = reinterpret_cast<dyibf_ticket*>(t);
    return DYIBF_OK;
}

/* loadReads + clasifyLoadedReads from the TEXT of the reads file(s), as a ticket: the text goes up on the ticket's stream,
 * is tokenised on the device (two short waits for the record / base counts, which size the launches that follow), and the
 * query + compaction are enqueued; the lists come back through dyibf_classify_end.  While this batch's query runs, the next
 * dyibf_classify_text_begin uploads and parses its text on another ticket's stream: the double-buffered reader of
 * src/file_prefetched.h:137-205, with the GPU as the consumer. */
int dyibf_classify_text_begin(dyibf_filter* f, const char* reads_text, uint64_t reads_bytes, const char* mates_text, uint64_t mates_bytes,
                              int format, int e_percent, uint32_t* n_reads, uint32_t* threshold, dyibf_ticket** ticket)
{
    if (!ticket) return fail(DYIBF_ERR_ARG, "ticket is NULL");
    *ticket = nullptr;
    if (!f || (reads_bytes && !reads_text) || (mates_bytes && !mates_text)) return fail(DYIBF_ERR_ARG, "dyibf_classify_text: NULL argument");
    if (format != DYIBF_FMT_FASTA && format != DYIBF_FMT_FASTQ) return fail(DYIBF_ERR_ARG, "dyibf_classify_text: bad format");
    const bool paired = mates_text != nullptr;
    DeviceGuard guard(f->prm.device);
    Ticket* t = nullptr;
    int rc = acquire_ticket(f, &t);
    if (rc) return rc;
    const uint32_t bins = f->prm.bins, W = f->prm.words_per_block;
    cudaStream_t st = t->sl.st;
    DrainOnError drain(st);
    t->n_rows = 0; t->batch.valid = false; t->ids_cap = 0;
    CU(t->csr_off.reserve((uint64_t)(bins + 1) * 8));
    // the text goes to the device once; the parsed forward read set ([reads] or [mates 1 | mates 2]) is the ticket's batch
    const uint64_t off2 = (reads_bytes + 15) & ~15ull, all_bytes = off2 + mates_bytes;
    CU(t->text.reserve(all_bytes + 32));
    char* d_text = static_cast<char*>(t->text.p);
    {
        Scope sc(f, st, K_COPY, 0);
        const char* src[2] = {reads_text, mates_text}; const uint64_t len[2] = {reads_bytes, mates_bytes}; char* dst[2] = {d_text, d_text + off2};
        for (int i = 0; i < 2; ++i) {
            if (!len[i]) continue;
            if (len[i] >= (1u << 20) && is_pageable(src[i])) { rc = staged_h2d(dst[i], src[i], len[i], st); if (rc) return rc; }
            else CU(cudaMemcpyAsync(dst[i], src[i], len[i], cudaMemcpyHostToDevice, st));
        }
    }
    // a record is at least two bytes of text (FASTA ">\n"), four in FASTQ ("@\n+\n")
    const uint64_t cap_seqs = (reads_bytes + mates_bytes) / (format == DYIBF_FMT_FASTQ ? 4 : 2) + 2;
    CU(t->bytes.reserve(reads_bytes + mates_bytes + 32));
    CU(t->offsets.reserve((cap_seqs + 2) * 8));
    uint8_t* d_seqs = static_cast<uint8_t*>(t->bytes.p);
    uint64_t* d_lim = static_cast<uint64_t*>(t->offsets.p);
    uint64_t n1 = 0, b1 = 0, n2 = 0, b2 = 0;
    rc = dyibf_parse_records(f->prm.device, d_text, reads_bytes, format, DYIBF_ALPHA_DNA5, d_seqs, reads_bytes + 16, d_lim, cap_seqs, 0,
                             &n1, &b1, DYIBF_MEM_DEVICE, st);
    if (rc) return rc;
    if (paired) {
        rc = dyibf_parse_records(f->prm.device, d_text + off2, mates_bytes, format, DYIBF_ALPHA_DNA5, d_seqs + b1, mates_bytes + 16,
                                 d_lim + n1, cap_seqs - n1, b1, &n2, &b2, DYIBF_MEM_DEVICE, st);
        if (rc) return rc;
        if (n2 != n1) return fail(DYIBF_ERR_ARG, "dyibf_classify_text: %llu reads but %llu mates", (unsigned long long)n1, (unsigned long long)n2);
    }
    f->kernel_launches += paired ? 10 : 5;
    const uint64_t n_fwd = n1 + n2, total_bytes = b1 + b2;
    if (2 * n_fwd > 0xffffffffull) return fail(DYIBF_ERR_ARG, "too many sequences in one batch");
    // the batch's threshold, d_mapper.h:313-315: average length over the whole read set (the rc half doubles both sums)
    const uint16_t avg = (uint16_t)(n_fwd ? total_bytes / n_fwd : 0);
    const uint32_t thr = dyibf_threshold(e_percent, f->prm.kmer, avg);
    const uint32_t group = paired ? 4 : 2;
    const uint32_t n_rows = (uint32_t)(2 * n_fwd / group);
    if (n_reads) *n_reads = n_rows;
    if (threshold) *threshold = thr;
    t->n_rows = n_rows;
    if (n_rows == 0) {
        CU(cudaMemsetAsync(t->csr_off.p, 0, (uint64_t)(bins + 1) * 8, st));
    } else {
        CU(t->rows.reserve((uint64_t)n_rows * W * 8 + 8));
        CU(t->sl.ctr.reserve(256));
        unsigned long long* d_max = reinterpret_cast<unsigned long long*>(static_cast<char*>(t->sl.ctr.p) + 128);
        CU(cudaMemsetAsync(d_max, 0, 8, st));
        CU(launch_max_len(d_lim, n_fwd, d_max, st));
        f->kernel_launches += 1;
        unsigned long long max_len = 0;
        CU(cudaMemcpyAsync(&max_len, d_max, 8, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        if (max_len >= f->hp.k && max_len - f->hp.k + 1 > 65535)
            return fail(DYIBF_ERR_ARG, "a sequence of %llu bases has more than 65535 k-mers (uint16_t in the reference)", max_len);
        PackSource src; memset(&src, 0, sizeof src);
        src.n_ranges = 1; src.cnt = (uint32_t)n_fwd; src.bytes[0] = d_seqs; src.offsets[0] = d_lim;
        rc = query_on_device(f, t->sl, src, total_bytes, DYIBF_ENC_ASCII, group, 1, thr, (uint32_t)std::max<unsigned long long>(max_len, 1),
                             (uint64_t*)t->rows.p, nullptr, st);
        if (rc) return rc;
        rc = enqueue_compaction(t, n_rows);
        if (rc) return rc;
        t->batch.valid = true; t->batch.n_fwd = (uint32_t)n_fwd; t->batch.group = group; t->batch.encoding = DYIBF_ENC_ASCII;
        t->batch.first_off = 0;
    }
    rc = enqueue_finish(t);
    if (rc) return rc;
    drain.release();
    *ticket = reinterpret_cast<dyibf_ticket*>(t);
    return DYIBF_OK;
}

/* the synchronous form: one ticket, begun and ended */
int dyibf_classify_text(dyibf_filter* f, const char* reads_text, uint64_t reads_bytes, const char* mates_text, uint64_t mates_bytes,
                        int format, int e_percent, uint64_t* bin_offsets, uint32_t* read_ids, uint64_t capacity, uint64_t* total,
                        uint32_t* n_reads, uint32_t* threshold)
{
    if (!f || !bin_offsets || (capacity && !read_ids)) return fail(DYIBF_ERR_ARG, "dyibf_classify_text: NULL argument");
    dyibf_ticket* t = nullptr;
    int rc = dyibf_classify_text_begin(f, reads_text, reads_bytes, mates_text, mates_bytes, format, e_percent, n_reads, threshold, &t);
    if (rc) return rc;
    rc = dyibf_classify_end(t, bin_offsets, read_ids, capacity, total);
    if (rc == DYIBF_ERR_CAPACITY) {
        // the lists stay in HBM for dyibf_classify_fetch: hand the ticket's batch to the handle with a sizing call
        std::string keep = last_error_string();
        Ticket* tk = reinterpret_cast<Ticket*>(t);
        std::vector<uint32_t> tmp((size_t)std::max<uint64_t>(tk->h_off[f->prm.bins], 1));
        std::vector<uint64_t> off((size_t)f->prm.bins + 1);
        uint64_t tot = 0;
        int rc2 = dyibf_classify_end(t, off.data(), tmp.data(), tmp.size(), &tot);
        if (rc2) return rc2;
        set_last_error(keep);
    }
    return rc;
}

int dyibf_classify_poll(dyibf_ticket* ticket, int* ready)
{
    Ticket* t = reinterpret_cast<Ticket*>(ticket);
    if (!t || !t->busy || !ready) return fail(DYIBF_ERR_ARG, "dyibf_classify_poll: not a ticket in flight");
    cudaError_t e = cudaEventQuery(t->done);
    if (e != cudaSuccess && e != cudaErrorNotReady) return fail(DYIBF_ERR_CUDA, "dyibf_classify_poll: %s", cudaGetErrorString(e));
    *ready = e == cudaSuccess;
    return DYIBF_OK;
}

int dyibf_classify_end(dyibf_ticket* ticket, uint64_t* bin_offsets, uint32_t* read_ids, uint64_t capacity, uint64_t* total)
{
    Ticket* t = reinterpret_cast<Ticket*>(ticket);
    if (!t || !t->busy) return fail(DYIBF_ERR_ARG, "dyibf_classify_end: not a ticket in flight");
    if (!bin_offsets || (capacity && !read_ids)) return fail(DYIBF_ERR_ARG, "NULL argument");
    dyibf_filter* f = t->f;
    DeviceGuard guard(f->prm.device);
    const uint32_t bins = f->prm.bins;
    cudaStream_t st = t->sl.st;
    cudaError_t e = cudaEventSynchronize(t->done);
    if (e != cudaSuccess) { t->busy = false; return fail(DYIBF_ERR_CUDA, "dyibf_classify_end: %s", cudaGetErrorString(e)); }
    memcpy(bin_offsets, t->h_off, (size_t)(bins + 1) * 8);
    const uint64_t tot = t->h_off[bins];
    if (total) *total = tot;
    if (tot > capacity)                                               // the ticket stays in flight: call again with room
        return fail(DYIBF_ERR_CAPACITY, "%llu (read,bin) pairs exceed the capacity of %llu", (unsigned long long)tot, (unsigned long long)capacity);
    if (tot > t->ids_cap) {                                           // rare: more lists than the first fill had room for
        CU(t->csr_ids.reserve(tot * 4));
        t->ids_cap = t->csr_ids.cap / 4;
        Scope sc(f, st, K_COMPACT);
        CU(launch_compact_fill((const uint64_t*)t->rows.p, t->n_rows, bins, (const uint64_t*)t->compact_ws.p,
                               (uint32_t*)t->csr_ids.p, t->ids_cap, 0, st));
    }
    if (tot) CU(cudaMemcpyAsync(read_ids, t->csr_ids.p, tot * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    // the ticket's batch becomes the handle's resident batch (buffers change hands, nothing is copied)
    f->h_csr_off.assign(t->h_off, t->h_off + bins + 1);
    f->csr_off.swap(t->csr_off); f->csr_ids.swap(t->csr_ids);
    f->batch_bytes.swap(t->bytes); f->batch_offsets.swap(t->offsets);
    f->batch = t->batch;
    t->busy = false;
    return DYIBF_OK;
}

void dyibf_classify_cancel(dyibf_ticket* ticket)
{
    Ticket* t = reinterpret_cast<Ticket*>(ticket);
    if (!t || !t->busy) return;
    DeviceGuard guard(t->f->prm.device);
    cudaStreamSynchronize(t->sl.st);
    t->busy = false;
}

} // extern "C"
