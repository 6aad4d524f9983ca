// dyibf_io.cu — the .filter loader and writer: disk <-> pinned host memory <-> HBM, several reader threads.
//
// Reference code replaced: the file constructors of SeqAnBloomFilter (src/d_bloom_filter.h:59-89: sdsl::load_from_file
// into host RAM, then _getMetadata) and save (:150-154: _setMetadata + sdsl::store_to_file).  The mapper times this stage
// as "Filter loading time" (src/d_mapper.h:913-921) and it is half of the reference's own test run.
//
// On-disk format (sdsl::int_vector<1>::serialize, include/sdsl-lite/include/sdsl/int_vector.hpp:1463-1472,
// write_header :534-544; trailer d_bloom_filter.h:255-277):
//   u64 (1<<56 | nbits) ; ceil(nbits/64) little-endian u64 words ; the last four words are [bins, h, k, 0].
// In HBM only the P*W data words are kept; the (all-zero) gap before the trailer and the trailer itself are
// synthesised when the filter is downloaded or saved, so a saved file is byte-identical to the reference's.
//
// Pipeline: T file threads (default: half of the host's cores, 4..16) pread chunks (4 MiB) of the data words into a ring of 2 T pinned buffers (allocated
// once per process); ONE thread issues the copies in file order: cudaMemcpyAsync(H2D) -> (replicas: cudaMemcpyPeerAsync
// device 0 -> the others over NVLink, on the same stream), so that file reads, PCIe and NVLink all overlap.  A bin shard
// (device r of G owns word columns [r W/G, (r+1) W/G) of every block) is a pitched copy out of the same pinned chunk.
// Saving runs the pipeline backwards: the one thread issues the downloads, the file threads wait for a chunk's event and pwrite it.
#include "dyibf_handle.h"

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <thread>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/statvfs.h>
#include <unistd.h>

using namespace dyibf;

namespace dyibf {

static std::atomic<int> g_io_threads{0};            // 0 = not set: half of the host's cores, between 4 and 16

// Pinned staging for the transfer threads: allocated once per process and kept (cudaMallocHost costs ~0.6 s per GiB, as
// much as the transfer of several GiB).  2 buffers x IO_CHUNK_BYTES per thread slot.
struct StagePool {
    std::mutex m;
    std::vector<void*> buf;       // 2 per slot
    size_t each = 0;
    ~StagePool() {}               // process exit frees it; the CUDA runtime may already be gone
    // buffers [2 t, 2 t + 1] of thread slot t, at least `bytes` each
    int get(int slots, size_t bytes, std::vector<void*>* out) {
        std::lock_guard<std::mutex> l(m);
        if (bytes > each) {
            for (void* p : buf) cudaFreeHost(p);
            buf.clear(); each = bytes;
        }
        while ((int)buf.size() < 2 * slots) {
            void* p = nullptr;
            if (cudaMallocHost(&p, each) != cudaSuccess) return DYIBF_ERR_NOMEM;
            buf.push_back(p);
        }
        out->assign(buf.begin(), buf.begin() + 2 * slots);
        return DYIBF_OK;
    }
};
static StagePool g_stage;
static std::mutex g_io_mutex;     // one transfer at a time uses the pool
static thread_local dyibf_io_stats g_last_io = {0, 0.0, 0.0, 0.0, 0, 0, 0.0, 0.0, 0.0, 0};

static double now_s()
{
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

int read_file_header(FILE* fp, const char* path, FileHeader* fh)
{
    uint64_t header = 0;
    if (fread(&header, 8, 1, fp) != 1) return fail(DYIBF_ERR_IO, "File \"%s\" could not be read.", path);
    if ((header >> 56) != 1) return fail(DYIBF_ERR_FORMAT, "File \"%s\" is not an sdsl bit_vector (width byte %u)", path, (unsigned)(header >> 56));
    fh->nbits = header & ((1ULL << 56) - 1);
    if (fh->nbits < META_BITS + 64) return fail(DYIBF_ERR_FORMAT, "File \"%s\": %llu bits is too small for a filter", path, (unsigned long long)fh->nbits);
    if (fseeko(fp, 0, SEEK_END) != 0) return fail(DYIBF_ERR_IO, "File \"%s\" could not be read.", path);
    fh->file_size = (uint64_t)ftello(fp);
    uint64_t n_words = (fh->nbits + 63) >> 6;
    if (fh->file_size < 8 + n_words * 8) return fail(DYIBF_ERR_FORMAT, "File \"%s\" is truncated: %llu bytes for %llu bits", path,
                                                     (unsigned long long)fh->file_size, (unsigned long long)fh->nbits);
    // _getMetadata, d_bloom_filter.h:255-265: three 64-bit reads at bit nbits-256 (+64, +128); they are
    // word-aligned iff nbits is a multiple of 64, otherwise they straddle words (sdsl bits::read_int)
    uint64_t m = fh->nbits - META_BITS;
    uint64_t w0 = m >> 6, sh = m & 63;
    uint64_t buf[5] = {0, 0, 0, 0, 0};
    size_t want = (size_t)std::min<uint64_t>(5, n_words - w0);
    if (fseeko(fp, (off_t)(8 + w0 * 8), SEEK_SET) != 0 || fread(buf, 8, want, fp) != want)
        return fail(DYIBF_ERR_IO, "File \"%s\" could not be read.", path);
    auto get = [&](int i) { return sh ? (buf[i] >> sh) | (buf[i + 1] << (64 - sh)) : buf[i]; };
    fh->bins = (uint32_t)get(0);           // truncation as the members have it (u32, u8, u8)
    fh->h = (uint8_t)get(1);
    fh->k = (uint8_t)get(2);
    return DYIBF_OK;
}

// One destination (or source) of the word stream: the words [col0, col0 + Wr) of every W-word block of the file live
// contiguously (Wr words per block) at d_words on `device`.  Wr == W: the whole table.
struct IoTarget { int device; uint64_t* d_words; uint32_t col0, Wr; };

struct IoResult { int rc = DYIBF_OK; std::string err; double file_s = 0; };

// blocks per chunk so that a chunk is about `want` bytes and holds whole blocks
static uint64_t blocks_per_chunk(uint32_t W, uint64_t want) { return std::max<uint64_t>(1, want / ((uint64_t)W * 8)); }

static std::atomic<int> g_io_chunk_mib{0};          // 0 = not set: DYIBF_IO_CHUNK_MIB from the environment, else 4
static uint64_t io_chunk_bytes()
{
    int mib = g_io_chunk_mib.load();
    if (mib <= 0) {
        const char* e = getenv("DYIBF_IO_CHUNK_MIB");
        mib = e ? atoi(e) : 0;
        if (mib < 1 || mib > 1024) mib = 4;
    }
    return (uint64_t)mib << 20;
}

// The ring of pinned chunks between the file threads and the one thread that talks to CUDA.  Slot c % N serves chunk c:
//   load:  a reader waits until the slot is free for chunk c, preads into it, marks it filled; the submitter (the calling
//          thread) takes the chunks in order, issues the copies, and frees the slot again for chunk c + N once the copies
//          of chunk c have finished;
//   save:  the submitter issues the download of chunk c into the slot and marks it filled; a writer waits for the copy's
//          event, pwrites and frees the slot.
// One submitting thread, because copies issued from many threads at once serialise inside the driver (measured: 8 threads
// each spent 2.7 ms per 8 MiB chunk inside cudaMemcpyAsync / cudaEventRecord, the copy engine idle most of the time).
struct ChunkRing {
    std::mutex m;
    std::condition_variable cv;
    std::vector<int64_t> filled;     // filled[slot] == c + 1: chunk c is in the slot (load: read; save: download issued)
    std::vector<int64_t> free_for;   // the slot may be used for chunk free_for[slot]
    bool abort = false;
    explicit ChunkRing(int n) : filled(n, 0), free_for(n) { for (int i = 0; i < n; ++i) free_for[i] = i; }
    bool wait_free(int slot, int64_t c) {
        std::unique_lock<std::mutex> l(m);
        cv.wait(l, [&] { return abort || free_for[slot] == c; });
        return !abort;
    }
    bool wait_filled(int slot, int64_t c) {
        std::unique_lock<std::mutex> l(m);
        cv.wait(l, [&] { return abort || filled[slot] == c + 1; });
        return !abort;
    }
    void set_filled(int slot, int64_t c) { { std::lock_guard<std::mutex> l(m); filled[slot] = c + 1; } cv.notify_all(); }
    void set_free(int slot, int64_t next_chunk) { { std::lock_guard<std::mutex> l(m); free_for[slot] = next_chunk; } cv.notify_all(); }
    void stop() { { std::lock_guard<std::mutex> l(m); abort = true; } cv.notify_all(); }
};

// file bytes [8, 8 + n_blocks * W * 8) <-> the targets.  to_device: load; otherwise save (exactly one whole-table target
// per column range: every word of the file is written once).
// fd < 0: the other end is host memory (`host`, laid out like the file's data words) instead of a file.
static int stream_words(int fd, const char* path, char* host, bool to_device, uint64_t n_blocks, uint32_t W,
                        const std::vector<IoTarget>& targets, bool replicas, dyibf_io_stats* stats)
{
    std::lock_guard<std::mutex> io_lock(g_io_mutex);
    const double t_begin = now_s();
    const uint64_t chunk_bytes = io_chunk_bytes();
    const uint64_t bpc = blocks_per_chunk(W, chunk_bytes);
    const int64_t n_chunks = (int64_t)((n_blocks + bpc - 1) / bpc);
    int T = g_io_threads.load();
    if (T < 1) T = (int)std::min(16u, std::max(4u, std::thread::hardware_concurrency() / 2));   // 8 on a 16-core box: 50.6 GB/s (16 threads: 43.6)
    if ((int64_t)T > n_chunks) T = (int)std::max<int64_t>(n_chunks, 1);
    const int N = 2 * T;                                              // slots: T being read or written, T in the copy queue
    const size_t buf_bytes = (size_t)(bpc * W * 8);
    std::vector<void*> pool;
    if (g_stage.get(T, std::max<size_t>(buf_bytes, (size_t)chunk_bytes), &pool) != DYIBF_OK) return fail(DYIBF_ERR_NOMEM, "filter I/O: pinned staging allocation failed");
    const double t_pool = now_s();
    const size_t nt = targets.size();
    // replicas: one stream on the first device drives the H2D and the peer copies; shards: one stream per device
    const size_t n_streams = replicas ? 1 : nt;
    std::vector<cudaStream_t> st(n_streams, nullptr);
    std::vector<cudaEvent_t> ev((size_t)N * n_streams, nullptr);       // ev[slot * n_streams + i]: the copies of the slot's chunk on stream i
    int prev = -1; cudaGetDevice(&prev);
    IoResult sub;                                                     // the submitter's own result
    auto bad = [](IoResult& r, int code, const std::string& what) { if (!r.rc) { r.rc = code; r.err = what; } };
    for (size_t i = 0; i < n_streams && !sub.rc; ++i) {
        if (cudaSetDevice(targets[i].device) != cudaSuccess || cudaStreamCreateWithFlags(&st[i], cudaStreamNonBlocking) != cudaSuccess)
            bad(sub, DYIBF_ERR_CUDA, "filter I/O: stream creation failed");
        for (int s = 0; s < N && !sub.rc; ++s)
            if (cudaEventCreateWithFlags(&ev[(size_t)s * n_streams + i], cudaEventDisableTiming) != cudaSuccess)
                bad(sub, DYIBF_ERR_CUDA, "filter I/O: event creation failed");
    }
    ChunkRing ring(N);
    std::atomic<int64_t> next{0};
    std::vector<IoResult> res(T);
    auto chunk_geom = [&](int64_t c, uint64_t* blk0, uint64_t* nb) { *blk0 = (uint64_t)c * bpc; *nb = std::min(bpc, n_blocks - *blk0); };

    // ---- file side: T threads ----
    auto file_worker = [&](int t) {
        IoResult& r = res[t];
        for (;;) {
            const int64_t c = next.fetch_add(1);
            if (c >= n_chunks) break;
            const int slot = (int)(c % N);
            uint64_t blk0, nb; chunk_geom(c, &blk0, &nb);
            const size_t bytes = (size_t)(nb * W * 8);
            const off_t file_off = (off_t)(8 + blk0 * W * 8);
            char* buf = static_cast<char*>(pool[slot]);
            if (to_device) {
                if (!ring.wait_free(slot, c)) break;
                const double t0 = now_s();
                size_t got = 0;
                if (fd < 0) { memcpy(buf, host + blk0 * W * 8, bytes); got = bytes; }
                while (got < bytes) {
                    ssize_t n = pread(fd, buf + got, bytes - got, file_off + (off_t)got);
                    if (n <= 0) { bad(r, DYIBF_ERR_IO, std::string("File \"") + path + "\" could not be read."); break; }
                    got += (size_t)n;
                }
                r.file_s += now_s() - t0;
                if (r.rc) { ring.stop(); break; }
                ring.set_filled(slot, c);
            } else {
                if (!ring.wait_filled(slot, c)) break;
                for (size_t i = 0; i < n_streams; ++i)
                    if (cudaEventSynchronize(ev[(size_t)slot * n_streams + i]) != cudaSuccess) bad(r, DYIBF_ERR_CUDA, "filter download: copy failed");
                const double t0 = now_s();
                size_t put = 0;
                if (fd < 0) { memcpy(host + blk0 * W * 8, buf, bytes); put = bytes; }
                while (put < bytes && !r.rc) {
                    ssize_t n = pwrite(fd, buf + put, bytes - put, file_off + (off_t)put);
                    if (n <= 0) { bad(r, DYIBF_ERR_IO, std::string("File \"") + path + "\" could not be written."); break; }
                    put += (size_t)n;
                }
                r.file_s += now_s() - t0;
                if (r.rc) { ring.stop(); break; }
                ring.set_free(slot, c + N);
            }
        }
    };
    std::vector<std::thread> th;
    if (!sub.rc) for (int t = 0; t < T; ++t) th.emplace_back(file_worker, t);

    // ---- CUDA side: this thread ----
    double wait_s = 0, submit_s = 0, starve_s = 0;
    auto copy_chunk = [&](int64_t c) -> cudaError_t {
        const int slot = (int)(c % N);
        uint64_t blk0, nb; chunk_geom(c, &blk0, &nb);
        const size_t bytes = (size_t)(nb * W * 8);
        char* buf = static_cast<char*>(pool[slot]);
        cudaError_t e = cudaSuccess;
        if (replicas) {
            const IoTarget& a = targets[0];
            e = cudaSetDevice(a.device);
            char* dst0 = (char*)a.d_words + blk0 * W * 8;
            if (e == cudaSuccess) e = cudaMemcpyAsync(dst0, buf, bytes, cudaMemcpyHostToDevice, st[0]);
            for (size_t i = 1; i < nt && e == cudaSuccess; ++i)                         // NVLink fan-out of the chunk
                e = cudaMemcpyPeerAsync((char*)targets[i].d_words + blk0 * W * 8, targets[i].device, dst0, a.device, bytes, st[0]);
            // the slot is reusable once the H2D has read it, but the event also orders the end of the call: after the fan-out
            if (e == cudaSuccess) e = cudaEventRecord(ev[(size_t)slot * n_streams], st[0]);
            return e;
        }
        for (size_t i = 0; i < nt && e == cudaSuccess; ++i) {
            const IoTarget& a = targets[i];
            e = cudaSetDevice(a.device);
            if (e != cudaSuccess) break;
            const cudaMemcpyKind kind = to_device ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost;
            char* dev_p = (char*)a.d_words + blk0 * a.Wr * 8;
            if (a.Wr == W)
                e = to_device ? cudaMemcpyAsync(dev_p, buf, bytes, kind, st[i]) : cudaMemcpyAsync(buf, dev_p, bytes, kind, st[i]);
            else if (to_device)
                e = cudaMemcpy2DAsync(dev_p, (size_t)a.Wr * 8, buf + (size_t)a.col0 * 8, (size_t)W * 8, (size_t)a.Wr * 8, (size_t)nb, kind, st[i]);
            else
                e = cudaMemcpy2DAsync(buf + (size_t)a.col0 * 8, (size_t)W * 8, dev_p, (size_t)a.Wr * 8, (size_t)a.Wr * 8, (size_t)nb, kind, st[i]);
            if (e == cudaSuccess) e = cudaEventRecord(ev[(size_t)slot * n_streams + i], st[i]);
        }
        return e;
    };
    for (int64_t c = 0; c < n_chunks && !sub.rc; ++c) {
        const int slot = (int)(c % N);
        double t0 = now_s();
        const bool go = to_device ? ring.wait_filled(slot, c) : ring.wait_free(slot, c);
        starve_s += now_s() - t0;
        if (!go) break;
        t0 = now_s();
        cudaError_t e = copy_chunk(c);
        submit_s += now_s() - t0;
        if (e != cudaSuccess) { bad(sub, DYIBF_ERR_CUDA, std::string(to_device ? "filter upload: " : "filter download: ") + cudaGetErrorString(e)); ring.stop(); break; }
        if (!to_device) { ring.set_filled(slot, c); continue; }
        // load: the slot of chunk c - T goes back to the readers once its copies are done
        const int64_t done = c - T;
        if (done >= 0) {
            t0 = now_s();
            for (size_t i = 0; i < n_streams; ++i) cudaEventSynchronize(ev[(size_t)(done % N) * n_streams + i]);
            wait_s += now_s() - t0;
            ring.set_free((int)(done % N), done + N);
        }
    }
    if (sub.rc) ring.stop();
    for (auto& x : th) x.join();
    double t0 = now_s();
    for (size_t i = 0; i < n_streams; ++i)
        if (st[i]) {
            cudaSetDevice(targets[i].device);
            if (cudaStreamSynchronize(st[i]) != cudaSuccess) bad(sub, DYIBF_ERR_CUDA, "filter I/O: copy failed");
            cudaStreamDestroy(st[i]);
        }
    wait_s += now_s() - t0;
    for (cudaEvent_t e : ev) if (e) cudaEventDestroy(e);
    if (prev >= 0) cudaSetDevice(prev);
    double file_s = 0;
    for (auto& r : res) file_s = std::max(file_s, r.file_s);
    if (stats) {
        stats->bytes = n_blocks * W * 8; stats->seconds = now_s() - t_begin; stats->file_seconds = file_s;
        stats->copy_wait_seconds = wait_s; stats->threads = T; stats->devices = (int)targets.size();
        stats->submit_seconds = submit_s; stats->starve_seconds = starve_s; stats->pool_seconds = t_pool - t_begin;
        stats->chunk_bytes = buf_bytes;
    }
    if (sub.rc) return fail(sub.rc, "%s", sub.err.c_str());
    for (auto& r : res) if (r.rc) return fail(r.rc, "%s", r.err.c_str());
    return DYIBF_OK;
}

// load the data words of an open, validated .filter file into the targets
int io_load_words(const char* path, uint64_t n_blocks, uint32_t W, const int* devices, uint64_t* const* d_words, int n,
                  bool bin_sharded)
{
    int fd = open(path, O_RDONLY);
    if (fd < 0) return fail(DYIBF_ERR_IO, "File \"%s\" could not be read.", path);
#ifdef POSIX_FADV_SEQUENTIAL
    posix_fadvise(fd, 0, 0, POSIX_FADV_SEQUENTIAL);
#endif
    std::vector<IoTarget> tg;
    const uint32_t Wr = bin_sharded ? W / (uint32_t)n : W;
    for (int i = 0; i < n; ++i) tg.push_back({devices[i], d_words[i], bin_sharded ? (uint32_t)i * Wr : 0u, Wr});
    dyibf_io_stats st; memset(&st, 0, sizeof st);
    int rc = stream_words(fd, path, nullptr, true, n_blocks, W, tg, !bin_sharded, &st);
    close(fd);
    g_last_io = st;
    return rc;
}

// write a complete .filter file: header, data words gathered from the sources, zero gap + trailer.
// The data words go through a shared mapping of the file where that is possible: write(2) on one file is serialised by the
// inode lock (tmpfs and ext4 alike: 8 pwrite threads moved 3.2 GB/s in total), while page faults on a mapping are served in
// parallel, so T threads copying out of the pinned chunks scale.  A mapping cannot report ENOSPC (the process would get
// SIGBUS), so the free space of the file system is checked first; DYIBF_IO_MMAP=0 forces pwrite.
int io_save_words(const char* path, const dyibf_filter* whole_params, uint64_t n_blocks, uint32_t W, const int* devices,
                  uint64_t* const* d_words, int n, bool bin_sharded)
{
    int fd = open(path, O_RDWR | O_CREAT | O_TRUNC, 0644);
    if (fd < 0) return fail(DYIBF_ERR_IO, "File \"%s\" could not be written.", path);
    const uint64_t header = (1ULL << 56) | whole_params->prm.nbits;
    const uint64_t data_words = n_blocks * W;
    const uint64_t tail_n = whole_params->prm.n_words - data_words;
    const uint64_t file_size = 8 + whole_params->prm.n_words * 8;
    std::vector<uint64_t> tail(tail_n);
    synth_tail(whole_params, data_words, tail_n, tail.data());
    bool ok = pwrite(fd, &header, 8, 0) == 8;
    ok = ok && (tail_n == 0 || pwrite(fd, tail.data(), tail_n * 8, (off_t)(8 + data_words * 8)) == (ssize_t)(tail_n * 8));
    int rc = DYIBF_OK;
    dyibf_io_stats st; memset(&st, 0, sizeof st);
    if (ok && data_words) {
        std::vector<IoTarget> tg;
        const uint32_t Wr = bin_sharded ? W / (uint32_t)n : W;
        const int n_src = bin_sharded ? n : 1;                       // replicas: every device holds the same words
        for (int i = 0; i < n_src; ++i) tg.push_back({devices[i], d_words[i], bin_sharded ? (uint32_t)i * Wr : 0u, Wr});
        char* map = nullptr;
        const char* env = getenv("DYIBF_IO_MMAP");
        struct statvfs vfs;
        if (!(env && atoi(env) == 0) && data_words * 8 >= (64ull << 20) && fstatvfs(fd, &vfs) == 0 &&
            (uint64_t)vfs.f_bavail * (uint64_t)vfs.f_frsize > file_size + (64ull << 20) && ftruncate(fd, (off_t)file_size) == 0) {
            void* m = mmap(nullptr, (size_t)file_size, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
            if (m != MAP_FAILED) {
                map = static_cast<char*>(m);
#ifdef MADV_HUGEPAGE
                madvise(m, (size_t)file_size, MADV_HUGEPAGE);          // where the file system honours it (tmpfs with shmem_enabled=advise): 512 x fewer faults
#endif
            }
        }
        rc = stream_words(map ? -1 : fd, path, map ? map + 8 : nullptr, false, n_blocks, W, tg, false, &st);
        if (map && munmap(map, (size_t)file_size) != 0) ok = false;
    }
    ok = (close(fd) == 0) && ok;
    g_last_io = st;
    if (rc) return rc;
    if (!ok) return fail(DYIBF_ERR_IO, "File \"%s\" could not be written.", path);
    return DYIBF_OK;
}

// words [first_block * W, (first_block + n_blocks) * W) of one table <-> host memory, through the threaded pinned pipeline
int io_copy_words(dyibf_filter* f, uint64_t* host_words, uint64_t first_block, uint64_t n_blocks, bool to_device)
{
    const uint32_t W = f->prm.words_per_block;
    std::vector<IoTarget> tg;
    tg.push_back({f->prm.device, f->d_words + first_block * W, 0u, W});
    dyibf_io_stats st; memset(&st, 0, sizeof st);
    int rc = stream_words(-1, "", reinterpret_cast<char*>(host_words), to_device, n_blocks, W, tg, false, &st);
    g_last_io = st;
    return rc;
}

static int load_impl(const char* path, bool expect, uint32_t bins, uint32_t n_hash, uint32_t kmer, uint64_t nbits,
                     int device, dyibf_filter** out)
{
    if (!path || !out) return fail(DYIBF_ERR_ARG, "NULL argument");
    *out = nullptr;
    FILE* fp = fopen(path, "rb");
    if (!fp) return fail(DYIBF_ERR_IO, "File \"%s\" could not be read.", path);
    FileHeader fh; int rc = read_file_header(fp, path, &fh);
    fclose(fp);
    if (rc) return rc;
    if (expect) {
        if (fh.nbits != nbits) return fail(DYIBF_ERR_SIZE_MISMATCH, "Size mismatch: \n\t Loaded file \t%llu bits\n\t expected\t%llu",
                                           (unsigned long long)fh.nbits, (unsigned long long)nbits);
    } else { bins = fh.bins; n_hash = fh.h; kmer = fh.k; nbits = fh.nbits; }
    dyibf_filter* f = nullptr;
    rc = new_handle(bins, n_hash, kmer, nbits, device, nullptr, false, &f);
    if (rc) return rc;
    uint64_t* dw = f->d_words;
    rc = io_load_words(path, f->prm.n_blocks, f->prm.words_per_block, &device, &dw, 1, false);
    if (rc) { std::string keep = last_error_string(); dyibf_destroy(f); set_last_error(keep); return rc; }
    *out = f;
    return DYIBF_OK;
}

} // namespace dyibf

extern "C" {

int dyibf_load_file(const char* path, int device, dyibf_filter** out)
{
    return load_impl(path, false, 0, 0, 0, 0, device, out);
}

int dyibf_load_file_expect(const char* path, uint32_t bins, uint32_t n_hash, uint32_t kmer, uint64_t nbits, int device,
                           dyibf_filter** out)
{
    return load_impl(path, true, bins, n_hash, kmer, nbits, device, out);
}

int dyibf_save_file(dyibf_filter* f, const char* path)
{
    if (!f || !path) return fail(DYIBF_ERR_ARG, "NULL argument");
    DeviceGuard guard(f->prm.device);
    CU(cudaDeviceSynchronize());                                      // builds still in flight on caller streams
    trim_build_scratch(f);
    uint64_t* dw = f->d_words; int dev = f->prm.device;
    return io_save_words(path, f, f->prm.n_blocks, f->prm.words_per_block, &dev, &dw, 1, false);
}

int dyibf_set_io_threads(int n)
{
    if (n < 0 || n > 64) return fail(DYIBF_ERR_ARG, "I/O threads must be 1..64 (0 = default)");
    g_io_threads.store(n);
    return DYIBF_OK;
}

int dyibf_set_io_chunk_mib(int mib)
{
    if (mib < 0 || mib > 1024) return fail(DYIBF_ERR_ARG, "I/O chunk size must be 0 (default) .. 1024 MiB");
    g_io_chunk_mib.store(mib);
    return DYIBF_OK;
}

int dyibf_last_io_stats(dyibf_io_stats* out)
{
    if (!out) return fail(DYIBF_ERR_ARG, "NULL argument");
    *out = g_last_io;
    return DYIBF_OK;
}

} // extern "C"
