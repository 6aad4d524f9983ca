// dyibf_hostcopy.cu — getting a caller's PAGEABLE memory onto the device at more than the driver's ~10 GB/s.
//
// cudaMemcpyAsync from pageable memory is staged by the driver on the calling thread: ~10 GB/s on the B200 box, and the
// call holds the host for the whole copy.  The reference's containers (SeqAn strings, src/d_bloom_filter.h:103 addKmers'
// text, the mapper's read batches src/mapper.h:109) are all pageable, so the drop-in boundary sees nothing else.  Here a
// few helper threads copy the range into pinned staging (memory-bandwidth bound, several times faster) and the DMA engine
// takes it from there at PCIe rate.
//
//   host_copy_streaming : memcpy with non-temporal stores.  The destination is pinned staging that only the DMA engine
//                         reads next; lines left dirty in the private caches of several cores made that DMA three times
//                         slower (17 instead of 51 GB/s, measured: tools/debug/async_probe.py), so the stores bypass them.
//   host_copy_parallel  : the same, split over up to 8 threads (persistent helpers; the caller copies a share too).
//   staged_h2d          : a large pageable range -> device memory through a process-wide ring of pinned pieces; piece i+1
//                         is being copied by the threads while piece i is on the bus.
#include "dyibf_handle.h"

#include <condition_variable>
#include <thread>

#include <emmintrin.h>

namespace dyibf {

void host_copy_streaming(void* dst, const void* src, size_t n)
{
    char* d = static_cast<char*>(dst); const char* s = static_cast<const char*>(src);
    while (n && (reinterpret_cast<uintptr_t>(d) & 15)) { *d++ = *s++; --n; }
    for (; n >= 64; n -= 64, d += 64, s += 64) {
        const __m128i a = _mm_loadu_si128(reinterpret_cast<const __m128i*>(s));
        const __m128i b = _mm_loadu_si128(reinterpret_cast<const __m128i*>(s + 16));
        const __m128i c = _mm_loadu_si128(reinterpret_cast<const __m128i*>(s + 32));
        const __m128i e = _mm_loadu_si128(reinterpret_cast<const __m128i*>(s + 48));
        _mm_stream_si128(reinterpret_cast<__m128i*>(d), a);
        _mm_stream_si128(reinterpret_cast<__m128i*>(d + 16), b);
        _mm_stream_si128(reinterpret_cast<__m128i*>(d + 32), c);
        _mm_stream_si128(reinterpret_cast<__m128i*>(d + 48), e);
    }
    while (n) { *d++ = *s++; --n; }
    _mm_sfence();
}

namespace {

class HostCopier {
public:
    static const int N = 7;                                          // helpers; the caller copies a share too
    static HostCopier& get() { static HostCopier* c = new HostCopier(); return *c; }   // never destroyed: its threads outlive main
    void copy(void* dst, const void* src, size_t n, int threads) {
        const int helpers = std::max(0, std::min(N, threads - 1));
        if (n < (1u << 20) || helpers == 0) { host_copy_streaming(dst, src, n); return; }
        std::unique_lock<std::mutex> call(call_m_);                  // one parallel copy at a time
        const size_t part = (n / (helpers + 1) + 63) & ~size_t(63);
        {
            std::lock_guard<std::mutex> l(m_);
            for (int i = 0; i < N; ++i) {
                const size_t a = std::min(n, part * (i + 1)), b = i < helpers ? (i + 1 == helpers ? n : std::min(n, part * (i + 2))) : a;
                job_[i] = {static_cast<char*>(dst) + a, static_cast<const char*>(src) + a, b - a};
            }
            pending_ = N; ++gen_;
        }
        cv_.notify_all();
        host_copy_streaming(dst, src, std::min(n, part));
        std::unique_lock<std::mutex> l(m_);
        done_.wait(l, [this] { return pending_ == 0; });
    }
private:
    struct Job { char* d; const char* s; size_t n; };
    HostCopier() { for (int i = 0; i < N; ++i) th_[i] = std::thread([this, i] { loop(i); }); for (auto& t : th_) t.detach(); }
    void loop(int i) {
        uint64_t seen = 0;
        for (;;) {
            Job j;
            {
                std::unique_lock<std::mutex> l(m_);
                cv_.wait(l, [&] { return gen_ != seen; });
                seen = gen_; j = job_[i];
            }
            if (j.n) host_copy_streaming(j.d, j.s, j.n);
            { std::lock_guard<std::mutex> l(m_); if (--pending_ == 0) done_.notify_all(); }
        }
    }
    std::thread th_[N]; Job job_[N];
    std::mutex m_, call_m_; std::condition_variable cv_, done_;
    int pending_ = 0; uint64_t gen_ = 0;
};

// The ring of staged_h2d: PIECES pinned buffers of PIECE_BYTES, portable across devices, allocated on first use and kept
// (cudaMallocHost costs ~0.6 s per GiB).  An event per piece says when the DMA that reads it has finished; it belongs to
// the device it was created on and is re-created when another device's stream uses the piece.
struct StageRing {
    static const int PIECES = 4;
    static const size_t PIECE_BYTES = 32u << 20;
    std::mutex m;
    void* p[PIECES] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ev[PIECES] = {nullptr, nullptr, nullptr, nullptr};
    int ev_dev[PIECES] = {-1, -1, -1, -1};
    bool in_flight[PIECES] = {false, false, false, false};
    unsigned next = 0;
};
StageRing* ring() { static StageRing* r = new StageRing(); return r; }

}  // namespace

void host_copy_parallel(void* dst, const void* src, size_t n, int threads)
{
    HostCopier::get().copy(dst, src, n, threads);
}

bool is_pageable(const void* p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return true; }
    return a.type == cudaMemoryTypeUnregistered;
}

int staged_h2d(void* d_dst, const void* h_src, size_t n, cudaStream_t st)
{
    StageRing* r = ring();
    std::lock_guard<std::mutex> lock(r->m);
    int dev = 0;
    CU(cudaGetDevice(&dev));
    for (size_t off = 0; off < n;) {
        const size_t len = std::min(n - off, StageRing::PIECE_BYTES);
        const int i = (int)(r->next++ % StageRing::PIECES);
        if (!r->p[i]) CU(cudaHostAlloc(&r->p[i], StageRing::PIECE_BYTES, cudaHostAllocPortable));
        if (r->in_flight[i]) { CU(cudaEventSynchronize(r->ev[i])); r->in_flight[i] = false; }
        if (r->ev_dev[i] != dev) {
            if (r->ev[i]) cudaEventDestroy(r->ev[i]);
            r->ev[i] = nullptr; r->ev_dev[i] = -1;
            CU(cudaEventCreateWithFlags(&r->ev[i], cudaEventDisableTiming));
            r->ev_dev[i] = dev;
        }
        host_copy_parallel(r->p[i], static_cast<const char*>(h_src) + off, len, 8);
        CU(cudaMemcpyAsync(static_cast<char*>(d_dst) + off, r->p[i], len, cudaMemcpyHostToDevice, st));
        CU(cudaEventRecord(r->ev[i], st));
        r->in_flight[i] = true;
        off += len;
    }
    return DYIBF_OK;
}

}  // namespace dyibf
