// dyibf_group.cu — one process, several B200s: the dyibf_group_* entry points of include/dyibf.h.
//
// The reference's callers are single processes: spawnDisMapper constructs ONE filter (src/d_mapper.h:911-922) and
// clasifyLoadedReads splits ONE read batch over its workers (src/d_mapper.h:338-375); build_filter splits the bins over its
// threads (src/d_build_filter.cpp:227-260).  A group is that, with GPUs for workers:
//
//   read-sharded (DYIBF_GROUP_READ_SHARDED; filter fits one GPU): every device holds a replica of the filter (loaded once
//     and fanned out over NVLink peer copies); device r filters rows [r n/G, (r+1) n/G) of the caller's ONE host batch on
//     its own streams; the per-bin lists concatenate in device order, which keeps ids ascending (the --threads 1 order).
//   bin-sharded (DYIBF_GROUP_BIN_SHARDED; filter larger than one GPU's HBM): device r owns the word columns
//     [r W/G, (r+1) W/G) of EVERY hash block as a self-contained IBF with bins/G bins and the same P (hash positions depend
//     only on the k-mer); every device sees every read; the per-bin lists of the bin ranges concatenate.  No reduction:
//     the threshold is per bin.
//   build: sequences are routed to the device that owns their bin's word column, so no two devices ever OR into the same
//     word; replicas are then made identical again by a column exchange over NVLink peer memory (exchange_columns_kernel:
//     every device pulls the columns it does not own straight out of their owners' HBM).
//
// One host thread per device drives that device's member handle; the members are ordinary dyibf_filter handles.
#include "dyibf_handle.h"

#include <condition_variable>
#include <functional>
#include <thread>

using namespace dyibf;

namespace dyibf {
int io_load_words(const char* path, uint64_t n_blocks, uint32_t W, const int* devices, uint64_t* const* d_words, int n, bool bin_sharded);
int io_save_words(const char* path, const dyibf_filter* whole_params, uint64_t n_blocks, uint32_t W, const int* devices,
                  uint64_t* const* d_words, int n, bool bin_sharded);

static const int MAX_GROUP = 16;

struct PeerTables { uint64_t* words[MAX_GROUP]; };

// Replica consistency after a bin-routed build: device `me` owns word columns [me Wr, (me+1) Wr) of every block; all other
// columns are pulled from their owners' tables over NVLink.  One thread moves one 16-byte (VEC = 2 words) or 8-byte piece;
// consecutive threads move consecutive pieces of the table, so local stores are coalesced and the remote loads of one owner
// are contiguous runs of Wr words.
template <int VEC>
__global__ void __launch_bounds__(256) exchange_columns_kernel(PeerTables peers, uint32_t me, uint32_t W, uint32_t Wr, uint64_t n_pieces)
{
    const uint32_t ppb = W / VEC;                                    // pieces per block
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_pieces; i += stride) {
        const uint32_t w = (uint32_t)(i % ppb) * VEC;               // first word of the piece inside its block
        const uint32_t owner = w / Wr;
        if (owner == me) continue;
        if (VEC == 2) {
            const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(peers.words[owner] + i * 2);
            *reinterpret_cast<ulonglong2*>(peers.words[me] + i * 2) = v;
        } else {
            peers.words[me][i] = peers.words[owner][i];
        }
    }
}

// Read-sharded lists -> the caller's CSR when the caller's buffer is pinned (device-accessible) host memory: bin b's list of
// this member goes to out[pre[b] ...] straight over PCIe, one warp per bin, instead of a download into a staging buffer and
// 1024 host memcpys per member (the members' host threads all write one array: ~20 ms per 80 M-read batch on 8 GPUs).
__global__ void __launch_bounds__(256) scatter_lists_kernel(const uint32_t* __restrict__ ids, const uint64_t* __restrict__ off,
                                                            const uint64_t* __restrict__ pre, uint32_t bins, uint32_t* __restrict__ out)
{
    const uint32_t lane = threadIdx.x & 31, n_warps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; b < bins; b += n_warps) {
        const uint64_t a = off[b], c = off[b + 1] - a;
        uint32_t* dst = out + pre[b];
        for (uint64_t k = lane; k < c; k += 32) dst[k] = ids[a + k];
    }
}

class Pool {                                                          // one persistent host thread per device
public:
    explicit Pool(int n) : n_(n), busy_(0), gen_(0), stop_(false) {
        for (int i = 0; i < n; ++i) th_.emplace_back([this, i] { loop(i); });
    }
    ~Pool() {
        { std::lock_guard<std::mutex> l(m_); stop_ = true; ++gen_; }
        cv_.notify_all();
        for (auto& t : th_) t.join();
    }
    void run(const std::function<void(int)>& fn) {
        std::unique_lock<std::mutex> l(m_);
        fn_ = &fn; busy_ = n_; ++gen_;
        cv_.notify_all();
        done_.wait(l, [this] { return busy_ == 0; });
        fn_ = nullptr;
    }
private:
    void loop(int i) {
        uint64_t seen = 0;
        for (;;) {
            const std::function<void(int)>* fn;
            {
                std::unique_lock<std::mutex> l(m_);
                cv_.wait(l, [&] { return gen_ != seen; });
                seen = gen_;
                if (stop_) return;
                fn = fn_;
            }
            (*fn)(i);
            { std::lock_guard<std::mutex> l(m_); if (--busy_ == 0) done_.notify_all(); }
        }
    }
    int n_, busy_; uint64_t gen_; bool stop_;
    const std::function<void(int)>* fn_ = nullptr;
    std::mutex m_; std::condition_variable cv_, done_;
    std::vector<std::thread> th_;
};

} // namespace dyibf

struct dyibf_group {
    int n = 0, mode = DYIBF_GROUP_READ_SHARDED;
    std::vector<int> devices;
    std::vector<dyibf_filter*> member;
    dyibf_params whole;                       // the whole filter's parameters (device = devices[0])
    uint32_t Wr = 0;                          // word columns per device (bin-sharded; and the routing unit of a replicated build)
    bool peer_ok = false;                     // every device can address every other device's HBM
    bool fetched_ready = false;
    std::vector<uint32_t> first_bin, n_bins;  // bin range a member owns (bin-sharded) / builds (read-sharded)
    std::vector<void*> h_ids; std::vector<size_t> h_ids_cap;   // pinned list staging per member
    std::vector<int> rc; std::vector<std::string> err;
    std::unique_ptr<Pool> pool;
    int next_async = 0;
};

namespace dyibf {

static int first_error(dyibf_group* g)
{
    for (int i = 0; i < g->n; ++i)
        if (g->rc[i]) { set_last_error(g->err[i]); return g->rc[i]; }
    return DYIBF_OK;
}

static void run_members(dyibf_group* g, const std::function<int(int)>& fn)
{
    std::function<void(int)> job = [&](int i) {
        DeviceGuard guard(g->devices[i]);
        g->rc[i] = fn(i);
        g->err[i] = g->rc[i] ? last_error_string() : std::string();
    };
    if (g->n == 1) job(0); else g->pool->run(job);
}

static int group_alloc(const int* devices, int n, int mode, uint32_t bins, uint64_t nbits, dyibf_group** out, uint32_t W)
{
    if (!out) return fail(DYIBF_ERR_ARG, "out is NULL");
    *out = nullptr;
    if (!devices || n < 1 || n > MAX_GROUP) return fail(DYIBF_ERR_ARG, "a group needs 1..%d devices", MAX_GROUP);
    if (mode != DYIBF_GROUP_READ_SHARDED && mode != DYIBF_GROUP_BIN_SHARDED) return fail(DYIBF_ERR_ARG, "unknown group mode %d", mode);
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < i; ++j)
            if (devices[i] == devices[j]) return fail(DYIBF_ERR_ARG, "device %d listed twice", devices[i]);
    if (mode == DYIBF_GROUP_BIN_SHARDED) {
        if (W % (uint32_t)n) return fail(DYIBF_ERR_ARG, "bin-sharding needs the %u words per hash block divisible by the %d devices", W, n);
        if ((uint64_t)(n - 1) * (W / n) * 64 >= bins) return fail(DYIBF_ERR_ARG, "device %d would own no bins (%u bins over %d devices)", n - 1, bins, n);
    }
    int ndev = 0;
    CU(cudaGetDeviceCount(&ndev));
    for (int i = 0; i < n; ++i)
        if (devices[i] < 0 || devices[i] >= ndev) return fail(DYIBF_ERR_CUDA, "CUDA device %d not present (%d devices)", devices[i], ndev);
    dyibf_group* g = new dyibf_group();
    g->n = n; g->mode = mode; g->devices.assign(devices, devices + n);
    g->member.assign(n, nullptr); g->rc.assign(n, 0); g->err.assign(n, std::string());
    g->h_ids.assign(n, nullptr); g->h_ids_cap.assign(n, 0);
    g->first_bin.assign(n, 0); g->n_bins.assign(n, bins);
    g->Wr = (W % (uint32_t)n == 0) ? W / (uint32_t)n : 0;
    if (g->Wr) {
        for (int i = 0; i < n; ++i) {
            const uint64_t first = (uint64_t)i * g->Wr * 64;
            g->first_bin[i] = (uint32_t)std::min<uint64_t>(first, bins);
            g->n_bins[i] = (uint32_t)(std::min<uint64_t>(bins, first + (uint64_t)g->Wr * 64) - g->first_bin[i]);
        }
    }
    // peer access between every pair (NVLink / NVSwitch on a B200 box); "already enabled" is fine
    g->peer_ok = true;
    int prev = -1; cudaGetDevice(&prev);
    for (int i = 0; i < n; ++i) {
        cudaSetDevice(devices[i]);
        for (int j = 0; j < n; ++j) {
            if (i == j) continue;
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, devices[i], devices[j]) != cudaSuccess || !can) { g->peer_ok = false; continue; }
            cudaError_t e = cudaDeviceEnablePeerAccess(devices[j], 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
            else if (e != cudaSuccess) { cudaGetLastError(); g->peer_ok = false; }
        }
    }
    if (prev >= 0) cudaSetDevice(prev);
    if (n > 1) g->pool.reset(new Pool(n));
    (void)nbits;
    *out = g;
    return DYIBF_OK;
}

static int make_members(dyibf_group* g, uint32_t bins, uint32_t h, uint32_t k, uint64_t nbits, bool zero)
{
    HashParams hp;
    int rc = init_params(bins, h, k, nbits, &g->whole, &hp);
    if (rc) return rc;
    g->whole.device = g->devices[0];
    const uint64_t P = g->whole.n_blocks;
    run_members(g, [&](int i) -> int {
        if (g->mode == DYIBF_GROUP_BIN_SHARDED)
            return new_handle(g->n_bins[i], h, k, P * 64ull * g->Wr + META_BITS, g->devices[i], nullptr, zero, &g->member[i]);
        return new_handle(bins, h, k, nbits, g->devices[i], nullptr, zero, &g->member[i]);
    });
    rc = first_error(g);
    if (rc) return rc;
    for (int i = 0; i < g->n; ++i)
        if (g->member[i]->prm.n_blocks != P) return fail(DYIBF_ERR_ARG, "internal: shard %d has %llu blocks, the filter %llu", i,
                                                          (unsigned long long)g->member[i]->prm.n_blocks, (unsigned long long)P);
    return DYIBF_OK;
}

static int ensure_h_ids(dyibf_group* g, int i, size_t bytes)
{
    if (bytes <= g->h_ids_cap[i]) return DYIBF_OK;
    if (g->h_ids[i]) { cudaFreeHost(g->h_ids[i]); g->h_ids[i] = nullptr; g->h_ids_cap[i] = 0; }
    const size_t want = bytes + bytes / 4 + 4096;
    CU(cudaMallocHost(&g->h_ids[i], want));
    g->h_ids_cap[i] = want;
    return DYIBF_OK;
}

// the lists of the members' last classify_range -> the caller's CSR
static int group_fetch(dyibf_group* g, uint64_t* bin_offsets, uint32_t* read_ids, uint64_t capacity, uint64_t* total)
{
    if (!g->fetched_ready) return fail(DYIBF_ERR_ARG, "no dyibf_group_classify result is resident");
    const uint32_t bins = g->whole.bins;
    const int n = g->n;
    uint64_t tot = 0;
    if (g->mode == DYIBF_GROUP_BIN_SHARDED) {
        std::vector<uint64_t> base(n + 1, 0);
        for (int i = 0; i < n; ++i) base[i + 1] = base[i] + g->member[i]->h_csr_off[g->n_bins[i]];
        tot = base[n];
        for (int i = 0; i < n; ++i)
            for (uint32_t b = 0; b < g->n_bins[i]; ++b) bin_offsets[g->first_bin[i] + b] = base[i] + g->member[i]->h_csr_off[b];
        bin_offsets[bins] = tot;
        if (total) *total = tot;
        if (tot > capacity) return fail(DYIBF_ERR_CAPACITY, "%llu (read,bin) pairs exceed the capacity of %llu", (unsigned long long)tot, (unsigned long long)capacity);
        run_members(g, [&](int i) -> int {                           // bin ranges are contiguous: straight into the caller's buffer
            dyibf_filter* f = g->member[i];
            const uint64_t mine = base[i + 1] - base[i];
            if (!mine) return DYIBF_OK;
            CU(cudaMemcpyAsync(read_ids + base[i], f->csr_ids.p, mine * 4, cudaMemcpyDeviceToHost, f->slot[0].st));
            CU(cudaStreamSynchronize(f->slot[0].st));
            return DYIBF_OK;
        });
        return first_error(g);
    }
    // read-sharded: bin b's list = member 0's ids | member 1's ids | ... (row ranges ascend with the member index)
    std::vector<uint64_t> pre((size_t)n * bins);                    // pre[i * bins + b]: where member i's part of bin b starts
    uint64_t pos = 0;
    for (uint32_t b = 0; b < bins; ++b) {
        bin_offsets[b] = pos;
        for (int i = 0; i < n; ++i) {
            pre[(size_t)i * bins + b] = pos;
            pos += g->member[i]->h_csr_off[b + 1] - g->member[i]->h_csr_off[b];
        }
    }
    bin_offsets[bins] = tot = pos;
    if (total) *total = tot;
    if (tot > capacity) return fail(DYIBF_ERR_CAPACITY, "%llu (read,bin) pairs exceed the capacity of %llu", (unsigned long long)tot, (unsigned long long)capacity);
    // a pinned output buffer is written by the devices themselves (mapped host memory under UVA)
    bool direct = false;
    if (tot) {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, read_ids) == cudaSuccess) direct = at.type == cudaMemoryTypeHost && at.devicePointer != nullptr;
        else cudaGetLastError();
        static const bool staged_lists = getenv("DYIBF_GROUP_STAGED_LISTS") != nullptr;   // A/B, read once
        if (staged_lists) direct = false;
    }
    run_members(g, [&](int i) -> int {
        dyibf_filter* f = g->member[i];
        const uint64_t mine = f->h_csr_off[bins];
        if (!mine) return DYIBF_OK;
        // the address THIS device reaches the pinned buffer under (it equals the host address only under unified addressing
        // with a portable registration); a device that cannot map it takes the staged path for its part
        void* dev_ids = nullptr;
        if (direct && (cudaHostGetDevicePointer(&dev_ids, read_ids, 0) != cudaSuccess || !dev_ids)) { cudaGetLastError(); dev_ids = nullptr; }
        if (dev_ids) {
            cudaStream_t st = f->slot[0].st;
            CU(f->misc.reserve((uint64_t)bins * 8 + 64));
            CU(cudaMemcpyAsync(f->misc.p, pre.data() + (size_t)i * bins, (uint64_t)bins * 8, cudaMemcpyHostToDevice, st));
            f->kernel_launches += 1;
            const unsigned blocks = (unsigned)std::min<uint32_t>((bins + 7) / 8, (uint32_t)f->sm_count * 8);
            scatter_lists_kernel<<<blocks, 256, 0, st>>>(static_cast<const uint32_t*>(f->csr_ids.p), static_cast<const uint64_t*>(f->csr_off.p),
                                                         static_cast<const uint64_t*>(f->misc.p), bins, static_cast<uint32_t*>(dev_ids));
            CU(cudaGetLastError());
            CU(cudaStreamSynchronize(st));
            return DYIBF_OK;
        }
        int rc = ensure_h_ids(g, i, mine * 4);
        if (rc) return rc;
        CU(cudaMemcpyAsync(g->h_ids[i], f->csr_ids.p, mine * 4, cudaMemcpyDeviceToHost, f->slot[0].st));
        CU(cudaStreamSynchronize(f->slot[0].st));
        const uint32_t* src = static_cast<const uint32_t*>(g->h_ids[i]);
        for (uint32_t b = 0; b < bins; ++b) {
            const uint64_t a = f->h_csr_off[b], c = f->h_csr_off[b + 1] - a;
            if (c) memcpy(read_ids + pre[(size_t)i * bins + b], src + a, c * 4);
        }
        return DYIBF_OK;
    });
    return first_error(g);
}

// bring every replica up to date after a bin-routed build: each device pulls the columns it does not own
static int exchange_columns(dyibf_group* g)
{
    PeerTables pt; memset(&pt, 0, sizeof pt);
    for (int i = 0; i < g->n; ++i) pt.words[i] = g->member[i]->d_words;
    const uint32_t W = g->whole.words_per_block;
    const bool vec2 = (g->Wr % 2u) == 0;
    const uint64_t n_pieces = g->whole.n_blocks * (uint64_t)W / (vec2 ? 2 : 1);
    run_members(g, [&](int i) -> int {                               // all builds are finished before anybody reads a peer
        CU(cudaDeviceSynchronize());
        return DYIBF_OK;
    });
    int rc = first_error(g);
    if (rc) return rc;
    run_members(g, [&](int i) -> int {
        dyibf_filter* f = g->member[i];
        const unsigned blocks = (unsigned)std::min<uint64_t>((n_pieces + 255) / 256, (uint64_t)f->sm_count * 16);
        f->kernel_launches += 1;
        if (vec2) exchange_columns_kernel<2><<<blocks, 256, 0, f->slot[0].st>>>(pt, (uint32_t)i, W, g->Wr, n_pieces);
        else      exchange_columns_kernel<1><<<blocks, 256, 0, f->slot[0].st>>>(pt, (uint32_t)i, W, g->Wr, n_pieces);
        CU(cudaGetLastError());
        CU(cudaStreamSynchronize(f->slot[0].st));
        return DYIBF_OK;
    });
    return first_error(g);
}

} // namespace dyibf

extern "C" {

int dyibf_shard_bins(uint32_t bins, int rank, int nranks, uint32_t* first_bin, uint32_t* n_bins)
{
    if (nranks < 1 || rank < 0 || rank >= nranks || bins == 0) return fail(DYIBF_ERR_ARG, "dyibf_shard_bins: bad rank %d of %d", rank, nranks);
    const uint32_t W = (bins + 63) / 64;
    if (W % (uint32_t)nranks) return fail(DYIBF_ERR_ARG, "bin-sharding needs the %u words per hash block divisible by the %d ranks", W, nranks);
    const uint64_t first = (uint64_t)rank * (W / nranks) * 64;
    if (first >= bins) return fail(DYIBF_ERR_ARG, "rank %d would own no bins", rank);
    if (first_bin) *first_bin = (uint32_t)first;
    if (n_bins) *n_bins = (uint32_t)(std::min<uint64_t>(bins, first + (uint64_t)(W / nranks) * 64) - first);
    return DYIBF_OK;
}

int dyibf_shard_reads(uint32_t n_reads, int rank, int nranks, uint32_t* first, uint32_t* last)
{
    if (nranks < 1 || rank < 0 || rank >= nranks) return fail(DYIBF_ERR_ARG, "dyibf_shard_reads: bad rank %d of %d", rank, nranks);
    const uint64_t per = ((uint64_t)n_reads + nranks - 1) / nranks;
    const uint64_t a = std::min<uint64_t>((uint64_t)rank * per, n_reads);
    if (first) *first = (uint32_t)a;
    if (last) *last = (uint32_t)std::min<uint64_t>(a + per, n_reads);
    return DYIBF_OK;
}

int dyibf_group_create(uint32_t bins, uint32_t n_hash, uint32_t kmer, uint64_t nbits, const int* devices, int n_devices,
                       int mode, dyibf_group** out)
{
    dyibf_params prm; HashParams hp;
    int rc = init_params(bins, n_hash, kmer, nbits, &prm, &hp);
    if (rc) return rc;
    dyibf_group* g = nullptr;
    rc = group_alloc(devices, n_devices, mode, bins, nbits, &g, prm.words_per_block);
    if (rc) return rc;
    rc = make_members(g, bins, n_hash, kmer, nbits, true);
    if (rc) { std::string keep = last_error_string(); dyibf_group_destroy(g); set_last_error(keep); return rc; }
    *out = g;
    return DYIBF_OK;
}

// expect: the caller's bins / h / k / nbits describe the file (only the bit count is checked against it), as
// dyibf_load_file_expect does for one device; else all four come from the trailer
static int group_load(const char* path, bool expect, uint32_t bins, uint32_t h, uint32_t k, uint64_t nbits, const int* devices,
                      int n_devices, int mode, dyibf_group** out)
{
    if (!path || !out) return fail(DYIBF_ERR_ARG, "NULL argument");
    *out = nullptr;
    uint32_t fbins = 0, fh = 0, fk = 0; uint64_t fnbits = 0;
    int rc = dyibf_peek_file(path, &fbins, &fh, &fk, &fnbits);
    if (rc) return rc;
    if (expect) {
        if (fnbits != nbits) return fail(DYIBF_ERR_SIZE_MISMATCH, "Size mismatch: \n\t Loaded file \t%llu bits\n\t expected\t%llu",
                                         (unsigned long long)fnbits, (unsigned long long)nbits);
    } else { bins = fbins; h = fh; k = fk; nbits = fnbits; }
    dyibf_params prm; HashParams hp;
    rc = init_params(bins, h, k, nbits, &prm, &hp);
    if (rc) return rc;
    dyibf_group* g = nullptr;
    rc = group_alloc(devices, n_devices, mode, bins, nbits, &g, prm.words_per_block);
    if (rc) return rc;
    rc = make_members(g, bins, h, k, nbits, false);
    if (!rc) {
        std::vector<uint64_t*> dw(g->n);
        for (int i = 0; i < g->n; ++i) dw[i] = g->member[i]->d_words;
        const bool sharded = mode == DYIBF_GROUP_BIN_SHARDED;
        if (!sharded && g->n > 1 && !g->peer_ok) {
            // no peer access: every replica takes its own copy over PCIe
            for (int i = 0; i < g->n && !rc; ++i) rc = io_load_words(path, prm.n_blocks, prm.words_per_block, &g->devices[i], &dw[i], 1, false);
        } else
            rc = io_load_words(path, prm.n_blocks, prm.words_per_block, g->devices.data(), dw.data(), g->n, sharded);
    }
    if (rc) { std::string keep = last_error_string(); dyibf_group_destroy(g); set_last_error(keep); return rc; }
    *out = g;
    return DYIBF_OK;
}

int dyibf_group_load_file(const char* path, const int* devices, int n_devices, int mode, dyibf_group** out)
{
    return group_load(path, false, 0, 0, 0, 0, devices, n_devices, mode, out);
}

int dyibf_group_load_file_expect(const char* path, uint32_t bins, uint32_t n_hash, uint32_t kmer, uint64_t nbits,
                                 const int* devices, int n_devices, int mode, dyibf_group** out)
{
    return group_load(path, true, bins, n_hash, kmer, nbits, devices, n_devices, mode, out);
}

void dyibf_group_destroy(dyibf_group* g)
{
    if (!g) return;
    g->pool.reset();
    for (int i = 0; i < g->n; ++i) {
        if (g->member[i]) dyibf_destroy(g->member[i]);
        if (g->h_ids[i]) cudaFreeHost(g->h_ids[i]);
    }
    delete g;
}

int dyibf_group_get_params(const dyibf_group* g, dyibf_params* whole, int* n_devices, int* mode)
{
    if (!g) return fail(DYIBF_ERR_ARG, "group is NULL");
    if (whole) *whole = g->whole;
    if (n_devices) *n_devices = g->n;
    if (mode) *mode = g->mode;
    return DYIBF_OK;
}

int dyibf_group_member(dyibf_group* g, int index, dyibf_filter** member, uint32_t* first_bin, uint32_t* n_bins)
{
    if (!g || index < 0 || index >= g->n) return fail(DYIBF_ERR_ARG, "dyibf_group_member: index %d out of range", index);
    if (member) *member = g->member[index];
    if (first_bin) *first_bin = g->mode == DYIBF_GROUP_BIN_SHARDED ? g->first_bin[index] : 0;
    if (n_bins) *n_bins = g->mode == DYIBF_GROUP_BIN_SHARDED ? g->n_bins[index] : g->whole.bins;
    return DYIBF_OK;
}

int dyibf_group_set_option(dyibf_group* g, int option, int64_t value)
{
    if (!g) return fail(DYIBF_ERR_ARG, "group is NULL");
    for (int i = 0; i < g->n; ++i) { int rc = dyibf_set_option(g->member[i], option, value); if (rc) return rc; }
    return DYIBF_OK;
}

int dyibf_group_save_file(dyibf_group* g, const char* path)
{
    if (!g || !path) return fail(DYIBF_ERR_ARG, "NULL argument");
    run_members(g, [&](int) -> int { CU(cudaDeviceSynchronize()); return DYIBF_OK; });
    int rc = first_error(g);
    if (rc) return rc;
    dyibf_filter whole;                                              // parameters only: the trailer is made from them
    whole.prm = g->whole;
    std::vector<uint64_t*> dw(g->n);
    for (int i = 0; i < g->n; ++i) dw[i] = g->member[i]->d_words;
    return io_save_words(path, &whole, g->whole.n_blocks, g->whole.words_per_block, g->devices.data(), dw.data(), g->n,
                         g->mode == DYIBF_GROUP_BIN_SHARDED);
}

int dyibf_broadcast_filter(dyibf_group* g, int src_member)
{
    if (!g || src_member < 0 || src_member >= g->n) return fail(DYIBF_ERR_ARG, "dyibf_broadcast_filter: bad source member");
    if (g->mode != DYIBF_GROUP_READ_SHARDED) return fail(DYIBF_ERR_ARG, "dyibf_broadcast_filter: only replicas are broadcast (read-sharded groups)");
    if (g->n == 1) return DYIBF_OK;
    dyibf_filter* src = g->member[src_member];
    { DeviceGuard guard(src->prm.device); CU(cudaDeviceSynchronize()); }
    const uint64_t bytes = src->data_words * 8, piece = 256ull << 20;
    run_members(g, [&](int i) -> int {                               // every replica pulls from the source over NVLink
        if (i == src_member) return DYIBF_OK;
        dyibf_filter* f = g->member[i];
        for (uint64_t a = 0; a < bytes; a += piece)
            CU(cudaMemcpyPeerAsync((char*)f->d_words + a, f->prm.device, (const char*)src->d_words + a, src->prm.device,
                                   std::min(piece, bytes - a), f->slot[0].st));
        CU(cudaStreamSynchronize(f->slot[0].st));
        return DYIBF_OK;
    });
    return first_error(g);
}

int dyibf_group_classify(dyibf_group* g, const uint8_t* seqs, const uint64_t* offsets, uint32_t n_seqs, int encoding, int group,
                         int derive_rc, uint32_t threshold, uint64_t* bin_offsets, uint32_t* read_ids, uint64_t capacity,
                         uint64_t* total)
{
    if (!g) return fail(DYIBF_ERR_ARG, "group is NULL");
    if (!bin_offsets || (capacity && !read_ids)) return fail(DYIBF_ERR_ARG, "NULL argument");
    int rc = check_query_args(g->member[0], seqs, offsets, n_seqs, encoding, group, derive_rc, threshold);
    if (rc) return rc;
    g->fetched_ready = false;
    const uint64_t n_total = derive_rc ? 2ull * n_seqs : n_seqs;
    const uint32_t n_rows = (uint32_t)(n_total / group);
    run_members(g, [&](int i) -> int {
        uint32_t a = 0, b = n_rows;
        if (g->mode == DYIBF_GROUP_READ_SHARDED) dyibf_shard_reads(n_rows, i, g->n, &a, &b);
        return classify_range(g->member[i], seqs, offsets, n_seqs, encoding, group, derive_rc, threshold, a, b);
    });
    rc = first_error(g);
    if (rc) return rc;
    g->fetched_ready = true;
    return group_fetch(g, bin_offsets, read_ids, capacity, total);
}

int dyibf_group_classify_fetch(dyibf_group* g, uint64_t* bin_offsets, uint32_t* read_ids, uint64_t capacity, uint64_t* total)
{
    if (!g || !bin_offsets || (capacity && !read_ids)) return fail(DYIBF_ERR_ARG, "NULL argument");
    return group_fetch(g, bin_offsets, read_ids, capacity, total);
}

int dyibf_group_classify_begin(dyibf_group* g, const uint8_t* seqs, const uint64_t* offsets, uint32_t n_seqs, int encoding,
                               int group, int derive_rc, uint32_t threshold, dyibf_ticket** ticket)
{
    if (!g) return fail(DYIBF_ERR_ARG, "group is NULL");
    if (g->mode != DYIBF_GROUP_READ_SHARDED)
        return fail(DYIBF_ERR_ARG, "dyibf_group_classify_begin: whole batches go to one replica each (read-sharded groups only)");
    // successive batches go to successive replicas; a replica with all tickets in flight is skipped
    int rc = DYIBF_ERR_BUSY;
    for (int tries = 0; tries < g->n; ++tries) {
        const int i = g->next_async;
        g->next_async = (g->next_async + 1) % g->n;
        rc = dyibf_classify_begin(g->member[i], seqs, offsets, n_seqs, encoding, group, derive_rc, threshold, ticket);
        if (rc != DYIBF_ERR_BUSY) return rc;
    }
    return rc;
}

int dyibf_group_add_sequences(dyibf_group* g, const uint8_t* seqs, const uint64_t* offsets, uint32_t n_seqs,
                              const uint32_t* bin_ids, int encoding)
{
    if (!g) return fail(DYIBF_ERR_ARG, "group is NULL");
    if (n_seqs == 0) return DYIBF_OK;
    if (!seqs || !offsets || !bin_ids) return fail(DYIBF_ERR_ARG, "NULL argument");
    const uint32_t bins = g->whole.bins;
    for (uint32_t s = 0; s < n_seqs; ++s)
        if (bin_ids[s] >= bins) return fail(DYIBF_ERR_ARG, "bin %u of sequence %u out of range (%u bins)", bin_ids[s], s, bins);
    const bool sharded = g->mode == DYIBF_GROUP_BIN_SHARDED;
    // Route every sequence to the device that owns its bin's word column.  Replicas without such an ownership (W not
    // divisible by the devices, or no peer access for the exchange afterwards) each add everything.
    const bool routed = sharded || (g->n > 1 && g->Wr != 0 && g->peer_ok);
    if (!routed) {
        run_members(g, [&](int i) -> int {
            return dyibf_add_sequences(g->member[i], seqs, offsets, n_seqs, bin_ids, encoding, DYIBF_MEM_HOST, nullptr);
        });
        return first_error(g);
    }
    auto owner = [&](uint32_t bin) { return (int)(bin / (g->Wr * 64u)); };
    bool ascending = true;
    for (uint32_t s = 1; s < n_seqs && ascending; ++s) ascending = owner(bin_ids[s]) >= owner(bin_ids[s - 1]);
    // per member: the sequences it adds, as (concat, limits, bins); contiguous ranges of the caller's set when the bins ascend
    struct Part { const uint8_t* seqs; const uint64_t* off; uint32_t n; std::vector<uint8_t> own_seqs; std::vector<uint64_t> own_off; std::vector<uint32_t> ids; };
    std::vector<Part> part(g->n);
    if (ascending) {
        uint32_t s = 0;
        for (int i = 0; i < g->n; ++i) {
            const uint32_t a = s;
            while (s < n_seqs && owner(bin_ids[s]) == i) ++s;
            part[i].seqs = seqs; part[i].off = offsets + a; part[i].n = s - a;
            part[i].ids.assign(bin_ids + a, bin_ids + s);
        }
    } else {
        for (uint32_t s = 0; s < n_seqs; ++s) {
            Part& p = part[owner(bin_ids[s])];
            if (p.own_off.empty()) p.own_off.push_back(0);
            p.own_seqs.insert(p.own_seqs.end(), seqs + offsets[s], seqs + offsets[s + 1]);
            p.own_off.push_back(p.own_seqs.size());
            p.ids.push_back(bin_ids[s]);
        }
        for (int i = 0; i < g->n; ++i) {
            Part& p = part[i];
            p.n = (uint32_t)p.ids.size();
            if (p.own_seqs.empty()) p.own_seqs.push_back(0);
            p.seqs = p.own_seqs.data(); p.off = p.own_off.data();
        }
    }
    if (sharded)
        for (int i = 0; i < g->n; ++i)
            for (uint32_t& b : part[i].ids) b -= g->first_bin[i];      // a shard numbers its bins from 0
    run_members(g, [&](int i) -> int {
        const Part& p = part[i];
        if (!p.n) return DYIBF_OK;
        return dyibf_add_sequences(g->member[i], p.seqs, p.off, p.n, p.ids.data(), encoding, DYIBF_MEM_HOST, nullptr);
    });
    int rc = first_error(g);
    if (rc || sharded) return rc;
    return exchange_columns(g);
}

int dyibf_group_clear_bins(dyibf_group* g, const uint32_t* bins, uint32_t n)
{
    if (!g || (n && !bins)) return fail(DYIBF_ERR_ARG, "NULL argument");
    for (uint32_t j = 0; j < n; ++j)
        if (bins[j] >= g->whole.bins) return fail(DYIBF_ERR_ARG, "bin %u out of range (%u bins)", bins[j], g->whole.bins);
    run_members(g, [&](int i) -> int {
        if (g->mode != DYIBF_GROUP_BIN_SHARDED) return dyibf_clear_bins(g->member[i], bins, n);
        std::vector<uint32_t> mine;
        for (uint32_t j = 0; j < n; ++j)
            if (bins[j] >= g->first_bin[i] && bins[j] < g->first_bin[i] + g->n_bins[i]) mine.push_back(bins[j] - g->first_bin[i]);
        return mine.empty() ? DYIBF_OK : dyibf_clear_bins(g->member[i], mine.data(), (uint32_t)mine.size());
    });
    return first_error(g);
}

} // extern "C"
