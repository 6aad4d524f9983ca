// compact_static.cu — dyibf_compact_static: the threshold compaction of compact_kernel.cu for callers that must not
// allocate.
//
// Who that is.  dyibf_compact keeps its per-(bin, chunk) counters in a workspace it cudaMallocs on first use and grows on
// demand.  That is the wrong thing to do for a host program that has handed the whole card to something else before the
// mapper stage starts: a bin-sharded cfg4 member holds a 64 GiB shard and the rest of the HBM may sit in another
// component's pool (a framework's caching allocator, an MPS client with a fixed memory limit), so a cudaMalloc in the
// middle of the run can fail for reasons this library cannot see — and even when it succeeds it stalls every stream of the
// device while the driver maps the pages, which a latency-bound service thread that feeds 100 000-read batches cannot
// afford.  Such callers bring the inputs and outputs from their own arena; the only other memory the compaction needs is
// the counter workspace, and here it is a fixed array in the library's own image: present on every GPU that loads
// libdyibf.so from module load until the process ends, never resized, never freed.
//
// No counterpart in the reference (its push_back lists allocate on the host heap, src/d_mapper.h:360-368).
// The kernels are the ones dyibf_compact runs (launch_compact_* in compact_kernel.cu); only where the workspace lives
// differs, so the lists are identical.
#include "dyibf_handle.h"

namespace dyibf {

// The workspace.  8 MiB per GPU: room for every batch up to about two million rows of a 1024-bin filter, or about
// 260 000 rows of an 8192-bin one (see dyibf.h for the exact rule).
static const uint64_t STATIC_WS_WORDS = DYIBF_COMPACT_STATIC_BYTES / sizeof(uint64_t);
static __device__ __align__(256) uint64_t g_compact_static_ws[STATIC_WS_WORDS];

// In host code `g_compact_static_ws` is only a handle the runtime maps to the real array; the array has a different
// address on every device.  The address is looked up once per device, with that device current, and kept here.
static const int STATIC_MAX_DEVICES = 64;
static std::mutex g_static_setup_mtx;
static uint64_t*  g_static_ws_addr[STATIC_MAX_DEVICES];           // zero-initialised: not looked up yet

// the caller has made `device` current
static int static_workspace(int device, uint64_t** ws)
{
    std::lock_guard<std::mutex> lk(g_static_setup_mtx);
    if (!g_static_ws_addr[device]) {
        uint64_t* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_compact_static_ws);
        if (e != cudaSuccess || !p)
            return fail(DYIBF_ERR_CUDA, "dyibf_compact_static: cudaGetSymbolAddress on device %d: %s", device,
                        e != cudaSuccess ? cudaGetErrorString(e) : "null address");
        g_static_ws_addr[device] = p;
    }
    *ws = g_static_ws_addr[device];
    return DYIBF_OK;
}

} // namespace dyibf

using namespace dyibf;

extern "C" int dyibf_compact_static(int device, uint32_t bins, const uint64_t* rows, uint32_t n_rows, uint64_t* bin_offsets,
                                    uint32_t* read_ids, uint64_t capacity, uint64_t* total, void* stream)
{
    if (bins == 0) return fail(DYIBF_ERR_ARG, "dyibf_compact_static: bins is 0");
    if (!bin_offsets || (n_rows && !rows) || (capacity && !read_ids)) return fail(DYIBF_ERR_ARG, "dyibf_compact_static: NULL argument");
    if (device < 0 || device >= STATIC_MAX_DEVICES)
        return fail(DYIBF_ERR_ARG, "dyibf_compact_static: device %d outside 0..%d", device, STATIC_MAX_DEVICES - 1);
    // size check first: nothing is launched, and no pointer is touched, for a batch the array cannot take
    const size_t need = compact_workspace_bytes(bins, n_rows);
    if (need > DYIBF_COMPACT_STATIC_BYTES)
        return fail(DYIBF_ERR_NOMEM, "dyibf_compact_static: %u rows x %u bins need %llu bytes of counters, the built-in workspace "
                                     "holds %llu (DYIBF_COMPACT_STATIC_BYTES); split the batch or use dyibf_compact",
                    n_rows, bins, (unsigned long long)need, (unsigned long long)DYIBF_COMPACT_STATIC_BYTES);
    DeviceGuard guard(device);
    if (!guard.ok) return fail(DYIBF_ERR_CUDA, "dyibf_compact_static: cannot select device %d", device);
    uint64_t* ws = nullptr;
    int rc = static_workspace(device, &ws);
    if (rc) return rc;

    cudaStream_t st = (cudaStream_t)stream;
    // the array still holds the counters of whichever call used it last: everything this batch reads starts at zero
    CU(cudaMemsetAsync(ws, 0, need, st));
    CU(launch_compact_count(rows, n_rows, bins, ws, st));
    CU(launch_compact_scan(ws, bins, n_rows, bin_offsets, st));
    uint64_t tot = 0;
    CU(cudaMemcpyAsync(&tot, bin_offsets + bins, 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (total) *total = tot;
    if (tot > capacity) return fail(DYIBF_ERR_CAPACITY, "%llu (read,bin) pairs exceed the capacity of %llu",
                                    (unsigned long long)tot, (unsigned long long)capacity);
    CU(launch_compact_fill(rows, n_rows, bins, ws, read_ids, capacity, 0, st));
    return DYIBF_OK;
}
