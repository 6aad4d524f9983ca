"""dyibf_compact_static — the compaction for callers that may not allocate device memory.

Some users of this library hand the whole GPU to other owners before the mapper stage runs (a 64 GiB filter shard next to
another component's memory pool, or a process under a fixed MPS memory limit) and feed batches from a latency-bound service
thread: for them the cudaMalloc inside dyibf_compact can fail, or stall every stream of the device, at a moment they do
not choose.  dyibf_compact_static keeps its counters in an 8 MiB array built into libdyibf.so instead and takes inputs and
outputs from the caller.  This file checks, through the C ABI, that it returns exactly the lists of dyibf_compact (integer
work: bit-exact) on two calls in a row — the second on a smaller batch, so stale counters of the first would show — that free
device memory does not move across a call, and that a batch too large for the array is refused before anything runs.

The file sorts last on purpose: `pytest -x` stops at the first failure, and a problem here must not hide the parity tests.
It also has to pass alone in a fresh process, where its first call is the one that looks the array up on the device.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rows(rng, n_rows, bins, density):
    """n_rows x W selection rows with about `density` of the bins set per row and no bit beyond `bins`."""
    W = (bins + 63) // 64
    bits = rng.random((n_rows, W * 64), dtype=np.float32) < density
    bits[:, bins:] = False
    return np.packbits(bits, axis=1, bitorder="little").view(np.uint64).reshape(n_rows, W)


def _expected(rows, bins):
    """Per bin the ascending row numbers, straight from the bits (what `--threads 1` leaves in origReadIdMap)."""
    bits = np.unpackbits(rows.view(np.uint8), axis=1, bitorder="little")[:, :bins]
    r, b = np.nonzero(bits.T)                      # r = bin, b = row; bins ascending, rows ascending inside a bin
    offs = np.zeros(bins + 1, dtype=np.uint64)
    offs[1:] = np.cumsum(np.bincount(r, minlength=bins))
    return offs, b.astype(np.uint32)


def test_compact_static_matches_compact_on_two_calls_in_a_row(dy):
    import torch
    from dream_yara_b200._lib import ERR_CAPACITY, ERR_NOMEM, MEM_DEVICE, OK, check
    L = dy.lib()
    dev = torch.device("cuda:0")
    bins, W = 1000, 16                                    # not a multiple of 64: the last word is partly unused
    f = dy.SeqAnBloomFilter(bins, 3, 18, (1 << 22) + 256)  # only dyibf_compact needs a handle
    rng = np.random.default_rng(20261106)
    st = torch.cuda.Stream(device=dev)
    sp = C.c_void_p(st.cuda_stream)

    # every device buffer of both calls exists before the first call: nothing below allocates between the calls
    cases = [(30011, 0.02), (777, 0.3)]
    host_rows = [_rows(rng, n, bins, d) for n, d in cases]
    cap = max(int(np.unpackbits(r.view(np.uint8)).sum()) for r in host_rows)
    d_rows = [torch.from_numpy(r.view(np.int64)).to(dev) for r in host_rows]
    d_off_ref = torch.zeros(bins + 1, dtype=torch.int64, device=dev)
    d_off_new = torch.full((bins + 1,), -1, dtype=torch.int64, device=dev)
    d_ids_ref = torch.zeros(cap, dtype=torch.int32, device=dev)
    d_ids_new = torch.full((cap,), -1, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()

    for call, (rows, dr) in enumerate(zip(host_rows, d_rows)):
        n_rows = rows.shape[0]
        t_ref, t_new = C.c_uint64(0), C.c_uint64(0)
        check(L.dyibf_compact(f._h, C.c_void_p(dr.data_ptr()), n_rows, C.c_void_p(d_off_ref.data_ptr()),
                              C.c_void_p(d_ids_ref.data_ptr()), cap, C.byref(t_ref), MEM_DEVICE, sp))
        st.synchronize()
        free_before = torch.cuda.mem_get_info(dev)[0]
        check(L.dyibf_compact_static(0, bins, C.c_void_p(dr.data_ptr()), n_rows, C.c_void_p(d_off_new.data_ptr()),
                                     C.c_void_p(d_ids_new.data_ptr()), cap, C.byref(t_new), sp))
        st.synchronize()
        free_after = torch.cuda.mem_get_info(dev)[0]
        if call > 0:            # the first call may make the driver load device code; from then on nothing may move
            assert free_after == free_before, "dyibf_compact_static changed the free device memory"
        exp_off, exp_ids = _expected(rows, bins)
        assert t_new.value == t_ref.value == int(exp_off[bins])
        tot = t_new.value
        off_ref = d_off_ref.cpu().numpy().view(np.uint64); off_new = d_off_new.cpu().numpy().view(np.uint64)
        ids_ref = d_ids_ref.cpu().numpy().view(np.uint32)[:tot]; ids_new = d_ids_new.cpu().numpy().view(np.uint32)[:tot]
        assert np.array_equal(off_new, off_ref), "call %d: bin offsets differ from dyibf_compact" % call
        assert np.array_equal(ids_new, ids_ref), "call %d: read ids differ from dyibf_compact" % call
        assert np.array_equal(off_new, exp_off) and np.array_equal(ids_new, exp_ids)

    # capacity too small: offsets and total are still delivered, as with dyibf_compact
    t = C.c_uint64(0)
    d_off_new.fill_(-1); torch.cuda.synchronize()
    rc = L.dyibf_compact_static(0, bins, C.c_void_p(d_rows[1].data_ptr()), host_rows[1].shape[0],
                                C.c_void_p(d_off_new.data_ptr()), C.c_void_p(d_ids_new.data_ptr()), 5, C.byref(t), sp)
    assert rc == ERR_CAPACITY and t.value == int(_expected(host_rows[1], bins)[0][bins])
    assert np.array_equal(d_off_new.cpu().numpy().view(np.uint64), _expected(host_rows[1], bins)[0])

    # an empty batch is fine; a batch whose counters do not fit the built-in array is refused with a message that says so
    assert L.dyibf_compact_static(0, bins, None, 0, C.c_void_p(d_off_new.data_ptr()), None, 0, C.byref(t), sp) == OK
    st.synchronize()
    assert t.value == 0 and not d_off_new.cpu().numpy().any()
    rc = L.dyibf_compact_static(0, 8192, C.c_void_p(d_rows[0].data_ptr()), 4_000_000, C.c_void_p(d_off_new.data_ptr()),
                                C.c_void_p(d_ids_new.data_ptr()), cap, C.byref(t), sp)
    assert rc == ERR_NOMEM
    msg = L.dyibf_last_error().decode()
    assert "DYIBF_COMPACT_STATIC_BYTES" in msg and "4000000 rows" in msg
    assert L.dyibf_compact_static(0, 0, None, 0, C.c_void_p(d_off_new.data_ptr()), None, 0, None, sp) == -1
    f.close()
