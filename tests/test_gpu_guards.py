"""GPU (B200): no kernel writes outside the buffers its caller gave it.  compute-sanitizer is not available on the GPU pool, so
the device-pointer entry points are run on outputs embedded in larger allocations whose surroundings hold a canary pattern;
sizes are ragged (no multiple of any tile, chunk or vector width) so that every tail path runs.  Results are checked
against the host-buffer path, which the parity tests pin to the oracle and the golden vectors."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GUARD = 4096          # bytes either side
CANARY = 0xA5


class Guarded:
    """n_bytes of device memory with GUARD canary bytes in front and behind (payload 256-byte aligned like cudaMalloc's)."""

    def __init__(self, n_bytes, fill=0):
        import torch
        self.n = int(n_bytes)
        self.buf = torch.full((GUARD + self.n + GUARD,), CANARY, dtype=torch.uint8, device="cuda")
        self.buf[GUARD:GUARD + self.n] = fill
        self.ptr = self.buf.data_ptr() + GUARD

    def payload(self, dtype):
        return self.buf[GUARD:GUARD + self.n].cpu().numpy().view(dtype)

    def intact(self):
        import torch
        torch.cuda.synchronize()
        return bool((self.buf[:GUARD] == CANARY).all()) and bool((self.buf[GUARD + self.n:] == CANARY).all())


def _ragged_reads(rng, n, lo, hi):
    lens = rng.integers(lo, hi, size=n)
    off = np.zeros(n + 1, dtype=np.uint64)
    off[1:] = np.cumsum(lens)
    codes = rng.integers(0, 4, size=int(off[-1]), dtype=np.uint8)
    codes[rng.integers(0, codes.size, size=codes.size // 200)] = 4          # some N
    return codes, off


@pytest.mark.parametrize("bins,group", [(64, 2), (130, 4), (1000, 2)])
def test_query_and_compact_device_outputs_stay_in_bounds(dy, bins, group):
    import torch
    from dream_yara_b200._lib import ENC_DNA5, MEM_DEVICE, check
    L = dy.lib()
    rng = np.random.default_rng(bins)
    k, h, nbits = 18, 3, (1 << 26) + 256
    f = dy.SeqAnBloomFilter(bins, h, k, nbits)
    g, goff = _ragged_reads(rng, bins, 300, 900)
    f.addSequences((g, goff), np.arange(bins, dtype=np.uint32))
    n_rows = 3001 if group == 2 else 1503                                    # forward sequences: n_rows * group / 2
    n_fwd = n_rows * group // 2
    # reads drawn from the genomes so that some bins are selected, lengths 1..250 (shorter than k included)
    codes, off = _ragged_reads(rng, n_fwd, 1, 251)
    for i in range(0, n_fwd, 3):
        b = int(rng.integers(0, bins)); ln = int(off[i + 1] - off[i]); gl = int(goff[b + 1] - goff[b])
        if ln <= gl:
            s = int(goff[b]) + int(rng.integers(0, gl - ln + 1))
            codes[int(off[i]):int(off[i + 1])] = g[s:s + ln]
    W = (bins + 63) // 64
    thr = 5
    want_rows, want_cnt = f.whichBinsBatch((codes, off), thr, group=group, derive_rc=True, counts=True, encoding=ENC_DNA5)
    want_bo, want_ids = f.classify((codes, off), thr, group=group, derive_rc=True, encoding=ENC_DNA5)

    d_codes = torch.from_numpy(codes).cuda(); d_off = torch.from_numpy(off.view(np.int64)).cuda()
    rows = Guarded(n_rows * W * 8, fill=0xFF)                                 # rows are overwritten, not OR-ed into garbage
    counts = Guarded(2 * n_fwd * bins * 2, fill=0xFF)
    st = torch.cuda.current_stream().cuda_stream
    check(L.dyibf_query_batch(f._h, C.c_void_p(d_codes.data_ptr()), C.c_void_p(d_off.data_ptr()), n_fwd, ENC_DNA5, group, 1, thr, 250,
                              C.c_void_p(rows.ptr), C.c_void_p(counts.ptr), MEM_DEVICE, C.c_void_p(st)))
    assert rows.intact() and counts.intact()
    assert np.array_equal(rows.payload(np.uint64).reshape(n_rows, W), want_rows)
    assert np.array_equal(counts.payload(np.uint16).reshape(2 * n_fwd, bins), want_cnt)

    total = int(want_bo[-1])
    assert total > 0
    boff = Guarded((bins + 1) * 8, fill=0xFF)
    ids = Guarded(total * 4, fill=0xFF)                                       # exact capacity: nothing to spare behind the lists
    tot = C.c_uint64(0)
    check(L.dyibf_compact(f._h, C.c_void_p(rows.ptr), n_rows, C.c_void_p(boff.ptr), C.c_void_p(ids.ptr), total, C.byref(tot),
                          MEM_DEVICE, C.c_void_p(st)))
    assert boff.intact() and ids.intact() and tot.value == total
    assert np.array_equal(boff.payload(np.uint64), want_bo) and np.array_equal(ids.payload(np.uint32), want_ids)
    if total > 1:                                                             # one id too few: DYIBF_ERR_CAPACITY, nothing written past the end
        small = Guarded((total - 1) * 4, fill=0xFF)
        rc = L.dyibf_compact(f._h, C.c_void_p(rows.ptr), n_rows, C.c_void_p(boff.ptr), C.c_void_p(small.ptr), total - 1, C.byref(tot),
                             MEM_DEVICE, C.c_void_p(st))
        assert rc == -6 and tot.value == total and small.intact() and boff.intact()
    f.close()


@pytest.mark.parametrize("fmt", [0, 1])
def test_record_reader_device_outputs_stay_in_bounds(dy, fmt):
    import torch
    from dream_yara_b200.ingest import parseRecordsDevice
    rng = np.random.default_rng(fmt)
    recs = []
    for i in range(2777):
        ln = int(rng.integers(0, 400))
        s = bytes(rng.choice(np.frombuffer(b"ACGTNacgtn", dtype=np.uint8), size=ln))
        if fmt == 0:
            w = int(rng.integers(1, 90))
            recs.append(b">r%d some text\n" % i + b"\n".join(s[j:j + w] for j in range(0, ln, w)) + b"\n")
        else:
            recs.append(b"@r%d\n%s\n+\n%s\n" % (i, s, b"I" * ln))
    text = b"".join(recs)
    want_seqs, want_limits = dy.parseRecords(text, fmt)
    n, nb = want_limits.size - 1, want_seqs.size
    d_text = Guarded(len(text))
    d_text.buf[GUARD:GUARD + len(text)] = torch.frombuffer(bytearray(text), dtype=torch.uint8).cuda()
    seqs = Guarded(nb, fill=0xFF)                                             # exact capacities
    lim = Guarded((n + 1) * 8, fill=0xFF)
    got = parseRecordsDevice(d_text.ptr, len(text), fmt, seqs.ptr, nb, lim.ptr, n)
    assert got == (n, nb)
    assert seqs.intact() and lim.intact() and d_text.intact()
    assert np.array_equal(seqs.payload(np.uint8), want_seqs) and np.array_equal(lim.payload(np.uint64), want_limits)
