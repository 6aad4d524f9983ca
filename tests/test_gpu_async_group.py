"""GPU (B200): the round-2 entry points of the C ABI — ticketed classify (dyibf_classify_begin / _end), the one-process
multi-GPU group (dyibf_group_*), the threaded .filter loader / writer, the device-side error word — against the oracle and
against the single-device path that the reference vectors already pin.

Reference behaviour being preserved: clasifyLoadedReads (src/d_mapper.h:308-387) called once per batch
(src/mapper.h:109) by ONE process that owns ONE filter (src/d_mapper.h:911-922); build_filter
(src/d_build_filter.cpp:217-272); the file constructors and save (src/d_bloom_filter.h:59-89,150-154).
The multi-device tests need 2 GPUs (gpurun --gpus 2) and skip on one."""
import ctypes as C
import os

import numpy as np
import pytest

from oracle_py import full_layout

pytestmark = pytest.mark.gpu


def _n_gpus(dy):
    n = C.c_int(0)
    assert dy.lib().dyibf_device_count(C.byref(n)) == 0
    return n.value


def _case(seed, bins, n_reads, k=18, L=100, paired=False, genome=3000):
    rng = np.random.default_rng(seed)
    contigs = [rng.integers(0, 4, genome).astype(np.uint8) for _ in range(bins)]
    for b in range(0, bins, 7):
        contigs[b][0] = 4                                            # leading-N contigs (SeqAn's hashInit quirk)
    n_seq = n_reads * (2 if paired else 1)
    reads = []
    for i in range(n_seq):
        c = contigs[int(rng.integers(0, bins))]
        st = int(rng.integers(0, len(c) - L))
        r = c[st:st + L].copy()
        m = rng.random(L) < 0.01
        r[m] = rng.integers(0, 5, int(m.sum()))
        reads.append(r if i % 9 else rng.integers(0, 4, L).astype(np.uint8))
    return contigs, reads


def _oracle_lists(oracle, contigs, reads, bins, h, k, nbits, thr, paired):
    p = oracle.params(bins, h, k, nbits)
    words = oracle.new_words(p)
    for b, c in enumerate(contigs):
        oracle.add_kmers(words, p, c, b)
    cat = np.concatenate(reads)
    off = np.zeros(len(reads) + 1, dtype=np.uint64); off[1:] = np.cumsum([len(r) for r in reads])
    fcat, foff = full_layout(cat, off, oracle)
    rows = oracle.classify(words, p, fcat, foff, paired, thr)
    return words, p, rows, oracle.compact(rows, bins)


# ---- ADVICE (round 1): the zero fill of a fresh filter must be complete before anything else touches the table ----------
def test_create_then_add_immediately_large_filter(dy, oracle):
    bins, h, k = 64, 3, 18
    nbits = (8 << 33) + 256                                           # 8 GiB: the fill takes milliseconds
    rng = np.random.default_rng(5)
    seq = rng.integers(0, 4, 400).astype(np.uint8)
    f = dy.SeqAnBloomFilter(bins, h, k, nbits)
    f.addSequences([seq], [17])                                       # straight after create
    rows = f.whichBinsBatch([seq], 400 - k + 1, group=1)
    assert rows[0, 0] == np.uint64(1 << 17)
    # every bit the oracle sets is set, and nothing else, anywhere in the 8 GiB
    p = oracle.params(bins, h, k, nbits)
    import torch
    from dream_yara_b200 import dist as _d
    t = _d._tensor_from_ptr(f.device_words(), p.P * p.W, torch.device("cuda", 0))
    nz = torch.nonzero(t).flatten().cpu().numpy()
    want = set()
    for hv in oracle.kmer_hashes(seq, k):
        for blk in oracle.positions(p, int(hv)):                       # block indices; bin 17 lives in word 0 of a block
            want.add(int(blk) * p.W)
    assert set(int(x) for x in nz) == want
    assert bool((t[torch.from_numpy(nz).cuda()] == (1 << 17)).all())
    f.close()


def test_device_error_word(dy):
    """device-resident bin ids / sequence lengths cannot be validated on the host: the kernels skip and flag"""
    import torch
    from dream_yara_b200._lib import DEVERR_BIN_RANGE, DEVERR_SEQ_TOO_LONG, ENC_DNA5, MEM_DEVICE
    bins, h, k, nbits = 128, 3, 18, (1 << 24) + 256
    L = dy.lib()
    f = dy.SeqAnBloomFilter(bins, h, k, nbits)
    rng = np.random.default_rng(1)
    seqs = rng.integers(0, 4, 3 * 500).astype(np.uint8)
    d_seq = torch.from_numpy(seqs).cuda()
    d_off = torch.tensor([0, 500, 1000, 1500], dtype=torch.int64).cuda()
    d_bin = torch.tensor([5, 128 + 64, 7], dtype=torch.int32).cuda()   # the middle one is out of range
    st = torch.cuda.current_stream().cuda_stream
    dy._lib.check(L.dyibf_add_sequences(f._h, C.c_void_p(d_seq.data_ptr()), C.c_void_p(d_off.data_ptr()), 3,
                                        C.c_void_p(d_bin.data_ptr()), ENC_DNA5, MEM_DEVICE, C.c_void_p(st)))
    assert f.device_errors() == DEVERR_BIN_RANGE
    assert f.device_errors() == 0                                      # reset by the read above
    g = dy.SeqAnBloomFilter(bins, h, k, nbits)
    g.addSequences([seqs[:500], seqs[1000:]], [5, 7])
    assert np.array_equal(f.download_words(), g.download_words())     # the bad sequence added nothing, anywhere
    # a 500-base sequence under a promised bound of 100 bases: 8-plane counters would wrap; it selects nothing + flag
    d_rows = torch.zeros(3 * 2, dtype=torch.int64).cuda()
    f.query_device(d_seq.data_ptr(), d_off.data_ptr(), 3, 10, d_rows.data_ptr(), group=1, derive_rc=False, max_seq_len=100, stream=st)
    assert f.device_errors() == DEVERR_SEQ_TOO_LONG
    assert int(d_rows.abs().sum()) == 0
    f.query_device(d_seq.data_ptr(), d_off.data_ptr(), 3, 10, d_rows.data_ptr(), group=1, derive_rc=False, max_seq_len=500, stream=st)
    assert f.device_errors() == 0 and int(d_rows[0]) == 1 << 5
    f.close(); g.close()


# ---- ticketed classify ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("paired", [False, True])
def test_classify_tickets_equal_synchronous(dy, oracle, paired):
    bins, h, k, nbits, thr = 192, 3, 18, (1 << 27) + 256, 20
    contigs, reads = _case(11, bins, 6 * 700, paired=paired)
    f = dy.SeqAnBloomFilter(bins, h, k, nbits)
    f.addSequences(contigs, np.arange(bins))
    group = 4 if paired else 2
    per = 700 * (2 if paired else 1)
    batches = [reads[i * per:(i + 1) * per] for i in range(6)]
    want = [f.classify(b, thr, group=group, derive_rc=True) for b in batches]
    # against the oracle for one batch (the synchronous path is pinned elsewhere; this pins the comparison itself)
    _, _, _, (wo, wi) = _oracle_lists(oracle, contigs, batches[0], bins, h, k, nbits, thr, paired)
    assert np.array_equal(want[0][0], wo) and np.array_equal(want[0][1], wi)
    # three tickets in flight, ended in order; then as many as the handle has (eight) + one too many
    tickets = [f.classifyBegin(b, thr, group=group, derive_rc=True) for b in batches[:3]]
    got = []
    for i in range(6):
        got.append(f.classifyEnd(tickets.pop(0)))
        if i + 3 < 6:
            tickets.append(f.classifyBegin(batches[i + 3], thr, group=group, derive_rc=True))
    for (o, i), (wo_, wi_) in zip(got, want):
        assert np.array_equal(o, wo_) and np.array_equal(i, wi_)
    four = [f.classifyBegin(b, thr, group=group, derive_rc=True) for b in batches[:4]]
    more = [f.classifyBegin(batches[i % 6], thr, group=group, derive_rc=True) for i in range(4, 8)]
    with pytest.raises(dy.DyibfError) as ei:
        f.classifyBegin(batches[4], thr, group=group, derive_rc=True)
    assert ei.value.code == -8                                         # DYIBF_ERR_BUSY
    for j, t in enumerate(more):
        o, i = f.classifyEnd(t)
        assert np.array_equal(o, want[(4 + j) % 6][0]) and np.array_equal(i, want[(4 + j) % 6][1])
    # ended out of order; a too-small buffer keeps the ticket alive
    o3, i3 = f.classifyEnd(four[3])
    assert np.array_equal(o3, want[3][0]) and np.array_equal(i3, want[3][1])
    offs = np.zeros(bins + 1, dtype=np.uint64); tot = C.c_uint64(0); small = np.zeros(4, dtype=np.uint32)
    rc = dy.lib().dyibf_classify_end(four[1].handle, C.c_void_p(offs.ctypes.data), C.c_void_p(small.ctypes.data), 4, C.byref(tot))
    assert rc == -6 and tot.value == want[1][1].size and np.array_equal(offs, want[1][0])
    o1, i1 = f.classifyEnd(four[1])
    assert np.array_equal(i1, want[1][1])
    # the ended ticket's batch is the resident batch: gatherReads works on it without a second upload
    f2 = dy.SeqAnBloomFilter(bins, h, k, nbits); f2.addSequences(contigs, np.arange(bins))
    f2.classify(batches[1], thr, group=group, derive_rc=True)
    b = int(np.argmax(np.diff(want[1][0])))
    for x, y in zip(f.gatherReads(b), f2.gatherReads(b)):
        assert np.array_equal(x, y)
    for t in (four[0], four[2]):
        f.classifyEnd(t)
    f.close(); f2.close()


def test_classify_ticket_list_overflow_refill(dy):
    """threshold 1 on a saturated filter: far more than 4 lists per row — the end call re-fills"""
    bins, h, k, nbits = 256, 2, 12, (1 << 20) + 256
    contigs, reads = _case(3, bins, 300, k=12, genome=20000)
    f = dy.SeqAnBloomFilter(bins, h, k, nbits)
    f.addSequences(contigs, np.arange(bins))
    want = f.classify(reads, 1, group=2, derive_rc=True)
    assert want[1].size > 4 * 300
    got = f.classifyEnd(f.classifyBegin(reads, 1, group=2, derive_rc=True))
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    f.close()


# ---- loader / writer -----------------------------------------------------------------------------------------------------
def test_threaded_loader_and_writer_roundtrip(dy, oracle, tmp_path):
    from dream_yara_b200._lib import IoStats
    bins, h, k = 130, 3, 18
    nbits = 3 * (1 << 30) + 777                                        # 384 MiB, unaligned trailer, W = 3 (odd), several chunks
    contigs, reads = _case(21, bins, 400)
    f = dy.SeqAnBloomFilter(bins, h, k, nbits)
    f.addSequences(contigs, np.arange(bins))
    words = f.download_words()
    p = oracle.params(bins, h, k, nbits)
    ow = oracle.new_words(p)
    for b, c in enumerate(contigs):
        oracle.add_kmers(ow, p, c, b)
    assert np.array_equal(words[: p.P * p.W], ow[: p.P * p.W])
    L = dy.lib()
    paths = []
    for threads in (1, 4, 7):
        assert L.dyibf_set_io_threads(threads) == 0
        path = tmp_path / ("t%d.filter" % threads)
        assert f.save(path)
        st = IoStats(); assert L.dyibf_last_io_stats(C.byref(st)) == 0
        assert st.bytes == p.P * p.W * 8 and st.threads == threads and st.seconds > 0
        paths.append(path)
    ref_path = tmp_path / "oracle.filter"
    oracle.save(ow, p, ref_path)
    blob = open(ref_path, "rb").read()
    for path in paths:
        assert open(path, "rb").read() == blob                         # byte-identical to the reference layout
    for threads in (1, 5):
        L.dyibf_set_io_threads(threads)
        g = dy.SeqAnBloomFilter(str(paths[0]))
        st = IoStats(); L.dyibf_last_io_stats(C.byref(st))
        assert st.threads == threads and st.bytes == p.P * p.W * 8
        assert np.array_equal(g.download_words(), words)
        g.close()
    L.dyibf_set_io_threads(4)
    f.close()


# ---- one process, several GPUs ---------------------------------------------------------------------------------------------
def test_shard_helpers(dy):
    L = dy.lib()
    a, b = C.c_uint32(0), C.c_uint32(0)
    assert L.dyibf_shard_bins(8192, 3, 8, C.byref(a), C.byref(b)) == 0 and (a.value, b.value) == (3072, 1024)
    assert L.dyibf_shard_bins(1000, 3, 4, C.byref(a), C.byref(b)) == 0 and (a.value, b.value) == (768, 232)
    assert L.dyibf_shard_bins(130, 1, 2, C.byref(a), C.byref(b)) == -1            # W = 3 does not divide
    assert L.dyibf_shard_reads(10, 2, 4, C.byref(a), C.byref(b)) == 0 and (a.value, b.value) == (6, 9)
    assert L.dyibf_shard_reads(10, 3, 4, C.byref(a), C.byref(b)) == 0 and (a.value, b.value) == (9, 10)
    from dream_yara_b200 import dist
    for n, w in ((10, 4), (7, 8), (100000, 3)):
        for r in range(w):
            L.dyibf_shard_reads(n, r, w, C.byref(a), C.byref(b))
            assert (a.value, b.value) == dist.read_shard_range(n, r, w)


@pytest.mark.parametrize("mode", ["read", "bin"])
def test_group_of_one_equals_plain_handle(dy, oracle, tmp_path, mode):
    bins, h, k, nbits, thr = 256, 3, 18, (1 << 27) + 256, 20
    contigs, reads = _case(31, bins, 900)
    g = dy.FilterGroup(bins, h, k, nbits, devices=[0], mode=mode)
    g.addSequences(contigs, np.arange(bins))
    _, _, _, (wo, wi) = _oracle_lists(oracle, contigs, reads, bins, h, k, nbits, thr, False)
    o, i = g.classify(reads, thr)
    assert np.array_equal(o, wo) and np.array_equal(i, wi)
    o, i = g.classify(reads, thr, capacity=None)
    path = tmp_path / "g.filter"
    assert g.save(path)
    f = dy.SeqAnBloomFilter(str(path))
    o2, i2 = f.classify(reads, thr)
    assert np.array_equal(o2, wo) and np.array_equal(i2, wi)
    f.close(); g.close()


@pytest.mark.parametrize("paired", [False, True])
def test_group_two_gpus_read_and_bin_sharded(dy, oracle, tmp_path, paired):
    if _n_gpus(dy) < 2:
        pytest.skip("needs 2 GPUs")
    bins, h, k, nbits, thr = 512, 3, 18, (1 << 28) + 256, 25
    contigs, reads = _case(41 + paired, bins, 5001, paired=paired)    # odd row count: uneven split
    words, p, rows, (wo, wi) = _oracle_lists(oracle, contigs, reads, bins, h, k, nbits, thr, paired)
    group = 4 if paired else 2
    single = dy.SeqAnBloomFilter(bins, h, k, nbits)
    single.addSequences(contigs, np.arange(bins))
    spath = tmp_path / "single.filter"
    assert single.save(spath)
    for mode in ("read", "bin"):
        g = dy.FilterGroup(bins, h, k, nbits, devices=[0, 1], mode=mode)
        # build through the group: sequences in descending bin order = the unsorted (gathering) routing path
        order = np.arange(bins)[::-1]
        g.addSequences([contigs[b] for b in order], order)
        o, i = g.classify(reads, thr, group=group, derive_rc=True)
        assert np.array_equal(o, wo) and np.array_equal(i, wi), mode
        o, i = g.classify(reads, thr, group=group, derive_rc=True, capacity=None)
        # too small a buffer: offsets + total still come back, fetch completes without a second pass
        L = dy.lib()
        cat = np.concatenate(reads); off = np.zeros(len(reads) + 1, dtype=np.uint64); off[1:] = np.cumsum([len(r) for r in reads])
        offs = np.zeros(bins + 1, dtype=np.uint64); tot = C.c_uint64(0); small = np.zeros(8, dtype=np.uint32)
        rc = L.dyibf_group_classify(g._g, C.c_void_p(cat.ctypes.data), C.c_void_p(off.ctypes.data), len(reads), 0, group, 1, thr,
                                    C.c_void_p(offs.ctypes.data), C.c_void_p(small.ctypes.data), 8, C.byref(tot))
        assert rc == -6 and tot.value == wi.size and np.array_equal(offs, wo)
        ids = np.zeros(wi.size, dtype=np.uint32)
        assert L.dyibf_group_classify_fetch(g._g, C.c_void_p(offs.ctypes.data), C.c_void_p(ids.ctypes.data), ids.size, C.byref(tot)) == 0
        assert np.array_equal(ids, wi)
        # a pinned output buffer: the devices scatter their lists into it themselves (no staging, no host memcpy); a guard word
        # behind the lists stays untouched
        hp = C.c_void_p(); assert L.dyibf_host_alloc(C.byref(hp), (wi.size + 4) * 4) == 0
        pin = np.ctypeslib.as_array(C.cast(hp, C.POINTER(C.c_uint32)), shape=(wi.size + 4,)); pin[:] = 0xdeadbeef
        assert L.dyibf_group_classify_fetch(g._g, C.c_void_p(offs.ctypes.data), hp, wi.size, C.byref(tot)) == 0
        assert np.array_equal(pin[:wi.size], wi) and np.all(pin[wi.size:] == 0xdeadbeef) and np.array_equal(offs, wo), mode
        L.dyibf_host_free(hp)
        # every member's table: replicas are identical to the single-device build (column exchange over NVLink);
        # shards are the column slices of it
        for m in range(2):
            mf, fb, nb = g.member(m)
            mw = mf.download_words()
            if mode == "read":
                assert np.array_equal(mw, single.download_words()), (mode, m)
            else:
                from dream_yara_b200 import dist
                want = dist.extract_bin_shard(words, bins, nbits, m, 2)
                assert (fb, nb) == (m * 256, 256) and np.array_equal(mw[: want.size], want), (mode, m)
        # the group's file is the single device's file, byte for byte; loading it into a fresh group gives the same answers
        gpath = tmp_path / ("g_%s.filter" % mode)
        assert g.save(gpath)
        assert open(gpath, "rb").read() == open(spath, "rb").read()
        g2 = dy.FilterGroup(str(gpath), devices=[0, 1], mode=mode)
        o, i = g2.classify(reads, thr, group=group, derive_rc=True)
        assert np.array_equal(o, wo) and np.array_equal(i, wi), mode
        # clearBins through the group, then compare with the single device
        g2.clearBins([3, 300, 511])
        single2 = dy.SeqAnBloomFilter(str(spath)); single2.clearBins([3, 300, 511])
        o, i = g2.classify(reads, thr, group=group, derive_rc=True)
        so, si = single2.classify(reads, thr, group=group, derive_rc=True)
        assert np.array_equal(o, so) and np.array_equal(i, si)
        single2.close(); g2.close()
        if mode == "read":
            # whole batches to successive replicas
            per = 1000 * (2 if paired else 1)
            bt = [reads[j * per:(j + 1) * per] for j in range(4)]
            tk = [g.classifyBegin(b, thr, group=group, derive_rc=True) for b in bt]
            for b, t in zip(bt, tk):
                so, si = single.classify(b, thr, group=group, derive_rc=True)
                o, i = g.classifyEnd(t)
                assert np.array_equal(o, so) and np.array_equal(i, si)
            # broadcast: rebuild member 1 only through its own handle, then replicate it
            mf1, _, _ = g.member(1)
            mf1.clearBins([9]); g.broadcast(1)
            mf0, _, _ = g.member(0)
            assert np.array_equal(mf0.download_words(), mf1.download_words())
        g.close()
    single.close()


def test_group_load_file_expect_takes_the_callers_parameters(dy, oracle, tmp_path):
    """SeqAnBloomFilter(fileName, n_bins, n_hash_func, kmer_size, vec_size) over a group: like
    dyibf_load_file_expect and the reference (d_bloom_filter.h:59-77) the members hash with the CALLER's bins / h / k, only
    the bit count is checked against the file.  Runs on one GPU (a group of one); with two it covers both sharding modes."""
    contigs, reads = _case(77, 128, 600, k=16, L=80)
    nbits = (1 << 24) + 256
    f = dy.SeqAnBloomFilter(128, 2, 16, nbits)
    f.addSequences(contigs, np.arange(128))
    path = tmp_path / "g.filter"
    f.save(path)
    # a trailer that disagrees with what the caller will say: h = 3, k = 20 written over [bins, h, k, 0] at the file's end
    raw = bytearray(path.read_bytes())
    trailer = np.frombuffer(raw, dtype=np.uint64, count=4, offset=len(raw) - 32).copy()
    assert list(trailer[:3]) == [128, 2, 16]
    trailer[1], trailer[2] = 3, 20
    raw[len(raw) - 32:] = trailer.tobytes()
    path.write_bytes(bytes(raw))
    assert dy.readFilterMetadata(path) == (128, 20)
    thr = 30
    want_off, want_ids = f.classify(reads, thr, group=2, derive_rc=True)
    single = dy.SeqAnBloomFilter(path, 128, 2, 16, nbits)
    a, b = single.classify(reads, thr, group=2, derive_rc=True)
    assert np.array_equal(a, want_off) and np.array_equal(b, want_ids)
    single.close()
    devs = [0, 1] if _n_gpus(dy) >= 2 else [0]
    for mode in ("read", "bin"):
        g = dy.FilterGroup(path, 128, 2, 16, nbits, devices=devs, mode=mode)
        assert (g.getNumberOfBins(), g._noOfHashFunc, g.getKmerSize()) == (128, 2, 16)
        a, b = g.classify(reads, thr, group=2, derive_rc=True)
        assert np.array_equal(a, want_off) and np.array_equal(b, want_ids), mode
        g.close()
    h = C.c_void_p()
    d = (C.c_int * len(devs))(*devs)
    rc = dy.lib().dyibf_group_load_file_expect(str(path).encode(), 128, 2, 16, nbits + 64, d, len(devs), 0, C.byref(h))
    assert rc == -5 and b"Size mismatch" in dy.lib().dyibf_last_error() and not h.value
    f.close()
