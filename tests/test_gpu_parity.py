"""GPU (B200): the CUDA path, called through the C ABI (libdyibf.so), against the oracle and the golden vectors.

Bit-exact bar: identical .filter bytes, identical per-read bin sets, identical per-bin k-mer counts,
identical per-bin read-id lists.
"""
import hashlib
import os

import numpy as np
import pytest

import ka
from oracle_py import full_layout

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def genomes(golden):
    return ka.records(golden, "genomes")


@pytest.fixture(scope="module")
def reads(golden):
    return ka.records(golden, "reads")


def _gpu_build(dy, records_per_bin, h, k, nbits=ka.GIB_BITS):
    f = dy.SeqAnBloomFilter(len(records_per_bin), h, k, nbits)
    seqs, bins = ka.flat(records_per_bin)
    keep = [(s, b) for s, b in zip(seqs, bins) if len(s) >= k]
    f.addSequences([s for s, _ in keep], [b for _, b in keep])
    return f


@pytest.fixture(scope="module")
def ka1_filter(dy, genomes):
    return _gpu_build(dy, [[g] for g in genomes], 2, 19)


def test_ka1_build_file_md5_and_query(dy, golden, genomes, reads, ka1_filter, tmp_path):
    f = ka1_filter
    path = tmp_path / "gpu_ka1.filter"
    assert f.save(path)
    assert path.stat().st_size == 1073741864
    assert ka.md5_file(path) == str(golden["reference_gold_md5"])
    f.set_early_exit(False)                                        # the reference's every-k-mer pass first
    rows = f.whichBinsBatch(reads, 25, group=2, derive_rc=True)
    assert np.array_equal(rows, golden["ka1_rows"])
    f.set_early_exit(True)
    assert np.array_equal(f.whichBinsBatch(reads, 25, group=2, derive_rc=True), golden["ka1_rows"])
    f.set_early_exit(False)
    # per-bin k-mer counts of single sequences
    idx = [int(i) for i in golden["ka1_count_reads"]]
    _, cnt = f.whichBinsBatch([reads[i] for i in idx], 25, group=1, counts=True)
    assert np.array_equal(cnt, golden["ka1_counts"])
    # reload from the file: parameters come from the trailer
    g = dy.SeqAnBloomFilter(str(path))
    assert (g.getNumberOfBins(), g.getKmerSize(), g._noOfHashFunc, g._noOfBits) == (64, 19, 2, ka.GIB_BITS)
    assert np.array_equal(g.whichBinsBatch(reads[:500], 25, group=2, derive_rc=True), golden["ka1_rows"][:500])
    assert dy.readFilterMetadata(str(path)) == (64, 19)
    g.close()


def test_ka2_baseline_config1(dy, golden, genomes, reads):
    f = _gpu_build(dy, [[g] for g in genomes], 3, 18)
    words = f.download_words()
    assert ka.md5_filter_words(words, f._noOfBits) == str(golden["ka2_md5"])
    assert np.array_equal(f.whichBinsBatch(reads, 29, group=2, derive_rc=True), golden["ka2_rows"])
    idx = [int(i) for i in golden["ka1_count_reads"]]
    _, cnt = f.whichBinsBatch([reads[i] for i in idx], 29, group=1, counts=True)
    assert np.array_equal(cnt, golden["ka2_counts"])
    f.close()


def test_ka4_130_bins(dy, golden, genomes, reads):
    f = _gpu_build(dy, ka.ka4_bins(genomes), 3, 18)
    assert (f._binIntWidth, f._noOfHashPos) == (3, 44739242)
    assert ka.md5_filter_words(f.download_words(), f._noOfBits) == str(golden["ka4_md5"])
    f.set_early_exit(False)
    assert np.array_equal(f.whichBinsBatch(reads, 29, group=2, derive_rc=True), golden["ka4_rows"])
    f.set_early_exit(True)
    assert np.array_equal(f.whichBinsBatch(reads, 29, group=2, derive_rc=True), golden["ka4_rows"])
    f.close()


@pytest.mark.parametrize("tag,paired", [("se", False), ("pe", True)])
def test_ka5_clasify_loaded_reads(dy, golden, reads, ka1_filter, tag, paired):
    """The mapper-stage mirror against the reference's own clasifyLoadedReads output (SE and PE)."""
    if "ka5_se_threshold" not in golden.files:
        pytest.skip("fixture made without libref_mapper.so")
    opts = dy.DisOptions(kmerSize=19, numberOfBins=64)
    opts.setErrorPercent(3)
    rs = dy.ReadSet(reads, paired=paired, rc=False)
    thr = dy.clasifyLoadedReads(rs, ka1_filter, opts)
    assert thr == int(golden["ka5_%s_threshold" % tag])
    bo, ids = golden["ka5_%s_bin_offsets" % tag], golden["ka5_%s_read_ids" % tag]
    for b in range(64):
        assert np.array_equal(opts.origReadIdMap[b], ids[int(bo[b]):int(bo[b + 1])]), b


def test_random_cases_vs_reference_vectors(dy, golden):
    for ci in golden["random_cases"]:
        pre = "rc%d_" % ci
        bins, h, k, nbits = (int(x) for x in golden[pre + "params"])
        f = dy.SeqAnBloomFilter(bins, h, k, nbits)
        f.addSequences((golden[pre + "seqs"], golden[pre + "seq_offsets"]), golden[pre + "bin_ids"])
        words = f.download_words()
        data = words[: f._noOfHashPos * f._binIntWidth]
        assert hashlib.sha256(data.tobytes()).hexdigest() == str(golden[pre + "words_sha256"]), pre
        thr = int(golden[pre + "threshold"])
        q = (golden[pre + "queries"], golden[pre + "query_offsets"])
        rows = f.whichBinsBatch(q, thr, group=2, derive_rc=True)
        assert np.array_equal(rows, golden[pre + "rows"]), pre
        nq = q[1].size - 1
        _, cnt = f.whichBinsBatch(q, thr, group=1, counts=True)
        assert np.array_equal(cnt[:6], golden[pre + "counts6"]), pre
        assert cnt.shape == (nq, bins)
        f.close()


def _random_batch(rng, n, k, lo=None, hi=250, p_n=0.01):
    seqs = []
    for _ in range(n):
        L = int(rng.integers(lo if lo is not None else max(k - 3, 0), hi))
        c = rng.integers(0, 4, size=L).astype(np.uint8)
        c[rng.random(L) < p_n] = 4
        if L and rng.random() < 0.1:
            c[0] = 4
        if L and rng.random() < 0.1:
            c[-1] = 4
        seqs.append(c)
    return seqs


@pytest.mark.parametrize("bins,h,k,nbits", [
    (1, 2, 14, (1 << 16) + 256), (64, 3, 18, (1 << 22) + 256), (100, 2, 20, 3000001), (128, 3, 18, (1 << 22) + 256),
    (256, 4, 24, (1 << 23) + 256), (500, 5, 32, (1 << 23) + 256), (1024, 3, 18, (1 << 26) + 256),
    (2048, 2, 18, (1 << 24) + 256), (4000, 3, 15, 90000001), (8192, 3, 18, (1 << 27) + 256), (640, 1, 18, (1 << 22) + 256),
    (1024, 6, 18, (1 << 24) + 256),
])
def test_build_query_counts_vs_oracle(dy, oracle, bins, h, k, nbits):
    """Every kernel variant (chunk width, lanes per block, h, plane count) against the oracle: filter words,
    counts, rows for SE and PE layouts incl. sequences shorter than k, empty sequences and N runs."""
    rng = np.random.default_rng(bins * 131 + h * 7 + k)
    p = oracle.params(bins, h, k, nbits)
    words = oracle.new_words(p)
    contigs = _random_batch(rng, 60, k, lo=0, hi=3000)
    bin_ids = rng.integers(0, bins, size=len(contigs)).astype(np.uint32)
    bin_ids[:4] = [0, bins - 1, bins // 2, min(63, bins - 1)]
    for c, b in zip(contigs, bin_ids):
        oracle.add_kmers(words, p, c, int(b))
    f = dy.SeqAnBloomFilter(bins, h, k, nbits)
    f.addSequences(contigs, bin_ids)
    nd = p.P * p.W
    assert np.array_equal(f.download_words()[:nd], words[:nd])

    # queries: substrings of the contigs (hits) + random (misses) + degenerate
    qs = []
    for _ in range(200):
        c = contigs[int(rng.integers(0, len(contigs)))]
        if len(c) > 160 and rng.random() < 0.7:
            st = int(rng.integers(0, len(c) - 150)); q = c[st:st + int(rng.integers(k, 150))].copy()
            m = rng.random(len(q)) < 0.02
            q[m] = rng.integers(0, 4, int(m.sum()))
        else:
            q = _random_batch(rng, 1, k)[0]
        qs.append(q)
    qs[0] = np.zeros(0, dtype=np.uint8); qs[1] = qs[2][: k - 1]; qs[3] = np.full(40, 4, dtype=np.uint8)
    while len(qs) % 4:
        qs.append(qs[-1])
    cat = np.concatenate(qs); off = np.zeros(len(qs) + 1, dtype=np.uint64); off[1:] = np.cumsum([len(q) for q in qs])
    fcat, foff = full_layout(cat, off, oracle)
    for thr in (0, 1, 4, 30):
        for paired in (False, True):
            want = oracle.classify(words, p, fcat, foff, paired, thr)
            got = f.whichBinsBatch((cat, off), thr, group=4 if paired else 2, derive_rc=True)
            assert np.array_equal(got, want), (thr, paired)
            got2 = f.whichBinsBatch((fcat, foff), thr, group=4 if paired else 2, derive_rc=False)
            assert np.array_equal(got2, want), (thr, paired, "full layout")
    # early-exit mode (DYIBF_OPT_EARLY_EXIT): identical rows from fewer lookups, at every threshold
    f.profile_enable(True)
    for thr in (0, 1, 4, 30, 61, 120, 133, 200, 300):
        for paired in (False, True):
            want = oracle.classify(words, p, fcat, foff, paired, thr)
            f.set_early_exit(False); f.profile(reset=True)
            full = f.whichBinsBatch((cat, off), thr, group=4 if paired else 2, derive_rc=True)
            n_full = f.profile(reset=True)["query_lookups"]
            f.set_early_exit(True)
            got = f.whichBinsBatch((cat, off), thr, group=4 if paired else 2, derive_rc=True)
            n_ee = f.profile(reset=True)["query_lookups"]
            assert np.array_equal(full, want) and np.array_equal(got, want), (thr, paired, "early exit")
            assert n_ee <= n_full and (thr < 30 or n_ee < n_full), (thr, n_ee, n_full)
    n_tiles = -(-(p.W // 2 if p.W % 2 == 0 else p.W) // 32)
    total_kmers = sum(max(len(q) - k + 1, 0) for q in qs) * 2
    assert n_full == total_kmers * n_tiles                          # full mode looks every k-mer of both strands up
    f.set_early_exit(False); f.profile_enable(False)
    # counts per sequence (both strands)
    _, cnt = f.whichBinsBatch((cat, off), 1, group=2, derive_rc=True, counts=True)
    n = len(qs)
    for i in list(range(8)) + [n - 1]:
        assert np.array_equal(cnt[i], oracle.count_bins(words, p, qs[i])), i
        assert np.array_equal(cnt[n + i], oracle.count_bins(words, p, oracle.revcomp(qs[i]))), ("rc", i)
    # compaction
    rows = f.whichBinsBatch((cat, off), 1, group=2, derive_rc=True)
    wo, wi = oracle.compact(rows, bins)
    go, gi = f.compact(rows)
    assert np.array_equal(go, wo) and np.array_equal(gi, wi)
    co, cid = f.classify((cat, off), 1, group=2, derive_rc=True)
    assert np.array_equal(co, wo) and np.array_equal(cid, wi)
    f.close()


def test_long_reads_use_16bit_counters(dy, oracle):
    bins, h, k, nbits = 1024, 3, 18, (1 << 24) + 256
    rng = np.random.default_rng(5)
    p = oracle.params(bins, h, k, nbits)
    words = oracle.new_words(p)
    contig = rng.integers(0, 4, 5000).astype(np.uint8)
    oracle.add_kmers(words, p, contig, 77)
    f = dy.SeqAnBloomFilter(bins, h, k, nbits)
    f.addSequences([contig], [77])
    qs = [contig[100:1500], contig[:300], rng.integers(0, 4, 2000).astype(np.uint8), contig[4000:]]
    rows, cnt = f.whichBinsBatch(qs, 900, group=1, counts=True)
    for i, q in enumerate(qs):
        c = oracle.count_bins(words, p, q)
        assert np.array_equal(cnt[i], c)
        assert np.array_equal(ka.bits_from_rows(rows[i:i + 1], bins)[0].astype(bool), c >= 900)
    assert cnt[0][77] == 1400 - k + 1
    f.close()


def test_clear_bins_and_update(dy, oracle):
    """dream_yara_update_filter's core: clearBins then re-add (d_update_filter.cpp:188-252)."""
    bins, h, k, nbits = 200, 3, 18, (1 << 22) + 256
    rng = np.random.default_rng(9)
    p = oracle.params(bins, h, k, nbits)
    words = oracle.new_words(p)
    contigs = [rng.integers(0, 4, 2000).astype(np.uint8) for _ in range(bins)]
    f = dy.SeqAnBloomFilter(bins, h, k, nbits)
    f.addSequences(contigs, np.arange(bins))
    for b, c in enumerate(contigs):
        oracle.add_kmers(words, p, c, b)
    clear = [3, 64, 65, 199]
    oracle.clear_bins(words, p, clear)
    f.clearBins(clear, 4)
    nd = p.P * p.W
    assert np.array_equal(f.download_words()[:nd], words[:nd])
    new = rng.integers(0, 4, 1500).astype(np.uint8)
    oracle.add_kmers(words, p, new, 64); f.addKmers(new, 64)
    assert np.array_equal(f.download_words()[:nd], words[:nd])
    f.close()


def test_single_sequence_api_and_errors(dy, oracle, tmp_path):
    bins, h, k, nbits = 70, 2, 16, 70000                     # unaligned trailer
    f = dy.SeqAnBloomFilter(bins, h, k, nbits)
    p = oracle.params(bins, h, k, nbits)
    words = oracle.new_words(p)
    text = "ACGTTGCANNACGTAGCTAGCTAGCATCGATCGATCGATGCATGCATGCATCGATCGTAGCTAGCTAGCRYACGT"
    f.addKmers(text, 69); oracle.add_kmers(words, p, oracle.encode_ascii(text), 69)
    sel = f.whichBins(text[5:60], 3)
    want = oracle.which_bins(words, p, oracle.encode_ascii(text[5:60]), 3).astype(bool)
    assert np.array_equal(sel, want) and sel[69]
    acc = np.zeros(bins, dtype=bool); acc[1] = True
    f.whichBins(text[5:60], 3, acc)                          # OR-accumulates, never clears
    assert acc[1] and acc[69]
    path = tmp_path / "u.filter"
    assert f.save(path)
    oracle.save(words, p, tmp_path / "o.filter")
    assert path.read_bytes() == (tmp_path / "o.filter").read_bytes()
    g = dy.SeqAnBloomFilter(str(path))
    assert (g.getNumberOfBins(), g._noOfHashFunc, g.getKmerSize()) == (bins, h, k)
    g.close()
    with pytest.raises(SystemExit):
        dy.SeqAnBloomFilter(str(tmp_path / "missing.filter"))
    with pytest.raises(SystemExit):
        dy.SeqAnBloomFilter(str(path), bins, h, k, nbits + 64)   # size mismatch -> exit(1) in the reference
    fa = tmp_path / "b.fa"
    fa.write_text(">r1\n%s\n%s\n>short\nACGT\n>r2\nNNNNACGTACGTACGTTTGACCA\n" % (text[:40], text[40:]))
    f.addFastaFile(str(fa), 5)
    oracle.add_kmers(words, p, oracle.encode_ascii(text), 5)
    oracle.add_kmers(words, p, oracle.encode_ascii("NNNNACGTACGTACGTTTGACCA"), 5)
    nd = p.P * p.W
    assert np.array_equal(f.download_words()[:nd], words[:nd])
    f.close()


@pytest.mark.parametrize("tag,paired", [("se", False), ("pe", True)])
def test_gather_reads_and_sorted_bins_vs_reference(dy, golden, gather_golden, ka1_filter, tag, paired):
    """loadFilteredReads + sortedBins on the device against the reference's own output
    (Dna5Q bytes with qualities; SE and PE)."""
    if "ka5_se_threshold" not in golden.files:
        pytest.skip("fixture made without libref_mapper.so")
    g = gather_golden
    f = ka1_filter
    thr = int(golden["ka5_%s_threshold" % tag])
    offs, ids = f.classify((g[tag + "_fwd_bytes"], golden["reads_offsets"]), thr, group=4 if paired else 2, derive_rc=True,
                           encoding=dy.ENC_DNA5Q)
    assert np.array_equal(offs, golden["ka5_%s_bin_offsets" % tag])
    assert np.array_equal(ids, golden["ka5_%s_read_ids" % tag])
    assert np.array_equal(f.sortedBins(offs), g[tag + "_sorted_bins"])
    lim, ss = g[tag + "_limits"], g[tag + "_seq_start"]
    for q, b in enumerate(g["bins"]):
        ob, ol, oi = f.gatherReads(int(b))
        a, e = int(ss[q]), int(ss[q + 1])
        assert np.array_equal(oi, g[tag + "_ids"][a:e]), (tag, b)
        assert np.array_equal(ol, lim[a:e + 1] - lim[a]), (tag, b)
        assert np.array_equal(ob, g[tag + "_bytes"][int(lim[a]):int(lim[e])]), (tag, b)


@pytest.mark.parametrize("paired", [False, True])
def test_gather_reads_random_vs_oracle(dy, oracle, paired):
    """Ragged read lengths (incl. empty reads and reads shorter than k), every bin, capacity retry path."""
    bins, h, k, nbits = 96, 3, 18, (1 << 22) + 256
    rng = np.random.default_rng(77 + paired)
    contigs = [rng.integers(0, 4, 3000).astype(np.uint8) for _ in range(bins)]
    f = dy.SeqAnBloomFilter(bins, h, k, nbits)
    f.addSequences(contigs, np.arange(bins))
    reads = []
    for i in range(400):
        c = contigs[int(rng.integers(0, bins))]
        L = int(rng.integers(0, 160))
        st = int(rng.integers(0, len(c) - 160))
        q = (rng.integers(0, 63, L) << 2).astype(np.uint8) | c[st:st + L]
        q[rng.random(L) < 0.02] = 252
        reads.append(q.astype(np.uint8))
    cat = np.concatenate(reads); off = np.zeros(len(reads) + 1, dtype=np.uint64); off[1:] = np.cumsum([len(r) for r in reads])
    offs, ids = f.classify((cat, off), 5, group=4 if paired else 2, derive_rc=True, encoding=dy.ENC_DNA5Q, capacity=None)
    fcat, foff = oracle.full_layout_dna5q(cat, off)
    assert int(offs[-1]) > 100
    for b in range(bins):
        mine = ids[int(offs[b]):int(offs[b + 1])]
        wb, wl, wi = oracle.load_filtered_reads(fcat, foff, paired, mine)
        ob, ol, oi = f.gatherReads(b)
        assert np.array_equal(oi, wi) and np.array_equal(ol, wl) and np.array_equal(ob, wb), b
    sb = f.sortedBins(offs)
    sizes = np.diff(offs.astype(np.int64))
    assert sorted(sb.tolist()) == list(range(bins)) and np.all(np.diff(sizes[sb]) <= 0)
    f.close()


@pytest.mark.parametrize("log2_bits", [32, 33])
def test_bucketed_build_equals_direct_and_oracle(dy, oracle, log2_bits):
    """The bucketed build (DYIBF_OPT_BUILD_BUCKET_MIB; tables larger than the L2, >= 32 M updates per call) against the
    oracle and the direct atomic kernel: auto budget, a budget so small that most updates overflow into the direct
    fallback, and the device-pointer entry with several scatter/apply passes.  512 MiB table: 32 regions of 16 MiB, one
    split level; 1 GiB table: 64 regions, two split levels."""
    import torch
    from bench_support import synth
    bins, h, k, nbits = 1000, 3, 18, (1 << log2_bits) + 256
    glen = 13_000
    genomes = synth.genomes(bins, glen)                           # 13 M bases -> 39 M updates, N runs, leading-N bins
    genomes[5, :40] = 4
    off = np.arange(bins + 1, dtype=np.uint64) * glen
    ids = np.arange(bins, dtype=np.uint32)
    p = oracle.params(bins, h, k, nbits)
    words = oracle.new_words(p)
    for b in range(bins):
        oracle.add_kmers(words, p, genomes[b], b)
    nd = p.P * p.W
    want = words[:nd]

    def built(mib, device_mode=False, chunk_mib=None):
        f = dy.SeqAnBloomFilter(bins, h, k, nbits)
        f.set_build_bucket_mib(mib)
        if chunk_mib:
            f.set_build_chunk_mib(chunk_mib)
        f.profile_enable(True)
        if device_mode:
            d_seq = torch.from_numpy(genomes.reshape(-1)).cuda()
            d_off = torch.from_numpy(off.view(np.int64)).cuda()
            d_ids = torch.from_numpy(ids.view(np.int32)).cuda()
            from dream_yara_b200._lib import check
            check(dy.lib().dyibf_add_sequences(f._h, d_seq.data_ptr(), d_off.data_ptr(), bins, d_ids.data_ptr(), dy.ENC_DNA5, 1, None))
            torch.cuda.synchronize()
        else:
            f.addSequences((genomes.reshape(-1), off), ids)
        got = f.download_words()[:nd]
        prof = f.profile()
        f.close()
        return got, prof

    two_level = log2_bits > 32
    direct, p0 = built(0)
    assert np.array_equal(direct, want) and p0["build_launches"] == 1
    auto, p1 = built(-1)                                           # everything in one sweep over the table
    assert np.array_equal(auto, want) and p1["build_apply_launches"] == 1
    assert (p1["build_split2_launches"] >= 1) == two_level
    tiny, p2 = built(32)                                           # 32 MiB of records for 39 M updates: overflow path
    assert np.array_equal(tiny, want) and p2["build_apply_launches"] >= 1
    dev, p3 = built(96, device_mode=True)                          # device entry, several sweeps
    assert np.array_equal(dev, want) and p3["build_apply_launches"] >= 2
    many, p4 = built(-1, chunk_mib=1)                              # 13 host chunks: split launches overlap on two streams
    assert np.array_equal(many, want) and p4["build_launches"] >= 14 and p4["build_apply_launches"] == 1


@pytest.mark.parametrize("bins,nbits,h,k", [(1000, (1 << 32) + 256, 3, 18), (130, 3 * (1 << 30) + 777, 2, 19), (64, (1 << 30) + 256, 3, 12),
                                            (200, (1 << 31) + 256, 5, 20)])
@pytest.mark.parametrize("max_cols", [0, 2, 4])
def test_sliced_build_equals_oracle(dy, oracle, bins, nbits, h, k, max_cols, monkeypatch):
    """The bit-sliced build (DYIBF_OPT_BUILD_MODE = 3: per-bin Bloom bit vectors in the L2, then interleaved with bit-matrix
    transposes) against the oracle: W = 16 (all columns in one pass: whole table rows), W = 3 (odd, non-power-of-two P,
    unaligned trailer), W = 1, W = 4 with bins that are not a multiple of 64 and the generic hash count; the same with the
    narrower column groups a small scratch would get (max_cols = 2 / 4: the interleave of one group overlaps the scatter of
    the next through two shadow buffers and a second stream); sequences in bin order
    and shuffled, host buffers in several chunks and device buffers, leading-N and short sequences, onto a table that already
    holds other bits (the interleave ORs)."""
    if max_cols:
        monkeypatch.setenv("DYIBF_SLICED_MAX_COLS", str(max_cols))
    else:
        monkeypatch.delenv("DYIBF_SLICED_MAX_COLS", raising=False)
    import torch
    from bench_support import synth
    from dream_yara_b200._lib import BUILD_SLICED, check
    glen = 9_000
    genomes = synth.genomes(bins, glen)
    genomes[min(5, bins - 1), :40] = 4
    rng = np.random.default_rng(bins)
    seqs, ids = [], []
    for b in range(bins):
        cut = int(rng.integers(1, glen - 1))
        seqs += [genomes[b, :cut], genomes[b, cut:]]; ids += [b, b]
        if b % 17 == 0:
            seqs.append(genomes[b, :k - 1]); ids.append(b)              # shorter than k: adds nothing
    p = oracle.params(bins, h, k, nbits)
    words = oracle.new_words(p)
    pre = rng.integers(0, 4, 5000).astype(np.uint8)
    oracle.add_kmers(words, p, pre, bins - 1)
    for s_, b in zip(seqs, ids):
        oracle.add_kmers(words, p, s_, b)
    nd = p.P * p.W
    order = rng.permutation(len(seqs))
    for variant in ("host", "host_chunks_shuffled", "device", "host_two_calls"):
        f = dy.SeqAnBloomFilter(bins, h, k, nbits)
        f.addSequences([pre], [bins - 1])                              # pre-existing content
        f.set_build_mode(BUILD_SLICED)
        f.profile_enable(True)
        if variant == "host":
            f.addSequences(seqs, ids)
        elif variant == "host_chunks_shuffled":
            f.set_build_chunk_mib(1)
            f.addSequences([seqs[i] for i in order], [ids[i] for i in order])
        elif variant == "host_two_calls":                              # the second call finds the shadow buffers zeroed by the first
            half = len(seqs) // 2
            f.addSequences(seqs[half:], ids[half:])
            assert f.last_build_mode() == BUILD_SLICED
            f.addSequences(seqs[:half], ids[:half])
        else:
            cat, off, enc = dy.as_batch(seqs)
            d_seq = torch.from_numpy(cat).cuda(); d_off = torch.from_numpy(off.view(np.int64)).cuda()
            d_ids = torch.from_numpy(np.asarray(ids, dtype=np.int32)).cuda()
            check(dy.lib().dyibf_add_sequences(f._h, d_seq.data_ptr(), d_off.data_ptr(), len(seqs), d_ids.data_ptr(), enc, 1, None))
            torch.cuda.synchronize()
        assert f.last_build_mode() == BUILD_SLICED, variant
        got = f.download_words()[:nd]
        prof = f.profile()
        assert np.array_equal(got, words[:nd]), (variant, int((got != words[:nd]).sum()))
        cols = min(p.W, max_cols) if max_cols else min(p.W, 16)
        assert prof["build_apply_launches"] >= (p.W + cols - 1) // cols, variant   # one interleave per column group at least
        assert f.device_errors() == 0
        f.close()


def test_sliced_build_auto_choice(dy):
    """auto mode: bin-clustered batches of >= 256 M updates into a table larger than the L2 take the sliced build; the same
    bases cut into 1 kbp pieces of alternating bins do not (their bins' vectors would not fit the L2 together)"""
    from bench_support import synth
    from dream_yara_b200._lib import BUILD_BUCKETED, BUILD_SLICED
    bins, h, k, nbits = 64, 3, 18, (1 << 33) + 256
    glen = 4_000_000                                                # runs of 4 Mbp: a few 16 MiB vectors in flight
    # no N: a piece that STARTS with N hashes its first k-mer differently (raw first character, shape_base.h:479-494), so the
    # subset property below would not hold
    genomes = np.minimum(synth.genomes(bins, glen), 3)
    f = dy.SeqAnBloomFilter(bins, h, k, nbits)
    f.addSequences((genomes.reshape(-1), np.arange(bins + 1, dtype=np.uint64) * glen), np.arange(bins, dtype=np.uint32))
    assert f.last_build_mode() == BUILD_SLICED
    a = f.download_words()
    g = dy.SeqAnBloomFilter(bins, h, k, nbits)
    pieces = genomes.reshape(bins, glen // 1000, 1000).transpose(1, 0, 2).reshape(-1, 1000)      # piece j of every bin, then j + 1
    g.addSequences((pieces.reshape(-1), np.arange(pieces.shape[0] + 1, dtype=np.uint64) * 1000),
                   np.tile(np.arange(bins, dtype=np.uint32), glen // 1000))
    assert g.last_build_mode() == BUILD_BUCKETED
    # the pieces lose the k-mers that span a cut: g's bits are a subset of a's
    b = g.download_words()
    assert np.array_equal(a | b, a) and not np.array_equal(a, b)
    f.close(); g.close()


def test_bucketed_build_wide_records_equals_direct(dy):
    """Tables above 16 GiB have regions of more than 2^21 words, so the level-1 records of the two-level split are 64 bit
    wide.  32 GiB table, GPU against GPU: the bucketed build must leave exactly the words the direct atomic kernel (pinned
    against the oracle at smaller sizes above) leaves."""
    import torch
    from bench_support import synth
    bins, h, k, nbits = 1024, 3, 18, (1 << 38) + 256
    glen = 40_000                                                 # 123 M updates: enough for the bucketed path on 1024 regions
    genomes = synth.genomes(bins, glen)
    off = np.arange(bins + 1, dtype=np.uint64) * glen
    ids = np.arange(bins, dtype=np.uint32)
    tables = []
    for mib in (0, -1):
        W = (bins + 63) // 64
        nd = ((nbits - 256) // (64 * W)) * W
        words = torch.zeros(nd, dtype=torch.int64, device="cuda")
        f = dy.SeqAnBloomFilter.attach(bins, h, k, nbits, words.data_ptr(), device=0, keep=words)
        f.set_build_bucket_mib(mib)
        f.profile_enable(True)
        f.addSequences((genomes.reshape(-1), off), ids)
        torch.cuda.synchronize()
        assert f.profile()["build_apply_launches"] == (0 if mib == 0 else 1)
        f.close()
        tables.append(words)
    assert int(torch.count_nonzero(tables[0])) > 100_000_000
    assert torch.equal(tables[0], tables[1])


def test_bin_sharded_build_equals_column_slices(dy, oracle):
    """SURVEY §8(e) "Build": what rank r of G builds — its own bin range into a self-contained filter with bins/G bins and
    the same P — is exactly its column slice of the one-process filter (here the G shards are built one after the other on
    one GPU; the all-gather interleave is covered by tests/test_dist_gloo.py)."""
    from dream_yara_b200 import dist as D
    rng = np.random.default_rng(77)
    bins, h, k, nbits, G = 512, 3, 18, (1 << 25) + 256, 4
    contigs = [rng.integers(0, 5 if b % 9 == 0 else 4, 3000).astype(np.uint8) for b in range(bins)]
    full = dy.SeqAnBloomFilter(bins, h, k, nbits)
    full.addSequences(contigs, np.arange(bins))
    words = full.download_words()
    p = oracle.params(bins, h, k, nbits)
    want = oracle.new_words(p)
    for b, c in enumerate(contigs):
        oracle.add_kmers(want, p, c, b)
    assert np.array_equal(words[: p.P * p.W], want[: p.P * p.W])
    for r in range(G):
        first, sb, snbits, Wr, P = D.bin_shard_params(bins, nbits, r, G)
        f = dy.SeqAnBloomFilter(sb, h, k, snbits)
        assert f._noOfHashPos == P == full._noOfHashPos
        mine = np.arange(first, first + sb)
        f.addSequences([contigs[int(b)] for b in mine], D.shard_bin_ids(mine, first))
        assert np.array_equal(f.download_words()[: P * Wr], D.extract_bin_shard(words, bins, nbits, r, G))
        f.close()
    full.close()


def test_bucketed_build_match_any_variant_in_subprocess(dy):
    """DYIBF_SPLIT_MATCH=1 (MATCH.ANY ranking of the radix split, read once per process) builds the same table as the
    direct atomic kernel: two split levels, overflow fallback included."""
    import subprocess
    import sys
    code = r'''
import hashlib, sys
import numpy as np
sys.path.insert(0, %r)
import dream_yara_b200 as dy
from bench_support import synth
bins, h, k, nbits, glen = 1000, 3, 18, (1 << 33) + 256, 13000
g = synth.genomes(bins, glen)
off = np.arange(bins + 1, dtype=np.uint64) * glen
out = []
for mib in (0, -1, 32):
    f = dy.SeqAnBloomFilter(bins, h, k, nbits)
    f.set_build_bucket_mib(mib)
    f.addSequences((g.reshape(-1), off), np.arange(bins, dtype=np.uint32))
    out.append(hashlib.md5(f.download_words().tobytes()).hexdigest())
    f.close()
print("MD5S", *out)
''' % ROOT
    env = dict(os.environ, DYIBF_SPLIT_MATCH="1")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env)
    assert r.returncode == 0, r.stderr[-2000:]
    md5s = [l for l in r.stdout.splitlines() if l.startswith("MD5S")][0].split()[1:]
    assert len(md5s) == 3 and md5s[0] == md5s[1] == md5s[2]


@pytest.mark.parametrize("paired", [False, True])
def test_query_chunking_does_not_change_results(dy, golden, reads, ka1_filter, paired):
    """DYIBF_OPT_QUERY_CHUNK_MIB only moves the chunk boundaries of the host-buffer pipeline: rows and per-bin lists are the
    same for the auto rule (a small batch is cut into chunks of at least 8192 rows), 1 MiB chunks (many chunks over both
    streams) and one large chunk — and equal to the reference's clasifyLoadedReads lists where the fixture holds them."""
    f = ka1_filter
    group = 4 if paired else 2
    many = reads * 6 if not paired else reads * 5                       # 38 400 / 32 000 sequences, [m1 | m2] halves for pairs
    out = []
    for mib in (0, 1, 512):
        f.set_query_chunk_mib(mib)
        rows = f.whichBinsBatch(many, 25, group=group, derive_rc=True)
        bo, ids = f.classify(many, 25, group=group, derive_rc=True)
        out.append((rows, bo, ids))
    f.set_query_chunk_mib(0)
    for rows, bo, ids in out[1:]:
        assert np.array_equal(rows, out[0][0]) and np.array_equal(bo, out[0][1]) and np.array_equal(ids, out[0][2])
    rows, bo, ids = out[0]
    bits = np.unpackbits(rows.view(np.uint8), axis=1, bitorder="little")[:, :64]
    assert np.array_equal(bits.sum(axis=0), np.diff(bo.astype(np.int64)))
    for b in (0, 7, 38, 63):
        assert np.array_equal(np.nonzero(bits[:, b])[0], ids[int(bo[b]):int(bo[b + 1])])
    if not paired:                                                      # six copies of KA1's reads: row r == KA1 row r % 6400
        assert np.array_equal(rows, np.tile(golden["ka1_rows"], (6, 1)))


def test_host_batch_argument_errors(dy):
    """Bad limits and over-long sequences are refused by the host-buffer calls (the length scan runs chunk by chunk inside the
    pipeline): DYIBF_ERR_ARG, and the handle stays usable afterwards."""
    import ctypes as C
    from dream_yara_b200._lib import ENC_DNA5, MEM_HOST
    L = dy.lib()
    f = dy.SeqAnBloomFilter(64, 2, 16, (1 << 24) + 256)
    rng = np.random.default_rng(3)
    n = 40_000                                                            # > 4 x 8192 rows: several chunks
    codes = rng.integers(0, 4, size=n * 50, dtype=np.uint8)
    off = np.arange(n + 1, dtype=np.uint64) * 50
    rows = np.zeros((n, 1), dtype=np.uint64)
    bo = np.zeros(65, dtype=np.uint64); ids = np.zeros(4 * n, dtype=np.uint32); tot = C.c_uint64(0)
    ptr = lambda a: C.c_void_p(a.ctypes.data)
    def query(o, max_len=0):
        return L.dyibf_query_batch(f._h, ptr(codes), ptr(o), n, ENC_DNA5, 2, 1, 3, max_len, ptr(rows), None, MEM_HOST, None)
    def classify(o):
        return L.dyibf_classify(f._h, ptr(codes), ptr(o), n, ENC_DNA5, 2, 1, 3, ptr(bo), ptr(ids), ids.size, C.byref(tot))
    assert query(off) == 0
    good = rows.copy()
    bad = off.copy(); bad[n - 5] = bad[n - 4] + 7                         # a decreasing pair in the last chunk
    assert query(bad) == -1 and "monotone" in L.dyibf_last_error().decode()
    assert classify(bad) == -1
    bad = off.copy(); bad[-1] = 0                                         # ends before it starts
    assert query(bad) == -1 and classify(bad) == -1
    assert query(off, max_len=70_000) == -1 and "65535" in L.dyibf_last_error().decode()
    long_codes = rng.integers(0, 4, size=70_000, dtype=np.uint8)          # one sequence with more than 65535 k-mers
    loff = np.array([0, 70_000], dtype=np.uint64)
    one = np.zeros((1, 1), dtype=np.uint64)
    rc = L.dyibf_query_batch(f._h, ptr(long_codes), ptr(loff), 1, ENC_DNA5, 1, 0, 3, 0, ptr(one), None, MEM_HOST, None)
    assert rc == -1 and "65535" in L.dyibf_last_error().decode()
    assert query(off) == 0 and np.array_equal(rows, good)                 # the handle still works
    assert classify(off) == 0 and int(bo[64]) == tot.value
    f.close()
