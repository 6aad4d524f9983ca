"""CPU: pin the C restatement (oracle/ibf_oracle.c) to vectors produced by the unmodified reference
(tests/golden/reference_vectors.npz, made by tests/golden/make_golden.py) and to the reference's own gold md5
(tests/gold/64-viral-binned-genomes.filter.md5, carried in the fixture as `reference_gold_md5`)."""
import hashlib

import numpy as np
import pytest

import ka
from oracle_py import Ref, full_layout


def _build(oracle, records_per_bin, h, k, nbits=ka.GIB_BITS):
    p = oracle.params(len(records_per_bin), h, k, nbits)
    words = oracle.new_words(p)
    for b, recs in enumerate(records_per_bin):
        for r in recs:
            if len(r) >= k:                                  # addFastaFile skips short records (:142)
                oracle.add_kmers(words, p, oracle.encode_ascii(r), b)
    return words, p


def _query(oracle, words, p, reads, thr):
    rows = np.zeros((len(reads), p.W), dtype=np.uint64)
    for i, r in enumerate(reads):
        c = oracle.encode_ascii(r)
        sel = oracle.which_bins(words, p, c, thr)
        oracle.which_bins(words, p, oracle.revcomp(c), thr, sel)
        bits = np.zeros(p.W * 64, dtype=np.uint8); bits[: p.bins] = sel
        rows[i] = np.packbits(bits, bitorder="little").view(np.uint64)
    return rows


@pytest.fixture(scope="module")
def genomes(golden):
    return ka.records(golden, "genomes")


@pytest.fixture(scope="module")
def reads(golden):
    return ka.records(golden, "reads")


def test_ka1_reference_test_config(oracle, golden, genomes, reads, tmp_path):
    """-b 64 -k 19 -nh 2 -bs 1: byte-identical file (gold md5 of the reference repo) and identical bin sets."""
    words, p = _build(oracle, [[g] for g in genomes], 2, 19)
    assert (p.W, p.P) == (1, 1 << 27)
    path = tmp_path / "ka1.filter"
    oracle.save(words, p, path)
    assert path.stat().st_size == int(golden["ka1_size"]) == 1073741864
    assert ka.md5_file(path) == str(golden["reference_gold_md5"]) == str(golden["ka1_md5"])
    assert list(words[-4:]) == [64, 2, 19, 0]
    assert int(np.unpackbits(words.view(np.uint8)).sum()) == int(golden["ka1_popcount"])
    rows = _query(oracle, words, p, reads, 25)
    assert np.array_equal(rows, golden["ka1_rows"])
    per_bin = ka.bits_from_rows(rows, 64).sum(axis=0)
    assert int(per_bin.sum()) == 6538 and list(per_bin[[0, 38, 43, 46]]) == [173, 106, 104, 155]
    # per-bin k-mer counts of a few reads
    for j, i in enumerate(golden["ka1_count_reads"]):
        assert np.array_equal(oracle.count_bins(words, p, oracle.encode_ascii(reads[int(i)])), golden["ka1_counts"][j])
    # load round trip
    w2, p2 = oracle.load(path)
    assert (p2.bins, p2.h, p2.k, p2.nbits) == (64, 2, 19, ka.GIB_BITS) and np.array_equal(w2, words)


def test_ka2_baseline_config1(oracle, golden, genomes, reads):
    words, p = _build(oracle, [[g] for g in genomes], 3, 18)
    oracle.set_metadata(words, p)
    assert ka.md5_filter_words(words, p.nbits) == str(golden["ka2_md5"])
    rows = _query(oracle, words, p, reads, 29)
    assert np.array_equal(rows, golden["ka2_rows"])
    assert int(ka.bits_from_rows(rows, 64).sum()) == 6532
    for j, i in enumerate(golden["ka1_count_reads"]):
        assert np.array_equal(oracle.count_bins(words, p, oracle.encode_ascii(reads[int(i)])), golden["ka2_counts"][j])


def test_ka3_hashes_leading_n(oracle, golden):
    codes = oracle.encode_ascii(golden["ka3_text"].tobytes())
    for k in (14, 18, 32):
        assert np.array_equal(oracle.kmer_hashes(codes, k), golden["ka3_hashes_k%d" % k])
    assert list(golden["ka3_hashes_k18"][:3]) == [70531860217, 282127440868, 1128509763474]


def test_ka4_130_bins_non_pow2(oracle, golden, genomes, reads):
    words, p = _build(oracle, ka.ka4_bins(genomes), 3, 18)
    assert (p.W, p.P) == (3, 44739242)
    oracle.set_metadata(words, p)
    assert ka.md5_filter_words(words, p.nbits) == str(golden["ka4_md5"])
    rows = _query(oracle, words, p, reads, 29)
    assert np.array_equal(rows, golden["ka4_rows"])
    assert int(ka.bits_from_rows(rows, 130).sum()) == 6659


def test_thresholds(oracle, golden):
    for a, e in enumerate(golden["thr_e"]):
        for b, k in enumerate(golden["thr_k"]):
            for c, L in enumerate(golden["thr_L"]):
                assert oracle.threshold(int(e), int(k), int(L)) == golden["thr_table"][a, b, c], (e, k, L)
    assert [oracle.threshold(e, 18, 150) for e in range(6)] == [133, 115, 79, 61, 25, 7]
    assert [oracle.threshold(e, 18, 100) for e in range(6)] == [83, 65, 47, 29, 11, 0]


@pytest.mark.parametrize("tag,paired", [("se", False), ("pe", True)])
def test_ka5_classify_matches_reference_mapper(oracle, golden, genomes, reads, tag, paired):
    """clasifyLoadedReads of the reference itself (SE and PE) vs oracle classify + compact."""
    if "ka5_se_threshold" not in golden.files:
        pytest.skip("fixture made without libref_mapper.so")
    words, p = _build(oracle, [[g] for g in genomes], 2, 19)
    thr = int(golden["ka5_%s_threshold" % tag])
    assert thr == oracle.threshold(3, 19, 100) == 25
    codes = [oracle.encode_ascii(r) for r in reads]
    cat = np.concatenate(codes); off = golden["reads_offsets"]
    fcat, foff = full_layout(cat, off, oracle)
    rows = oracle.classify(words, p, fcat, foff, paired, thr)
    offs, ids = oracle.compact(rows, 64)
    assert np.array_equal(offs, golden["ka5_%s_bin_offsets" % tag])
    assert np.array_equal(ids, golden["ka5_%s_read_ids" % tag])


def test_random_cases(oracle, golden):
    for ci in golden["random_cases"]:
        pre = "rc%d_" % ci
        bins, h, k, nbits = (int(x) for x in golden[pre + "params"])
        p = oracle.params(bins, h, k, nbits)
        words = oracle.new_words(p)
        s_cat, s_off, b_ids = golden[pre + "seqs"], golden[pre + "seq_offsets"], golden[pre + "bin_ids"]
        for i in range(s_off.size - 1):
            oracle.add_kmers(words, p, s_cat[int(s_off[i]):int(s_off[i + 1])], int(b_ids[i]))
        data = words[: p.P * p.W]
        assert hashlib.sha256(data.tobytes()).hexdigest() == str(golden[pre + "words_sha256"]), pre
        q_cat, q_off = golden[pre + "queries"], golden[pre + "query_offsets"]
        thr = int(golden[pre + "threshold"])
        rows = np.zeros((q_off.size - 1, p.W), dtype=np.uint64)
        for i in range(q_off.size - 1):
            c = q_cat[int(q_off[i]):int(q_off[i + 1])]
            sel = oracle.which_bins(words, p, c, thr)
            oracle.which_bins(words, p, oracle.revcomp(c), thr, sel)
            bits = np.zeros(p.W * 64, dtype=np.uint8); bits[:bins] = sel
            rows[i] = np.packbits(bits, bitorder="little").view(np.uint64)
            if i < 6:
                assert np.array_equal(oracle.count_bins(words, p, c), golden[pre + "counts6"][i]), (pre, i)
        assert np.array_equal(rows, golden[pre + "rows"]), pre


def test_clear_bins_and_unaligned_metadata(oracle):
    p = oracle.params(130, 2, 16, 70000)                    # nbits not a multiple of 64: trailer straddles words
    words = oracle.new_words(p)
    rng = np.random.default_rng(1)
    for b in (0, 5, 64, 129):
        oracle.add_kmers(words, p, rng.integers(0, 4, 300).astype(np.uint8), b)
    before = words.copy()
    oracle.clear_bins(words, p, [5, 129])
    bits_b = np.unpackbits(before[: p.P * p.W].view(np.uint8), bitorder="little").reshape(p.P, p.W * 64)
    bits_a = np.unpackbits(words[: p.P * p.W].view(np.uint8), bitorder="little").reshape(p.P, p.W * 64)
    assert not bits_a[:, [5, 129]].any() and bits_b[:, [5, 129]].any()
    keep = np.ones(p.W * 64, dtype=bool); keep[[5, 129]] = False
    assert np.array_equal(bits_a[:, keep], bits_b[:, keep])


@pytest.mark.skipif(not Ref.available(), reason="oracle/_ref/libref_ibf.so not built")
def test_oracle_vs_compiled_reference_direct(oracle):
    """Where the compiled reference is present, compare with it directly on fresh random input."""
    ref = Ref()
    rng = np.random.default_rng(7)
    bins, h, k, nbits = 300, 3, 21, (1 << 22) + 256
    f = ref.new(bins, h, k, nbits)
    p = oracle.params(bins, h, k, nbits)
    words = oracle.new_words(p)
    seqs = []
    for _ in range(50):
        c = rng.integers(0, 5, int(rng.integers(k, 500))).astype(np.uint8)
        b = int(rng.integers(0, bins))
        ref.add_kmers(f, c, b); oracle.add_kmers(words, p, c, b); seqs.append(c)
    assert np.array_equal(ref.words(f)[: p.P * p.W], words[: p.P * p.W])
    for c in seqs[:20]:
        q = c[: 120]
        assert np.array_equal(ref.kmer_hashes(q, k), oracle.kmer_hashes(q, k))
        assert np.array_equal(ref.revcomp(q), oracle.revcomp(q))
        for thr in (1, 3, 50):
            assert np.array_equal(ref.which_bins(f, q, thr, bins), oracle.which_bins(words, p, q, thr))
    ref.free(f)


@pytest.mark.parametrize("tag,paired", [("se", False), ("pe", True)])
def test_load_filtered_reads_vs_reference(oracle, golden, gather_golden, tag, paired):
    """loadFilteredReads restatement (d_mapper.h:393-443) and the Dna5Q reverse complement against the per-bin read
    sets the reference's own loadFilteredReads produced (qualities set, SE and PE)."""
    if "ka5_se_threshold" not in golden.files:
        pytest.skip("fixture made without libref_mapper.so")
    g = gather_golden
    fwd = g[tag + "_fwd_bytes"]
    fcat, foff = oracle.full_layout_dna5q(fwd, golden["reads_offsets"])
    bo, ids = golden["ka5_%s_bin_offsets" % tag], golden["ka5_%s_read_ids" % tag]
    lim, ss = g[tag + "_limits"], g[tag + "_seq_start"]
    for q, b in enumerate(g["bins"]):
        mine = ids[int(bo[b]):int(bo[b + 1])]
        ob, ol, oi = oracle.load_filtered_reads(fcat, foff, paired, mine)
        a, e = int(ss[q]), int(ss[q + 1])
        assert e - a == mine.size * (4 if paired else 2)
        assert np.array_equal(oi, g[tag + "_ids"][a:e]), (tag, b)
        assert np.array_equal(ol, lim[a:e + 1] - lim[a]), (tag, b)
        assert np.array_equal(ob, g[tag + "_bytes"][int(lim[a]):int(lim[e])]), (tag, b)
    # sortedBins: sizes non-increasing along the reference's order
    sizes = np.diff(bo.astype(np.int64))
    sb = g[tag + "_sorted_bins"]
    assert sorted(sb.tolist()) == list(range(64)) and np.all(np.diff(sizes[sb]) <= 0)


@pytest.mark.skipif(not Ref.available(), reason="oracle/_ref/libref_ibf.so not built")
@pytest.mark.parametrize("bins,h,k,nbits", [
    (1, 1, 4, 64 * 50 + 256),                 # one bin, one hash, tiny k
    (63, 2, 12, (1 << 16) + 256),             # W = 1, last word bit unused
    (64, 3, 18, (1 << 18) + 256),             # W = 1 exactly (BASELINE configs[0] shape)
    (65, 4, 19, 128 * 1001 + 256),            # W = 2, P = 1001 (odd)
    (130, 5, 31, 192 * 777 + 256 + 64),       # W = 3, (N - 256) % S != 0: unused words in front of the trailer
    (1024, 3, 32, 1024 * 257 + 256),          # k = 32: the leading-N term is gone from the first k-mer on
    (200, 2, 25, 256 * 4096 + 256),           # power-of-two P with W = 4
])
def test_oracle_vs_compiled_reference_shapes(oracle, bins, h, k, nbits):
    """The restatement against the compiled reference over the corners of the parameter space: W = 1..16, odd and power-of-two P,
    unaligned trailers, k from 4 to 32, 1..5 hash functions; sequences with N runs, leading N, length k exactly."""
    ref = Ref()
    rng = np.random.default_rng(bins * 131 + k)
    f = ref.new(bins, h, k, nbits)
    p = oracle.params(bins, h, k, nbits)
    words = oracle.new_words(p)
    seqs = []
    for i in range(40):
        n = k if i % 7 == 0 else int(rng.integers(k, 300))
        c = rng.integers(0, 4, n).astype(np.uint8)
        if i % 3 == 0:
            c[: int(rng.integers(1, min(n, 6) + 1))] = 4              # leading N (the hashInit quirk)
        if i % 4 == 0:
            a = int(rng.integers(0, n)); c[a:a + int(rng.integers(1, 20))] = 4
        b = int(rng.integers(0, bins))
        ref.add_kmers(f, c, b); oracle.add_kmers(words, p, c, b); seqs.append(c)
    assert np.array_equal(ref.words(f)[: p.P * p.W], words[: p.P * p.W])
    for c in seqs[:25]:
        q = c[: max(k, min(c.size, 150))]
        assert np.array_equal(ref.kmer_hashes(q, k), oracle.kmer_hashes(q, k))
        for thr in (1, 2, max(1, (q.size - k + 1) // 2), q.size - k + 1):
            assert np.array_equal(ref.which_bins(f, q, thr, bins), oracle.which_bins(words, p, q, thr)), (thr, q.size)
    ref.free(f)
