"""GPU (B200): the kmer_direct filter (`-ft kmer_direct`, reference src/d_kdx_filter.h) through the
C ABI (a filter with zero hash functions) against the vectors of the unmodified reference class and the oracle."""
import os

import numpy as np
import pytest

import ka

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def kdx_golden():
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "kdx_vectors.npz"))


def test_kdx_ka1_file_rows_counts_and_add(dy, golden, kdx_golden, tmp_path):
    genomes, reads = ka.records(golden, "genomes"), ka.records(golden, "reads")
    reads = [reads[int(i)] for i in kdx_golden["read_idx"]]
    k, bins, nbits = int(kdx_golden["kx1_k"]), int(kdx_golden["kx1_bins"]), int(kdx_golden["kx1_nbits"])
    assert dy.SeqAnKDXFilter.vectorSize(bins, k) == nbits
    f = dy.SeqAnKDXFilter(bins, k, nbits)
    assert (f._noOfHashFunc, f._noOfHashPos) == (0, 1 << (2 * k))
    f.addSequences(genomes, np.arange(bins))
    path = tmp_path / "kx1.filter"
    assert f.save(path)
    assert ka.md5_file(path) == str(kdx_golden["kx1_md5"])
    assert np.array_equal(f.download_words(f.n_words - 4, 4), kdx_golden["kx1_trailer"])
    f.set_early_exit(False)
    assert np.array_equal(f.whichBinsBatch(reads, 53, group=2, derive_rc=True), kdx_golden["kx1_rows"])
    f.set_early_exit(True)
    assert np.array_equal(f.whichBinsBatch(reads, 53, group=2, derive_rc=True), kdx_golden["kx1_rows"])
    f.set_early_exit(False)
    assert np.array_equal(f.whichBinsBatch(reads[:800], 30, group=2, derive_rc=True), kdx_golden["kx1_rows_t30"])
    _, cnt = f.whichBinsBatch([reads[int(i)] for i in kdx_golden["kx1_count_reads"]], 53, group=1, counts=True)
    assert np.array_equal(cnt, kdx_golden["kx1_counts"])
    offs, ids = f.classify(reads, 53, group=2, derive_rc=True)
    want = ka.bits_from_rows(kdx_golden["kx1_rows"], bins)
    assert np.array_equal(np.diff(offs.astype(np.int64)), want.sum(axis=0))
    # the file constructor recognises the kind by the zero hash count in the trailer
    g = dy.SeqAnKDXFilter(str(path))
    assert (g.getNumberOfBins(), g.getKmerSize(), g._noOfHashFunc, g._noOfBits) == (bins, k, 0, nbits)
    assert np.array_equal(g.whichBinsBatch(reads[:600], 53, group=2, derive_rc=True), kdx_golden["kx1_rows"][:600])
    g.close()
    assert dy.readFilterMetadata(str(path)) == (bins, k)
    # incremental add
    f.addKmers(genomes[5][::-1], 3)
    assert ka.md5_filter_words(f.download_words(), nbits) == str(kdx_golden["kx1u_md5"])
    assert np.array_equal(f.whichBinsBatch(reads[:1200], 53, group=2, derive_rc=True), kdx_golden["kx1u_rows"])
    f.close()


def test_kdx_two_words_per_block(dy, golden, kdx_golden):
    genomes, reads = ka.records(golden, "genomes"), ka.records(golden, "reads")
    reads = [reads[int(i)] for i in kdx_golden["read_idx"]]
    k, bins, nbits = int(kdx_golden["kx2_k"]), int(kdx_golden["kx2_bins"]), int(kdx_golden["kx2_nbits"])
    f = dy.SeqAnKDXFilter(bins, k, nbits)
    f.addSequences([genomes[b % 64] if b < 64 else genomes[b % 64][::-1] for b in range(bins)], np.arange(bins))
    assert ka.md5_filter_words(f.download_words(), nbits) == str(kdx_golden["kx2_md5"])
    assert np.array_equal(f.whichBinsBatch(reads[:2000], 57, group=2, derive_rc=True), kdx_golden["kx2_rows"])
    f.close()


@pytest.mark.parametrize("bins,k,blocks", [(64, 9, None), (130, 8, None), (64, 10, 50_000), (1024, 7, None)])
def test_kdx_vs_oracle_random_incl_out_of_table_kmers(dy, oracle, bins, k, blocks):
    """random filters incl. what the reference cannot do: sequences that start with N (k-mer values beyond 4^k), a table
    with fewer than 4^k blocks, bins not a multiple of 64 — such k-mers add nothing and match nothing; clearBins"""
    rng = np.random.default_rng(bins * 100 + k)
    W = (bins + 63) // 64
    nbits = (blocks if blocks else (1 << (2 * k))) * 64 * W + 256 + (64 if bins == 130 else 0)
    p = oracle.kdx_params(bins, k, nbits)
    words = oracle.new_words(p)
    f = dy.SeqAnKDXFilter(bins, k, nbits)
    assert f._noOfHashPos == p.P
    seqs = []
    for b in range(bins):
        c = rng.integers(0, 4, int(rng.integers(k - 2, 600))).astype(np.uint8)
        c[rng.random(c.size) < 0.03] = 4
        if b % 5 == 0:
            c[0] = 4
        seqs.append(c)
        oracle.kdx_add_kmers(words, p, c, b)
    f.addSequences(seqs, np.arange(bins))
    nd = p.P * p.W
    assert np.array_equal(f.download_words()[:nd], words[:nd])
    reads = []
    for i in range(300):
        src = seqs[int(rng.integers(0, bins))]
        if src.size > 80 and i % 3:
            st = int(rng.integers(0, src.size - 60)); c = src[st:st + 60].copy()
        else:
            c = rng.integers(0, 5, int(rng.integers(k, 120))).astype(np.uint8)
        if i % 7 == 0:
            c[0] = 4
        if i % 11 == 0:
            c[-1] = 4
        reads.append(c)
    for thr in (1, 5, 20):
        assert np.array_equal(f.whichBinsBatch(reads, thr, group=2, derive_rc=True), oracle.kdx_rows(words, p, reads, thr))
    _, cnt = f.whichBinsBatch(reads[:40], 1, group=1, counts=True)
    assert np.array_equal(cnt, np.stack([oracle.kdx_count_bins(words, p, c) for c in reads[:40]]))
    clear = [0, bins - 1, bins // 2]
    f.clearBins(clear)
    oracle.clear_bins(words, p, clear)
    assert np.array_equal(f.download_words()[:nd], words[:nd])
    f.close()
