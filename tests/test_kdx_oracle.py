"""CPU: the oracle's restatement of the kmer_direct filter (oracle/ibf_oracle.c kdx_*; reference src/d_kdx_filter.h)
against the vectors of the unmodified reference class (tests/golden/make_golden_kdx.py)."""
import numpy as np
import pytest

import ka
from oracle_py import RefKdx


@pytest.fixture(scope="module")
def kdx_golden():
    import os
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "kdx_vectors.npz"))


def _codes(oracle, golden, kdx_golden):
    genomes = [oracle.encode_ascii(g) for g in ka.records(golden, "genomes")]
    reads = ka.records(golden, "reads")
    reads = [oracle.encode_ascii(reads[int(i)]) for i in kdx_golden["read_idx"]]
    return genomes, reads


def _file_md5(oracle, words, p):
    w = words.copy()
    oracle.set_metadata(w, p)                      # h == 0: the trailer is [bins, 0, k, 0], as SeqAnKDXFilter writes it
    return ka.md5_filter_words(w, p.nbits)


def test_kdx_oracle_ka(oracle, golden, kdx_golden):
    g, r = _codes(oracle, golden, kdx_golden)
    k, bins, nbits = int(kdx_golden["kx1_k"]), int(kdx_golden["kx1_bins"]), int(kdx_golden["kx1_nbits"])
    p = oracle.kdx_params(bins, k, nbits)
    assert p.P * 64 == int(kdx_golden["kx1_hashpos"])            # the reference's _noOfHashPos is 64 x the real block count
    words = oracle.new_words(p)
    for b, c in enumerate(g):
        oracle.kdx_add_kmers(words, p, c, b)
    assert _file_md5(oracle, words, p) == str(kdx_golden["kx1_md5"])
    assert np.array_equal(oracle.kdx_rows(words, p, r[:1500], 53), kdx_golden["kx1_rows"][:1500])
    assert np.array_equal(oracle.kdx_rows(words, p, r[:800], 30), kdx_golden["kx1_rows_t30"])
    cnt = np.stack([oracle.kdx_count_bins(words, p, r[int(i)]) for i in kdx_golden["kx1_count_reads"]])
    assert np.array_equal(cnt, kdx_golden["kx1_counts"])
    oracle.kdx_add_kmers(words, p, g[5][::-1].copy(), 3)
    assert _file_md5(oracle, words, p) == str(kdx_golden["kx1u_md5"])
    assert np.array_equal(oracle.kdx_rows(words, p, r[:1200], 53), kdx_golden["kx1u_rows"])


def test_kdx_oracle_two_words_per_block(oracle, golden, kdx_golden):
    g, r = _codes(oracle, golden, kdx_golden)
    k, bins, nbits = int(kdx_golden["kx2_k"]), int(kdx_golden["kx2_bins"]), int(kdx_golden["kx2_nbits"])
    p = oracle.kdx_params(bins, k, nbits)
    words = oracle.new_words(p)
    for b in range(bins):
        oracle.kdx_add_kmers(words, p, g[b % 64] if b < 64 else g[b % 64][::-1].copy(), b)
    assert _file_md5(oracle, words, p) == str(kdx_golden["kx2_md5"])
    assert np.array_equal(oracle.kdx_rows(words, p, r[:2000], 57), kdx_golden["kx2_rows"])


@pytest.mark.skipif(not RefKdx.available(), reason="oracle/_ref without the kmer_direct shim")
def test_kdx_oracle_vs_compiled_reference_random(oracle):
    rk = RefKdx()
    rng = np.random.default_rng(5)
    for bins, k in ((64, 8), (192, 7), (64, 10)):
        nbits = (1 << (2 * k)) * bins + 256
        f = rk.new(bins, k, nbits)
        p = oracle.kdx_params(bins, k, nbits)
        words = oracle.new_words(p)
        for b in range(bins):
            c = rng.integers(0, 4, int(rng.integers(k, 400))).astype(np.uint8)
            c[rng.random(c.size) < 0.02] = 4
            c[0] = min(c[0], 3)                                  # no leading N: out of bounds in the reference
            rk.add_kmers(f, c, b); oracle.kdx_add_kmers(words, p, c, b)
        nd = p.P * p.W
        assert np.array_equal(rk.words(f)[:nd], words[:nd])
        for _ in range(50):
            c = rng.integers(0, 4, int(rng.integers(k, 200))).astype(np.uint8)
            thr = int(rng.integers(1, 20))
            assert np.array_equal(rk.which_bins(f, c, thr, bins), oracle.kdx_which_bins(words, p, c, thr))
        rk.free(f)
