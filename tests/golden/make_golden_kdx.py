"""Golden vectors for the kmer_direct filter: the UNMODIFIED seqan::SeqAnKDXFilter
(src/d_kdx_filter.h, through oracle/_ref/libref_ibf.so) on the reference's 64 viral genomes and 6400 test reads.
Run in the build container:   python tests/golden/make_golden_kdx.py   -> tests/golden/kdx_vectors.npz
Sizes as dream_yara_build_filter computes them for `-ft kmer_direct` (src/d_build_filter.cpp:305-313):
vec_size = 4^k * bins + 256.  Reads with an N at either end are left out (the reference indexes out of bounds for a strand that starts with N);
all indices below are positions in `read_idx`."""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import ka                                  # noqa: E402
from oracle_py import Ref, RefKdx          # noqa: E402


def rows_of(rk, ref, f, reads_codes, thr, bins):
    W = (bins + 63) // 64
    rows = np.zeros((len(reads_codes), W), dtype=np.uint64)
    for i, c in enumerate(reads_codes):
        sel = rk.which_bins(f, c, thr, bins)
        rk.which_bins(f, ref.revcomp(c), thr, bins, sel)
        bits = np.zeros(W * 64, dtype=np.uint8); bits[:bins] = sel
        rows[i] = np.packbits(bits, bitorder="little").view(np.uint64)
    return rows


def counts_by_sweep(rk, f, codes, bins, k):
    counts = np.zeros(bins, dtype=np.uint16)
    for t in range(1, len(codes) - k + 2):
        sel = rk.which_bins(f, codes, t, bins)
        counts[sel != 0] = t
        if not sel.any():
            break
    return counts


def main():
    ref, rk = Ref(), RefKdx()
    golden = np.load(os.path.join(HERE, "reference_vectors.npz"))
    genomes, reads = ka.records(golden, "genomes"), ka.records(golden, "reads")
    g_codes = [ref.ascii_to_dna5(g) for g in genomes]
    all_codes = [ref.ascii_to_dna5(r) for r in reads]
    assert not any(c[0] == 4 for c in g_codes)
    # reads that begin or end with N are left out: one of their strands starts with N and the reference then reads
    # beyond its bit vector (it segfaults here); the vectors carry the indices of the reads they cover
    keep = np.array([i for i, c in enumerate(all_codes) if c[0] != 4 and c[-1] != 4], dtype=np.uint32)
    r_codes = [all_codes[i] for i in keep]
    tmp = tempfile.mkdtemp(prefix="kdx_")
    out = {"read_idx": keep}

    # KX1: 64 bins, k = 12 -> 128 MiB; threshold getThreshold(e=3, k=12, 100 bp) = 53
    k, bins = 12, 64
    nbits = (1 << (2 * k)) * bins + 256
    f = rk.new(bins, k, nbits)
    for b, g in enumerate(g_codes):
        rk.add_kmers(f, g, b)
    path = os.path.join(tmp, "kx1.filter"); assert rk.save(f, path) == 0
    w = rk.words(f)
    out.update(kx1_k=k, kx1_bins=bins, kx1_nbits=np.uint64(nbits), kx1_md5=np.array(ka.md5_file(path)),
               kx1_trailer=w[-4:].copy(), kx1_hashpos=np.uint64(rk.params(f)[3]),
               kx1_popcount=np.uint64(int(np.unpackbits(w.view(np.uint8)).sum())),
               kx1_rows=rows_of(rk, ref, f, r_codes, 53, bins), kx1_rows_t30=rows_of(rk, ref, f, r_codes[:800], 30, bins))
    idx = np.array([0, 1, 2, 99, 3199, len(r_codes) - 1])
    out.update(kx1_count_reads=idx, kx1_counts=np.stack([counts_by_sweep(rk, f, r_codes[i], bins, k) for i in idx]))
    # incremental add on the same filter.  (clearBins, d_kdx_filter.h:71-97, walks _noOfHashPos = 64 x the real number of
    # blocks and writes far beyond the vector — dream_yara_update_filter -ft kmer_direct segfaults in the reference — so
    # no vector pins it; the device clear covers the real blocks, tested against the oracle only.)
    rk.add_kmers(f, g_codes[5][::-1].copy(), 3)
    path2 = os.path.join(tmp, "kx1u.filter"); assert rk.save(f, path2) == 0
    out.update(kx1u_md5=np.array(ka.md5_file(path2)), kx1u_rows=rows_of(rk, ref, f, r_codes[:1200], 53, bins))
    rk.free(f)
    # the file constructor reads bins and k from the trailer (:57-66, :219-232)
    g = rk.load(path)
    assert rk.params(g)[:3] == (bins, k, nbits)
    rk.free(g)

    # KX2: 128 bins (two words per block), k = 11 -> 64 MiB; bins 64.. hold the reversed genomes; threshold 57
    k, bins = 11, 128
    nbits = (1 << (2 * k)) * bins + 256
    f = rk.new(bins, k, nbits)
    for b in range(bins):
        rk.add_kmers(f, g_codes[b % 64] if b < 64 else g_codes[b % 64][::-1].copy(), b)
    path = os.path.join(tmp, "kx2.filter"); assert rk.save(f, path) == 0
    out.update(kx2_k=k, kx2_bins=bins, kx2_nbits=np.uint64(nbits), kx2_md5=np.array(ka.md5_file(path)),
               kx2_rows=rows_of(rk, ref, f, r_codes[:2000], 57, bins))
    rk.free(f)
    np.savez_compressed(os.path.join(HERE, "kdx_vectors.npz"), **out)
    print({k_: (v.shape if hasattr(v, "shape") and v.shape else v) for k_, v in out.items()})


if __name__ == "__main__":
    main()
