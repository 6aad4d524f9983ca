#!/usr/bin/env python
"""Generate tests/golden/*.npz from the UNMODIFIED reference (run in the build container only).

Needs /root/reference (fixtures tests/input/*.fa and tests/gold/*.md5) and the reference compiled by
`make -C oracle ref _ref/libref_mapper.so` (oracle/ref_shim.cpp, oracle/ref_mapper_shim.cpp).  Every output
value below comes from the reference's own code: SeqAnBloomFilter (build, save, whichBins), SeqAn's Shape hash,
reverseComplement and Iupac->Dna5 conversion, DisOptions::getThreshold and clasifyLoadedReads.

    python tests/golden/make_golden.py
"""
import ctypes as C
import hashlib
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle_py import REF_MAPPER_SO, Ref  # noqa: E402

REFROOT = "/root/reference"
GIB_BITS = 1 << 33


def md5_file(path):
    h = hashlib.md5()
    with open(path, "rb") as fh:
        for blk in iter(lambda: fh.read(1 << 24), b""):
            h.update(blk)
    return h.hexdigest()


def read_fasta(path):
    recs = []
    for chunk in open(path, "rb").read().split(b">")[1:]:
        nl = chunk.find(b"\n")
        recs.append((chunk[:nl].decode(), chunk[nl + 1:].replace(b"\n", b"")))
    return recs


def pack_records(seqs):
    offs = np.zeros(len(seqs) + 1, dtype=np.uint64)
    offs[1:] = np.cumsum([len(s) for s in seqs], dtype=np.uint64)
    return np.frombuffer(b"".join(seqs), dtype=np.uint8).copy(), offs


def ka4_bins(genomes):
    """130 bins, multi-record files, leading-N contigs. Returns list of record lists."""
    out = []
    for b in range(130):
        seq = genomes[b % 64]
        if 64 <= b < 128:
            seq = seq[::-1]
        if b == 128:
            seq = b"N" + seq[1:4000]
        if b == 129:
            seq = b"NN" + (b"N" + seq[1:4000])[2:3000]
        if b % 10 == 3:
            out.append([seq[:3000], seq[3000:], b"ACGTACGT"])
        else:
            out.append([seq])
    return out


def build_with_ref(ref, bins_records, h, k, tmpdir, tag, nbits=GIB_BITS + 256):
    """dream_yara_build_filter semantics: addFastaFile per bin file (reference reader + conversion), save."""
    f = ref.new(len(bins_records), h, k, nbits)
    for b, recs in enumerate(bins_records):
        path = os.path.join(tmpdir, "%s_%d.fa" % (tag, b))
        with open(path, "wb") as fh:
            for i, r in enumerate(recs):
                fh.write(b">%d_%d\n" % (b, i))
                for j in range(0, len(r), 80):
                    fh.write(r[j:j + 80] + b"\n")
        assert ref.add_fasta(f, path, b) == 0
        os.unlink(path)
    out = os.path.join(tmpdir, tag + ".filter")
    assert ref.save(f, out) == 0
    words = ref.words(f)
    info = dict(md5=md5_file(out), size=os.path.getsize(out), popcount=int(np.unpackbits(words.view(np.uint8)).sum()))
    os.unlink(out)
    return f, info


def query_reads(ref, f, reads_codes, thr, bins):
    """fwd OR rc per read with the reference's whichBins and reverseComplement -> rows [n x W] u64."""
    W = (bins + 63) // 64
    rows = np.zeros((len(reads_codes), W), dtype=np.uint64)
    for i, c in enumerate(reads_codes):
        sel = np.zeros(bins, dtype=np.uint8)
        ref.which_bins(f, c, thr, bins, sel)
        ref.which_bins(f, ref.revcomp(c), thr, bins, sel)
        bits = np.zeros(W * 64, dtype=np.uint8); bits[:bins] = sel
        rows[i] = np.packbits(bits, bitorder="little").view(np.uint64)
    return rows


def counts_by_sweep(ref, f, codes, bins, k):
    """counts[b] = max{t : b in whichBins(read, t)} — recovers the private `counts` exactly."""
    possible = len(codes) - k + 1
    counts = np.zeros(bins, dtype=np.uint16)
    for t in range(1, possible + 1):
        sel = ref.which_bins(f, codes, t, bins)
        counts[sel != 0] = t
        if not sel.any():
            break
    return counts


def main():
    assert Ref.available(), "run `make -C oracle ref` first"
    ref = Ref()
    tmpdir = tempfile.mkdtemp(prefix="golden_")
    genomes = [s for _, s in read_fasta(os.path.join(REFROOT, "tests/input/64-viral-genomes.fa"))]
    reads = [s for _, s in read_fasta(os.path.join(REFROOT, "tests/input/64-viral-reads.fa"))]
    assert len(genomes) == 64 and len(reads) == 6400
    gold_md5 = open(os.path.join(REFROOT, "tests/gold/64-viral-binned-genomes.filter.md5")).read().split()[0]
    g_cat, g_off = pack_records(genomes)
    r_cat, r_off = pack_records(reads)
    reads_codes = [ref.ascii_to_dna5(r) for r in reads]
    out = dict(genomes_ascii=g_cat, genomes_offsets=g_off, reads_ascii=r_cat, reads_offsets=r_off,
               reference_gold_md5=np.array(gold_md5))

    # KA1: the reference's own test configuration (tests/run_tests.py:116-122): -b 64 -k 19 -nh 2 -bs 1
    f, info = build_with_ref(ref, [[g] for g in genomes], 2, 19, tmpdir, "ka1")
    assert info["md5"] == gold_md5, (info, gold_md5)
    out.update(ka1_md5=np.array(info["md5"]), ka1_size=info["size"], ka1_popcount=info["popcount"],
               ka1_rows=query_reads(ref, f, reads_codes, 25, 64))
    # per-bin counts of a few reads (fwd strand), recovered by threshold sweep
    idx = np.array([0, 1, 2, 99, 100, 3199, 4600, 6399])
    out.update(ka1_count_reads=idx, ka1_counts=np.stack([counts_by_sweep(ref, f, reads_codes[i], 64, 19) for i in idx]))
    ka1_file = os.path.join(tmpdir, "ka1_keep.filter"); ref.save(f, ka1_file)
    ref.free(f)

    # KA2: BASELINE config 1 (k=18, h=3)
    f, info = build_with_ref(ref, [[g] for g in genomes], 3, 18, tmpdir, "ka2")
    out.update(ka2_md5=np.array(info["md5"]), ka2_popcount=info["popcount"], ka2_rows=query_reads(ref, f, reads_codes, 29, 64))
    out.update(ka2_counts=np.stack([counts_by_sweep(ref, f, reads_codes[i], 64, 18) for i in idx]))
    ref.free(f)

    # KA3: SeqAn hashes incl. the leading-N term
    s = b"NACGTNNACGTACGTTGCAGGCTAACGTAGCTAGCTAGGATCGATCGNACGT"
    out.update(ka3_text=np.frombuffer(s, dtype=np.uint8).copy(), ka3_hashes_k18=ref.kmer_hashes(ref.ascii_to_dna5(s), 18),
               ka3_hashes_k32=ref.kmer_hashes(ref.ascii_to_dna5(s), 32), ka3_hashes_k14=ref.kmer_hashes(ref.ascii_to_dna5(s), 14))

    # KA4: 130 bins (W=3, P not a power of two), multi-record bins, leading-N contigs
    f, info = build_with_ref(ref, ka4_bins(genomes), 3, 18, tmpdir, "ka4")
    out.update(ka4_md5=np.array(info["md5"]), ka4_popcount=info["popcount"], ka4_rows=query_reads(ref, f, reads_codes, 29, 130))
    ref.free(f)

    # KA5 + thresholds: the reference's own clasifyLoadedReads / getThreshold (oracle/ref_mapper_shim.cpp)
    if os.path.exists(REF_MAPPER_SO):
        m = C.CDLL(REF_MAPPER_SO)
        m.refm_threshold.restype = C.c_uint16
        m.refm_threshold.argtypes = [C.c_int, C.c_uint32, C.c_uint16]
        es, ks, Ls = np.arange(0, 11), np.arange(14, 33), np.array([36, 50, 75, 100, 101, 125, 150, 151, 200, 250, 300, 1000])
        thr = np.zeros((es.size, ks.size, Ls.size), dtype=np.uint16)
        for a, e in enumerate(es):
            for b, k in enumerate(ks):
                for c, L in enumerate(Ls):
                    thr[a, b, c] = m.refm_threshold(int(e), int(k), int(L))
        out.update(thr_e=es, thr_k=ks, thr_L=Ls, thr_table=thr)
        m.refm_classify.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_int, C.c_uint32,
                                    C.POINTER(C.c_uint16), C.c_void_p, C.c_void_p, C.c_uint64]
        codes = np.concatenate(reads_codes); offs = r_off.copy()
        for tag, paired in (("se", 0), ("pe", 1)):          # PE: mates = first/second half of the read file (KA5)
            bo = np.zeros(65, dtype=np.uint32); ids = np.zeros(1 << 20, dtype=np.uint32); t = C.c_uint16(0)
            m.refm_classify(ka1_file.encode(), codes.ctypes.data, offs.ctypes.data, 6400, paired, 3, 1, C.byref(t),
                            bo.ctypes.data, ids.ctypes.data, ids.size)
            out["ka5_%s_threshold" % tag] = int(t.value)
            out["ka5_%s_bin_offsets" % tag] = bo.astype(np.uint64)
            out["ka5_%s_read_ids" % tag] = ids[: bo[64]].copy()
    os.unlink(ka1_file)

    # small random cases straight through the reference class (odd bin counts, all h, k up to 32, tiny vectors)
    rng = np.random.default_rng(20261104)
    cases = []
    for ci, (bins, h, k, nbits) in enumerate([(1, 2, 14, 4096 + 256), (63, 3, 20, 100000), (64, 4, 32, (1 << 20) + 256),
                                               (65, 5, 31, 999999), (200, 3, 18, (1 << 22) + 256), (1024, 3, 18, (1 << 24) + 256),
                                               (1000, 2, 25, 5000000), (8192, 3, 18, (1 << 25) + 256), (130, 2, 16, 70000)]):
        f = ref.new(bins, h, k, nbits)
        seqs, binids = [], []
        for _ in range(40):
            L = int(rng.integers(k - 2, 400))
            c = rng.integers(0, 4, size=L).astype(np.uint8)
            c[rng.random(L) < 0.02] = 4
            if rng.random() < 0.3 and L:
                c[0] = 4
            if rng.random() < 0.2 and L:
                c[-1] = 4
            seqs.append(c); binids.append(int(rng.integers(0, bins)))
        for c, b in zip(seqs, binids):
            if len(c) >= k:
                ref.add_kmers(f, c, b)
        W_ = (bins + 63) // 64
        words = ref.words(f)[: ((nbits - 256) // (64 * W_)) * W_].copy()      # the P*W data words
        qs = []
        for j in range(30):
            src = seqs[int(rng.integers(0, len(seqs)))]
            if len(src) > 60 and rng.random() < 0.8:
                st = int(rng.integers(0, len(src) - 50)); q = src[st:st + int(rng.integers(k, 120))].copy()
                m_ = rng.random(len(q)) < 0.02
                q[m_] = rng.integers(0, 4, size=int(m_.sum())).astype(np.uint8)
            else:
                q = rng.integers(0, 5, size=int(rng.integers(k, 150))).astype(np.uint8)
            qs.append(q)
        thr = int(rng.integers(1, 6))
        rows = query_reads(ref, f, qs, thr, bins)
        cnts = np.stack([counts_by_sweep(ref, f, q, bins, k) for q in qs[:6]])
        ref.free(f)
        s_cat = np.concatenate(seqs); s_off = np.zeros(len(seqs) + 1, dtype=np.uint64); s_off[1:] = np.cumsum([len(x) for x in seqs])
        q_cat = np.concatenate(qs); q_off = np.zeros(len(qs) + 1, dtype=np.uint64); q_off[1:] = np.cumsum([len(x) for x in qs])
        p = "rc%d_" % ci
        out.update({p + "params": np.array([bins, h, k, nbits], dtype=np.uint64), p + "seqs": s_cat, p + "seq_offsets": s_off,
                    p + "bin_ids": np.array(binids, dtype=np.uint32), p + "words_sha256": np.array(hashlib.sha256(words.tobytes()).hexdigest()),
                    p + "words_popcount": int(np.unpackbits(words.view(np.uint8)).sum()),
                    p + "queries": q_cat, p + "query_offsets": q_off, p + "threshold": thr, p + "rows": rows, p + "counts6": cnts})
        cases.append(ci)
    out["random_cases"] = np.array(cases)
    np.savez_compressed(os.path.join(HERE, "reference_vectors.npz"), **out)
    print("wrote", os.path.join(HERE, "reference_vectors.npz"), os.path.getsize(os.path.join(HERE, "reference_vectors.npz")), "bytes")
    for kname in ("ka1_md5", "ka2_md5", "ka4_md5", "ka1_popcount", "ka2_popcount", "ka4_popcount"):
        print(kname, out[kname])


if __name__ == "__main__":
    main()
