#!/usr/bin/env python
"""Generate tests/golden/gather_vectors.npz from the UNMODIFIED reference (run in the build container only).

Pins SURVEY §8(f) row 2 — loadFilteredReads (src/d_mapper.h:393-443) and sortedBins (:843-854) — with the reference's
own code (oracle/ref_mapper_shim.cpp -> oracle/_ref/libref_mapper.so): the KA1 filter (k=19, h=2, 64 bins), the 6400
test reads with synthetic qualities, single-end and paired-end (mates = first/second half of the file, as KA5), -e 3.
For a few bins it stores the per-bin read set the child mapper receives (Dna5Q bytes incl. reverse complements, limits)
and origReadIdMap[bin] as loadFilteredReads leaves it; plus the bin order sortedBins returns.

    make -C oracle ref _ref/libref_mapper.so && python tests/golden/make_golden_gather.py
"""
import ctypes as C
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
from oracle_py import REF_MAPPER_SO, Ref  # noqa: E402
import make_golden as mg  # noqa: E402

BINS = np.array([0, 5, 38, 46, 63], dtype=np.uint32)


def main():
    assert Ref.available() and os.path.exists(REF_MAPPER_SO), "run `make -C oracle ref _ref/libref_mapper.so` first"
    ref = Ref()
    tmpdir = tempfile.mkdtemp(prefix="golden_gather_")
    genomes = [s for _, s in mg.read_fasta(os.path.join(mg.REFROOT, "tests/input/64-viral-genomes.fa"))]
    reads = [s for _, s in mg.read_fasta(os.path.join(mg.REFROOT, "tests/input/64-viral-reads.fa"))]
    f, info = mg.build_with_ref(ref, [[g] for g in genomes], 2, 19, tmpdir, "ka1")
    ka1_file = os.path.join(tmpdir, "ka1.filter"); ref.save(f, ka1_file); ref.free(f)
    codes = np.concatenate([ref.ascii_to_dna5(r) for r in reads])
    _, offs = mg.pack_records(reads)
    quals = ((np.arange(codes.size, dtype=np.uint64) * 7 + 11) % 63).astype(np.uint8)

    m = C.CDLL(REF_MAPPER_SO)
    V = C.c_void_p
    m.refm_gather.argtypes = [C.c_char_p, V, V, V, C.c_uint32, C.c_int, C.c_int, V, C.c_uint32, V, V, C.c_uint64, V, V,
                              C.c_uint64, V, V]
    out = dict(quals=quals, bins=BINS, ka1_md5=np.array(info["md5"]))
    for tag, paired in (("se", 0), ("pe", 1)):
        fwd = np.zeros(codes.size, dtype=np.uint8)
        cap_b, cap_s = 8 << 20, 1 << 16
        ob = np.zeros(cap_b, dtype=np.uint8); ol = np.zeros(cap_s + 1, dtype=np.uint64); oi = np.zeros(cap_s, dtype=np.uint32)
        ss = np.zeros(BINS.size + 1, dtype=np.uint64); sb = np.zeros(64, dtype=np.uint32)
        m.refm_gather(ka1_file.encode(), codes.ctypes.data, quals.ctypes.data, offs.ctypes.data, 6400, paired, 3,
                      BINS.ctypes.data, BINS.size, fwd.ctypes.data, ob.ctypes.data, cap_b, ol.ctypes.data, oi.ctypes.data,
                      cap_s, ss.ctypes.data, sb.ctypes.data)
        n = int(ss[-1]); nb = int(ol[n])
        assert n < cap_s and nb < cap_b
        out.update({tag + "_fwd_bytes": fwd, tag + "_bytes": ob[:nb].copy(), tag + "_limits": ol[:n + 1].copy(),
                    tag + "_ids": oi[:n].copy(), tag + "_seq_start": ss, tag + "_sorted_bins": sb})
        print(tag, "sequences", n, "bytes", nb, "sorted bins head", sb[:8])
    os.unlink(ka1_file)
    path = os.path.join(HERE, "gather_vectors.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
