"""From-disk measurement of the filtering stage: a FASTQ file of --reads synthetic
150 bp reads and the cfg2 filter (1024 bins, 16 GiB, k=18, h=3) as a .filter file, both on a file system of this box, through
tools/bin/dyibf_filter_reads (file -> reader thread -> text tickets -> per-bin lists), with the tool's own timers — the same
three lines the reference mapper prints (src/d_mapper.h:947-958: "Filter loading time", "Reads loading time", "Reads filtering
time").  Prints one JSON object.

    python bench_support/from_disk.py [--reads 10000000] [--rb 100000,1000000] [--dir /dev/shm]
"""
import argparse
import json
import os
import re
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def write_fastq(path, codes):
    """[n, L] Dna5 codes -> four-line FASTQ with fixed-width ids (vectorised: one 2-D byte array)."""
    n, L = codes.shape
    rec = np.empty((n, 1 + 9 + 1 + L + 3 + L + 1), dtype=np.uint8)
    rec[:, 0] = ord("@")
    digits = np.arange(n, dtype=np.int64)
    for d in range(9):
        rec[:, 9 - d] = ord("0") + (digits // 10 ** d) % 10
    rec[:, 10] = ord("\n")
    rec[:, 11:11 + L] = np.frombuffer(b"ACGTN", dtype=np.uint8)[codes]
    rec[:, 11 + L:14 + L] = np.frombuffer(b"\n+\n", dtype=np.uint8)
    rec[:, 14 + L:14 + 2 * L] = ord("I")
    rec[:, 14 + 2 * L] = ord("\n")
    with open(path, "wb") as fh:
        fh.write(rec.tobytes())
    return rec.size


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reads", type=int, default=10_000_000)
    ap.add_argument("--rb", default="100000,1000000")
    ap.add_argument("--dir", default="/dev/shm")
    ap.add_argument("--filter-gib", type=int, default=16)
    ap.add_argument("--genome-len", type=int, default=4_000_000)
    ap.add_argument("--inflight", type=int, default=3)
    a = ap.parse_args()
    import dream_yara_b200 as dy
    from bench_support import synth
    exe = os.path.join(ROOT, "tools", "bin", "dyibf_filter_reads")
    bins, k, h, RL = 1024, 18, 3, 150
    nbits = a.filter_gib * (1 << 33) + 256
    gen = synth.genomes(bins, a.genome_len)
    f = dy.SeqAnBloomFilter(bins, h, k, nbits)
    f.addSequences((gen.reshape(-1), np.arange(bins + 1, dtype=np.uint64) * a.genome_len), np.arange(bins, dtype=np.uint32))
    fpath = os.path.join(a.dir, "dyibf_from_disk_%d.filter" % os.getpid())
    rpath = os.path.join(a.dir, "dyibf_from_disk_%d.fq" % os.getpid())
    out = {"reads": a.reads, "read_len": RL, "where": a.dir, "filter_gib": a.filter_gib, "runs": []}
    try:
        assert f.save(fpath)
        f.close()
        codes = synth.reads(gen, 2, 0, a.reads, RL)
        out["fastq_bytes"] = write_fastq(rpath, codes)
        del codes, gen
        for rb in [int(x) for x in a.rb.split(",")]:
            t0 = time.perf_counter()
            r = subprocess.run([exe, "-fi", fpath, "-e", "3", "-rb", str(rb), "-v", "--inflight", str(a.inflight), rpath],
                               capture_output=True, text=True)
            wall = time.perf_counter() - t0
            if r.returncode != 0:
                out["runs"].append({"rb": rb, "error": r.stderr[-400:]})
                continue
            t = {m.group(1): float(m.group(2)) for m in re.finditer(r"(Filter loading|Reads loading|Reads filtering) time:\s+([0-9.e+-]+) sec", r.stderr)}
            avg = re.search(r"Avg reads per bin:\s+([0-9.e+-]+)", r.stderr)
            out["runs"].append({"rb": rb, "inflight": a.inflight, "wall_s": wall, "filter_loading_s": t.get("Filter loading"),
                                "reads_loading_s_reader_thread": t.get("Reads loading"), "reads_filtering_s": t.get("Reads filtering"),
                                "reads_per_s_filtering_stage": a.reads / t["Reads filtering"], "reads_per_s_whole_program": a.reads / wall,
                                "text_gbs_filtering_stage": out["fastq_bytes"] / t["Reads filtering"] / 1e9,
                                "avg_reads_per_bin": float(avg.group(1)) if avg else None})
    finally:
        for p in (fpath, rpath):
            try:
                os.unlink(p)
            except OSError:
                pass
    out["note"] = ("tools/bin/dyibf_filter_reads -fi F -e 3 -rb RB -v reads.fq: the FASTQ text is read by a reader thread, cut into batches "
                   "of RB records, uploaded and tokenised on the device per batch (dyibf_classify_text_begin), filtered, and the per-bin "
                   "lists come back; wall_s includes process start, CUDA context creation and the filter load")
    print(json.dumps(out))


if __name__ == "__main__":
    main()
