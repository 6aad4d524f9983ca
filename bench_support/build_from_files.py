"""From-disk measurement of the builder (BASELINE.json configs[4] as a user runs it): --bins FASTA files of --genome-len bases
each (80 columns per line) in a directory of this box, through tools/bin/dyibf_build_filter (the reference's
dream_yara_build_filter command line): files -> reader threads -> text batches -> device tokeniser (dyibf_add_fasta_texts) ->
sliced build -> .filter file.  The file written must be byte-identical to the one the library writes for the same contigs
added as sequences.  Prints one JSON object.

    python bench_support/build_from_files.py [--bins 1024] [--genome-len 4000000] [--threads 16] [--dir /dev/shm]
"""
import argparse
import json
import os
import re
import shutil
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def write_fasta(path, codes, width=80):
    """one record, `width` characters per line (vectorised)"""
    L = int(codes.size)
    rows = (L + width - 1) // width
    pad = rows * width - L
    padded = np.full(rows * width, ord("\n"), dtype=np.uint8)
    padded[:L] = np.frombuffer(b"ACGTN", dtype=np.uint8)[codes]
    body = np.full((rows, width + 1), ord("\n"), dtype=np.uint8)
    body[:, :width] = padded.reshape(rows, width)
    flat = body.reshape(-1)
    if pad:                                                            # the last line is shorter: drop its padding
        flat = np.concatenate([flat[: (rows - 1) * (width + 1) + (width - pad)], np.array([ord("\n")], dtype=np.uint8)])
    with open(path, "wb") as fh:
        fh.write(b">contig of " + os.path.basename(path).encode() + b"\n")
        fh.write(flat.tobytes())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bins", type=int, default=1024)
    ap.add_argument("--genome-len", type=int, default=4_000_000)
    ap.add_argument("--filter-gib", type=int, default=16)
    ap.add_argument("--threads", type=int, default=min(16, os.cpu_count() or 8))
    ap.add_argument("--dir", default="/dev/shm")
    a = ap.parse_args()
    import dream_yara_b200 as dy
    import ka
    from bench_support import synth
    exe = os.path.join(ROOT, "tools", "bin", "dyibf_build_filter")
    k, h = 18, 3
    d = os.path.join(a.dir, "dyibf_build_files_%d" % os.getpid())
    os.makedirs(d)
    out = {"bins": a.bins, "genome_len": a.genome_len, "filter_gib": a.filter_gib, "threads": a.threads, "where": a.dir}
    try:
        gen = synth.genomes(a.bins, a.genome_len)
        t0 = time.perf_counter()
        for b in range(a.bins):
            write_fasta(os.path.join(d, "%d.fa" % b), gen[b])
        out["fasta_bytes"] = sum(os.path.getsize(os.path.join(d, f)) for f in os.listdir(d))
        out["write_files_s"] = time.perf_counter() - t0
        filt = os.path.join(a.dir, "dyibf_build_files_%d.filter" % os.getpid())
        want = os.path.join(a.dir, "dyibf_build_files_%d.want.filter" % os.getpid())
        t0 = time.perf_counter()
        r = subprocess.run([exe, "-b", str(a.bins), "-t", str(a.threads), "-k", str(k), "-nh", str(h), "-bs", str(a.filter_gib), "-v",
                            d + "/", "-o", filt], capture_output=True, text=True)
        wall = time.perf_counter() - t0
        if r.returncode != 0:
            out["error"] = r.stderr[-600:]
            print(json.dumps(out)); return
        m_add = re.search(r"All bins are done adding kmers!\s+([0-9.e+-]+) sec", r.stderr)
        m_fin = re.search(r"Finshed in\s+([0-9.e+-]+)", r.stderr)
        kmers = a.bins * (a.genome_len - k + 1)
        out.update({"wall_s": wall, "adding_kmers_s": float(m_add.group(1)) if m_add else None, "tool_total_s": float(m_fin.group(1)) if m_fin else None,
                    "kmers": kmers})
        if m_add:
            out["kmers_per_s_adding_stage"] = kmers / float(m_add.group(1))
            out["text_gbs_adding_stage"] = out["fasta_bytes"] / float(m_add.group(1)) / 1e9
        # the same contigs as sequences through the library -> must give the same file
        f = dy.SeqAnBloomFilter(a.bins, h, k, a.filter_gib * (1 << 33) + 256)
        f.addSequences((gen.reshape(-1), np.arange(a.bins + 1, dtype=np.uint64) * a.genome_len), np.arange(a.bins, dtype=np.uint32))
        assert f.save(want)
        f.close()
        from bench_support.cfg5_fullsize import files_equal
        same, detail = files_equal(filt, want)
        out["file_identical_to_library_build"] = same
        out["detail"] = detail
        out["tail"] = [l for l in r.stderr.strip().splitlines() if not l.startswith("[bin ")][-6:]
        for p in (filt, want):
            os.unlink(p)
    finally:
        shutil.rmtree(d, ignore_errors=True)
    out["note"] = ("tools/bin/dyibf_build_filter -b B -t T -k 18 -nh 3 -bs G -v DIR -o F: reader threads read whole FASTA files, the text goes "
                   "to the device in batches (dyibf_add_fasta_texts: tokeniser + sliced build), then dyibf_save_file; wall_s includes process "
                   "start, CUDA context creation and the 16 GiB file write")
    print(json.dumps(out))


if __name__ == "__main__":
    main()
