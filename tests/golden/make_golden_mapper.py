#!/usr/bin/env python
"""Golden vectors for the drop-in proof of the whole programs (run in the build container only).

Runs the UNMODIFIED reference programs on the reference's own test data and records what they print:

  * tests/cpp/_bin/ref_dream_yara_mapper (the reference mapper as it is, CPU filter; tests/cpp/Makefile `ref_mapper`) on
    tests/gold/64-viral-binned-indices + tests/input/64-viral-reads.fa with the KA1 filter, for the three invocations of
    the reference's test suite (tests/run_tests.py:129-150: `-e 3 --threads 1`, `... -sm record -s 10`, `... -sm tag -s 10`)
    plus `-vv`: the per-bin list sizes it prints (`bin <b>\t<n>`), "Avg reads per bin", and a sha256 of the SAM records
    (header lines starting with @PG carry the command line and are left out);
  * the reference's own filter update semantics (oracle/_ref: clearBins + addFastaFile + save) for the update round trip
    the patched dream_yara_update_filter is checked against: md5 of the updated file.

The GPU test (tests/test_cpp_reference_mains.py) runs the same command lines with the reference's translation units
compiled against include/ of this repository and -ldyibf, and compares.

    make -C tests/cpp ref_mapper data && python tests/golden/make_golden_mapper.py
"""
import hashlib
import json
import os
import re
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
TESTS = os.path.dirname(HERE)
sys.path.insert(0, TESTS)
from oracle_py import Ref  # noqa: E402

REFROOT = "/root/reference"
BIN = os.path.join(TESTS, "cpp", "_bin")
ARGS = {"t1": ["-e", "3", "--threads", "1"],
        "rec.t1": ["-e", "3", "--threads", "1", "-sm", "record", "-s", "10"],
        "tag.t1": ["-e", "3", "--threads", "1", "-sm", "tag", "-s", "10"]}


def split_bins(genomes_fa, out_dir):
    """tests/manage_bins.py `split`: 101-line slices of the multi-FASTA -> 0.fa .. 63.fa"""
    os.makedirs(out_dir, exist_ok=True)
    lines = open(genomes_fa).read().split("\n")
    for b in range(64):
        with open(os.path.join(out_dir, "%d.fa" % b), "w") as fh:
            fh.write("\n".join(lines[b * 101:(b + 1) * 101]))


def sam_digest(path):
    h = hashlib.sha256(); n = 0
    for line in open(path, "rb"):
        if line.startswith(b"@PG"):
            continue
        h.update(line); n += 1
    return h.hexdigest(), n


def md5_file(path):
    h = hashlib.md5()
    with open(path, "rb") as fh:
        for blk in iter(lambda: fh.read(1 << 24), b""):
            h.update(blk)
    return h.hexdigest()


def main():
    out = {}
    ref = Ref()
    with tempfile.TemporaryDirectory() as tmp:
        bins_dir = os.path.join(tmp, "bins")
        split_bins(os.path.join(REFROOT, "tests/input/64-viral-genomes.fa"), bins_dir)
        # dream_yara_build_filter -b 64 -k 19 -nh 2 -bs 1 (KA1), by the reference class itself
        f = ref.new(64, 2, 19, (1 << 33) + 256)
        for b in range(64):
            assert ref.add_fasta(f, os.path.join(bins_dir, "%d.fa" % b), b) == 0
        filt = os.path.join(tmp, "ka1.filter")
        assert ref.save(f, filt) == 0
        out["ka1_md5"] = md5_file(filt)
        assert out["ka1_md5"] == open(os.path.join(REFROOT, "tests/gold/64-viral-binned-genomes.filter.md5")).read().split()[0]
        for tag, args in ARGS.items():
            sam = os.path.join(tmp, tag + ".sam")
            cmd = [os.path.join(BIN, "ref_dream_yara_mapper"), os.path.join(REFROOT, "tests/gold/64-viral-binned-indices/"),
                   os.path.join(REFROOT, "tests/input/64-viral-reads.fa"), "-fi", filt, "-o", sam, "-vv"] + args
            r = subprocess.run(cmd, capture_output=True, text=True)
            assert r.returncode == 0, r.stderr[-2000:]
            counts = [int(m.group(2)) for m in re.finditer(r"^bin (\d+)\t(\d+)$", r.stderr, re.M)]
            avg = re.search(r"Avg reads per bin:\s+([0-9.]+)", r.stderr)
            dig, n = sam_digest(sam)
            out[tag] = {"args": args, "bin_counts": counts, "avg_reads_per_bin": avg.group(1) if avg else None,
                        "sam_sha256": dig, "sam_lines": n}
            print(tag, len(counts), sum(counts), avg.group(1) if avg else None, dig[:16], n, file=sys.stderr)
        # update round trip (dream_yara_update_filter semantics, src/d_update_filter.cpp:188-252): bins 3 and 40 are cleared
        # and re-filled from other genomes' files (5.fa and 41.fa under the names 3.fa and 40.fa)
        ref.clear_bins(f, [3, 40], 2)
        assert ref.add_fasta(f, os.path.join(bins_dir, "5.fa"), 3) == 0
        assert ref.add_fasta(f, os.path.join(bins_dir, "41.fa"), 40) == 0
        upd = os.path.join(tmp, "ka1_updated.filter")
        assert ref.save(f, upd) == 0
        out["update"] = {"bins": [3, 40], "sources": [5, 41], "md5": md5_file(upd)}
        ref.free(f)
    with open(os.path.join(HERE, "mapper_vectors.json"), "w") as fh:
        json.dump(out, fh, indent=1)
    print("wrote mapper_vectors.json", file=sys.stderr)


if __name__ == "__main__":
    main()
