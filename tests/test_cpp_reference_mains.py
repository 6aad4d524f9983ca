"""GPU (B200): the reference's OWN programs, compiled from its UNMODIFIED translation units against include/ of this
repository and linked with -ldyibf (tests/cpp/Makefile, tests/cpp/stage_reference.py), run on the reference's own test data
with the command lines of its own test suite (tests/run_tests.py:116-150).

  dream_yara_build_filter   src/d_build_filter.cpp:217-272, unmodified   -> the gold md5 of tests/gold/*.filter.md5
  dream_yara_update_filter  src/d_update_filter.cpp:188-252, unmodified  -> md5 of the file the reference class produces
  dream_yara_mapper         src/d_mapper.cpp unmodified + src/d_mapper.h with the one-call patch of INTEGRATION.md §3.1
                            -> `-vv` per-bin list sizes, "Avg reads per bin" and the SAM records of the reference mapper
                               (tests/golden/mapper_vectors.json, made by the unmodified reference mapper in the build container)

The binaries and the staged test data are built in the container that has /root/reference and travel to the GPU box as
built files; the tests skip where they are absent."""
import hashlib
import json
import os
import re
import subprocess

import pytest

import ka

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
BIN = os.path.join(HERE, "cpp", "_bin")
DATA = os.path.join(HERE, "cpp", "_data")


def _need(*names):
    for n in names:
        if not os.path.exists(os.path.join(BIN, n)):
            pytest.skip("%s not built (make -C tests/cpp %s needs the reference tree)" % (n, n))
    if not os.path.exists(os.path.join(DATA, "64-viral-genomes.fa")):
        pytest.skip("reference test data not staged (make -C tests/cpp data)")


@pytest.fixture(scope="module")
def vectors():
    with open(os.path.join(HERE, "golden", "mapper_vectors.json")) as fh:
        return json.load(fh)


@pytest.fixture(scope="module")
def bins_dir(tmp_path_factory):
    """tests/manage_bins.py `split`: 101-line slices of the multi-FASTA -> 0.fa .. 63.fa"""
    d = tmp_path_factory.mktemp("bins")
    lines = open(os.path.join(DATA, "64-viral-genomes.fa")).read().split("\n") if os.path.exists(DATA) else []
    for b in range(64):
        with open(d / ("%d.fa" % b), "w") as fh:
            fh.write("\n".join(lines[b * 101:(b + 1) * 101]))
    return d


@pytest.fixture(scope="module")
def ka1_filter_file(bins_dir, tmp_path_factory, vectors):
    _need("dream_yara_build_filter")
    out = tmp_path_factory.mktemp("filter") / "64-viral-binned-genomes.filter"
    # the reference test suite's own command line (tests/run_tests.py:116-122)
    r = subprocess.run([os.path.join(BIN, "dream_yara_build_filter"), str(bins_dir) + "/", "-o", str(out),
                        "-b", "64", "-t", "4", "-k", "19", "-nh", "2", "-bs", "1"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    return out


def test_reference_build_filter_main_gold_md5(ka1_filter_file, golden, vectors):
    assert ka1_filter_file.stat().st_size == 1073741864
    md5 = ka.md5_file(ka1_filter_file)
    assert md5 == str(golden["reference_gold_md5"]) == vectors["ka1_md5"]     # tests/gold/64-viral-binned-genomes.filter.md5


def test_reference_update_filter_main(ka1_filter_file, bins_dir, tmp_path, vectors):
    _need("dream_yara_update_filter")
    import shutil
    work = tmp_path / "upd.filter"
    shutil.copyfile(ka1_filter_file, work)
    newdir = tmp_path / "new"; newdir.mkdir()
    files = []
    for b, src in zip(vectors["update"]["bins"], vectors["update"]["sources"]):   # the file NAME is the bin number
        shutil.copyfile(bins_dir / ("%d.fa" % src), newdir / ("%d.fa" % b))
        files.append(str(newdir / ("%d.fa" % b)))
    r = subprocess.run([os.path.join(BIN, "dream_yara_update_filter"), str(work)] + files + ["-t", "2"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    assert ka.md5_file(work) == vectors["update"]["md5"]
    # a bin beyond the filter's: the reference's message and exit code (src/d_update_filter.cpp:196-200)
    shutil.copyfile(bins_dir / "1.fa", newdir / "64.fa")
    r = subprocess.run([os.path.join(BIN, "dream_yara_update_filter"), str(work), str(newdir / "64.fa")], capture_output=True, text=True)
    assert r.returncode == 1 and "has only 64 Bins" in r.stderr


def _sam_digest(path):
    h = hashlib.sha256(); n = 0
    for line in open(path, "rb"):
        if line.startswith(b"@PG"):
            continue
        h.update(line); n += 1
    return h.hexdigest(), n


@pytest.mark.parametrize("tag", ["t1", "rec.t1", "tag.t1"])
def test_patched_reference_mapper(ka1_filter_file, tmp_path, vectors, tag):
    _need("dream_yara_mapper")
    want = vectors[tag]
    sam = tmp_path / ("64-viral-reads.%s.sam" % tag)
    cmd = [os.path.join(BIN, "dream_yara_mapper"), os.path.join(DATA, "64-viral-binned-indices/"),
           os.path.join(DATA, "64-viral-reads.fa"), "-fi", str(ka1_filter_file), "-o", str(sam), "-vv"] + want["args"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    counts = [int(m.group(2)) for m in re.finditer(r"^bin (\d+)\t(\d+)$", r.stderr, re.M)]
    assert counts == want["bin_counts"]                                   # origReadIdMap sizes, KA1: 6538 pairs
    avg = re.search(r"Avg reads per bin:\s+([0-9.]+)", r.stderr)
    assert avg and avg.group(1) == want["avg_reads_per_bin"]
    assert "Filter loading time" in r.stderr and "Reads filtering time" in r.stderr
    dig, n = _sam_digest(sam)
    assert n == want["sam_lines"] and dig == want["sam_sha256"]          # the mapper downstream sees identical read sets


def test_patched_reference_mapper_two_gpus(ka1_filter_file, tmp_path, vectors, dy):
    """DYIBF_DEVICES=0,1: the unmodified spawnDisMapper constructs ONE filter object that spans both GPUs"""
    import ctypes as C
    _need("dream_yara_mapper")
    n = C.c_int(0); dy.lib().dyibf_device_count(C.byref(n))
    if n.value < 2:
        pytest.skip("needs 2 GPUs")
    want = vectors["t1"]
    for mode in ("read", "bin"):
        if mode == "bin":
            continue                                                      # 64 bins = one word column: not divisible
        sam = tmp_path / ("two.%s.sam" % mode)
        cmd = [os.path.join(BIN, "dream_yara_mapper"), os.path.join(DATA, "64-viral-binned-indices/"),
               os.path.join(DATA, "64-viral-reads.fa"), "-fi", str(ka1_filter_file), "-o", str(sam), "-vv"] + want["args"]
        r = subprocess.run(cmd, capture_output=True, text=True, env=dict(os.environ, DYIBF_DEVICES="0,1", DYIBF_GROUP_MODE=mode))
        assert r.returncode == 0, r.stderr[-3000:]
        counts = [int(m.group(2)) for m in re.finditer(r"^bin (\d+)\t(\d+)$", r.stderr, re.M)]
        assert counts == want["bin_counts"]
        dig, nl = _sam_digest(sam)
        assert nl == want["sam_lines"] and dig == want["sam_sha256"]
