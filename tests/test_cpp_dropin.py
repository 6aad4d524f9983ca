"""The C++ drop-in (include/d_bloom_filter.h, include/d_mapper_filter.h) with real SeqAn types.

CPU: the driver compiles against the SeqAn vendored in the reference tree (only where /root/reference exists).
GPU: the prebuilt driver (tests/cpp/_bin/dropin_test; SeqAn is needed at compile time only) builds the reference's
test filter from FASTA files, queries and classifies the test reads; outputs are compared with the vectors the
unmodified reference produced (tests/golden)."""
import os
import subprocess

import numpy as np
import pytest

import ka

HERE = os.path.dirname(os.path.abspath(__file__))
BIN = os.path.join(HERE, "cpp", "_bin", "dropin_test")
REF = "/root/reference"


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "include", "seqan")), reason="needs the reference's vendored SeqAn")
def test_dropin_headers_compile_against_seqan(dy):
    r = subprocess.run(["make", "-C", os.path.join(HERE, "cpp")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-4000:]
    assert os.path.exists(BIN)


def _write_fasta(path, records):
    with open(path, "wb") as fh:
        for name, seq in records:
            fh.write(b">" + name.encode() + b"\n")
            for j in range(0, len(seq), 80):
                fh.write(seq[j:j + 80] + b"\n")


@pytest.mark.gpu
def test_dropin_build_query_classify(golden, tmp_path):
    if not os.path.exists(BIN):
        pytest.skip("tests/cpp/_bin/dropin_test was not prebuilt (needs /root/reference at build time)")
    genomes, reads = ka.records(golden, "genomes"), ka.records(golden, "reads")
    bins_dir = tmp_path / "bins"; bins_dir.mkdir()
    for b, g in enumerate(genomes):
        _write_fasta(bins_dir / ("%d.fa" % b), [("bin%d" % b, g)])
    _write_fasta(tmp_path / "reads.fa", [("r%d" % i, r) for i, r in enumerate(reads)])
    filt = tmp_path / "f.filter"
    r = subprocess.run([BIN, "build", "64", "2", "19", str(ka.GIB_BITS), str(bins_dir), str(filt)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert ka.md5_file(filt) == str(golden["reference_gold_md5"])
    # per-read whichBins through the class, Dna5Q reads as in the mapper (first 400 reads: one PCIe round trip each)
    _write_fasta(tmp_path / "reads400.fa", [("r%d" % i, r) for i, r in enumerate(reads[:400])])
    r = subprocess.run([BIN, "query", str(filt), str(tmp_path / "reads400.fa"), "25", str(tmp_path / "q.txt")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "devices 1" in r.stderr
    want = ka.bits_from_rows(golden["ka1_rows"][:400], 64)
    for i, line in enumerate(open(tmp_path / "q.txt")):
        assert [int(x) for x in line.split()] == list(np.nonzero(want[i])[0]), i
    # the same through a MOVED multi-GPU object, bin-sharded and read-sharded (the move constructor dropped the device
    # count and the mode, so whichBins asked member 0 only); needs 2 GPUs
    import torch
    if torch.cuda.device_count() >= 2:
        # the gold filter has one word per hash block (read-sharded only); a 128-bin filter of its own for the bin shards
        rng = np.random.default_rng(5)
        d128 = tmp_path / "bins128"; d128.mkdir()
        contigs = [bytes(rng.choice(list(b"ACGT"), 2500).astype(np.uint8)) for _ in range(128)]
        for b, g in enumerate(contigs):
            _write_fasta(d128 / ("%d.fa" % b), [("bin%d" % b, g)])
        f128 = tmp_path / "f128.filter"
        r = subprocess.run([BIN, "build", "128", "2", "19", str((1 << 26) + 256), str(d128), str(f128)], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        some = [contigs[(7 * i) % 128][100 + i: 200 + i] for i in range(60)]
        _write_fasta(tmp_path / "reads60.fa", [("r%d" % i, x) for i, x in enumerate(some)])
        outs = {}
        for tag, env in (("one", {}), ("bin", dict(DYIBF_DEVICES="0,1", DYIBF_GROUP_MODE="bin")), ("read", dict(DYIBF_DEVICES="0,1", DYIBF_GROUP_MODE="read"))):
            r = subprocess.run([BIN, "query", str(f128), str(tmp_path / "reads60.fa"), "60", str(tmp_path / ("q_%s.txt" % tag))],
                               capture_output=True, text=True, env=dict(os.environ, **env))
            assert r.returncode == 0, r.stderr
            assert ("devices 2" if env else "devices 1") in r.stderr
            outs[tag] = open(tmp_path / ("q_%s.txt" % tag)).read()
        lines = outs["one"].split("\n")
        assert all(str((7 * i) % 128) in lines[i].split() for i in range(60))          # every read finds its own bin
        assert any(int(x) >= 64 for ln in lines for x in ln.split())                      # ... incl. bins of the second shard
        assert outs["bin"] == outs["one"] and outs["read"] == outs["one"]
    # the batched clasifyLoadedReads replacement, SE and PE, vs the reference's own clasifyLoadedReads
    for tag, paired in (("se", 0), ("pe", 1)):
        out = tmp_path / ("c_%s.txt" % tag)
        r = subprocess.run([BIN, "classify", str(filt), str(tmp_path / "reads.fa"), "3", str(paired), str(out)], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        bo, ids = golden["ka5_%s_bin_offsets" % tag], golden["ka5_%s_read_ids" % tag]
        lines = open(out).read().split("\n")
        for b in range(64):
            assert [int(x) for x in lines[b].split()] == list(ids[int(bo[b]):int(bo[b + 1])]), (tag, b)
        assert "bin 0\t173" in r.stderr                         # -vv dump, d_mapper.h:380-386
    # sortedBins + loadFilteredReads replacements vs the reference's own output (Dna5Q bytes incl. qualities)
    gg = np.load(os.path.join(HERE, "golden", "gather_vectors.npz"))
    for tag, paired in (("se", 0), ("pe", 1)):
        out = tmp_path / ("g_%s.txt" % tag)
        r = subprocess.run([BIN, "gather", str(filt), str(tmp_path / "reads.fa"), "3", str(paired), str(out)], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        lines = open(out).read().split("\n")
        order, at, per_bin = [], 0, {}
        while at < len(lines) and lines[at].startswith("bin "):
            _, b, n = lines[at].split()
            b, n = int(b), int(n)
            order.append(b)
            per_bin[b] = (lines[at + 1], lines[at + 2: at + 2 + n])
            at += 2 + n
        assert order == list(gg[tag + "_sorted_bins"])
        lim, ss = gg[tag + "_limits"], gg[tag + "_seq_start"]
        for q, b in enumerate(gg["bins"]):
            a, e = int(ss[q]), int(ss[q + 1])
            ids_line, seq_lines = per_bin[int(b)]
            assert [int(x) for x in ids_line.split()] == list(gg[tag + "_ids"][a:e]), (tag, b)
            assert len(seq_lines) == e - a
            for j, ln in enumerate(seq_lines):
                want = gg[tag + "_bytes"][int(lim[a + j]):int(lim[a + j + 1])]
                assert [int(x) for x in ln.split()] == list(want), (tag, b, j)
    # load failure: message + exit(1) as the reference
    r = subprocess.run([BIN, "query", str(tmp_path / "nope.filter"), str(tmp_path / "reads400.fa"), "25", str(tmp_path / "x")], capture_output=True, text=True)
    assert r.returncode == 1 and "could not be read" in r.stderr


@pytest.mark.gpu
def test_dropin_kdx_filter(golden, tmp_path):
    """include/d_kdx_filter.h (SeqAnKDXFilter, -ft kmer_direct) against the vectors of the unmodified reference class"""
    if not os.path.exists(BIN):
        pytest.skip("tests/cpp/_bin/dropin_test was not prebuilt (needs /root/reference at build time)")
    kg = np.load(os.path.join(HERE, "golden", "kdx_vectors.npz"))
    genomes, reads = ka.records(golden, "genomes"), ka.records(golden, "reads")
    reads = [reads[int(i)] for i in kg["read_idx"]][:300]
    bins_dir = tmp_path / "bins"; bins_dir.mkdir()
    for b, g in enumerate(genomes):
        _write_fasta(bins_dir / ("%d.fa" % b), [("bin%d" % b, g)])
    _write_fasta(tmp_path / "reads.fa", [("r%d" % i, r) for i, r in enumerate(reads)])
    filt, out = tmp_path / "kx.filter", tmp_path / "kx.txt"
    r = subprocess.run([BIN, "kdx", "64", str(int(kg["kx1_k"])), str(bins_dir), str(filt), str(tmp_path / "reads.fa"), "53", str(out)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert ka.md5_file(filt) == str(kg["kx1_md5"])
    want = ka.bits_from_rows(kg["kx1_rows"][:300], 64)
    for i, line in enumerate(open(out)):
        assert [int(x) for x in line.split()] == list(np.nonzero(want[i])[0]), i
