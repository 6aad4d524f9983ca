"""The standalone builder / updater (tools/dyibf_build_filter.cpp, tools/dyibf_update_filter.cpp): the reference's
dream_yara_build_filter / dream_yara_update_filter command lines on top of the C ABI."""
import gzip
import os
import subprocess

import numpy as np
import pytest

import ka

HERE = os.path.dirname(os.path.abspath(__file__))
TOOLS = os.path.join(os.path.dirname(HERE), "tools")
BUILD = os.path.join(TOOLS, "bin", "dyibf_build_filter")
UPDATE = os.path.join(TOOLS, "bin", "dyibf_update_filter")


@pytest.fixture(scope="module")
def tools(dy):
    r = subprocess.run(["make", "-C", TOOLS], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    return BUILD, UPDATE


def _write_fasta(path, records, width=80, gz=False):
    op = gzip.open if gz else open
    with op(path, "wb") as fh:
        for name, seq in records:
            fh.write(b">" + name.encode() + b" some description\n")
            for j in range(0, len(seq), width):
                fh.write(seq[j:j + width] + b"\n")


def test_cli_argument_checks(tools, tmp_path):
    build, update = tools
    r = subprocess.run([build, "-bs", "3", str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 1 and "should be a power of 2" in r.stderr          # d_build_filter.cpp:198-210
    r = subprocess.run([build, "-k", "40", str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 1
    r = subprocess.run([build, "-ft", "none", str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 1 and "bloom or kmer_direct" in r.stderr
    r = subprocess.run([build, "-ft", "kmer_direct", "-k", "16", str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 1 and "-k <= 15" in r.stderr                        # the reference computes 1u << 2k
    r = subprocess.run([update, str(tmp_path / "x.filter"), str(tmp_path / "abc.fna")], capture_output=True, text=True)
    assert r.returncode == 1 and "doesn't have a valid name" in r.stderr        # d_update_filter.cpp:158-162


@pytest.mark.gpu
def test_build_and_update_cli(tools, golden, oracle, tmp_path):
    build, update = tools
    genomes = ka.records(golden, "genomes")
    d = tmp_path / "bins"; d.mkdir()
    for b, g in enumerate(genomes):
        _write_fasta(d / ("%d.fna" % b), [("bin%d" % b, g)])
    filt = tmp_path / "cli.filter"
    # the reference's own test command line (tests/run_tests.py:116-122), incl. its "-nh ' 2'"
    r = subprocess.run([build, "-b", "64", "-t", "4", "-k", "19", "-nh", " 2", "-bs", "1", "-v", str(d), "-o", str(filt)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "Finshed in" in r.stderr and "[bin 63] Done adding kmers!" in r.stderr
    assert filt.stat().st_size == 1073741864
    assert ka.md5_file(filt) == str(golden["reference_gold_md5"])

    # update bins 5 and 40: multi-record gzip FASTA with a short record and lower-case / IUPAC characters
    rng = np.random.default_rng(3)
    new5 = [bytes(rng.choice(list(b"ACGTacgtNRY"), 3000).astype(np.uint8)), b"ACGT", bytes(rng.choice(list(b"ACGT"), 500).astype(np.uint8))]
    new40 = [genomes[7][::-1]]
    u = tmp_path / "upd"; u.mkdir()
    _write_fasta(u / "5.fa.gz", [("a", new5[0]), ("short", new5[1]), ("c", new5[2])], width=60, gz=True)
    with open(u / "40.fastq", "wb") as fh:                   # FASTQ whose quality line starts with '@' and contains '+'
        fh.write(b"@x desc\n" + new40[0] + b"\n+\n@" + b"+" * (len(new40[0]) - 1) + b"\n")
    r = subprocess.run([update, "-t", "2", "-v", str(filt), str(u / "5.fa.gz"), str(u / "40.fastq")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    p = oracle.params(64, 2, 19, ka.GIB_BITS)
    words = oracle.new_words(p)
    for b, g in enumerate(genomes):
        oracle.add_kmers(words, p, oracle.encode_ascii(g), b)
    oracle.clear_bins(words, p, [5, 40])
    for rec in new5:
        if len(rec) >= 19:
            oracle.add_kmers(words, p, oracle.encode_ascii(rec), 5)
    oracle.add_kmers(words, p, oracle.encode_ascii(new40[0]), 40)
    oracle.save(words, p, tmp_path / "want.filter")
    assert ka.md5_file(filt) == ka.md5_file(tmp_path / "want.filter")
    # a bin beyond the filter's range is refused like the reference does
    _write_fasta(u / "64.fna", [("x", genomes[0])])
    r = subprocess.run([update, str(filt), str(u / "64.fna")], capture_output=True, text=True)
    assert r.returncode == 1 and "has only 64 Bins" in r.stderr
    # SeqAn's reader rules (readRecord into an IupacString): a '>' in the middle of a FASTA line starts a new record, a character
    # outside the IUPAC alphabet aborts.  The device tokeniser (dyibf_parse_records, pinned against the compiled reference) says
    # what the records are; the tool must add exactly those.
    from dream_yara_b200.ingest import parseRecords
    text = b">r1 x\n" + genomes[3][:700] + b">mid line\n" + genomes[3][700:1500] + b"\n>short\nACGT\n"
    with open(u / "7.fa", "wb") as fh:
        fh.write(text)
    r = subprocess.run([update, str(filt), str(u / "7.fa")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    seqs, limits = parseRecords(text, 0, iupac=True)
    assert limits.size - 1 == 3
    oracle.clear_bins(words, p, [7])
    for i in range(limits.size - 1):
        rec = seqs[int(limits[i]):int(limits[i + 1])]
        if rec.size >= 19:
            oracle.add_kmers(words, p, oracle.encode_ascii(rec.tobytes()), 7)
    oracle.save(words, p, tmp_path / "want2.filter")
    assert ka.md5_file(filt) == ka.md5_file(tmp_path / "want2.filter")
    with open(u / "8.fa", "wb") as fh:
        fh.write(b">bad\nACGTACGTACGTACGTACGT-ACGTACGTACGTACGT\n")
    r = subprocess.run([update, str(filt), str(u / "8.fa")], capture_output=True, text=True)
    assert r.returncode == 1 and "ParseError" in r.stderr


@pytest.mark.gpu
def test_build_cli_many_threads_small_files(tools, tmp_path):
    """Every file must reach the filter whichever reader task finishes last (a task that was not the last to
    finish could push its files after the final flush): 64 reader tasks over 64 small files, several times, against the
    one-task build of the same directory."""
    build, _ = tools
    rng = np.random.default_rng(11)
    d = tmp_path / "bins"; d.mkdir()
    for b in range(64):
        _write_fasta(d / ("%d.fna" % b), [("bin%d" % b, bytes(rng.choice(list(b"ACGT"), 400 + 37 * b).astype(np.uint8)))])
    common = ["-b", "64", "-k", "19", "-nh", "3", "-bs", "1", str(d)]
    one = tmp_path / "one.filter"
    r = subprocess.run([build, "-t", "1", "-o", str(one)] + common, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    want = ka.md5_file(one)
    for rep in range(3):
        many = tmp_path / "many.filter"
        r = subprocess.run([build, "-t", "64", "-o", str(many)] + common, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        assert ka.md5_file(many) == want, "run %d with 64 reader tasks lost or changed bins" % rep


@pytest.mark.gpu
def test_build_cli_kmer_direct(tools, golden, oracle, dy, tmp_path):
    """dream_yara_build_filter -ft kmer_direct (d_build_filter.cpp:305-313): 4^k * bins + 256 bits, trailer [bins, 0, k, 0]"""
    build, update = tools
    genomes, reads = ka.records(golden, "genomes"), ka.records(golden, "reads")
    d = tmp_path / "bins"; d.mkdir()
    for b, g in enumerate(genomes):
        _write_fasta(d / ("%d.fa" % b), [("bin%d" % b, g)])
    filt = tmp_path / "kdx.filter"
    r = subprocess.run([build, "-b", "64", "-t", "2", "-ft", "kmer_direct", "-k", "14", str(d), "-o", str(filt)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    nbits = (1 << 28) * 64 + 256
    assert filt.stat().st_size == 8 + nbits // 8
    p = oracle.kdx_params(64, 14, nbits)
    words = oracle.new_words(p)
    for b, g in enumerate(genomes):
        oracle.kdx_add_kmers(words, p, oracle.encode_ascii(g), b)
    oracle.set_metadata(words, p)
    f = dy.SeqAnKDXFilter(str(filt))
    assert (f.getNumberOfBins(), f.getKmerSize(), f._noOfHashFunc) == (64, 14, 0)
    assert np.array_equal(f.download_words(), words)
    # update: clear + re-add bin 7 from a new file; the kind comes from the trailer
    _write_fasta(tmp_path / "7.fa", [("new7", genomes[7][::-1])])
    r = subprocess.run([update, "-ft", "kmer_direct", str(filt), str(tmp_path / "7.fa")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    oracle.clear_bins(words, p, [7])
    oracle.kdx_add_kmers(words, p, oracle.encode_ascii(genomes[7][::-1]), 7)
    oracle.set_metadata(words, p)
    g2 = dy.SeqAnKDXFilter(str(filt))
    assert np.array_equal(g2.download_words(), words)
    codes = [oracle.encode_ascii(x) for x in reads[:200]]
    assert np.array_equal(g2.whichBinsBatch(reads[:200], 45, group=2, derive_rc=True), oracle.kdx_rows(words, p, codes, 45))
    f.close(); g2.close()


@pytest.mark.gpu
def test_filter_reads_cli_vs_reference_mapper_output(tools, dy, golden, tmp_path):
    """tools/dyibf_filter_reads = the filtering stage of dream_yara_mapper: the `-vv` per-bin sizes, "Avg reads per bin"
    and the per-bin lists of the reference's own mapper run on its test data, single- and
    paired-end, FASTA and gzipped FASTQ input, one batch and several."""
    exe = os.path.join(TOOLS, "bin", "dyibf_filter_reads")
    genomes, reads = ka.records(golden, "genomes"), ka.records(golden, "reads")
    f = dy.SeqAnBloomFilter(64, 2, 19, ka.GIB_BITS)
    f.addSequences(genomes, np.arange(64))
    filt = tmp_path / "ka1.filter"
    assert f.save(filt)
    f.close()
    _write_fasta(tmp_path / "reads.fa", [("r%d" % i, r) for i, r in enumerate(reads)], width=60)
    out = tmp_path / "se.txt"
    r = subprocess.run([exe, "-fi", str(filt), "-e", "3", "-vv", "-o", str(out), str(tmp_path / "reads.fa")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    sizes = [int(l.split("\t")[1]) for l in r.stderr.splitlines() if l.startswith("bin ")]
    bo, ids = golden["ka5_se_bin_offsets"], golden["ka5_se_read_ids"]
    assert sizes == [int(bo[b + 1] - bo[b]) for b in range(64)] and sizes[0] == 173 and sum(sizes) == 6538
    assert "Avg reads per bin:\t\t102.156" in r.stderr                  # the reference mapper's own line
    lines = [l.split() for l in open(out)]
    assert [int(l[2]) for l in lines] == sorted(sizes, reverse=True)      # sortedBins order
    for l in lines:
        b = int(l[1])
        assert [int(x) for x in l[3:]] == list(ids[int(bo[b]):int(bo[b + 1])])
    # paired-end, gzipped FASTQ mates (KA5: pair j = reads j and j + 3200)
    for name, part, q in (("m1.fq.gz", reads[:3200], b"I"), ("m2.fq.gz", reads[3200:], b"@")):
        with gzip.open(tmp_path / name, "wb") as fh:
            for i, s in enumerate(part):
                fh.write(b"@p%d\n%s\n+\n%s\n" % (i, s, q * len(s)))
    r = subprocess.run([exe, "-fi", str(filt), "-e", "3", "-vv", str(tmp_path / "m1.fq.gz"), str(tmp_path / "m2.fq.gz")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    sizes = [int(l.split("\t")[1]) for l in r.stderr.splitlines() if l.startswith("bin ")]
    bo = golden["ka5_pe_bin_offsets"]
    assert sizes == [int(bo[b + 1] - bo[b]) for b in range(64)]
    assert "Avg reads per bin:\t\t204.312" in r.stderr
    # several batches (-rb 1000): ids are batch-local, sizes add up
    r = subprocess.run([exe, "-fi", str(filt), "-e", "3", "-rb", "1000", "-vv", str(tmp_path / "reads.fa")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    sizes = np.array([int(l.split("\t")[1]) for l in r.stderr.splitlines() if l.startswith("bin ")]).reshape(7, 64)
    assert int(sizes.sum()) == 6538
    want = ka.bits_from_rows(golden["ka1_rows"], 64)
    assert np.array_equal(sizes[2], want[2000:3000].sum(axis=0))
    # layouts the line cutter hands over whole (junk before the first record; trailing blank lines are plain white space):
    # same batches, same sizes
    with open(tmp_path / "junk.fa", "wb") as fh:
        fh.write(b"; a comment line\n" + open(tmp_path / "reads.fa", "rb").read() + b"\n\n  \n")
    r2 = subprocess.run([exe, "-fi", str(filt), "-e", "3", "-rb", "1000", "-vv", str(tmp_path / "junk.fa")], capture_output=True, text=True)
    assert r2.returncode == 0, r2.stderr
    assert [l for l in r2.stderr.splitlines() if l.startswith("bin ")] == [l for l in r.stderr.splitlines() if l.startswith("bin ")]
    with open(tmp_path / "tail.fq", "wb") as fh:
        for i, s_ in enumerate(reads):
            fh.write(b"@r%d\n%s\n+\n%s\n" % (i, s_, b"I" * len(s_)))
        fh.write(b"\n \n")
    r3 = subprocess.run([exe, "-fi", str(filt), "-e", "3", "-rb", "1000", "-vv", str(tmp_path / "tail.fq")], capture_output=True, text=True)
    assert r3.returncode == 0, r3.stderr
    assert [l for l in r3.stderr.splitlines() if l.startswith("bin ")] == [l for l in r.stderr.splitlines() if l.startswith("bin ")]
    assert "Reads count:\t\t\t6400" in r3.stderr
    # FASTQ with sequences and qualities wrapped at 50 columns and blank lines between some records: the head check sends the
    # file to the whole-text path, the device reader's general FASTQ parser tokenises it
    with open(tmp_path / "wrapped.fq", "wb") as fh:
        for i, s_ in enumerate(reads):
            q_ = b"@" + b"I" * (len(s_) - 1)                                   # every quality line group starts with '@'
            fh.write(b"@r%d\n" % i + b"\n".join(s_[j:j + 50] for j in range(0, len(s_), 50)) + b"\n+\n"
                     + b"\n".join(q_[j:j + 50] for j in range(0, len(q_), 50)) + b"\n" + (b"\n" if i % 7 == 0 else b""))
    r4 = subprocess.run([exe, "-fi", str(filt), "-e", "3", "-rb", "1000", "-vv", str(tmp_path / "wrapped.fq")], capture_output=True, text=True)
    assert r4.returncode == 0, r4.stderr
    assert [l for l in r4.stderr.splitlines() if l.startswith("bin ")] == [l for l in r.stderr.splitlines() if l.startswith("bin ")]
    # several GPUs of this one process (a replica each, successive batches to successive GPUs): same lists, through the text
    # tickets (regular layout) and through the parsed path (wrapped FASTQ)
    import ctypes as C
    ndev = C.c_int(0); dy.lib().dyibf_device_count(C.byref(ndev))
    if ndev.value >= 2:
        for name in ("reads.fa", "wrapped.fq"):
            r5 = subprocess.run([exe, "-fi", str(filt), "-e", "3", "-rb", "1000", "-vv", "--devices", "0,1", str(tmp_path / name)],
                                capture_output=True, text=True)
            assert r5.returncode == 0, r5.stderr
            assert [l for l in r5.stderr.splitlines() if l.startswith("bin ")] == [l for l in r.stderr.splitlines() if l.startswith("bin ")], name
    r = subprocess.run([exe, "-fi", str(tmp_path / "nope.filter"), str(tmp_path / "reads.fa")], capture_output=True, text=True)
    assert r.returncode == 1 and "could not be read" in r.stderr
