"""GPU (B200): dyibf_parse_records — SeqAn's FASTA / FASTQ record reader on the device — against the oracle's restatement
and the golden vectors of the unmodified reference, and the chain text -> parse -> filter."""
import ctypes as C

import numpy as np
import pytest

import ka
from ingest_cases import ASCII_TO_DNA5, cases

pytestmark = pytest.mark.gpu

CASES = cases()


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_device_reader_vs_oracle_and_reference(dy, oracle, ingest_golden, case):
    name, fmt, iupac, text, device_ok = case
    want = oracle.parse_records(text, fmt, iupac)
    if want is None or not device_ok:
        with pytest.raises(dy.DyibfError) as ei:
            dy.parseRecords(text, fmt, iupac=iupac)
        assert ei.value.code == -4                                   # DYIBF_ERR_FORMAT
        return
    seqs, limits = dy.parseRecords(text, fmt, iupac=iupac)
    assert np.array_equal(seqs, want[0]) and np.array_equal(limits, want[1])
    assert np.array_equal(ASCII_TO_DNA5[seqs], ingest_golden[name + "_codes"])
    assert np.array_equal(limits, ingest_golden[name + "_limits"])
    if text:
        from dream_yara_b200.ingest import countRecords
        assert countRecords(text, fmt, iupac=iupac) == (limits.size - 1, seqs.size)


def test_device_reader_random_texts(dy, oracle):
    """tile boundaries everywhere: random record lengths, line widths, CRLF, ids with '>' '@' '+', qualities with
    every printable character; sizes from a few bytes to several tiles"""
    rng = np.random.default_rng(99)
    for t in range(60):
        fq = t % 2
        crlf = b"\r\n" if t % 3 == 0 else b"\n"
        recs = []
        for i in range(int(rng.integers(1, 400 if t > 20 else 12))):
            n = int(rng.integers(0, 400))
            s = bytes(rng.choice(np.frombuffer(b"ACGTNacgtn", dtype=np.uint8), n).tobytes())
            if fq:
                q = bytes(rng.integers(33, 127, n, dtype=np.uint8).tobytes())
                recs.append(b"@r%d +@> \t x" % i + crlf + s + crlf + b"+" + (b"r%d" % i if i % 2 else b"") + crlf + q + crlf)
            else:
                w = int(rng.integers(1, 90))
                recs.append(b">r%d > @ + x" % i + crlf + crlf.join(s[j:j + w] for j in range(0, n, w)) + crlf)
        text = b"".join(recs)
        if t % 5 == 0:
            text = text.rstrip()                                     # no trailing newline
        want = oracle.parse_records(text, fq)
        assert want is not None
        seqs, limits = dy.parseRecords(text, fq)
        assert np.array_equal(limits, want[1]), "case %d" % t
        assert np.array_equal(seqs, want[0]), "case %d" % t


def test_device_reader_first_error_offset(dy):
    text = b">r1\nACGT\n>r2\nAC" + b"G" * 9000 + b"X" + b"ACGT" * 3000 + b"-\n"
    with pytest.raises(dy.DyibfError) as ei:
        dy.parseRecords(text, 0)
    assert "at byte %d" % text.index(b"X") in str(ei.value)
    with pytest.raises(dy.DyibfError) as ei:
        dy.parseRecords(b"@r\nACGT\n+\nIII\n@r2\nAC\n+\nII\n", 1)
    assert ei.value.code == -4 and "differ in length" in str(ei.value)


def test_device_reader_device_memory_limit_base_and_capacity(dy):
    import torch
    from dream_yara_b200.ingest import parseRecordsDevice
    L = dy.lib()
    text = b"".join(b"@r%d\n%s\n+\n%s\n" % (i, b"ACGTN" * (i % 40), b"I" * (5 * (i % 40))) for i in range(3000))
    want_seqs, want_limits = dy.parseRecords(text, 1)
    d_text = torch.frombuffer(bytearray(b"\0" + text), dtype=torch.uint8).cuda()[1:]      # a mis-aligned device pointer
    d_seqs = torch.zeros(want_seqs.size + 8, dtype=torch.uint8, device="cuda")
    d_lim = torch.zeros(want_limits.size, dtype=torch.int64, device="cuda")
    n, nb = parseRecordsDevice(d_text.data_ptr(), len(text), 1, d_seqs.data_ptr(), d_seqs.numel(), d_lim.data_ptr(),
                               d_lim.numel() - 1, limit_base=1000)
    assert (n, nb) == (want_limits.size - 1, want_seqs.size)
    assert np.array_equal(d_seqs.cpu().numpy()[:nb], want_seqs)
    assert np.array_equal(d_lim.cpu().numpy().view(np.uint64), want_limits + np.uint64(1000))
    ns, nbb = C.c_uint64(), C.c_uint64()
    rc = L.dyibf_parse_records(0, C.c_void_p(d_text.data_ptr()), len(text), 1, 0, C.c_void_p(d_seqs.data_ptr()), 10,
                               C.c_void_p(d_lim.data_ptr()), d_lim.numel() - 1, 0, C.byref(ns), C.byref(nbb), 1, None)
    assert rc == -6 and (ns.value, nbb.value) == (n, nb)              # DYIBF_ERR_CAPACITY, sizes reported


def test_text_to_bins_end_to_end(dy, golden, tmp_path):
    """FASTA text of the reference's 64 viral genomes and 6400 test reads -> device reader -> build + classify: the
    reference's own gold md5 and its per-read bin sets (KA1), and loadReads / addFastaFile of the host mirror."""
    genomes, reads = ka.records(golden, "genomes"), ka.records(golden, "reads")
    f = dy.SeqAnBloomFilter(64, 2, 19, ka.GIB_BITS)
    for b, g in enumerate(genomes):                                   # one FASTA file per bin, 70 columns, as build_filter reads them
        p = tmp_path / ("%d.fa" % b)
        p.write_bytes(b">genome %d\n" % b + b"\n".join(g[j:j + 70] for j in range(0, len(g), 70)) + b"\n")
        f.addFastaFile(str(p), b)
    assert ka.md5_filter_words(f.download_words(), f._noOfBits) == str(golden["reference_gold_md5"])
    rp = tmp_path / "reads.fa"
    rp.write_bytes(b"".join(b">read.%d\n%s\n" % (i, r) for i, r in enumerate(reads)))
    rs = dy.loadReads(str(rp))
    assert rs.getReadsCount() == 6400 and not rs.paired
    rows = f.whichBinsBatch((rs.concat, rs.offsets), 25, group=2, derive_rc=True, encoding=rs.encoding)
    assert np.array_equal(rows, golden["ka1_rows"])
    q1, q2 = tmp_path / "m1.fq", tmp_path / "m2.fq"
    q1.write_bytes(b"".join(b"@p%d/1\n%s\n+\n%s\n" % (i, r, b"I" * len(r)) for i, r in enumerate(reads[:3200])))
    q2.write_bytes(b"".join(b"@p%d/2\n%s\n+\n%s\n" % (i, r, b"@" * len(r)) for i, r in enumerate(reads[3200:])))
    pe = dy.loadReads(str(q1), str(q2))
    assert pe.paired and pe.getPairsCount() == 3200
    rows_pe = f.whichBinsBatch((pe.concat, pe.offsets), 25, group=4, derive_rc=True, encoding=pe.encoding)
    assert np.array_equal(rows_pe, golden["ka1_rows"][:3200] | golden["ka1_rows"][3200:])
    assert dy.read_fasta(rp, iupac=False)[:2] == reads[:2]
    f.close()


def test_classify_text_equals_parse_then_classify(dy, golden, gather_golden):
    """dyibf_classify_text (text up, lists back) against the reference's own clasifyLoadedReads lists (KA5, SE and PE) and
    against dyibf_classify on the parsed reads; the batch it leaves resident serves dyibf_gather_reads."""
    genomes, reads = ka.records(golden, "genomes"), ka.records(golden, "reads")
    f = dy.SeqAnBloomFilter(64, 2, 19, ka.GIB_BITS)
    f.addSequences(genomes, np.arange(64))
    fa = b"".join(b">r%d some text\n%s\n%s\n" % (i, r[:61], r[61:]) for i, r in enumerate(reads))
    offs, ids, n, thr = f.classifyText(fa, fmt=0, e_percent=3)
    assert (n, thr) == (6400, int(golden["ka5_se_threshold"]))
    assert np.array_equal(offs, golden["ka5_se_bin_offsets"]) and np.array_equal(ids, golden["ka5_se_read_ids"])
    q1 = b"".join(b"@p%d/1\n%s\n+\n%s\n" % (i, r, b"I" * len(r)) for i, r in enumerate(reads[:3200]))
    q2 = b"".join(b"@p%d/2\n%s\n+\n%s\n" % (i, r, b"+" * len(r)) for i, r in enumerate(reads[3200:]))
    offs, ids, n, thr = f.classifyText(q1, q2, fmt=1, e_percent=3)
    assert (n, thr) == (3200, int(golden["ka5_pe_threshold"]))
    assert np.array_equal(offs, golden["ka5_pe_bin_offsets"]) and np.array_equal(ids, golden["ka5_pe_read_ids"])
    o2, i2 = f.classify(reads, thr, group=4, derive_rc=True)
    assert np.array_equal(offs, o2) and np.array_equal(ids, i2)
    # the resident batch (ASCII) feeds loadFilteredReads: forward mates, then reverse complements
    b = int(np.argmax(np.diff(offs.astype(np.int64))))
    seq_bytes, limits, gids = f.gatherReads(b)
    m = int(offs[b + 1] - offs[b])
    assert limits.size - 1 == 4 * m and list(gids[:m]) == list(ids[int(offs[b]):int(offs[b + 1])])
    first = int(gids[0])
    assert seq_bytes[int(limits[0]):int(limits[1])].tobytes() == reads[first]
    comp = bytes.maketrans(b"ACGTN", b"TGCAN")
    assert seq_bytes[int(limits[2 * m]):int(limits[2 * m + 1])].tobytes() == reads[first][::-1].translate(comp)
    with pytest.raises(dy.DyibfError):
        f.classifyText(q1, q2[: len(q2) // 2 + 3], fmt=1, e_percent=3)
    offs, ids, n, thr = f.classifyText(b"", fmt=1, e_percent=3)
    assert n == 0 and ids.size == 0
    f.close()


def test_classify_text_tickets_in_flight(dy, golden):
    """dyibf_classify_text_begin: three text batches in flight on one handle (upload + tokenising of batch i+1 overlap the
    query of batch i) give, batch by batch, the lists of the synchronous call; the ticket ended last is the resident batch;
    a malformed text fails in begin and leaves the tickets usable."""
    import ctypes as C
    from dream_yara_b200._lib import check
    L = dy.lib()
    genomes, reads = ka.records(golden, "genomes"), ka.records(golden, "reads")
    f = dy.SeqAnBloomFilter(64, 2, 19, ka.GIB_BITS)
    f.addSequences(genomes, np.arange(64))
    batches = [reads[i:i + 800] for i in range(0, 6400, 800)]
    texts = [np.frombuffer(b"".join(b"@r%d\n%s\n+\n%s\n" % (i, r, b"F" * len(r)) for i, r in enumerate(b)), dtype=np.uint8) for b in batches]
    want = [f.classifyText(t.tobytes(), fmt=1, e_percent=3) for t in texts]
    q, got = [], []

    def end():
        t, n, thr = q.pop(0)
        offs = np.zeros(65, dtype=np.uint64); ids = np.zeros(8 * 800, dtype=np.uint32); tot = C.c_uint64(0)
        check(L.dyibf_classify_end(t, offs.ctypes.data, ids.ctypes.data, ids.size, C.byref(tot)))
        got.append((offs, ids[:tot.value].copy(), n, thr))

    for t in texts:
        if len(q) == 3:
            end()
        tk, n, thr = C.c_void_p(), C.c_uint32(0), C.c_uint32(0)
        check(L.dyibf_classify_text_begin(f._h, t.ctypes.data, t.size, None, 0, 1, 3, C.byref(n), C.byref(thr), C.byref(tk)))
        q.append((tk, n.value, thr.value))
    bad = np.frombuffer(b"@x\nACGT\n+\nIIII\n@y\nAC-T\n+\nIIII\n", dtype=np.uint8)     # a character outside the alphabet: ParseError
    tk = C.c_void_p()
    assert L.dyibf_classify_text_begin(f._h, bad.ctypes.data, bad.size, None, 0, 1, 3, None, None, C.byref(tk)) == -4 and not tk.value
    while q:
        end()
    for (o1, i1, n1, t1), (o0, i0, n0, t0) in zip(got, want):
        assert (n1, t1) == (n0, t0) and np.array_equal(o1, o0) and np.array_equal(i1, i0)
    # the batch ended last is resident: its first bin's reads come back
    offs = got[-1][0]
    b = int(np.argmax(np.diff(offs.astype(np.int64))))
    seq_bytes, limits, gids = f.gatherReads(b)
    first = int(gids[0])
    assert seq_bytes[int(limits[0]):int(limits[1])].tobytes() == batches[-1][first]
    f.close()


def test_add_fasta_texts_equals_add_sequences(dy):
    """dyibf_add_fasta_texts (addFastaFile for many files from their text, tokenised on the device) against the same records
    added through dyibf_add_sequences: the joined one-pass form (FASTA texts that start with '>') and the one-by-one form
    (FASTQ, text before the first '>'); a character outside the IUPAC alphabet names its text."""
    from dream_yara_b200.ingest import parseRecords
    rng = np.random.default_rng(11)
    bins, h, k = 130, 3, 18
    nbits = (1 << 30) + 256

    def rand(n, alphabet=b"ACGT"):
        return bytes(rng.choice(np.frombuffer(alphabet, dtype=np.uint8), n).tobytes())

    def wrap(s, w):
        return b"\n".join(s[j:j + w] for j in range(0, len(s), w))

    texts, ids = [], []
    for b in range(0, bins, 3):
        recs = [rand(int(rng.integers(0, 3000)), b"ACGTacgtNRYKM" if b % 9 == 0 else b"ACGT") for _ in range(int(rng.integers(1, 5)))]
        t = b"".join(b">c%d_%d text > more\n" % (b, i) + wrap(r, int(rng.integers(30, 90))) + b"\n" for i, r in enumerate(recs))
        if b % 12 == 0:
            t = t[:-1]                                               # no newline at the end of the file
        if b % 15 == 0:
            t += b">cut short by the end of the file"                # an id line without a newline
        if b % 21 == 0:
            t += b">x\nACGTACGTACGTACGTACGTAC>a record that starts in the middle of a line\nGGGGGGGGGGGGGGGGGGGGGGGGG\n"
        texts.append(t); ids.append(b)
    texts.append(b""); ids.append(1)                                 # an empty file

    def reference(texts_, ids_, fmts=None):
        g = dy.SeqAnBloomFilter(bins, h, k, nbits)
        seqs, sb = [], []
        for j, (t, b) in enumerate(zip(texts_, ids_)):
            if not t:
                continue
            s_, lim = parseRecords(t, 0 if fmts is None else fmts[j], iupac=True)
            for i in range(lim.size - 1):
                seqs.append(s_[int(lim[i]):int(lim[i + 1])]); sb.append(b)
        g.addSequences(seqs, sb, encoding=2)
        w = g.download_words(); g.close()
        return w

    f = dy.SeqAnBloomFilter(bins, h, k, nbits)
    f.addFastaTexts(texts, ids)
    assert np.array_equal(f.download_words(), reference(texts, ids)) and f.device_errors() == 0
    f.close()
    # one by one: a FASTQ file and a FASTA file with text before its first '>'
    mixed = [texts[0], b"@q1\n" + rand(400) + b"\n+\n" + b"I" * 400 + b"\n@q2\nACGTN\n+\nIIIII\n", b"; comment\n" + texts[2]]
    mids, fmts = [5, 77, 129], [0, 1, 0]
    f = dy.SeqAnBloomFilter(bins, h, k, nbits)
    f.addFastaTexts(mixed, mids, fmts)
    assert np.array_equal(f.download_words(), reference(mixed, mids, fmts))
    # ParseError: names the text, adds nothing
    before = f.download_words()
    with pytest.raises(dy.DyibfError) as ei:
        f.addFastaTexts([texts[0], b">bad\nACGTACGTACGTACGT-ACGTACGTACGT\n", texts[1]], [1, 2, 3])
    assert ei.value.code == -4 and ei.value.bad_index == 1 and "of text 1" in str(ei.value)
    assert np.array_equal(f.download_words(), before)
    f.close()
