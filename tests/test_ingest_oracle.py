"""CPU: the oracle's restatement of SeqAn's FASTA / FASTQ record reader (oracle/ibf_oracle.c ibf_parse_records) against
the golden vectors the unmodified reference produced (tests/golden/make_golden_ingest.py) and, where oracle/_ref is
present, against the reference reader itself."""
import hashlib

import numpy as np
import pytest

from ingest_cases import ASCII_TO_DNA5, cases
from oracle_py import Ref

CASES = cases()


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_oracle_reader_matches_reference_vectors(oracle, ingest_golden, case):
    name, fmt, iupac, text, _ = case
    if text:
        assert str(ingest_golden[name + "_text_md5"]) == hashlib.md5(text).hexdigest(), "ingest_cases.py changed: regenerate the vectors"
    got = oracle.parse_records(text, fmt, iupac)
    if not int(ingest_golden[name + "_ok"]):
        assert got is None, "the reference throws on this text"
        return
    assert got is not None
    assert np.array_equal(ASCII_TO_DNA5[got[0]], ingest_golden[name + "_codes"])
    assert np.array_equal(got[1], ingest_golden[name + "_limits"])


def test_oracle_reader_on_reference_reads(oracle, golden, ingest_golden):
    """the reference's own test reads (tests/input/64-viral-reads.fa), rebuilt as FASTA text from the golden records"""
    cat, off = golden["reads_ascii"], golden["reads_offsets"]
    text = b"".join(b">r%d\n%s\n" % (i, cat[int(off[i]):int(off[i + 1])].tobytes()) for i in range(off.size - 1))
    seqs, limits = oracle.parse_records(text, 0)
    assert limits.size - 1 == int(ingest_golden["viral_reads_n"])
    assert hashlib.md5(ASCII_TO_DNA5[seqs].tobytes()).hexdigest() == str(ingest_golden["viral_reads_md5"])
    assert hashlib.md5(limits.tobytes()).hexdigest() == str(ingest_golden["viral_reads_limits_md5"])


@pytest.mark.skipif(not Ref.available(), reason="oracle/_ref not built")
def test_oracle_reader_vs_compiled_reference(oracle, tmp_path):
    ref = Ref()
    if not hasattr(ref.l, "ref_read_records"):
        pytest.skip("oracle/_ref predates ref_read_records")
    rng = np.random.default_rng(7)
    for t in range(20):
        fq = t % 2
        recs = []
        for i in range(int(rng.integers(1, 60))):
            n = int(rng.integers(0, 300))
            s = bytes(rng.choice(np.frombuffer(b"ACGTNacgtn", dtype=np.uint8), n).tobytes())
            if fq:
                q = bytes(rng.integers(33, 127, n, dtype=np.uint8).tobytes())
                recs.append(b"@r%d\n%s\n+\n%s\n" % (i, s, q))
            else:
                w = int(rng.integers(1, 80))
                recs.append(b">r%d x\n" % i + b"\n".join(s[j:j + w] for j in range(0, n, w)) + b"\n")
        text = b"".join(recs)
        path = tmp_path / ("t%d.%s" % (t, "fq" if fq else "fa"))
        path.write_bytes(text)
        want = ref.read_records(path)
        got = oracle.parse_records(text, fq)
        assert (want is None) == (got is None)
        if want is not None:
            assert np.array_equal(ASCII_TO_DNA5[got[0]], want[0]) and np.array_equal(got[1], want[1])
