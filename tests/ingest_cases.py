"""FASTA / FASTQ texts for the record-reader parity tests (oracle vs the reference's SeqAn reader vs the device parser).

Each case: (name, format 0=FASTA/1=FASTQ, iupac alphabet?, text, device_ok) — device_ok False marks inputs the reference
accepts but the device parser deliberately rejects (a FASTQ record that starts right behind the previous record's qualities
without white space).  Multi-line FASTQ and blank lines between FASTQ records go through the general (speculative) parser of
csrc/ingest_general.cu."""
import numpy as np


def _rand_seq(rng, n, alphabet=b"ACGT"):
    return bytes(rng.choice(np.frombuffer(alphabet, dtype=np.uint8), n).tobytes())


def cases():
    rng = np.random.default_rng(1234)
    out = []
    # FASTA: reads alphabet
    out.append(("fa_simple", 0, False, b">r1\nACGTACGTNN\n>r2 desc\nacgtn\n", True))
    out.append(("fa_multiline_crlf", 0, False, b">r1\r\nACGT\r\nAC GT\r\n\r\n>r2\r\nTTTT\tGG\r\n", True))
    out.append(("fa_no_trailing_newline", 0, False, b">r1\nACGT\n>r2\nGGCC", True))
    out.append(("fa_leading_junk", 0, False, b"; comment\nACGT\n>r1\nAC\nGT\n", True))
    out.append(("fa_gt_in_id_and_empty_records", 0, False, b">r1 >not a record\nACGT\n>empty\n>r3\n\nA\n>last", True))
    out.append(("fa_gt_mid_line", 0, False, b">r1\nACGT>r2 starts mid line\nGGGG\n", True))
    out.append(("fa_lone_cr", 0, False, b">r1\rACGT\r>r2\rGG\r", True))
    out.append(("fa_bad_char", 0, False, b">r1\nACGTXACGT\n", True))          # ParseError in the reference
    out.append(("fa_u_in_reads", 0, False, b">r1\nACGUACGT\n", True))          # 'U' is not Dna5: ParseError
    out.append(("fa_no_record", 0, False, b"ACGT\nACGT\n", True))              # UnexpectedEnd
    out.append(("fa_empty", 0, False, b"", True))
    # FASTA: contigs through Iupac
    out.append(("fa_iupac", 0, True, b">c1\nACGTUMRWSYKVHDBNacgtumrwsykvhdbn\n>c2\nNNNNACGT\n", True))
    out.append(("fa_iupac_bad", 0, True, b">c1\nACGT-ACGT\n", True))
    big = []
    for i in range(300):
        s = _rand_seq(rng, int(rng.integers(0, 700)), b"ACGTN" if i % 7 else b"ACGTacgtNn")
        w = int(rng.integers(20, 90))
        big.append(b">contig_%d some > text\n" % i + b"\n".join(s[j:j + w] for j in range(0, len(s), w)) + b"\n")
    out.append(("fa_random_wrapped", 0, False, b"".join(big), True))
    out.append(("fa_long_single_line", 0, True, b">chr\n" + _rand_seq(rng, 300_000, b"ACGTNRYK") + b"\n>chr2\n" + _rand_seq(rng, 70_001) + b"\n", True))
    # FASTQ
    out.append(("fq_simple", 1, False, b"@r1\nACGTN\n+\nIIII#\n@r2\nGG\n+r2\n@+\n", True))
    out.append(("fq_crlf_no_trailing_newline", 1, False, b"@r1\r\nACGT\r\n+\r\n!!!!\r\n@r2\r\nTTGA\r\n+\r\n@@@@", True))
    out.append(("fq_trailing_blank_lines", 1, False, b"@r1\nACGT\n+\nIIII\n\n\n  \n", True))
    out.append(("fq_empty_read", 1, False, b"@r1\n\n+\n\n@r2\nAC\n+\nII\n", True))
    out.append(("fq_qual_too_long", 1, False, b"@r1\nACGT\n+\nIIIII\n@r2\nAC\n+\nII\n", True))     # ParseError
    out.append(("fq_qual_too_short_last", 1, False, b"@r1\nACGT\n+\nIII\n@r2\nAC\n+\nII\n", True))  # garbage -> ParseError
    out.append(("fq_bad_char", 1, False, b"@r1\nAC-T\n+\nIIII\n", True))
    out.append(("fq_missing_plus", 1, False, b"@r1\nACGT\n", True))
    out.append(("fq_empty", 1, False, b"", True))
    out.append(("fq_multiline", 1, False, b"@r1\nACGT\nACGT\n+\nIIII\nIIII\n@r2\nAC\n+\nII\n", True))
    out.append(("fq_blank_between", 1, False, b"@r1\nACGT\n+\nIIII\n\n@r2\nAC\n+\nII\n", True))
    # the general parser: qualities cut short by the end of the text are accepted; junk before the first '@'; '+' line with text;
    # an error deep inside a wrapped file; a record glued to the previous qualities (reference: fine, device: refused)
    out.append(("fq_multiline_short_quals_at_end", 1, False, b"@r1\nACGT\nAC\n+\nIIII\nII\n@r2\nACGTA\nCC\n+r2 again\nIII\nI", True))
    out.append(("fq_multiline_leading_junk", 1, False, b"junk line\n\n@r1\nAC\nGT\n+\nII\nII\n", True))
    out.append(("fq_multiline_bad_quals", 1, False, b"@r1\nACGT\nACGT\n+\nIIII\nIIIII\n@r2\nAC\n+\nII\n", True))      # ParseError
    out.append(("fq_multiline_bad_char", 1, False, b"@r1\nACGT\nAC-T\n+\nIIII\nIIII\n", True))                          # ParseError
    out.append(("fq_nospace", 1, False, b"@r1\nACGT\n+\nIIII@r2\nAC\n+\nII\n", False))
    recs = []
    for i in range(2000):
        n = int(rng.integers(1, 260))
        s = _rand_seq(rng, n, b"ACGT" if i % 11 else b"ACGTNacgtn")
        q = bytes(rng.integers(33, 127, n, dtype=np.uint8).tobytes())           # includes '@', '+', '>'
        recs.append(b"@read%d/1 x\n" % i + s + b"\n+\n" + q + b"\n")
    out.append(("fq_random", 1, False, b"".join(recs), True))
    # wrapped at random widths, blank lines between some records, qualities from the whole printable range: wrapped quality
    # lines that start with '@' or '+' are candidates that are not records
    recs = []
    for i in range(700):
        n = int(rng.integers(0, 400))
        s = _rand_seq(rng, n, b"ACGT" if i % 13 else b"ACGTNacgtn")
        q = bytearray(rng.integers(33, 127, n, dtype=np.uint8).tobytes())
        ws_, wq = int(rng.integers(7, 120)), int(rng.integers(5, 90))
        for j in range(0, n, wq):
            if rng.random() < 0.3:
                q[j] = ord("@") if rng.random() < 0.7 else ord("+")
        q = bytes(q)
        wrap = lambda x, w: b"\n".join(x[j:j + w] for j in range(0, len(x), w)) if x else b""
        recs.append(b"@read%d wrapped\n" % i + wrap(s, ws_) + b"\n+" + (b"read%d" % i if i % 3 == 0 else b"") + b"\n" + wrap(q, wq) + b"\n"
                    + (b"\n" * int(rng.integers(0, 3)) if i % 5 == 0 else b""))
    out.append(("fq_random_wrapped", 1, False, b"".join(recs), True))
    return out


ASCII_TO_DNA5 = np.full(256, 4, dtype=np.uint8)
for _c, _v in ((b"Aa", 0), (b"Cc", 1), (b"Gg", 2), (b"TtUu", 3)):
    for _b in _c:
        ASCII_TO_DNA5[_b] = _v
