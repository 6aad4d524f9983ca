"""FASTA / FASTQ text -> sequences on the GPU (reference: SeqAn's readRecords as loadReads and addFastaFile drive it,
src/mapper.h:451, src/store_seqs.h:175-188, src/d_bloom_filter.h:127-147; seqan seq_io/fasta_fastq.h:268-345).

The file is read (and, for .gz, inflated) on the host — that is I/O — and tokenised by libdyibf.so's
dyibf_parse_records on the device; there is no host tokenizer behind it.
"""
import ctypes as C
import gzip

import numpy as np

from ._lib import ALPHA_DNA5, ALPHA_IUPAC, ENC_ASCII, FMT_FASTA, FMT_FASTQ, MEM_DEVICE, MEM_HOST, check, lib

_FASTQ_EXT = (".fq", ".fastq")
_FASTA_EXT = (".fa", ".fasta", ".fna", ".ffn", ".faa", ".frn", ".fas")


def guessFormat(path, head=b""):
    """SeqFileIn picks the format from the file extension (seq_io/sequence_file.h); content decides when it says nothing."""
    name = str(path).lower()
    if name.endswith(".gz"):
        name = name[:-3]
    if name.endswith(_FASTQ_EXT):
        return FMT_FASTQ
    if name.endswith(_FASTA_EXT):
        return FMT_FASTA
    return FMT_FASTQ if head[:1] == b"@" else FMT_FASTA


def parseRecords(text, fmt, iupac=False, device=0, limit_base=0):
    """All records of `text` (bytes): (sequence characters uint8[n_bases], limits uint64[n_records + 1]).
    Raises DyibfError(ERR_FORMAT) where SeqAn throws a ParseError / UnexpectedEnd (and for FASTQ that is not in the
    four-line layout)."""
    buf = np.frombuffer(bytes(text), dtype=np.uint8) if not isinstance(text, np.ndarray) else np.ascontiguousarray(text, dtype=np.uint8)
    n = int(buf.size)
    ns, nb = C.c_uint64(0), C.c_uint64(0)
    alpha = ALPHA_IUPAC if iupac else ALPHA_DNA5
    ptr = C.c_void_p(buf.ctypes.data) if n else None
    # a record needs at least two bytes of text and a sequence character one: no sizing pass needed
    seqs = np.zeros(max(n, 1), dtype=np.uint8)
    limits = np.zeros(n // 2 + 2, dtype=np.uint64)
    check(lib().dyibf_parse_records(device, ptr, n, fmt, alpha, C.c_void_p(seqs.ctypes.data), seqs.size,
                                    C.c_void_p(limits.ctypes.data), limits.size - 1, limit_base, C.byref(ns), C.byref(nb),
                                    MEM_HOST, None))
    return seqs[: nb.value], limits[: ns.value + 1]


def countRecords(text, fmt, iupac=False, device=0):
    """(n_records, n_bases) without producing the sequences (the sizing call of dyibf_parse_records)."""
    buf = np.frombuffer(bytes(text), dtype=np.uint8)
    ns, nb = C.c_uint64(0), C.c_uint64(0)
    check(lib().dyibf_parse_records(device, C.c_void_p(buf.ctypes.data) if buf.size else None, int(buf.size), fmt,
                                    ALPHA_IUPAC if iupac else ALPHA_DNA5, None, 0, None, 0, 0, C.byref(ns), C.byref(nb),
                                    MEM_HOST, None))
    return ns.value, nb.value


def readRecords(path, iupac=False, device=0, fmt=None):
    """readRecords(seqs, SeqFileIn(path)): the sequences of a FASTA / FASTQ (.gz) file as (characters, limits), ASCII
    (use with encoding=ENC_ASCII)."""
    opener = gzip.open if str(path).endswith(".gz") else open
    with opener(path, "rb") as fh:
        data = fh.read()
    if fmt is None:
        fmt = guessFormat(path, data[:1])
    return parseRecords(data, fmt, iupac=iupac, device=device)


def parseRecordsDevice(text_ptr, n_bytes, fmt, out_seqs_ptr, capacity_bytes, out_limits_ptr, capacity_seqs, iupac=False,
                       device=0, limit_base=0, stream=None):
    """Device-resident variant: text and outputs are device pointers; returns (n_records, n_bases)."""
    ns, nb = C.c_uint64(0), C.c_uint64(0)
    check(lib().dyibf_parse_records(device, C.c_void_p(text_ptr), n_bytes, fmt, ALPHA_IUPAC if iupac else ALPHA_DNA5,
                                    C.c_void_p(out_seqs_ptr), capacity_bytes, C.c_void_p(out_limits_ptr), capacity_seqs,
                                    limit_base, C.byref(ns), C.byref(nb), MEM_DEVICE, C.c_void_p(stream) if stream else None))
    return ns.value, nb.value


__all__ = ["parseRecords", "countRecords", "readRecords", "parseRecordsDevice", "guessFormat", "ENC_ASCII"]
