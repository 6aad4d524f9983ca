"""ctypes access to the CPU checkers under oracle/ — TEST INFRASTRUCTURE.

    Oracle  : oracle/libibf_oracle.so   (plain-C restatement, oracle/ibf_oracle.c; built with `make -C oracle`)
    Ref     : oracle/_ref/libref_ibf.so (the UNMODIFIED reference class compiled from /root/reference;
              present only where `make -C oracle ref` ran — it travels to the GPU box as a built file)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "libibf_oracle.so")
REF_SO = os.path.join(ORACLE_DIR, "_ref", "libref_ibf.so")
REF_MAPPER_SO = os.path.join(ORACLE_DIR, "_ref", "libref_mapper.so")

MAX_HASH = 16


class IbfParams(C.Structure):
    _fields_ = [("bins", C.c_uint32), ("h", C.c_uint32), ("k", C.c_uint32), ("W", C.c_uint32),
                ("S", C.c_uint64), ("nbits", C.c_uint64), ("P", C.c_uint64), ("pre", C.c_uint64 * MAX_HASH)]


def build_oracle(force=False):
    src = os.path.join(ORACLE_DIR, "ibf_oracle.c")
    if force or not os.path.exists(ORACLE_SO) or os.path.getmtime(ORACLE_SO) < os.path.getmtime(src):
        subprocess.run(["make", "-C", ORACLE_DIR, "-B", "libibf_oracle.so"], check=True, capture_output=True)
    return ORACLE_SO


def _p(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


_V = C.c_void_p


class Oracle:
    """The C restatement. Filters are plain numpy uint64 word arrays (the sdsl word array incl. trailer)."""

    def __init__(self):
        self.l = C.CDLL(build_oracle())
        l = self.l
        l.ibf_params_init.argtypes = [C.POINTER(IbfParams), C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint64]
        l.ibf_nwords.restype = C.c_uint64; l.ibf_nwords.argtypes = [C.c_uint64]
        l.ibf_kmer_hashes.restype = C.c_uint64
        l.ibf_kmer_hashes.argtypes = [_V, C.c_uint64, C.c_uint32, _V]
        l.ibf_positions.argtypes = [C.POINTER(IbfParams), C.c_uint64, _V]
        l.ibf_add_kmers.argtypes = [_V, C.POINTER(IbfParams), _V, C.c_uint64, C.c_uint32]
        l.ibf_count_bins.argtypes = [_V, C.POINTER(IbfParams), _V, C.c_uint64, _V]
        l.ibf_which_bins.argtypes = [_V, C.POINTER(IbfParams), _V, C.c_uint64, C.c_uint16, _V]
        l.ibf_threshold.restype = C.c_uint16; l.ibf_threshold.argtypes = [C.c_int, C.c_uint32, C.c_uint16]
        l.ibf_classify.argtypes = [_V, C.POINTER(IbfParams), _V, _V, C.c_uint32, C.c_int, C.c_uint16, _V]
        l.ibf_compact.restype = C.c_uint64; l.ibf_compact.argtypes = [_V, C.c_uint32, C.c_uint32, _V, _V]
        l.ibf_clear_bins.argtypes = [_V, C.POINTER(IbfParams), _V, C.c_uint32]
        l.ibf_set_metadata.argtypes = [_V, C.POINTER(IbfParams)]
        l.ibf_save.argtypes = [_V, C.POINTER(IbfParams), C.c_char_p]
        l.ibf_file_nbits.restype = C.c_uint64; l.ibf_file_nbits.argtypes = [C.c_char_p]
        l.ibf_load.argtypes = [C.c_char_p, _V, C.c_uint64, C.POINTER(IbfParams)]
        l.ibf_encode_ascii.argtypes = [C.c_char_p, C.c_uint64, _V]
        l.ibf_revcomp.argtypes = [_V, C.c_uint64, _V]
        l.ibf_revcomp_dna5q.argtypes = [_V, C.c_uint64, _V]
        l.ibf_load_filtered_reads.restype = C.c_uint64
        l.ibf_load_filtered_reads.argtypes = [_V, _V, C.c_uint32, C.c_int, _V, C.c_uint32, _V, _V, _V]
        l.ibf_parse_records.restype = C.c_int64
        l.ibf_parse_records.argtypes = [C.c_char_p, C.c_uint64, C.c_int, C.c_int, _V, C.c_uint64, _V, C.c_uint64,
                                        C.POINTER(C.c_uint64)]

    def params(self, bins, h, k, nbits):
        p = IbfParams()
        if self.l.ibf_params_init(C.byref(p), bins, h, k, nbits) != 0:
            raise ValueError("bad filter parameters")
        return p

    def new_words(self, p):
        return np.zeros(self.l.ibf_nwords(p.nbits), dtype=np.uint64)

    def encode_ascii(self, text):
        if isinstance(text, str):
            text = text.encode("ascii")
        out = np.zeros(max(len(text), 1), dtype=np.uint8)
        self.l.ibf_encode_ascii(text, len(text), _p(out))
        return out[: len(text)]

    def revcomp(self, codes):
        codes = np.ascontiguousarray(codes, dtype=np.uint8)
        out = np.zeros(max(codes.size, 1), dtype=np.uint8)
        self.l.ibf_revcomp(_p(codes), codes.size, _p(out))
        return out[: codes.size]

    def kmer_hashes(self, codes, k):
        codes = np.ascontiguousarray(codes, dtype=np.uint8)
        n = max(codes.size - k + 1, 0)
        out = np.zeros(max(n, 1), dtype=np.uint64)
        got = self.l.ibf_kmer_hashes(_p(codes), codes.size, k, _p(out))
        return out[:got]

    def positions(self, p, kmer_hash):
        out = np.zeros(p.h, dtype=np.uint64)
        self.l.ibf_positions(C.byref(p), C.c_uint64(int(kmer_hash)), _p(out))
        return out

    def add_kmers(self, words, p, codes, bin_no):
        codes = np.ascontiguousarray(codes, dtype=np.uint8)
        self.l.ibf_add_kmers(_p(words), C.byref(p), _p(codes), codes.size, bin_no)

    def count_bins(self, words, p, codes):
        codes = np.ascontiguousarray(codes, dtype=np.uint8)
        counts = np.zeros(p.bins, dtype=np.uint16)
        self.l.ibf_count_bins(_p(words), C.byref(p), _p(codes), codes.size, _p(counts))
        return counts

    def which_bins(self, words, p, codes, threshold, selected=None):
        codes = np.ascontiguousarray(codes, dtype=np.uint8)
        sel = np.zeros(p.bins, dtype=np.uint8) if selected is None else selected
        self.l.ibf_which_bins(_p(words), C.byref(p), _p(codes), codes.size, threshold, _p(sel))
        return sel

    def threshold(self, e_percent, k, read_len):
        return self.l.ibf_threshold(e_percent, k, read_len)

    def classify(self, words, p, concat, offsets, paired, threshold):
        """concat/offsets: the FULL reference layout (incl. reverse complements). Returns rows [numReads x W]."""
        concat = np.ascontiguousarray(concat, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        n_seqs = offsets.size - 1
        n_rows = n_seqs // (4 if paired else 2)
        rows = np.zeros((n_rows, p.W), dtype=np.uint64)
        self.l.ibf_classify(_p(words), C.byref(p), _p(concat), _p(offsets), n_seqs, int(paired), threshold, _p(rows))
        return rows

    def compact(self, rows, bins):
        rows = np.ascontiguousarray(rows, dtype=np.uint64)
        n = rows.shape[0]
        offs = np.zeros(bins + 1, dtype=np.uint32)
        total = self.l.ibf_compact(_p(rows), n, bins, _p(offs), None)
        ids = np.zeros(max(total, 1), dtype=np.uint32)
        self.l.ibf_compact(_p(rows), n, bins, _p(offs), _p(ids))
        return offs.astype(np.uint64), ids[:total]

    def full_layout_dna5q(self, concat, offsets):
        """appendReverseComplement on raw Dna5Q bytes (quality kept, N stays 252): [fwd | rc] concat, offsets."""
        concat = np.ascontiguousarray(concat, dtype=np.uint8); offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        n = offsets.size - 1
        base = int(offsets[0]); total = int(offsets[-1]) - base
        cat = np.zeros(2 * total, dtype=np.uint8)
        cat[:total] = concat[base:base + total]
        off = np.zeros(2 * n + 1, dtype=np.uint64)
        off[: n + 1] = offsets - offsets[0]
        for i in range(n):
            a, b = int(off[i]), int(off[i + 1])
            out = np.zeros(max(b - a, 1), dtype=np.uint8)
            self.l.ibf_revcomp_dna5q(_p(np.ascontiguousarray(cat[a:b])), b - a, _p(out))
            cat[total + a: total + b] = out[: b - a]
            off[n + 1 + i] = total + b
        return cat, off

    def load_filtered_reads(self, full_concat, full_offsets, paired, ids):
        """loadFilteredReads over the FULL read set: (bytes, limits, extended ids)."""
        fc = np.ascontiguousarray(full_concat, dtype=np.uint8); fo = np.ascontiguousarray(full_offsets, dtype=np.uint64)
        ids = np.ascontiguousarray(ids, dtype=np.uint32)
        n_out = ids.size * (4 if paired else 2)
        longest = int(np.max(np.diff(fo.astype(np.int64)))) if fo.size > 1 else 0
        ob = np.zeros(max(n_out * longest, 1), dtype=np.uint8)
        ol = np.zeros(n_out + 1, dtype=np.uint64); oi = np.zeros(max(n_out, 1), dtype=np.uint32)
        n = self.l.ibf_load_filtered_reads(_p(fc), _p(fo), fo.size - 1, int(paired), _p(ids), ids.size, _p(ob), _p(ol), _p(oi))
        assert n == n_out
        return ob[: int(ol[n_out])], ol, oi[:n_out]

    def parse_records(self, text, fmt, iupac=False):
        """SeqAn's readRecords over FASTA (fmt 0) / FASTQ (fmt 1) text -> (sequence characters, limits), or None where
        the reference throws a ParseError / UnexpectedEnd."""
        text = bytes(text)
        seqs = np.zeros(max(len(text), 1), dtype=np.uint8)
        limits = np.zeros(len(text) + 2, dtype=np.uint64)
        nb = C.c_uint64(0)
        n = self.l.ibf_parse_records(text, len(text), fmt, int(iupac), _p(seqs), seqs.size, _p(limits), len(text) + 1, C.byref(nb))
        if n < 0:
            return None
        return seqs[: nb.value].copy(), limits[: n + 1].copy()

    # ---- kmer_direct filter (src/d_kdx_filter.h)
    def kdx_params(self, bins, k, nbits):
        l = self.l
        l.kdx_params_init.argtypes = [C.POINTER(IbfParams), C.c_uint32, C.c_uint32, C.c_uint64]
        l.kdx_add_kmers.argtypes = [_V, C.POINTER(IbfParams), _V, C.c_uint64, C.c_uint32]
        l.kdx_count_bins.argtypes = [_V, C.POINTER(IbfParams), _V, C.c_uint64, _V]
        l.kdx_which_bins.argtypes = [_V, C.POINTER(IbfParams), _V, C.c_uint64, C.c_uint16, _V]
        p = IbfParams()
        if l.kdx_params_init(C.byref(p), bins, k, nbits) != 0:
            raise ValueError("bad kmer_direct filter parameters")
        return p

    def kdx_add_kmers(self, words, p, codes, bin_no):
        codes = np.ascontiguousarray(codes, dtype=np.uint8)
        self.l.kdx_add_kmers(_p(words), C.byref(p), _p(codes), codes.size, bin_no)

    def kdx_count_bins(self, words, p, codes):
        codes = np.ascontiguousarray(codes, dtype=np.uint8)
        counts = np.zeros(p.bins, dtype=np.uint16)
        self.l.kdx_count_bins(_p(words), C.byref(p), _p(codes), codes.size, _p(counts))
        return counts

    def kdx_which_bins(self, words, p, codes, threshold, selected=None):
        codes = np.ascontiguousarray(codes, dtype=np.uint8)
        sel = np.zeros(p.bins, dtype=np.uint8) if selected is None else selected
        self.l.kdx_which_bins(_p(words), C.byref(p), _p(codes), codes.size, threshold, _p(sel))
        return sel

    def kdx_rows(self, words, p, reads_codes, threshold):
        """one row per read: whichBins(fwd) OR whichBins(revcomp), as clasifyLoadedReads combines them"""
        W = p.W
        rows = np.zeros((len(reads_codes), W), dtype=np.uint64)
        for i, c in enumerate(reads_codes):
            sel = self.kdx_which_bins(words, p, c, threshold)
            self.kdx_which_bins(words, p, self.revcomp(c), threshold, sel)
            bits = np.zeros(W * 64, dtype=np.uint8); bits[: p.bins] = sel
            rows[i] = np.packbits(bits, bitorder="little").view(np.uint64)
        return rows

    def clear_bins(self, words, p, bins):
        b = np.ascontiguousarray(bins, dtype=np.uint32)
        self.l.ibf_clear_bins(_p(words), C.byref(p), _p(b), b.size)

    def set_metadata(self, words, p):
        self.l.ibf_set_metadata(_p(words), C.byref(p))

    def save(self, words, p, path):
        if self.l.ibf_save(_p(words), C.byref(p), str(path).encode()) != 0:
            raise OSError("oracle save failed: %s" % path)

    def load(self, path):
        nbits = self.l.ibf_file_nbits(str(path).encode())
        if not nbits:
            raise OSError("not a filter file: %s" % path)
        words = np.zeros(self.l.ibf_nwords(nbits), dtype=np.uint64)
        p = IbfParams()
        if self.l.ibf_load(str(path).encode(), _p(words), words.size, C.byref(p)) != 0:
            raise OSError("oracle load failed: %s" % path)
        return words, p


def full_layout(concat, offsets, oracle, paired=False):
    """appendReverseComplement (store_seqs.h:271-290) on the host: forward set -> [fwd | rc] Dna5 codes."""
    offsets = np.asarray(offsets, dtype=np.uint64)
    n = offsets.size - 1
    parts = [concat[int(offsets[i]):int(offsets[i + 1])] for i in range(n)]
    parts += [oracle.revcomp(p) for p in parts]
    offs = np.zeros(2 * n + 1, dtype=np.uint64)
    offs[1:] = np.cumsum([len(p) for p in parts], dtype=np.uint64)
    cat = np.concatenate(parts) if parts else np.zeros(0, dtype=np.uint8)
    return cat, offs


class Ref:
    """The unmodified reference class behind oracle/ref_shim.cpp. `available()` says whether the .so exists."""

    @staticmethod
    def available():
        return os.path.exists(REF_SO)

    def __init__(self):
        self.l = C.CDLL(REF_SO)
        l = self.l
        l.ref_new.restype = _V; l.ref_new.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint64]
        l.ref_load.restype = _V; l.ref_load.argtypes = [C.c_char_p]
        l.ref_free.argtypes = [_V]
        l.ref_params.argtypes = [_V] + [C.POINTER(C.c_uint32)] * 3 + [C.POINTER(C.c_uint64), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64)]
        l.ref_words.restype = _V; l.ref_words.argtypes = [_V]
        l.ref_nwords.restype = C.c_uint64; l.ref_nwords.argtypes = [_V]
        l.ref_add_kmers.argtypes = [_V, _V, C.c_uint64, C.c_uint32]
        l.ref_add_fasta.argtypes = [_V, C.c_char_p, C.c_uint32]
        l.ref_clear_bins.argtypes = [_V, _V, C.c_uint32, C.c_uint32]
        l.ref_save.argtypes = [_V, C.c_char_p]
        l.ref_which_bins.argtypes = [_V, _V, C.c_uint64, C.c_uint16, _V]
        l.ref_kmer_hashes.restype = C.c_uint64; l.ref_kmer_hashes.argtypes = [_V, C.c_uint64, C.c_uint32, _V]
        l.ref_revcomp.argtypes = [_V, C.c_uint64, _V]
        l.ref_ascii_to_dna5.argtypes = [C.c_char_p, C.c_uint64, _V]
        l.ref_classify.argtypes = [_V, _V, _V, C.c_uint32, C.c_int, C.c_uint16, C.c_uint32, _V]
        if hasattr(l, "ref_read_records"):
            l.ref_read_records.restype = C.c_int64
            l.ref_read_records.argtypes = [C.c_char_p, C.c_int, _V, C.c_uint64, _V, C.c_uint64, C.POINTER(C.c_uint64)]

    def new(self, bins, h, k, nbits):
        return self.l.ref_new(bins, h, k, nbits)

    def load(self, path):
        return self.l.ref_load(str(path).encode())

    def free(self, f):
        self.l.ref_free(f)

    def words(self, f):
        n = self.l.ref_nwords(f)
        return np.ctypeslib.as_array(C.cast(self.l.ref_words(f), C.POINTER(C.c_uint64)), shape=(n,))

    def add_kmers(self, f, codes, bin_no):
        codes = np.ascontiguousarray(codes, dtype=np.uint8)
        self.l.ref_add_kmers(f, _p(codes), codes.size, bin_no)

    def add_fasta(self, f, path, bin_no):
        return self.l.ref_add_fasta(f, str(path).encode(), bin_no)

    def clear_bins(self, f, bins, threads=1):
        b = np.ascontiguousarray(bins, dtype=np.uint32)
        self.l.ref_clear_bins(f, _p(b), b.size, threads)

    def save(self, f, path):
        return self.l.ref_save(f, str(path).encode())

    def which_bins(self, f, codes, threshold, bins, selected=None):
        codes = np.ascontiguousarray(codes, dtype=np.uint8)
        sel = np.zeros(bins, dtype=np.uint8) if selected is None else selected
        self.l.ref_which_bins(f, _p(codes), codes.size, threshold, _p(sel))
        return sel

    def kmer_hashes(self, codes, k):
        codes = np.ascontiguousarray(codes, dtype=np.uint8)
        out = np.zeros(max(codes.size - k + 1, 1), dtype=np.uint64)
        n = self.l.ref_kmer_hashes(_p(codes), codes.size, k, _p(out))
        return out[:n]

    def revcomp(self, codes):
        codes = np.ascontiguousarray(codes, dtype=np.uint8)
        out = np.zeros(max(codes.size, 1), dtype=np.uint8)
        self.l.ref_revcomp(_p(codes), codes.size, _p(out))
        return out[: codes.size]

    def ascii_to_dna5(self, text):
        if isinstance(text, str):
            text = text.encode("ascii")
        out = np.zeros(max(len(text), 1), dtype=np.uint8)
        self.l.ref_ascii_to_dna5(text, len(text), _p(out))
        return out[: len(text)]

    def read_records(self, path, iupac=False):
        """SeqAn's own reader on a file (extension selects the format): (Dna5 codes, limits), or None where it throws."""
        size = os.path.getsize(path)
        codes = np.zeros(max(size, 1), dtype=np.uint8)
        limits = np.zeros(size + 2, dtype=np.uint64)
        nb = C.c_uint64(0)
        n = self.l.ref_read_records(str(path).encode(), int(iupac), _p(codes), codes.size, _p(limits), size + 1, C.byref(nb))
        if n == -3:
            raise OSError("cannot open %s" % path)
        if n < 0:
            return None
        return codes[: nb.value].copy(), limits[: n + 1].copy()

    def classify(self, f, concat, offsets, paired, threshold, threads, W):
        concat = np.ascontiguousarray(concat, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        n_seqs = offsets.size - 1
        rows = np.zeros((n_seqs // (4 if paired else 2), W), dtype=np.uint64)
        self.l.ref_classify(f, _p(concat), _p(offsets), n_seqs, int(paired), threshold, threads, _p(rows))
        return rows


class RefKdx:
    """The unmodified seqan::SeqAnKDXFilter behind oracle/ref_shim.cpp (ref_kdx_*)."""

    @staticmethod
    def available():
        return os.path.exists(REF_SO) and hasattr(C.CDLL(REF_SO), "ref_kdx_new")

    def __init__(self):
        self.l = C.CDLL(REF_SO)
        l = self.l
        l.ref_kdx_new.restype = _V; l.ref_kdx_new.argtypes = [C.c_uint32, C.c_uint32, C.c_uint64]
        l.ref_kdx_load.restype = _V; l.ref_kdx_load.argtypes = [C.c_char_p]
        l.ref_kdx_free.argtypes = [_V]
        l.ref_kdx_words.restype = _V; l.ref_kdx_words.argtypes = [_V]
        l.ref_kdx_nwords.restype = C.c_uint64; l.ref_kdx_nwords.argtypes = [_V]
        l.ref_kdx_params.argtypes = [_V, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        l.ref_kdx_add_kmers.argtypes = [_V, _V, C.c_uint64, C.c_uint32]
        l.ref_kdx_add_fasta.argtypes = [_V, C.c_char_p, C.c_uint32]
        l.ref_kdx_clear_bins.argtypes = [_V, _V, C.c_uint32, C.c_uint32]
        l.ref_kdx_save.argtypes = [_V, C.c_char_p]
        l.ref_kdx_which_bins.argtypes = [_V, _V, C.c_uint64, C.c_uint16, _V]

    def new(self, bins, k, nbits):
        return self.l.ref_kdx_new(bins, k, nbits)

    def load(self, path):
        return self.l.ref_kdx_load(str(path).encode())

    def free(self, f):
        self.l.ref_kdx_free(f)

    def words(self, f):
        n = self.l.ref_kdx_nwords(f)
        return np.ctypeslib.as_array(C.cast(self.l.ref_kdx_words(f), C.POINTER(C.c_uint64)), shape=(n,))

    def params(self, f):
        b, k, nb, hp = C.c_uint32(), C.c_uint32(), C.c_uint64(), C.c_uint64()
        self.l.ref_kdx_params(f, C.byref(b), C.byref(k), C.byref(nb), C.byref(hp))
        return b.value, k.value, nb.value, hp.value

    def add_kmers(self, f, codes, bin_no):
        codes = np.ascontiguousarray(codes, dtype=np.uint8)
        self.l.ref_kdx_add_kmers(f, _p(codes), codes.size, bin_no)

    def add_fasta(self, f, path, bin_no):
        return self.l.ref_kdx_add_fasta(f, str(path).encode(), bin_no)

    def clear_bins(self, f, bins, threads=1):
        b = np.ascontiguousarray(bins, dtype=np.uint32)
        self.l.ref_kdx_clear_bins(f, _p(b), b.size, threads)

    def save(self, f, path):
        return self.l.ref_kdx_save(f, str(path).encode())

    def which_bins(self, f, codes, threshold, bins, selected=None):
        codes = np.ascontiguousarray(codes, dtype=np.uint8)
        sel = np.zeros(bins, dtype=np.uint8) if selected is None else selected
        self.l.ref_kdx_which_bins(f, _p(codes), codes.size, threshold, _p(sel))
        return sel
