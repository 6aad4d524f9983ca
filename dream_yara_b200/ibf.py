"""Host-side mirror of seqan::SeqAnBloomFilter (reference: src/d_bloom_filter.h:41-331) over the C ABI.

Same member names, argument meaning and error behaviour as the reference class; every method forwards to
libdyibf.so (CUDA, sm_100a).  Batched entries (`whichBinsBatch`, `addSequences`, `compact`, `classify`) are
what a patched clasifyLoadedReads / build_filter call; the per-sequence `whichBins` / `addKmers` stay for API
parity and are batches of one.
"""
import ctypes as C
import gzip
import sys

import numpy as np

from . import _lib
from ._lib import ENC_ASCII, ENC_DNA5, ENC_DNA5Q, MEM_DEVICE, MEM_HOST, check, lib

filterMetadataSize = 256   # d_misc_options.h:40
INT_WIDTH = 0x40           # d_misc_options.h:41


def _ptr(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def as_batch(seqs, encoding=None):
    """Normalise sequence input to (concat uint8, offsets uint64, encoding).

    Accepts a str/bytes (one ASCII sequence), a 1-D uint8 array (one sequence of Dna5 codes), a list of those,
    or an already concatenated (concat, offsets) pair.
    """
    if isinstance(seqs, tuple) and len(seqs) == 2:
        concat = np.ascontiguousarray(seqs[0], dtype=np.uint8)
        offsets = np.ascontiguousarray(seqs[1], dtype=np.uint64)
        return concat, offsets, ENC_DNA5 if encoding is None else encoding
    if isinstance(seqs, (str, bytes, np.ndarray)):
        seqs = [seqs]
    parts, enc = [], encoding
    for s in seqs:
        if isinstance(s, str):
            s = s.encode("ascii")
        if isinstance(s, (bytes, bytearray)):
            this = ENC_ASCII
            arr = np.frombuffer(bytes(s), dtype=np.uint8)
        else:
            this = ENC_DNA5
            arr = np.ascontiguousarray(s, dtype=np.uint8)
        if enc is None:
            enc = this
        elif encoding is None and enc != this:
            raise ValueError("mixed ASCII and code sequences in one batch")
        parts.append(arr)
    offsets = np.zeros(len(parts) + 1, dtype=np.uint64)
    if parts:
        offsets[1:] = np.cumsum([len(p) for p in parts], dtype=np.uint64)
        concat = np.concatenate(parts) if len(parts) > 1 else np.ascontiguousarray(parts[0])
    else:
        concat = np.zeros(0, dtype=np.uint8)
    if concat.size == 0:
        concat = np.zeros(1, dtype=np.uint8)   # keep a valid address
    return concat, offsets, ENC_DNA5 if enc is None else enc


def read_fasta(path, iupac=True, device=0):
    """Records of a FASTA/FASTQ(.gz) file as ASCII byte strings (what SeqFileIn::readRecord yields); tokenised on the
    device (ingest.readRecords)."""
    from .ingest import readRecords
    cat, lim = readRecords(path, iupac=iupac, device=device)
    return [cat[int(lim[i]):int(lim[i + 1])].tobytes() for i in range(lim.size - 1)]


class SeqAnBloomFilter:
    """IBF resident in the HBM of one B200.  Constructors as in the reference:

        SeqAnBloomFilter(n_bins, n_hash_func, kmer_size, vec_size)             d_bloom_filter.h:48
        SeqAnBloomFilter(fileName, n_bins, n_hash_func, kmer_size, vec_size)   d_bloom_filter.h:59
        SeqAnBloomFilter(fileName)                                             d_bloom_filter.h:79
    """

    def __init__(self, *args, device=0, _handle=None, _keep=None):
        self._h = C.c_void_p()
        self._keep = _keep                 # object owning externally attached device memory
        L = lib()
        if _handle is not None:
            self._h = _handle
        elif len(args) == 4:
            n_bins, n_hash, kmer, vec_size = args
            check(L.dyibf_create(n_bins, n_hash, kmer, vec_size, device, C.byref(self._h)))
        elif len(args) == 1:
            rc = L.dyibf_load_file(str(args[0]).encode(), device, C.byref(self._h))
            self._exit_like_reference(rc)
        elif len(args) == 5:
            fileName, n_bins, n_hash, kmer, vec_size = args
            rc = L.dyibf_load_file_expect(str(fileName).encode(), n_bins, n_hash, kmer, vec_size, device, C.byref(self._h))
            self._exit_like_reference(rc)
        else:
            raise TypeError("SeqAnBloomFilter(n_bins, n_hash_func, kmer_size, vec_size) | (fileName) | "
                            "(fileName, n_bins, n_hash_func, kmer_size, vec_size)")
        p = _lib.Params()
        check(L.dyibf_get_params(self._h, C.byref(p)))
        self._noOfBins, self._noOfHashFunc, self._kmerSize = p.bins, p.n_hash, p.kmer
        self._binIntWidth, self._noOfBits, self._noOfHashPos = p.words_per_block, p.nbits, p.n_blocks
        self._blockBitSize = p.words_per_block * 64
        self.n_words, self.device = p.n_words, p.device

    @staticmethod
    def _exit_like_reference(rc):
        # the reference prints to std::cerr and calls exit(1) on load failure / size mismatch (:66-76,:81-85)
        if rc in (_lib.ERR_IO, _lib.ERR_SIZE_MISMATCH, _lib.ERR_FORMAT):
            sys.stderr.write(lib().dyibf_last_error().decode("utf-8", "replace") + "\n")
            raise SystemExit(1)
        check(rc)

    @classmethod
    def attach(cls, n_bins, n_hash, kmer, vec_size, device_ptr, device=0, keep=None):
        """Filter whose P*W data words already sit in device memory owned by `keep` (e.g. a torch tensor)."""
        h = C.c_void_p()
        check(lib().dyibf_attach(n_bins, n_hash, kmer, vec_size, device, C.c_void_p(device_ptr), C.byref(h)))
        return cls(device=device, _handle=h, _keep=keep)

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            if getattr(self, "_owned", True):                    # a FilterGroup member view does not own its handle
                lib().dyibf_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    # ---- reference surface ------------------------------------------------------------------------------
    def getNumberOfBins(self):
        return self._noOfBins

    def getKmerSize(self):
        return self._kmerSize

    def size_mb(self):
        return lib().dyibf_size_mb(self._h)

    def addKmers(self, text, binNo):
        """addKmers(text, binNo), :91-94. The reference hashes len<k texts out of bounds; here they add nothing."""
        self.addSequences([text], [binNo])

    def addFastaFile(self, fastaFile, binNo):
        """addFastaFile, :127-147: every record of the file (len >= k) goes to bin `binNo`."""
        from .ingest import readRecords
        try:
            cat, lim = readRecords(fastaFile, iupac=True, device=self.device)
        except OSError:
            raise RuntimeError("Unable to open contigs File: %s" % fastaFile)   # reference: throw const char*
        n = lim.size - 1
        if n:                                                # records shorter than k are skipped by the build (:142)
            self.addSequences((cat if cat.size else np.zeros(1, np.uint8), lim), np.full(n, binNo, dtype=np.uint32), encoding=ENC_ASCII)

    def clearBins(self, bins2clear, threadsCount=1):
        """clearBins, :99-123 (threadsCount is accepted and ignored: the GPU does the pass)."""
        b = np.ascontiguousarray(bins2clear, dtype=np.uint32)
        check(lib().dyibf_clear_bins(self._h, _ptr(b), b.size))

    def save(self, fileName):
        """save, :150-154 -> bool"""
        return lib().dyibf_save_file(self._h, str(fileName).encode()) == _lib.OK

    def whichBins(self, text, threshold, selected=None):
        """whichBins(selected, text, threshold) :161-224 OR-accumulates into `selected`;
        whichBins(text, threshold) :226-231 returns a fresh vector."""
        rows = self.whichBinsBatch([text], threshold, group=1)
        bits = np.unpackbits(rows.view(np.uint8), bitorder="little")[: self._noOfBins].astype(bool)
        if selected is None:
            return bits
        sel = np.asarray(selected)
        sel |= bits
        return selected

    # ---- batched entries --------------------------------------------------------------------------------
    def addSequences(self, seqs, bin_ids, encoding=None):
        concat, offsets, enc = as_batch(seqs, encoding)
        bins = np.ascontiguousarray(bin_ids, dtype=np.uint32)
        n = offsets.size - 1
        if bins.size != n:
            raise ValueError("bin_ids must have one entry per sequence")
        check(lib().dyibf_add_sequences(self._h, _ptr(concat), _ptr(offsets), n, _ptr(bins), enc, MEM_HOST, None))

    def whichBinsBatch(self, seqs, threshold, group=1, derive_rc=False, counts=False, encoding=None, max_seq_len=0):
        """Rows [n_rows x W] u64 (bit b of word w = bin 64w+b); with counts=True also u16 [n_total x bins]."""
        concat, offsets, enc = as_batch(seqs, encoding)
        n = offsets.size - 1
        n_total = 2 * n if derive_rc else n
        if n_total % group:
            raise ValueError("%d sequences do not divide into groups of %d" % (n_total, group))
        rows = np.zeros((n_total // group, self._binIntWidth), dtype=np.uint64)
        cnt = np.zeros((n_total, self._noOfBins), dtype=np.uint16) if counts else None
        check(lib().dyibf_query_batch(self._h, _ptr(concat), _ptr(offsets), n, enc, group, int(bool(derive_rc)),
                                      int(threshold), int(max_seq_len), _ptr(rows), _ptr(cnt), MEM_HOST, None))
        return (rows, cnt) if counts else rows

    def compact(self, rows):
        """Per-bin ascending row ids (CSR) from selection rows: (bin_offsets u64[bins+1], read_ids u32[total])."""
        rows = np.ascontiguousarray(rows, dtype=np.uint64).reshape(-1, self._binIntWidth)
        offs = np.zeros(self._noOfBins + 1, dtype=np.uint64)
        total = C.c_uint64(0)
        L = lib()
        rc = L.dyibf_compact(self._h, _ptr(rows), rows.shape[0], _ptr(offs), None, 0, C.byref(total), MEM_HOST, None)
        if rc not in (_lib.OK, _lib.ERR_CAPACITY):
            check(rc)
        ids = np.zeros(max(int(total.value), 1), dtype=np.uint32)
        if total.value:
            check(L.dyibf_compact(self._h, _ptr(rows), rows.shape[0], _ptr(offs), _ptr(ids), total.value,
                                  C.byref(total), MEM_HOST, None))
        return offs, ids[: total.value]

    def classify(self, seqs, threshold, group=2, derive_rc=True, encoding=None, capacity=None):
        """Query + compaction in one call (rows stay in HBM): (bin_offsets, read_ids)."""
        concat, offsets, enc = as_batch(seqs, encoding)
        n = offsets.size - 1
        n_rows = (2 * n if derive_rc else n) // group
        offs = np.zeros(self._noOfBins + 1, dtype=np.uint64)
        total = C.c_uint64(0)
        cap = int(capacity) if capacity is not None else max(4 * n_rows, 1024)
        L = lib()
        ids = np.zeros(cap, dtype=np.uint32)
        rc = L.dyibf_classify(self._h, _ptr(concat), _ptr(offsets), n, enc, group, int(bool(derive_rc)), int(threshold),
                              _ptr(offs), _ptr(ids), cap, C.byref(total))
        if rc == _lib.ERR_CAPACITY and capacity is None:          # lists are still in HBM: fetch, don't re-run
            cap = int(total.value)
            ids = np.zeros(cap, dtype=np.uint32)
            rc = L.dyibf_classify_fetch(self._h, _ptr(offs), _ptr(ids), cap, C.byref(total))
        check(rc)
        return offs, ids[: total.value]

    def classifyBegin(self, seqs, threshold, group=2, derive_rc=True, encoding=None):
        """dyibf_classify_begin: enqueue one batch and return a ticket at once (up to four in flight per filter).  The ticket
        keeps the host arrays alive until classifyEnd."""
        concat, offsets, enc = as_batch(seqs, encoding)
        n = offsets.size - 1
        t = C.c_void_p()
        check(self._begin(_ptr(concat), _ptr(offsets), n, enc, group, int(bool(derive_rc)), int(threshold), C.byref(t)))
        return ClassifyTicket(t, (concat, offsets), (2 * n if derive_rc else n) // group, self._noOfBins)

    def _begin(self, *a):
        return lib().dyibf_classify_begin(self._h, *a)

    def classifyEnd(self, ticket, capacity=None):
        """dyibf_classify_end: wait for the ticket, (bin_offsets, read_ids) exactly as classify returns them."""
        offs = np.zeros(ticket.bins + 1, dtype=np.uint64)
        total = C.c_uint64(0)
        cap = int(capacity) if capacity is not None else max(4 * ticket.n_rows, 1024)
        ids = np.zeros(cap, dtype=np.uint32)
        L = lib()
        rc = L.dyibf_classify_end(ticket.handle, _ptr(offs), _ptr(ids), cap, C.byref(total))
        if rc == _lib.ERR_CAPACITY and capacity is None:          # the ticket stays in flight: again, with room
            cap = int(total.value)
            ids = np.zeros(cap, dtype=np.uint32)
            rc = L.dyibf_classify_end(ticket.handle, _ptr(offs), _ptr(ids), cap, C.byref(total))
        check(rc)
        ticket.keep = None
        return offs, ids[: total.value]

    def launch_count(self):
        return int(lib().dyibf_launch_count(self._h))

    def device_errors(self, reset=True):
        fl = C.c_uint32(0)
        check(lib().dyibf_device_errors(self._h, C.byref(fl), int(reset)))
        return int(fl.value)

    def addFastaTexts(self, texts, bin_ids, formats=None):
        """addFastaFile for many files at once, from their text (bytes each; FASTA unless formats says 1 = FASTQ): tokenised on
        the device with SeqAn's rules, all records added in one build call.  Raises DyibfError(ERR_FORMAT) where SeqAn throws;
        the exception's `bad_index` is the offending text."""
        n = len(texts)
        keep = [np.frombuffer(bytes(t), dtype=np.uint8) for t in texts]
        ptrs = (C.c_void_p * n)(*[C.c_void_p(k.ctypes.data) if k.size else None for k in keep])
        sizes = np.array([k.size for k in keep], dtype=np.uint64)
        bins = np.ascontiguousarray(bin_ids, dtype=np.uint32)
        fm = None if formats is None else np.ascontiguousarray(formats, dtype=np.int32)
        bad = C.c_uint32(0)
        rc = lib().dyibf_add_fasta_texts(self._h, C.cast(ptrs, C.c_void_p), _ptr(sizes), _ptr(bins), None if fm is None else _ptr(fm), n, C.byref(bad))
        if rc:
            try:
                check(rc)
            except _lib.DyibfError as e:
                e.bad_index = bad.value
                raise

    def classifyText(self, reads_text, mates_text=None, fmt=1, e_percent=5, capacity=None):
        """loadReads + clasifyLoadedReads from the text of the reads file(s) (bytes; fmt 0 FASTA / 1 FASTQ): only text goes up,
        only the lists come back.  Returns (bin_offsets, read_ids, n_reads, threshold)."""
        t1 = np.frombuffer(bytes(reads_text), dtype=np.uint8)
        t2 = np.frombuffer(bytes(mates_text), dtype=np.uint8) if mates_text is not None else None
        offs = np.zeros(self._noOfBins + 1, dtype=np.uint64)
        total, n_reads, thr = C.c_uint64(0), C.c_uint32(0), C.c_uint32(0)
        cap = int(capacity) if capacity is not None else max(len(t1) // 16, 1024)
        ids = np.zeros(cap, dtype=np.uint32)
        L = lib()
        p2 = None if t2 is None else C.c_void_p(t2.ctypes.data if t2.size else np.zeros(1, np.uint8).ctypes.data)
        rc = L.dyibf_classify_text(self._h, C.c_void_p(t1.ctypes.data) if t1.size else None, t1.size, p2, 0 if t2 is None else t2.size,
                                   int(fmt), int(e_percent), _ptr(offs), _ptr(ids), cap, C.byref(total), C.byref(n_reads), C.byref(thr))
        if rc == _lib.ERR_CAPACITY and capacity is None:
            cap = int(total.value)
            ids = np.zeros(cap, dtype=np.uint32)
            rc = L.dyibf_classify_fetch(self._h, _ptr(offs), _ptr(ids), cap, C.byref(total))
        check(rc)
        return offs, ids[: total.value], n_reads.value, thr.value

    def gatherReads(self, binNo):
        """loadFilteredReads (d_mapper.h:393-443) for one bin of the last classify(derive_rc=True):
        (bytes u8, limits u64[n+1], ids u32[n]) — fwd reads, [second mates,] then all reverse complements."""
        L = lib()
        ns, nb = C.c_uint64(0), C.c_uint64(0)
        rc = L.dyibf_gather_reads(self._h, int(binNo), None, 0, None, None, 0, C.byref(ns), C.byref(nb))
        if rc not in (_lib.OK, _lib.ERR_CAPACITY):
            check(rc)
        out = np.zeros(max(nb.value, 1), dtype=np.uint8)
        lim = np.zeros(ns.value + 1, dtype=np.uint64)
        ids = np.zeros(max(ns.value, 1), dtype=np.uint32)
        if ns.value:
            check(L.dyibf_gather_reads(self._h, int(binNo), _ptr(out), nb.value, _ptr(lim), _ptr(ids), ns.value,
                                       C.byref(ns), C.byref(nb)))
        return out[: nb.value], lim, ids[: ns.value]

    @staticmethod
    def sortedBins(bin_offsets):
        """sortedBins (d_mapper.h:843-854): bins by descending list size."""
        offs = np.ascontiguousarray(bin_offsets, dtype=np.uint64)
        out = np.zeros(offs.size - 1, dtype=np.uint32)
        check(lib().dyibf_sorted_bins(_ptr(offs), offs.size - 1, _ptr(out)))
        return out

    # ---- raw words / instrumentation ---------------------------------------------------------------------
    def download_words(self, first=0, n=None):
        n = self.n_words - first if n is None else n
        out = np.zeros(n, dtype=np.uint64)
        check(lib().dyibf_download_words(self._h, _ptr(out), first, n))
        return out

    def upload_words(self, words, first=0):
        w = np.ascontiguousarray(words, dtype=np.uint64)
        check(lib().dyibf_upload_words(self._h, _ptr(w), first, w.size))

    def device_words(self):
        return lib().dyibf_device_words(self._h)

    def set_early_exit(self, on=True):
        """DYIBF_OPT_EARLY_EXIT: stop a sequence once its bin set is decided (rows identical, fewer lookups)."""
        check(lib().dyibf_set_option(self._h, _lib.OPT_EARLY_EXIT, int(bool(on))))

    def set_build_bucket_mib(self, mib):
        """DYIBF_OPT_BUILD_BUCKET_MIB: scratch size of the bucketed build (-1 auto, 0 = direct atomics only)."""
        check(lib().dyibf_set_option(self._h, _lib.OPT_BUILD_BUCKET_MIB, int(mib)))

    def set_build_mode(self, mode):
        """DYIBF_OPT_BUILD_MODE: 0 auto, 1 direct atomics, 2 bucketed (radix split + apply), 3 sliced (per-bin bit vectors in the L2)."""
        check(lib().dyibf_set_option(self._h, _lib.OPT_BUILD_MODE, int(mode)))

    def last_build_mode(self):
        v = C.c_int64(0)
        check(lib().dyibf_get_option(self._h, _lib.OPT_LAST_BUILD_MODE, C.byref(v)))
        return int(v.value)

    def set_build_chunk_mib(self, mib):
        """DYIBF_OPT_BUILD_CHUNK_MIB: bases per staged chunk of addSequences on host buffers."""
        check(lib().dyibf_set_option(self._h, _lib.OPT_BUILD_CHUNK_MIB, int(mib)))

    def set_query_chunk_mib(self, mib):
        """DYIBF_OPT_QUERY_CHUNK_MIB: staged bytes per chunk of the host-buffer query (0 = auto)."""
        check(lib().dyibf_set_option(self._h, _lib.OPT_QUERY_CHUNK_MIB, int(mib)))

    def profile_enable(self, on=True):
        check(lib().dyibf_profile_enable(self._h, int(on)))

    def profile(self, reset=False):
        p = _lib.Profile()
        check(lib().dyibf_profile_get(self._h, C.byref(p), int(reset)))
        return {k: getattr(p, k) for k, _ in p._fields_}

    # device-pointer entry used by bench.py / multi-GPU code (torch tensors' data_ptr())
    def query_device(self, seqs_ptr, offsets_ptr, n_seqs, threshold, rows_ptr, group=2, derive_rc=True,
                     encoding=ENC_DNA5, max_seq_len=0, counts_ptr=None, stream=None):
        check(lib().dyibf_query_batch(self._h, C.c_void_p(seqs_ptr), C.c_void_p(offsets_ptr), n_seqs, encoding, group,
                                      int(bool(derive_rc)), int(threshold), int(max_seq_len), C.c_void_p(rows_ptr),
                                      C.c_void_p(counts_ptr) if counts_ptr else None, MEM_DEVICE,
                                      C.c_void_p(stream) if stream else None))


class ClassifyTicket:
    """One asynchronous classify in flight (dyibf_ticket)."""

    def __init__(self, handle, keep, n_rows, bins):
        self.handle, self.keep, self.n_rows, self.bins = handle, keep, n_rows, bins

    def ready(self):
        r = C.c_int(0)
        check(lib().dyibf_classify_poll(self.handle, C.byref(r)))
        return bool(r.value)


class FilterGroup:
    """One filter over several GPUs of ONE process (dyibf_group, include/dyibf.h): mode "read" = a replica per device, the
    rows of every batch split over the devices; mode "bin" = device r owns the bins dyibf_shard_bins(r) of a filter larger
    than one GPU's HBM.  What spawnDisMapper's single filter object (src/d_mapper.h:911-922) becomes on an 8-GPU box.

        FilterGroup(n_bins, n_hash_func, kmer_size, vec_size, devices=[0, 1], mode="read")
        FilterGroup(fileName, devices=[...], mode=...)
        FilterGroup(fileName, n_bins, n_hash_func, kmer_size, vec_size, devices=[...], mode=...)   the caller's parameters win
    """

    def __init__(self, *args, devices=(0,), mode="read"):
        L = lib()
        self._g = C.c_void_p()
        devs = (C.c_int * len(devices))(*[int(d) for d in devices])
        m = {"read": _lib.GROUP_READ_SHARDED, "bin": _lib.GROUP_BIN_SHARDED}[mode]
        if len(args) == 4:
            check(L.dyibf_group_create(args[0], args[1], args[2], args[3], devs, len(devices), m, C.byref(self._g)))
        elif len(args) == 1:
            rc = L.dyibf_group_load_file(str(args[0]).encode(), devs, len(devices), m, C.byref(self._g))
            SeqAnBloomFilter._exit_like_reference(rc)
        elif len(args) == 5:                                      # (fileName, n_bins, n_hash_func, kmer_size, vec_size), :59-77
            rc = L.dyibf_group_load_file_expect(str(args[0]).encode(), args[1], args[2], args[3], args[4], devs, len(devices), m,
                                                C.byref(self._g))
            SeqAnBloomFilter._exit_like_reference(rc)
        else:
            raise TypeError("FilterGroup(n_bins, n_hash_func, kmer_size, vec_size) | (fileName) | (fileName, n_bins, n_hash_func, kmer_size, vec_size)")
        p = _lib.Params(); nd, md = C.c_int(0), C.c_int(0)
        check(L.dyibf_group_get_params(self._g, C.byref(p), C.byref(nd), C.byref(md)))
        self._noOfBins, self._noOfHashFunc, self._kmerSize = p.bins, p.n_hash, p.kmer
        self._binIntWidth, self._noOfBits, self._noOfHashPos = p.words_per_block, p.nbits, p.n_blocks
        self.n_words, self.n_devices, self.mode = p.n_words, nd.value, mode
        self.devices = list(devices)

    def close(self):
        if getattr(self, "_g", None) is not None and self._g:
            lib().dyibf_group_destroy(self._g)
            self._g = C.c_void_p()

    __del__ = close

    def getNumberOfBins(self):
        return self._noOfBins

    def getKmerSize(self):
        return self._kmerSize

    def member(self, i):
        """(SeqAnBloomFilter view of device i's handle — not owning —, first_bin, n_bins)"""
        h = C.c_void_p(); fb, nb = C.c_uint32(0), C.c_uint32(0)
        check(lib().dyibf_group_member(self._g, int(i), C.byref(h), C.byref(fb), C.byref(nb)))
        f = SeqAnBloomFilter(_handle=h)
        f._owned = False                                          # the group owns the handle
        f._owner = self
        return f, fb.value, nb.value

    def addSequences(self, seqs, bin_ids, encoding=None):
        concat, offsets, enc = as_batch(seqs, encoding)
        bins = np.ascontiguousarray(bin_ids, dtype=np.uint32)
        n = offsets.size - 1
        if bins.size != n:
            raise ValueError("bin_ids must have one entry per sequence")
        check(lib().dyibf_group_add_sequences(self._g, _ptr(concat), _ptr(offsets), n, _ptr(bins), enc))

    def clearBins(self, bins2clear, threadsCount=1):
        b = np.ascontiguousarray(bins2clear, dtype=np.uint32)
        check(lib().dyibf_group_clear_bins(self._g, _ptr(b), b.size))

    def save(self, fileName):
        return lib().dyibf_group_save_file(self._g, str(fileName).encode()) == _lib.OK

    def broadcast(self, src_member=0):
        check(lib().dyibf_broadcast_filter(self._g, int(src_member)))

    def set_early_exit(self, on=True):
        check(lib().dyibf_group_set_option(self._g, _lib.OPT_EARLY_EXIT, int(bool(on))))

    def classify(self, seqs, threshold, group=2, derive_rc=True, encoding=None, capacity=None):
        """clasifyLoadedReads over all devices of the group: (bin_offsets, read_ids), identical to one device's."""
        concat, offsets, enc = as_batch(seqs, encoding)
        n = offsets.size - 1
        n_rows = (2 * n if derive_rc else n) // group
        offs = np.zeros(self._noOfBins + 1, dtype=np.uint64)
        total = C.c_uint64(0)
        cap = int(capacity) if capacity is not None else max(4 * n_rows, 1024)
        ids = np.zeros(cap, dtype=np.uint32)
        L = lib()
        rc = L.dyibf_group_classify(self._g, _ptr(concat), _ptr(offsets), n, enc, group, int(bool(derive_rc)), int(threshold),
                                    _ptr(offs), _ptr(ids), cap, C.byref(total))
        if rc == _lib.ERR_CAPACITY and capacity is None:
            cap = int(total.value)
            ids = np.zeros(cap, dtype=np.uint32)
            rc = L.dyibf_group_classify_fetch(self._g, _ptr(offs), _ptr(ids), cap, C.byref(total))
        check(rc)
        return offs, ids[: total.value]

    def classifyBegin(self, seqs, threshold, group=2, derive_rc=True, encoding=None):
        """whole batches to successive replicas (read-sharded groups); end with classifyEnd"""
        concat, offsets, enc = as_batch(seqs, encoding)
        n = offsets.size - 1
        t = C.c_void_p()
        check(lib().dyibf_group_classify_begin(self._g, _ptr(concat), _ptr(offsets), n, enc, group, int(bool(derive_rc)),
                                               int(threshold), C.byref(t)))
        return ClassifyTicket(t, (concat, offsets), (2 * n if derive_rc else n) // group, self._noOfBins)

    classifyEnd = SeqAnBloomFilter.classifyEnd


class SeqAnKDXFilter(SeqAnBloomFilter):
    """The kmer_direct filter (`-ft kmer_direct`; reference: seqan::SeqAnKDXFilter, src/d_kdx_filter.h:41-311): same layout,
    file format and members as the IBF, but the block of a k-mer is its value — no hashing.  Constructors as in the reference:

        SeqAnKDXFilter(n_bins, kmer_size, vec_size)     d_kdx_filter.h:49-56   (vec_size = 4^k * n_bins + 256 in build_filter)
        SeqAnKDXFilter(fileName)                        d_kdx_filter.h:58-67

    On the C ABI it is a filter with zero hash functions.  Where the reference indexes out of bounds (a k-mer value that is
    not a block of the table: sequences that start with N, tables smaller than 4^k blocks; clearBins walking 64 x the
    blocks there are) this class skips the k-mer / covers the real blocks."""

    def __init__(self, *args, device=0):
        if len(args) == 3:
            n_bins, kmer, vec_size = args
            super().__init__(n_bins, 0, kmer, vec_size, device=device)
        elif len(args) == 1:
            super().__init__(args[0], device=device)
            if self._noOfHashFunc != 0:
                raise ValueError("%s is a bloom filter file (%d hash functions), not kmer_direct" % (args[0], self._noOfHashFunc))
        else:
            raise TypeError("SeqAnKDXFilter(n_bins, kmer_size, vec_size) | (fileName)")

    @staticmethod
    def vectorSize(n_bins, kmer_size):
        """vec_size as dream_yara_build_filter computes it, d_build_filter.cpp:307-309 (u32 shift: k <= 15)"""
        return ((1 << (2 * kmer_size)) & 0xFFFFFFFF) * n_bins + filterMetadataSize
