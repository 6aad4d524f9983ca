"""Host-side mirror of the mapper's filtering stage (reference: src/d_mapper.h).

    DisOptions.getThreshold      d_mapper.h:83-92
    clasifyLoadedReads           d_mapper.h:308-387
    readFilterMetadata           d_mapper.cpp:532-548
    loadFilteredReads            d_mapper.h:393-443
    sortedBins                   d_mapper.h:843-854

The read set has the layout loadReads + appendReverseComplement leave behind (src/mapper.h:444-469,
src/store_seqs.h:271-290): single-end [fwd | rc], paired-end [m1 | m2 | m1 rc | m2 rc].  `ReadSet` holds the
forward half only (rc=False) — the reverse complements are then derived on the GPU — or the full set (rc=True).
"""
import ctypes as C
import sys
import time

import numpy as np

from ._lib import ENC_DNA5, lib, check
from .ibf import SeqAnBloomFilter, as_batch


def readFilterMetadata(filterFile):
    """(numberOfBins, kmerSize) from the trailer of a .filter file, d_mapper.cpp:532-548."""
    bins, h, k, nbits = C.c_uint32(), C.c_uint32(), C.c_uint32(), C.c_uint64()
    check(lib().dyibf_peek_file(str(filterFile).encode(), C.byref(bins), C.byref(h), C.byref(k), C.byref(nbits)))
    return bins.value, k.value


class ReadSet:
    """reads.seqs of the main mapper: concatenated bases + limits (ConcatDirect StringSet, store_seqs.h:54-61)."""

    def __init__(self, seqs, paired=False, rc=False, encoding=None):
        self.concat, self.offsets, self.encoding = as_batch(seqs, encoding)
        self.paired, self.rc = bool(paired), bool(rc)
        n = self.offsets.size - 1
        per_read = (4 if paired else 2) if rc else (2 if paired else 1)
        if n % per_read:
            raise ValueError("%d sequences do not form a %s read set" % (n, "paired-end" if paired else "single-end"))

    def __len__(self):                              # length(reads.seqs) as the reference sees it (incl. rc)
        n = self.offsets.size - 1
        return n if self.rc else 2 * n

    def lengthSum(self):
        s = int(self.offsets[-1] - self.offsets[0])
        return s if self.rc else 2 * s

    def getReadsCount(self):                        # bits_reads.h:57-62
        return len(self) // 2

    def getPairsCount(self):                        # bits_reads.h:64-69
        return len(self) // 4


def loadReads(readsFile, matesFile=None, device=0):
    """loadReads (src/mapper.h:444-469) for the filter's purposes: the forward read set of one or two FASTA / FASTQ
    (.gz) files, tokenised on the device (SeqAn's readRecords, ingest.py); paired-end files become [mates 1 | mates 2]
    as readRecords(records, pair) leaves them (src/file_pair.h:82-88).  The reverse complements the reference appends
    next (appendReverseComplement) are derived on the GPU by the query (ReadSet rc=False)."""
    from .ingest import readRecords
    from ._lib import ENC_ASCII
    cat, lim = readRecords(readsFile, device=device)
    if matesFile is not None:
        cat2, lim2 = readRecords(matesFile, device=device)
        if lim2.size != lim.size:
            raise ValueError("the two read files hold %d and %d records" % (lim.size - 1, lim2.size - 1))
        cat = np.concatenate([cat, cat2])
        lim = np.concatenate([lim, lim2[1:] + lim[-1]])
    if cat.size == 0:
        cat = np.zeros(1, dtype=np.uint8)
    return ReadSet((cat, lim), paired=matesFile is not None, rc=False, encoding=ENC_ASCII)


class DisOptions:
    """The fields of DisOptions/Options the filtering stage reads (d_mapper.h:49-94, mapper.h:49-131)."""

    def __init__(self, errorRate=0.05, kmerSize=20, numberOfBins=64, threadsCount=1, verbose=0):
        self.errorRate = np.float32(errorRate)      # `float errorRate` (mapper.h:68)
        self.kmerSize = int(kmerSize)
        self.numberOfBins = int(numberOfBins)
        self.threadsCount = int(threadsCount)
        self.verbose = int(verbose)
        self.origReadIdMap = []
        self.filterReads = 0.0
        self.loadFilter = 0.0

    def setErrorPercent(self, e):
        """-e <int percent>: errorRate = e / 100.0 stored as float (d_mapper.cpp:340-342)."""
        self.errorRate = np.float32(e / 100.0)

    def getThreshold(self, readLen):
        return int(lib().dyibf_threshold_rate(C.c_float(float(self.errorRate)), self.kmerSize, int(readLen) & 0xFFFF))


def clasifyLoadedReads(reads, filter, disOptions, stderr=None):
    """Fill disOptions.origReadIdMap (one ascending uint32 id list per bin) for the loaded read batch.

    Mirrors d_mapper.h:308-387 including its quirks: one threshold for the whole batch from the *average*
    read length (:313-315), and for threshold 0 every id is pushed to every bin *and then* the normal loop still
    runs, so ids appear twice (:320-333, no early return).
    """
    err = stderr if stderr is not None else sys.stderr
    t0 = time.perf_counter()
    numReads = reads.getReadsCount()
    avgReadLen = (reads.lengthSum() // (numReads * 2)) & 0xFFFF if numReads else 0
    threshold = disOptions.getThreshold(avgReadLen)
    bins = disOptions.numberOfBins
    pre = None
    if threshold == 0:
        err.write("[WARNING!] 0 k-mer is required to filter a read!\n"
                  "All reads will pass filteration and be mapped everywhere.\n "
                  "This will be extremly slow.\n "
                  "Choose an approprate error rate based on kmer size and read length\n")
        pre = np.arange(numReads, dtype=np.uint32)
    if reads.paired:
        numReads = reads.getPairsCount()
    group = 4 if reads.paired else 2
    offs, ids = filter.classify((reads.concat, reads.offsets), threshold, group=group, derive_rc=not reads.rc,
                                encoding=reads.encoding)
    disOptions.binOffsets = offs
    disOptions.origReadIdMap = []
    for b in range(bins):
        lst = ids[int(offs[b]):int(offs[b + 1])]
        disOptions.origReadIdMap.append(np.concatenate([pre, lst]) if pre is not None else lst.copy())
    disOptions.filterReads += time.perf_counter() - t0
    if disOptions.verbose > 1:
        for b in range(bins):
            err.write("bin %d\t%d\n" % (b, len(disOptions.origReadIdMap[b])))
    return threshold


def sortedBins(disOptions):
    """Bin numbers by descending number of reads (d_mapper.h:843-854)."""
    sizes = np.array([len(x) for x in disOptions.origReadIdMap], dtype=np.uint64)
    offs = np.zeros(sizes.size + 1, dtype=np.uint64)
    offs[1:] = np.cumsum(sizes)
    return SeqAnBloomFilter.sortedBins(offs)


def loadFilteredReads(filter, disOptions, binNo):
    """The read set the mapper of bin `binNo` receives (d_mapper.h:393-443), built on the device from the batch the last
    clasifyLoadedReads left resident: returns ReadSet-style (bytes, limits) and extends disOptions.origReadIdMap[binNo]
    exactly as the reference does (mate ids, then reverse-strand ids appended).  Not valid after threshold == 0
    batches (the reference's double-push quirk makes the list differ from the filter's)."""
    t0 = time.perf_counter()
    seq_bytes, limits, ids = filter.gatherReads(binNo)
    disOptions.origReadIdMap[binNo] = ids
    disOptions.copyReads = getattr(disOptions, "copyReads", 0.0) + time.perf_counter() - t0
    return seq_bytes, limits
