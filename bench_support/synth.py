"""ctypes wrapper of bench_support/synth.c: deterministic synthetic binned genomes and reads."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "libsynth.so")
_l = None


def _lib():
    global _l
    if _l is None:
        src = os.path.join(HERE, "synth.c")
        if not os.path.exists(SO) or os.path.getmtime(SO) < os.path.getmtime(src):
            subprocess.run(["make", "-C", HERE, "-B"], check=True, capture_output=True)
        _l = C.CDLL(SO)
        _l.synth_genomes.argtypes = [C.c_uint32, C.c_uint64, C.c_void_p]
        _l.synth_genome.argtypes = [C.c_uint32, C.c_uint64, C.c_void_p]
        _l.synth_reads.argtypes = [C.c_void_p, C.c_uint32, C.c_uint64, C.c_uint32, C.c_uint64, C.c_uint64, C.c_uint32,
                                   C.c_void_p, C.c_void_p]
        _l.synth_pairs.argtypes = [C.c_void_p, C.c_uint32, C.c_uint64, C.c_uint32, C.c_uint64, C.c_uint64, C.c_uint32,
                                   C.c_void_p, C.c_void_p]
    return _l


def genomes(bins, length, out=None, first_bin=0):
    """uint8 [bins, length] Dna5 codes of bins first_bin .. first_bin+bins-1."""
    if out is None:
        out = np.empty((bins, length), dtype=np.uint8)
    l = _lib()
    for b in range(bins):
        l.synth_genome(first_bin + b, length, out[b].ctypes.data)
    return out


def reads(genomes_arr, cfg, first, n, L, out=None, origin=False):
    """uint8 [n, L] Dna5 codes of reads first .. first+n-1 of configuration cfg (forward strand as sequenced)."""
    bins, glen = genomes_arr.shape
    if out is None:
        out = np.empty((n, L), dtype=np.uint8)
    org = np.empty(n, dtype=np.uint32) if origin else None
    _lib().synth_reads(genomes_arr.ctypes.data, bins, glen, cfg, first, n, L, out.ctypes.data,
                       org.ctypes.data if origin else None)
    return (out, org) if origin else out


def pairs(genomes_arr, cfg, first, n, L, out=None):
    """uint8 [2, n, L]: mates 1 and mates 2 of pairs first .. first+n-1 (the [m1 | m2] forward read set of the mapper)."""
    bins, glen = genomes_arr.shape
    if out is None:
        out = np.empty((2, n, L), dtype=np.uint8)
    _lib().synth_pairs(genomes_arr.ctypes.data, bins, glen, cfg, first, n, L, out[0].ctypes.data, out[1].ctypes.data)
    return out
