"""Shared helpers for the known-answer configurations KA1..KA5."""
import hashlib

import numpy as np

GIB_BITS = (1 << 33) + 256


def records(golden, what):
    cat, off = golden[what + "_ascii"], golden[what + "_offsets"]
    return [cat[int(off[i]):int(off[i + 1])].tobytes() for i in range(off.size - 1)]


def ka4_bins(genomes):
    out = []
    for b in range(130):
        seq = genomes[b % 64]
        if 64 <= b < 128:
            seq = seq[::-1]
        if b == 128:
            seq = b"N" + seq[1:4000]
        if b == 129:
            seq = b"NN" + (b"N" + seq[1:4000])[2:3000]
        out.append([seq[:3000], seq[3000:], b"ACGTACGT"] if b % 10 == 3 else [seq])
    return out


def md5_file(path):
    h = hashlib.md5()
    with open(path, "rb") as fh:
        for blk in iter(lambda: fh.read(1 << 24), b""):
            h.update(blk)
    return h.hexdigest()


def md5_filter_words(words, nbits):
    """md5 of the .filter file a word array serialises to (8-byte sdsl header + words)."""
    h = hashlib.md5()
    h.update(np.array([(1 << 56) | nbits], dtype=np.uint64).tobytes())
    h.update(np.ascontiguousarray(words).tobytes())
    return h.hexdigest()


def flat(records_per_bin):
    """[(record bytes, bin)] for every record with its bin."""
    seqs, bins = [], []
    for b, recs in enumerate(records_per_bin):
        for r in recs:
            seqs.append(r); bins.append(b)
    return seqs, bins


def bits_from_rows(rows, bins):
    return np.unpackbits(np.ascontiguousarray(rows).view(np.uint8), axis=1, bitorder="little")[:, :bins]
