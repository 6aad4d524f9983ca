"""Golden vectors for the record reader: what the UNMODIFIED reference's SeqAn reader
(readRecords as loadReads drives it / the readRecord loop of addFastaFile, through oracle/_ref/libref_ibf.so) returns for
the texts of tests/ingest_cases.py and for the reference's own test reads.  Run in the build container (needs
/root/reference and `make -C oracle ref`):   python tests/golden/make_golden_ingest.py
Writes tests/golden/ingest_vectors.npz: per case <name>_ok (0 = SeqAn threw), <name>_codes (Dna5 codes), <name>_limits."""
import hashlib
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from ingest_cases import cases            # noqa: E402
from oracle_py import Ref                 # noqa: E402

REFROOT = "/root/reference"


def main():
    ref = Ref()
    out = {}
    td = tempfile.mkdtemp()
    for name, fmt, iupac, text, _ in cases():
        if not text:                      # SeqFileIn cannot guess the format of an empty file; readRecords: no records
            out[name + "_ok"] = np.uint8(1); out[name + "_codes"] = np.zeros(0, np.uint8); out[name + "_limits"] = np.zeros(1, np.uint64)
            continue
        path = os.path.join(td, name + (".fq" if fmt else ".fa"))
        with open(path, "wb") as fh:
            fh.write(text)
        res = ref.read_records(path, iupac)
        out[name + "_ok"] = np.uint8(res is not None)
        if res is not None:
            out[name + "_codes"], out[name + "_limits"] = res
        out[name + "_text_md5"] = np.array(hashlib.md5(text).hexdigest())
    codes, limits = ref.read_records(os.path.join(REFROOT, "tests/input/64-viral-reads.fa"))
    out["viral_reads_md5"] = np.array(hashlib.md5(codes.tobytes()).hexdigest())
    out["viral_reads_limits_md5"] = np.array(hashlib.md5(limits.tobytes()).hexdigest())
    out["viral_reads_n"] = np.uint64(limits.size - 1)
    np.savez_compressed(os.path.join(HERE, "ingest_vectors.npz"), **out)
    print("wrote", len(out), "arrays")


if __name__ == "__main__":
    main()
