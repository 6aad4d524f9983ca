"""Device record reader (dyibf_parse_records) on a synthetic FASTQ / FASTA text resident in HBM: time per call, text GB/s.
    python tools/ingest_bench.py [--reads N] [--read-len L] [--fasta] [--calls K]
Run under `ncu --metrics gpu__time_duration.sum` for the per-kernel launch list."""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reads", type=int, default=2_000_000)
    ap.add_argument("--read-len", type=int, default=150)
    ap.add_argument("--fasta", action="store_true")
    ap.add_argument("--calls", type=int, default=5)
    a = ap.parse_args()
    import torch
    from dream_yara_b200.ingest import parseRecordsDevice
    n, L = a.reads, a.read_len
    rng = np.random.default_rng(1)
    codes = rng.integers(0, 4, (n, L), dtype=np.uint8)
    seq = np.frombuffer(b"ACGT", dtype=np.uint8)[codes]
    if a.fasta:
        rec = np.empty((n, L + 4), dtype=np.uint8)
        rec[:, 0:3] = np.frombuffer(b">r\n", dtype=np.uint8); rec[:, 3:3 + L] = seq; rec[:, 3 + L] = 10
    else:
        rec = np.empty((n, 2 * L + 7), dtype=np.uint8)
        rec[:, 0:3] = np.frombuffer(b"@r\n", dtype=np.uint8); rec[:, 3:3 + L] = seq
        rec[:, 3 + L:6 + L] = np.frombuffer(b"\n+\n", dtype=np.uint8); rec[:, 6 + L:6 + 2 * L] = ord("I"); rec[:, 6 + 2 * L] = 10
    d_text = torch.from_numpy(rec.reshape(-1)).cuda()
    d_seq = torch.empty(n * L + 16, dtype=torch.uint8, device="cuda")
    d_lim = torch.empty(n + 1, dtype=torch.int64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    call = lambda: parseRecordsDevice(d_text.data_ptr(), d_text.numel(), 0 if a.fasta else 1, d_seq.data_ptr(), d_seq.numel(),
                                      d_lim.data_ptr(), n, stream=st)
    call()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(a.calls):
        got = call()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.calls
    assert got == (n, n * L)
    assert bool(torch.equal(d_seq[:n * L], torch.from_numpy(seq.reshape(-1)).cuda()))
    print("format %s reads %d text %.1f MB: %.3f ms per call, %.1f GB/s of text, %.1f M reads/s"
          % ("fasta" if a.fasta else "fastq", n, d_text.numel() / 1e6, ms, d_text.numel() / ms / 1e6, n / ms / 1e3))


if __name__ == "__main__":
    main()
