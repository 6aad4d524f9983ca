#!/usr/bin/env python
"""bench.py — IBF bin-filter throughput (BASELINE.json metric) on N B200s of one node.

Workload (BASELINE.json configs[1], "cfg2"): synthetic 1024-bin, 16 GiB IBF (k=18, h=3) built from 1024 x 4 Mbp
synthetic genomes; 10 M single-end 150 bp reads per GPU at e=3 (threshold 61).  One "step" = the whole hot path
over one batch of 10 M reads: pack (2-bit + reverse complement on the device) + IBF query kernel -> per-read bin
bitmasks.  N > 1 is read-sharded (filter replicated by an NCCL broadcast over NVLink, every rank gets its own
10 M reads: weak scaling, no data-path collective).

    python bench.py --gpus 1 --steps 5 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
           bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # the reference's own CPU whichBins path on the host cores

Prints ONE JSON line (rank 0).  `value`: reads/s with the read bytes resident in HBM (CUDA events, max over
ranks).  `e2e`: the same through the public host-buffer call (pinned host memory in, bitmask rows out, copies
inside the timed region).  `roofline`: the query kernel against the measured HBM peak.  `cpu_baseline`: the
reference's own code (oracle/_ref) on a bounded sample of the same reads, all host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

K, H, E = 18, 3, 3
SECTOR = 32


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--bins", type=int, default=1024)
    ap.add_argument("--filter-gib", type=int, default=16)
    ap.add_argument("--genome-len", type=int, default=4_000_000)
    ap.add_argument("--reads", type=int, default=10_000_000, help="reads per GPU per step")
    ap.add_argument("--read-len", type=int, default=150)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def algorithmic_bytes_per_read(bins, read_len):
    """per k-mer lookup h * 32*ceil(W*8/32) sector bytes; 2 strands per single-end read."""
    W = (bins + 63) // 64
    per_lookup = H * SECTOR * ((W * 8 + SECTOR - 1) // SECTOR)
    return (read_len - K + 1) * 2 * per_lookup


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.tmp = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=self.tmp, stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        self.tmp.flush(); self.tmp.seek(0)
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.tmp.read().splitlines():
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 8:
                continue
            try:
                sm.append(float(parts[1])); mx.append(float(parts[2])); power.append(float(parts[3]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        os.unlink(self.tmp.name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "power_w_max": float(max(power)), "samples": len(sm)}


def full_layout_codes(fwd):
    """appendReverseComplement on a [n, L] Dna5 code array -> concat [fwd | rc], offsets."""
    rc = fwd[:, ::-1]
    rc = np.where(rc >= 4, 4, 3 - rc).astype(np.uint8)
    n, L = fwd.shape
    cat = np.concatenate([fwd.reshape(-1), rc.reshape(-1)])
    off = np.arange(2 * n + 1, dtype=np.uint64) * L
    return cat, off
