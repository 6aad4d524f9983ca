#!/usr/bin/env python
"""bench.py — IBF bin-filter throughput (BASELINE.json metric) on N B200s of one node.

Workload (BASELINE.json configs[1], "cfg2"): synthetic 1024-bin, 16 GiB IBF (k=18, h=3) built from 1024 x 4 Mbp
synthetic genomes; 10 M single-end 150 bp reads per GPU at e=3 (threshold 61).  One "step" = the whole hot path
over one batch of 10 M reads: pack (2-bit + reverse complement on the device) + IBF query kernel -> per-read bin
bitmasks + compaction of the bitmasks into per-bin read-id lists.  N > 1 is read-sharded (filter replicated by an NCCL broadcast over NVLink, every rank gets its own
10 M reads: weak scaling, no data-path collective).

    python bench.py --gpus 1 --steps 5 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
           bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # the reference's own CPU whichBins path on the host cores

Prints ONE JSON line (rank 0).  `value`: reads/s with the read bytes resident in HBM (CUDA events, max over
ranks).  `e2e`: the same through the reference-facing host-buffer call dyibf_classify (pinned host reads in, per-bin
read-id lists out, copies inside the timed region); `e2e.rows_call` times dyibf_query_batch (bitmask rows out) beside it.  `roofline`: the query kernel against the measured HBM peak.  `cpu_baseline`: the
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
    ap.add_argument("--workload", default="cfg2", choices=["cfg2", "cfg3", "cfg4", "cfg5"],
                    help="BASELINE.json configs[1] (default, the headline), configs[2] (paired-end, e sweep), configs[3] "
                         "(bin-sharded, 1024 bins and a 64 GiB shard per GPU) or configs[4] (IBF build: one step = addKmers "
                         "over 1024 bins x --genome-len bases, successive steps add new contigs to the same filter)")
    ap.add_argument("--bins", type=int, default=None, help="bins (cfg4: per GPU); default 1024")
    ap.add_argument("--filter-gib", type=int, default=None, help="filter size in GiB (cfg4: per GPU shard); default 16 / 64")
    ap.add_argument("--genome-len", type=int, default=4_000_000)
    ap.add_argument("--reads", type=int, default=10_000_000,
                    help="reads (cfg3: pairs) per GPU per step; cfg4: reads per step, seen by every GPU")
    ap.add_argument("--read-len", type=int, default=None)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-compact", action="store_true", help="leave the per-bin list compaction out of the timed step")
    ap.add_argument("--no-dropin", action="store_true", help="skip the extra 100 000-reads-per-call dyibf_classify measurement")
    ap.add_argument("--dropin-batch", type=int, default=100_000, help="reads (pairs) per drop-in call; the reference's -rb default")
    ap.add_argument("--no-ingest", action="store_true", help="skip the extra device FASTQ-reader measurement (cfg2)")
    ap.add_argument("--no-early-exit", action="store_true", help="skip the extra DYIBF_OPT_EARLY_EXIT measurement")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the extras that ride in the cfg2 line: filter load/save, cfg3 sweep, cfg5 build (N=1); multi-GPU "
                         "parity, cfg4 bin-sharded, one-process group (N>1)")
    ap.add_argument("--extras", default="all", help="comma list out of io,cfg3,cfg5,parity,cfg4,single,async (default all)")
    a = ap.parse_args()
    a.paired = a.workload == "cfg3"
    a.bin_sharded = a.workload == "cfg4"
    a.bins = a.bins or 1024
    a.filter_gib = a.filter_gib or (64 if a.bin_sharded else 16)
    a.read_len = a.read_len or (100 if a.bin_sharded else 150)
    a.unit = "pairs/s" if a.paired else ("k-mers/s" if a.workload == "cfg5" else "reads/s")
    a.metric = {"cfg5": "IBF build k-mers/s (addKmers, 1024 bins, k=18, h=3)",
                "cfg2": "IBF bin-filter reads/s (150bp,e=3)", "cfg3": "IBF bin-filter pairs/s (2x150bp,e=3)",
                "cfg4": "IBF bin-filter reads/s (100bp,e=3), bin-sharded"}[a.workload]
    return a


def algorithmic_bytes_per_read(bins, read_len, paired=False):
    """per k-mer lookup h * 32*ceil(W*8/32) sector bytes; 2 strands per read, 4 sequences per pair.
    `bins` = the bins resident on ONE GPU (cfg4: the shard's)."""
    W = (bins + 63) // 64
    per_lookup = H * SECTOR * ((W * 8 + SECTOR - 1) // SECTOR)
    return (read_len - K + 1) * (4 if paired else 2) * per_lookup


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
    """appendReverseComplement on a [n, L] (single-end) or [2, n, L] (paired-end: [m1 | m2]) Dna5 code array ->
    concat [fwd | rc], offsets."""
    fwd = fwd.reshape(-1, fwd.shape[-1])
    rc = fwd[:, ::-1]
    rc = np.where(rc >= 4, 4, 3 - rc).astype(np.uint8)
    n, L = fwd.shape
    cat = np.concatenate([fwd.reshape(-1), rc.reshape(-1)])
    off = np.arange(2 * n + 1, dtype=np.uint64) * L
    return cat, off


# ---------------------------------------------------------------------------------------------------------------
# reference arm: the reference's own CPU code on the host cores (no GPU code on this path)
# ---------------------------------------------------------------------------------------------------------------
class CpuFilter:
    """The reference's SeqAnBloomFilter compiled from /root/reference (oracle/_ref/libref_ibf.so, kind
    "reference"), or the C restatement oracle/ibf_oracle.c (kind "port") where that file is absent."""

    def __init__(self, bins, h, k, nbits):
        from oracle_py import Oracle, Ref
        self.bins, self.h, self.k, self.nbits = bins, h, k, nbits
        self.W = (bins + 63) // 64
        self.cores = os.cpu_count() or 1
        if Ref.available():
            self.kind, self.ref = "reference", Ref()
            self.f = self.ref.new(bins, h, k, nbits)
            self.words = self.ref.words(self.f)
        else:
            self.kind, self.o = "port", Oracle()
            self.p = self.o.params(bins, h, k, nbits)
            self.words = self.o.new_words(self.p)

    def build(self, genomes):
        """build_filter (d_build_filter.cpp:217-272): bins over threads, addKmers per bin; a thread owns whole 64-bin
        columns so that the reference's non-atomic `word |= mask` never races (see bench_support/build_bench.py)."""
        from bench_support.build_bench import _cpu_build
        _cpu_build(self, genomes, np.arange(genomes.shape[0], dtype=np.uint32))

    def classify(self, layout, threshold, paired=False):
        """the filtering loop of clasifyLoadedReads over a prepared [fwd | rc] read set (full_layout_codes)"""
        cat, off = layout
        if self.kind == "reference":
            return self.ref.classify(self.f, cat, off, paired, threshold, self.cores, self.W)
        os.environ.setdefault("OMP_NUM_THREADS", str(self.cores))
        return self.o.classify(self.words, self.p, cat, off, paired, threshold)

    def timed_sample(self, reads_fn, threshold, target_s, paired=False):
        """Time the filtering loop on a prefix of the workload sized for ~target_s seconds."""
        n = 20_000
        fwd = reads_fn(n); lay = full_layout_codes(fwd)
        t = time.perf_counter(); rows = self.classify(lay, threshold, paired); dt = time.perf_counter() - t
        rate = n / dt
        n2 = int(min(max(rate * target_s, n), 2_000_000))
        if n2 > 2 * n:
            fwd = reads_fn(n2); lay = full_layout_codes(fwd)
            t = time.perf_counter(); rows = self.classify(lay, threshold, paired); dt = time.perf_counter() - t
            n = n2
        return n / dt, n, dt, rows, fwd


def synth_fwd(synth, genomes, args, first, n, out=None):
    """forward read set of the workload: [n, L] reads (cfg2/cfg4) or [2, n, L] = [m1 | m2] pairs (cfg3)"""
    if args.paired:
        return synth.pairs(genomes, 3, first, n, args.read_len, out=out)
    return synth.reads(genomes, 4 if args.bin_sharded else 2, first, n, args.read_len, out=out)


def run_reference(args):
    """The reference's own CPU path (oracle/_ref) on the host cores, same config/metric as the B200 arm; cfg4: one
    GPU's 1024-bin shard (the whole 8192-bin table does not fit host RAM; the per-shard work is what a GPU does)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from bench_support import synth
    nbits = args.filter_gib * (1 << 33) + 256
    cf = CpuFilter(args.bins, H, K, nbits)
    from oracle_py import Oracle
    thr = Oracle().threshold(E, K, args.read_len)
    t0 = time.perf_counter()
    genomes = synth.genomes(args.bins, args.genome_len)
    cf.build(genomes)
    build_s = time.perf_counter() - t0
    # size a step for ~4 s of CPU work from a probe, then W warm-up + K timed steps on successive samples
    rate, n_probe, _, _, _ = cf.timed_sample(lambda n: synth_fwd(synth, genomes, args, 0, n), thr, 4.0, args.paired)
    n_step = int(min(max(rate * 4.0, 20_000), 2_000_000))
    times = []
    for s in range(args.warmup + args.steps):
        lay = full_layout_codes(synth_fwd(synth, genomes, args, (s % 4) * n_step, n_step))
        t = time.perf_counter(); cf.classify(lay, thr, args.paired); dt = time.perf_counter() - t
        if s >= args.warmup:
            times.append(dt)
    total = sum(times)
    value = n_step * args.steps / total
    what = "pair" if args.paired else "read"
    line = {
        "impl": "reference", "metric": args.metric, "value": value, "unit": args.unit,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": workload_config(args, thr, args.gpus),
        "cpu_baseline": {"value": value, "unit": args.unit, "cores": cf.cores, "kind": cf.kind,
                         "sample": "%d-%s prefix samples of the workload per step; filter built on the host "
                                   "by the same code in %.1f s" % (n_step, what, build_s)},
        "e2e": {"value": value, "unit": args.unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(args, thr, world):
    per_unit = algorithmic_bytes_per_read(args.bins, args.read_len, args.paired)
    gather_gb = round(args.reads * per_unit / 1e9)
    if args.workload == "cfg2":
        desc = ("cfg2: synthetic %d-bin %d GiB IBF (k=%d, h=%d), %d single-end %d bp reads per GPU, e=%d (threshold %d), "
                "read-sharded" % (args.bins, args.filter_gib, K, H, args.reads, args.read_len, E, thr))
        par = "read-sharded x%d (filter replicated, NCCL broadcast)" % world
    elif args.workload == "cfg3":
        desc = ("cfg3: synthetic %d-bin %d GiB IBF (k=%d, h=%d), %d paired-end 2x%d bp pairs per GPU per step, e=%d "
                "(threshold %d) + e=0..5 sweep, read-sharded" % (args.bins, args.filter_gib, K, H, args.reads, args.read_len, E, thr))
        par = "read-sharded x%d (filter replicated, NCCL broadcast)" % world
    else:
        desc = ("cfg4: synthetic %d-bin %d GiB IBF (k=%d, h=%d) bin-sharded over %d GPU(s) (%d bins, %d GiB per GPU), "
                "%d single-end %d bp reads per step seen by every GPU, e=%d (threshold %d)"
                % (args.bins * world, args.filter_gib * world, K, H, world, args.bins, args.filter_gib, args.reads, args.read_len, E, thr))
        par = "bin-sharded x%d (each GPU owns a contiguous bin range of every hash block; reads all-gathered over NCCL; no reduction)" % world
    return {"workload": desc, "bins": args.bins * (world if args.bin_sharded else 1),
            "filter_gib": args.filter_gib * (world if args.bin_sharded else 1), "kmer": K, "hashes": H,
            "reads_per_gpu": args.reads, "read_len": args.read_len, "paired": args.paired, "threshold": thr,
            "genome_bp_per_bin": args.genome_len,
            "l2": "no flush needed: every step gathers ~%d GB at random from a %d GiB table per GPU (>> 126 MB L2)"
                  % (gather_gb, args.filter_gib),
            "parallelism": par}


# ---------------------------------------------------------------------------------------------------------------
# B200 arm
# ---------------------------------------------------------------------------------------------------------------
def run_b200(args):
    import ctypes as C
    import torch
    import torch.distributed as dist
    import dream_yara_b200 as dy
    from dream_yara_b200._lib import ENC_DNA5, MEM_DEVICE, MEM_HOST, check
    from bench_support import synth

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the IBF path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    L = dy.lib()
    bins, n, RL = args.bins, args.reads, args.read_len
    nbits = args.filter_gib * (1 << 33) + 256
    W = (bins + 63) // 64
    P = (nbits - 256) // (64 * W)
    thr = int(L.dyibf_threshold(E, K, RL))

    # ---- setup (untimed): genomes, filter, this rank's reads
    paired, sharded = args.paired, args.bin_sharded
    group = 4 if paired else 2
    n_seq = n * (2 if paired else 1)                                     # forward sequences per step ([m1 | m2] for pairs)
    units_world = 1 if sharded else world                                # cfg4: every GPU sees every read
    genomes = synth.genomes(bins, args.genome_len, first_bin=rank * bins if sharded else 0)
    words = torch.zeros(P * W, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()          # the fill runs on torch's stream; the library's streams are not ordered behind it
    f = dy.SeqAnBloomFilter.attach(bins, H, K, nbits, words.data_ptr(), device=local, keep=words)
    f.set_early_exit(False)            # headline = the reference's every-k-mer pass (fixed bytes per read); the early-exit default is an extra below
    t_build = time.perf_counter()
    if rank == 0 or sharded:                                             # cfg4: every rank builds its own bin shard
        goff = np.arange(bins + 1, dtype=np.uint64) * args.genome_len
        f.addSequences((genomes.reshape(-1), goff), np.arange(bins, dtype=np.uint32))
    torch.cuda.synchronize()
    t_build = time.perf_counter() - t_build
    replication = None
    if world > 1 and not sharded:                                        # read-sharded: replicate over NVLink
        step = 1 << 27                                                   # 1 GiB pieces
        dist.broadcast(words[:1024], src=0)                              # communicator set-up is not the transfer
        dist.barrier(); torch.cuda.synchronize()
        t_rep = time.perf_counter()
        for a in range(0, words.numel(), step):
            dist.broadcast(words[a:a + step], src=0)
        torch.cuda.synchronize()
        t_rep = time.perf_counter() - t_rep
        trep = torch.tensor([t_rep], dtype=torch.float64, device=dev)
        dist.all_reduce(trep, op=dist.ReduceOp.MAX)
        replication = {"seconds": float(trep.item()), "bytes": int(words.numel()) * 8, "gbs": words.numel() * 8 / float(trep.item()) / 1e9,
                       "how": "ncclBroadcast of the P*W data words from rank 0 in 1 GiB pieces into memory every rank adopts with "
                              "dyibf_attach (NVLink / NVSwitch); max over ranks, wall clock"}

    hp = C.c_void_p(); check(L.dyibf_host_alloc(C.byref(hp), n_seq * RL))
    h_reads = np.ctypeslib.as_array(C.cast(hp, C.POINTER(C.c_uint8)), shape=(2, n, RL) if paired else (n, RL))
    if sharded and world > 1:
        # every rank draws n/world reads from its own bins; one NCCL all-gather gives every GPU the whole batch
        per = n // world
        assert per * world == n, "--reads must be divisible by the number of GPUs for cfg4"
        mine = torch.from_numpy(synth_fwd(synth, genomes, args, rank * per, per)).to(dev)
        allr = torch.empty((n, RL), dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(allr, mine)
        h_reads[:] = allr.cpu().numpy()
        del mine, allr
    else:
        synth_fwd(synth, genomes, args, rank * n, n, out=h_reads)
    op = C.c_void_p(); check(L.dyibf_host_alloc(C.byref(op), (n_seq + 1) * 8))
    h_off = np.ctypeslib.as_array(C.cast(op, C.POINTER(C.c_uint64)), shape=(n_seq + 1,))
    h_off[:] = np.arange(n_seq + 1, dtype=np.uint64) * RL
    rp = C.c_void_p(); check(L.dyibf_host_alloc(C.byref(rp), n * W * 8))
    h_rows = np.ctypeslib.as_array(C.cast(rp, C.POINTER(C.c_uint64)), shape=(n, W))
    d_reads = torch.from_numpy(h_reads.reshape(-1)).to(dev)
    d_off = torch.from_numpy(h_off.view(np.int64)).to(dev)
    d_rows = torch.zeros(n * W, dtype=torch.int64, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    cur_thr = [thr]

    # north_star (c): the rows are compacted into per-bin read-id lists (origReadIdMap, d_mapper.h:360-368) inside the step
    ids_cap = 4 * n
    d_boff = torch.zeros(bins + 1, dtype=torch.int64, device=dev)
    d_ids = torch.zeros(ids_cap, dtype=torch.int32, device=dev)
    n_pairs = C.c_uint64(0)

    enabled = set(args.extras.split(",")) if args.extras != "all" else {"io", "cfg3", "cfg5", "parity", "cfg4", "single", "async"}

    def want(name):
        return not args.no_extras and name in enabled and args.workload == "cfg2"

    _cpu = {}

    def cpu_filter():
        """the reference's filter object on the host with the words of THIS GPU's table (one 16 GiB download, cached)"""
        if "cf" not in _cpu:
            cf_ = CpuFilter(bins, H, K, nbits)
            check(L.dyibf_download_words(f._h, C.c_void_p(cf_.words.ctypes.data), 0, P * W))
            _cpu["cf"] = cf_
        return _cpu["cf"]

    def mk_ctx():
        from types import SimpleNamespace
        return SimpleNamespace(L=L, dy=dy, f=f, words=words, dev=dev, local=local, rank=rank, world=world, args=args, genomes=genomes,
                               h_reads=h_reads, d_reads=d_reads, d_off=d_off, d_rows=d_rows, n=n, RL=RL, bins=bins, W=W, P=P, thr=thr,
                               nbits=nbits, stream=stream, barrier=barrier, torch=torch, dist=dist, units_world=units_world,
                               peak=measured_peak()[0], cpu_filter=cpu_filter)

    def mk_ctx_light():
        from types import SimpleNamespace
        return SimpleNamespace(L=L, dy=dy, dev=dev, local=local, rank=rank, world=world, args=args, stream=stream, barrier=barrier,
                               torch=torch, dist=dist, peak=measured_peak()[0])

    def device_step():
        f.query_device(d_reads.data_ptr(), d_off.data_ptr(), n_seq, cur_thr[0], d_rows.data_ptr(), group=group, derive_rc=True,
                       encoding=ENC_DNA5, max_seq_len=RL, stream=stream)
        if not args.no_compact:
            rc = L.dyibf_compact(f._h, C.c_void_p(d_rows.data_ptr()), n, C.c_void_p(d_boff.data_ptr()), C.c_void_p(d_ids.data_ptr()),
                                 ids_cap, C.byref(n_pairs), MEM_DEVICE, C.c_void_p(stream))
            if rc not in (0, -6):                                      # -6: more pairs than ids_cap (tiny thresholds); offsets are complete
                check(rc)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- value: reads resident in HBM, CUDA events on the launching stream
    for _ in range(args.warmup):
        device_step()
    f.profile_enable(True); f.profile(reset=True)
    sampler = ClockSampler(local); 
    barrier()
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = f.launch_count()
    e0.record()
    for _ in range(args.steps):
        device_step()
    e1.record()
    barrier()
    gpu_launches = f.launch_count() - launches0                          # counted by the library at every kernel launch
    clocks = sampler.stop() if rank == 0 else None
    ms = e0.elapsed_time(e1)
    prof = f.profile(reset=True); f.profile_enable(False)
    t = torch.tensor([ms, prof["query_ms"] / max(prof["query_launches"], 1), prof["pack_ms"] / max(prof["pack_launches"], 1),
                      prof["compact_ms"] / args.steps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, query_ms, pack_ms, compact_ms = (float(x) for x in t.tolist())
    value = units_world * n * args.steps / (ms_total / 1e3)

    # ---- e2e: the reference-facing host-buffer call (clasifyLoadedReads = dyibf_classify): pinned reads in, per-bin read-id lists
    # out, copies inside the timed region.  The bitmask-rows call (dyibf_query_batch, 1.28 GB of rows back) is timed beside it.
    e2e = None
    ip = C.c_void_p()
    if not args.no_e2e:
        check(L.dyibf_host_alloc(C.byref(ip), ids_cap * 4))
        h_ids = np.ctypeslib.as_array(C.cast(ip, C.POINTER(C.c_uint32)), shape=(ids_cap,))
        h_boff = np.zeros(bins + 1, dtype=np.uint64)
        tot_h = C.c_uint64(0)
        def host_step():
            check(L.dyibf_query_batch(f._h, hp, op, n_seq, ENC_DNA5, group, 1, thr, RL, rp, None, MEM_HOST, None))
        def classify_step():
            check(L.dyibf_classify(f._h, hp, op, n_seq, ENC_DNA5, group, 1, thr, C.c_void_p(h_boff.ctypes.data), ip, ids_cap, C.byref(tot_h)))
        def host_timed(fn):
            for _ in range(max(1, min(args.warmup, 2))):
                fn()
            barrier()
            t0 = time.perf_counter()
            for _ in range(args.steps):
                fn()
            barrier()
            dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(dt, op=dist.ReduceOp.MAX)
            return float(dt.item())
        dt_rows = host_timed(host_step)
        dt_cls = host_timed(classify_step)
        h2d = world * (n_seq * RL + (n_seq + 1) * 8)
        e2e = {"value": units_world * n * args.steps / dt_cls, "unit": args.unit,
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": world * ((bins + 1) * 8 + int(tot_h.value) * 4),
               "timing": "host wall clock around K synchronous dyibf_classify(host) calls (reads in, per-bin read-id lists out), max over ranks",
               "rows_call": {"value": units_world * n * args.steps / dt_rows, "unit": args.unit, "h2d_bytes_per_step": h2d,
                             "d2h_bytes_per_step": world * n * W * 8, "call": "dyibf_query_batch(host): bitmask rows out"}}
        # device-resident and host paths must agree bit for bit
        assert np.array_equal(h_rows.reshape(-1).view(np.int64), d_rows.cpu().numpy()), "host-mode rows != device-mode rows"
        if not args.no_compact:
            tot = int(tot_h.value)
            assert tot == int(n_pairs.value) and np.array_equal(h_boff.view(np.int64), d_boff.cpu().numpy()) and \
                np.array_equal(h_ids[:tot].view(np.int32), d_ids[:tot].cpu().numpy()), "host-mode lists != device-mode lists"

    # ---- extra: drop-in batching: one dyibf_classify call per 100 000 reads, as clasifyLoadedReads is called
    # (src/mapper.h:109), pageable host buffers in, per-bin id lists out — what the patched reference mapper would see
    dropin = None
    if not args.no_dropin and not sharded:
        rb = min(args.dropin_batch, n)
        n_calls = max(1, min(n // rb, 20))
        per_call = rb * (2 if paired else 1)
        offs_b = np.arange(per_call + 1, dtype=np.uint64) * RL
        boff = np.zeros(bins + 1, dtype=np.uint64); ids_b = np.zeros(4 * rb, dtype=np.uint32); tot_b = C.c_uint64(0)
        def batch_of(i):                                                   # [reads] or [mates 1 | mates 2] of batch i, pageable
            if paired:
                return np.ascontiguousarray(h_reads[:, i * rb:(i + 1) * rb]).reshape(-1)
            return np.array(h_reads[i * rb:(i + 1) * rb]).reshape(-1)
        batches = [batch_of(i) for i in range(n_calls)]
        def dropin_call_on(srcs, i):
            rc = L.dyibf_classify(f._h, C.c_void_p(srcs[i].ctypes.data), C.c_void_p(offs_b.ctypes.data), per_call, ENC_DNA5, group, 1, thr,
                                  C.c_void_p(boff.ctypes.data), C.c_void_p(ids_b.ctypes.data), ids_b.size, C.byref(tot_b))
            check(rc)
        def dropin_run(srcs):
            dropin_call_on(srcs, 0); dropin_call_on(srcs, 0)
            barrier()
            t0 = time.perf_counter()
            pairs = 0
            for i in range(n_calls):
                dropin_call_on(srcs, i)
                pairs += tot_b.value
            dt_b = time.perf_counter() - t0
            barrier()
            dtt = torch.tensor([dt_b], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(dtt, op=dist.ReduceOp.MAX)
            return float(dtt.item()), pairs
        dt_auto, pairs_b = dropin_run(batches)
        # the lists of the last batch must be the rows of the device pass, transposed
        lastrows = d_rows.view(n, W)[(n_calls - 1) * rb:n_calls * rb].cpu().numpy().view(np.uint64)
        bits = np.unpackbits(lastrows.view(np.uint8), axis=1, bitorder="little")[:, :bins]
        assert int(bits.sum()) == int(boff[bins]) and np.array_equal(bits.sum(axis=0).astype(np.uint64), np.diff(boff)), \
            "drop-in lists differ from the device rows"
        f.profile_enable(True); f.profile(reset=True)                      # a separate, untimed pass: kernel time inside one call
        for i in range(n_calls):
            dropin_call_on(batches, i)
        pb = f.profile(reset=True); f.profile_enable(False)
        kern_ms = {k: pb[k + "_ms"] / n_calls for k in ("pack", "query", "compact")}
        f.set_query_chunk_mib(1024)                                        # A/B: the whole batch as one chunk (no overlap inside a call)
        dt_one, _ = dropin_run(batches)
        f.set_query_chunk_mib(0)
        dt_pin = None
        if not paired:                                                     # A/B: the same calls on pinned host memory
            dt_pin, _ = dropin_run([h_reads[i * rb:(i + 1) * rb].reshape(-1) for i in range(n_calls)])
        rate = lambda dt: units_world * rb * n_calls / dt
        dropin = {"value": rate(dt_auto), "unit": args.unit, "reads_per_call": rb, "calls": n_calls,
                  "ms_per_call": dt_auto / n_calls * 1e3, "kernel_ms_per_call": kern_ms,
                  "kernel_ms_note": "summed over the two stream slots of a call (their kernels overlap), so the sum may exceed ms_per_call",
                  "read_bin_pairs": int(pairs_b),
                  "one_chunk_per_call": rate(dt_one), "pinned_input": rate(dt_pin) if dt_pin else None,
                  "note": "dyibf_classify per batch of the reference's default size: pageable host reads in, per-bin read-id lists "
                          "out, synchronous calls, host wall clock, max over ranks; not part of value/e2e"}
        if want("async"):
            from bench_support import extras as X
            pinned_b = None if paired else [h_reads[i * rb:(i + 1) * rb].reshape(-1) for i in range(n_calls)]
            dropin["asynchronous"] = X.dropin_async(mk_ctx(), batches, pinned_b, per_call, rb, group)
            # the same tickets as a caller gets them by default (early exit on): identical lists, fewer lookups
            f.set_early_exit(True)
            try:
                ee = X.dropin_async(mk_ctx(), batches, pinned_b, per_call, rb, group)
            finally:
                f.set_early_exit(False)
            dropin["asynchronous_early_exit_default"] = {"value": ee["value"], "pinned_input": ee.get("pinned_input"), "unit": args.unit,
                                                         "lists_equal_synchronous": ee["lists_equal_synchronous"],
                                                         "note": "the same ticketed batches with the library's default DYIBF_OPT_EARLY_EXIT = 1"}

    # ---- extra (not the headline): the same step with DYIBF_OPT_EARLY_EXIT — identical rows, fewer lookups
    early = None
    if not args.no_early_exit:
        full_rows = d_rows.clone()
        full_lookups = prof["query_lookups"] / max(prof["query_launches"], 1)
        f.set_early_exit(True)
        for _ in range(2):
            device_step()
        f.profile_enable(True); f.profile(reset=True)
        barrier()
        e0.record()
        for _ in range(args.steps):
            device_step()
        e1.record()
        barrier()
        p2 = f.profile(reset=True); f.profile_enable(False)
        same = bool(torch.equal(full_rows, d_rows))
        t2 = torch.tensor([e0.elapsed_time(e1), p2["query_ms"] / max(p2["query_launches"], 1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t2, op=dist.ReduceOp.MAX)
        early = {"value": units_world * n * args.steps / (float(t2[0]) / 1e3), "unit": args.unit, "kernel_ms": float(t2[1]),
                 "lookups_frac": (p2["query_lookups"] / max(p2["query_launches"], 1)) / max(full_lookups, 1),
                 "rows_equal_full": same,
                 "roofline": {"bound": "hbm", "unit": "GB/s", "peak": measured_peak()[0],
                              "achieved": (p2["query_lookups"] / max(p2["query_launches"], 1)) * algorithmic_bytes_per_read(bins, RL, paired)
                                          / ((RL - K + 1) * (4 if paired else 2)) / (float(t2[1]) / 1e3) / 1e9,
                              "definition": "bytes touched in 32 B sectors (lookups performed x h x sector bytes per lookup) / kernel time"},
                 "note": "DYIBF_OPT_EARLY_EXIT=1, the library's default since round 2 (the mapper never asks for counts): a sequence stops "
                         "once no remaining k-mer can change its bin set; rows identical.  The headline value / e2e / roofline switch it "
                         "off (the reference's every-k-mer pass, fixed algorithmic bytes per read)"}
        early["roofline"]["frac"] = early["roofline"]["achieved"] / early["roofline"]["peak"]
        if not same:
            raise AssertionError("early-exit rows differ from the full pass")
        if not args.no_e2e:
            # e2e of the early-exit mode through the mapper-facing call (lists out, 36 MB back instead of 1.28 GB of rows)
            classify_step()
            barrier()
            t0 = time.perf_counter()
            for _ in range(args.steps):
                classify_step()
            barrier()
            dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(dt, op=dist.ReduceOp.MAX)
            early["e2e"] = units_world * n * args.steps / float(dt.item())
            early["e2e_over_value"] = early["e2e"] / early["value"]
            early["e2e_h2d_gbs_per_gpu"] = (n_seq * RL + (n_seq + 1) * 8) * args.steps / float(dt.item()) / 1e9
            early["e2e_call"] = "dyibf_classify(host): pinned reads in, per-bin lists out"
            tot = int(tot_h.value)
            assert tot == int(n_pairs.value) and np.array_equal(h_ids[:tot].view(np.int32), d_ids[:tot].cpu().numpy()), "early-exit host lists differ"
        f.set_early_exit(False)
        del full_rows

    # ---- extra: the step in front of the filter — FASTQ text of the first reads -> sequences, on the device
    ingest = None
    if args.workload == "cfg2" and not args.no_ingest:
        from dream_yara_b200.ingest import parseRecordsDevice
        n_ing = min(n, 2_000_000)
        rec = np.empty((n_ing, 2 * RL + 7), dtype=np.uint8)                  # "@r\n" seq "\n+\n" qual "\n"
        rec[:, 0:3] = np.frombuffer(b"@r\n", dtype=np.uint8)
        rec[:, 3:3 + RL] = np.frombuffer(b"ACGTN", dtype=np.uint8)[np.minimum(h_reads[:n_ing], 4)]
        rec[:, 3 + RL:6 + RL] = np.frombuffer(b"\n+\n", dtype=np.uint8)
        rec[:, 6 + RL:6 + 2 * RL] = ord("I")
        rec[:, 6 + 2 * RL] = ord("\n")
        d_text = torch.from_numpy(rec.reshape(-1)).to(dev)
        d_seq_out = torch.empty(n_ing * RL + 16, dtype=torch.uint8, device=dev)
        d_lim_out = torch.empty(n_ing + 1, dtype=torch.int64, device=dev)
        def ingest_step():
            return parseRecordsDevice(d_text.data_ptr(), d_text.numel(), 1, d_seq_out.data_ptr(), d_seq_out.numel(),
                                      d_lim_out.data_ptr(), n_ing, device=local, stream=stream)
        for _ in range(2):
            got = ingest_step()
        barrier()
        e0.record()
        for _ in range(args.steps):
            ingest_step()
        e1.record()
        barrier()
        ing_ms = e0.elapsed_time(e1) / args.steps
        lut = torch.full((256,), 4, dtype=torch.uint8, device=dev)
        for ch, v in ((b"Aa", 0), (b"Cc", 1), (b"Gg", 2), (b"Tt", 3)):
            for b_ in ch:
                lut[b_] = v
        same = got == (n_ing, n_ing * RL) and bool(torch.equal(lut[d_seq_out[:n_ing * RL].long()], d_reads[:n_ing * RL].clamp(max=4))) \
            and bool(torch.equal(d_lim_out, torch.arange(n_ing + 1, dtype=torch.int64, device=dev) * RL))
        if not same:
            raise AssertionError("device FASTQ reader output differs from the reads it was written from")
        ingest = {"value": n_ing / (ing_ms / 1e3), "unit": "reads/s", "text_gbs": d_text.numel() / (ing_ms / 1e3) / 1e9,
                  "ms": ing_ms, "reads": n_ing, "text_bytes": int(d_text.numel()), "output_equal_reads": True,
                  "note": "dyibf_parse_records (FASTQ, device-resident text of the first reads of the workload -> sequence "
                          "characters + limits; two reads of the text: summary + class counts, scans, emit); not part of value/e2e"}
        del d_text, d_seq_out, d_lim_out, rec

    # ---- roofline of the dominant kernel (query) against the measured HBM peak
    bytes_per_read = algorithmic_bytes_per_read(bins, RL, paired)
    peak, peak_src = measured_peak()
    achieved = n * bytes_per_read / (query_ms / 1e3) / 1e9
    vec = 16 if W % 2 == 0 else 8
    tl2 = 0
    while tl2 < 5 and (1 << tl2) < (W * 8) // vec:
        tl2 += 1
    planes = 8 if (RL - K + 1) <= 255 else 16
    kernel_name = "query_kernel<%d,%d,%d,%d>" % (vec, tl2, H if planes == 8 else 0, planes)       # launch_query's dispatch
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": None, "kernel": kernel_name, "kernel_ms": query_ms, "pack_kernel_ms": pack_ms, "compact_kernels_ms": compact_ms,
                "read_bin_pairs_per_step": int(n_pairs.value),
                "algorithmic_bytes_per_launch": n * bytes_per_read, "peak_source": peak_src,
                "frac_of_nominal_8000": achieved / 8000.0}
    traffic_file = os.path.join(ROOT, "profiles", "query_kernel_traffic.json")
    if os.path.exists(traffic_file) and args.workload == "cfg2" and bins == 1024 and RL == 150:
        try:
            tj = json.load(open(traffic_file))
            if tj.get("reads_per_launch") and tj.get("kernel") == kernel_name:
                roofline["traffic"] = tj["dram_bytes_per_launch"] * (n / tj["reads_per_launch"])
                roofline["traffic_source"] = "DERIVED, not measured in this run: " + str(tj.get("source")) + \
                    "; scaled by reads per launch; query_kernel.cu commit of the capture: " + str(tj.get("kernel_commit"))
        except Exception:
            pass

    # ---- CPU baseline: the reference's own code on a bounded sample of the same reads (rank 0, N = 1)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cf = cpu_filter()                                                  # same filter words as the GPU (downloaded from HBM)
        prefix = (lambda m: h_reads[:, :m]) if paired else (lambda m: h_reads[:m])
        if paired:
            # the GPU rows of a prefix of pairs need that prefix queried on its own ([m1 | m2] halves move with n)
            def gpu_rows(m):
                sub = np.ascontiguousarray(h_reads[:, :m]).reshape(-1)
                off = np.arange(2 * m + 1, dtype=np.uint64) * RL
                return f.whichBinsBatch((sub, off), thr, group=4, derive_rc=True, encoding=ENC_DNA5, max_seq_len=RL)
        rate, n_s, dt_s, rows_cpu, _ = cf.timed_sample(prefix, thr, 15.0, paired)
        if paired:
            rows_gpu = gpu_rows(n_s)
        else:
            rows_gpu = h_rows[:n_s] if e2e else d_rows.cpu().numpy().view(np.uint64).reshape(n, W)[:n_s]
        same = bool(np.array_equal(rows_cpu, rows_gpu))
        cpu = {"value": rate, "unit": args.unit, "cores": cf.cores, "kind": cf.kind,
               "sample": "first %d %s of the workload, %.1f s, same filter words (downloaded from HBM)"
                         % (n_s, "pairs" if paired else "reads", dt_s),
               "build": "oracle/Makefile: g++ -O3 -march=x86-64-v3 (built in the CPU container, run here; not -march=native)",
               "rows_equal_gpu": same}
        if not same:
            raise AssertionError("GPU rows differ from the CPU reference on the sampled prefix")

    # ---- cfg3: the e = 0..5 sweep BASELINE.json names (thresholds 133,115,79,61,25,7), full pass and early exit
    sweep = None
    if args.workload == "cfg3":
        sweep = {}
        for e in range(6):
            cur_thr[0] = int(L.dyibf_threshold(e, K, RL))
            res = {"threshold": cur_thr[0]}
            for mode in ("full", "early_exit"):
                f.set_early_exit(mode == "early_exit")
                device_step()
                barrier()
                e0.record()
                for _ in range(2):
                    device_step()
                e1.record()
                barrier()
                tt = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
                if world > 1:
                    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                res[mode] = units_world * n * 2 / (float(tt.item()) / 1e3)
            sweep["e%d" % e] = res
        f.set_early_exit(False); cur_thr[0] = thr

    # ---- extras that ride in the cfg2 line (bench_support/extras.py): the other BASELINE configs, multi-GPU parity against the
    # reference's CPU code, the stages either side of the kernel.  Order matters: cfg3 overwrites the read buffer, cfg4 needs
    # the cfg2 filter's memory.
    extras = {}
    if args.workload == "cfg2" and not args.no_extras:
        from bench_support import extras as X
        ctx = mk_ctx()
        if world > 1 and want("parity"):
            extras["multi_gpu_parity"] = X.multi_gpu_parity(ctx)
        if world > 1 and want("single"):
            extras["single_process"] = X.single_process(ctx)
        if world == 1 and want("io"):
            extras["filter_io"] = X.filter_io(ctx)
        if want("cfg3"):
            extras["cfg3"] = X.cfg3(ctx)
        _cpu.clear()
        if world == 1 and want("cfg5"):
            extras["cfg5"] = X.cfg5(ctx)
        if world > 1 and want("cfg4"):
            f.close(); f = None
            del words, d_reads, d_rows, d_ids, ctx
            torch.cuda.empty_cache()
            ctx4 = mk_ctx_light()
            extras["cfg4"] = X.cfg4(ctx4)

    if rank == 0:
        line = {"metric": args.metric, "value": value, "unit": args.unit, "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
                "config": workload_config(args, thr, world), "clocks": clocks, "e2e": e2e, "gpu_launches": int(gpu_launches),
                "roofline": roofline, "cpu_baseline": cpu, "filter_build_s": t_build, "early_exit": early, "dropin_batches": dropin, "ingest": ingest, "impl": "b200"}
        if replication is not None:
            line["filter_replication"] = replication
        line.update(extras)
        if sweep is not None:
            line["e_sweep"] = sweep
        # the verdicts and rates of the extras once more, LAST and small: a reader who keeps only the tail of the line still sees them
        summary = {}
        try:
            pick = lambda d, keys: {k: d.get(k) for k in keys} if isinstance(d, dict) else None       # noqa: E731
            if extras.get("multi_gpu_parity"):
                summary["multi_gpu_parity"] = pick(extras["multi_gpu_parity"], ("read_sharded", "bin_sharded"))
            sp = extras.get("single_process")
            if sp:
                summary["single_process"] = pick(sp, ("value", "lists_equal_reference", "lists_consistent_with_rank0"))
                summary["single_process"]["bin_sharded_group"] = pick(sp.get("bin_sharded_group"), ("lists_equal_reference", "words_equal_reference"))
                summary["group_build"] = pick(sp.get("group_build"), ("kmers_per_s", "words_equal_reference", "words_equal_single_gpu_build"))
            if replication is not None:
                summary["filter_replication_gbs"] = replication["gbs"]
            if dropin:
                summary["dropin_batches"] = {"value": dropin["value"], "pinned_input": dropin.get("pinned_input"),
                                             "asynchronous": (dropin.get("asynchronous") or {}).get("value")}
            c4 = extras.get("cfg4")
            if c4:
                summary["cfg4"] = {"value": c4.get("value"), "frac": (c4.get("roofline") or {}).get("frac"), "skipped": c4.get("skipped"),
                                   "shard_vs_reference": pick(c4.get("shard_vs_reference"), ("rows_equal_reference", "lists_equal_reference", "skipped"))}
            if extras.get("cfg3"):
                summary["cfg3"] = pick(extras["cfg3"], ("value",))
            c5 = extras.get("cfg5")
            if c5:
                summary["cfg5"] = {"value": c5.get("value"), "e2e": (c5.get("e2e") or {}).get("value"),
                                   "words_equal_reference": (c5.get("cpu_baseline") or {}).get("filter_words_equal_gpu")}
            summary["headline"] = {"value": value, "e2e": e2e["value"] if e2e else None, "frac": roofline["frac"], "n_gpus": world}
        except Exception as ex:                                            # a summary must never cost the line
            summary["error"] = repr(ex)[:200]
        line["summary"] = summary
        print(json.dumps(line), flush=True)
    if f is not None:
        f.close()
    for p_ in (hp, op, rp, ip):
        L.dyibf_host_free(p_)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.workload == "cfg5":
        from bench_support import build_bench
        (build_bench.run_reference if a.impl == "reference" else build_bench.run_b200)(a, sys.modules[__name__])
    elif a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
