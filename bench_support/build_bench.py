"""bench.py --workload cfg5: IBF build (BASELINE.json configs[4]).

One step = dyibf_add_sequences (addKmers, d_bloom_filter.h:91-94,293-315) over `bins` contigs of `genome_len`
bases, one per bin; successive steps add NEW contigs to the same 16 GiB filter (chunk c = synthetic bins
c*bins .. (c+1)*bins-1), so K steps model K/25 of the ~100 Gbp job BASELINE names when genome_len = 4 Mbp.
`value` = k-mers/s with the contig bytes resident in HBM; `e2e` = the same through the host-buffer call (pinned
host bytes in) plus ONE download of the finished filter words at the end of the K steps.
"""
import ctypes as C
import json
import os
import time

import numpy as np

K, H = 18, 3
MAX_CHUNKS = 8          # distinct synthetic chunks kept resident (host generation is ~10 s per 4 Gbp chunk); more steps cycle


def _config(args, world):
    kmers = args.bins * (args.genome_len - K + 1)
    return {"workload": "cfg5: IBF build, %d bins x %d bp new synthetic contigs per step (%d k-mers, k=%d, h=%d) into a %d GiB "
                        "filter per GPU; %d steps = %.1f Gbp of the ~100 Gbp job" % (
                            args.bins, args.genome_len, kmers, K, H, args.filter_gib, args.steps,
                            args.steps * args.bins * args.genome_len / 1e9),
            "bins": args.bins, "filter_gib": args.filter_gib, "kmer": K, "hashes": H, "genome_bp_per_bin": args.genome_len,
            "kmers_per_step": kmers,
            "l2": "no flush needed: every step ORs %.1f G single bits into random words of a %d GiB table (>> 126 MB L2)"
                  % (kmers * H / 1e9, args.filter_gib),
            "parallelism": "replicas x%d (each GPU builds its own filter from its own contigs; no collective)" % world}


def _spread_bins(n_seqs, bins):
    """bin of sequence i when a SAMPLE of n_seqs contigs is spread over all 64-bin columns: every host thread gets work
    and no two threads ever share a 64-bit word"""
    ncol = bins // 64
    i = np.arange(n_seqs, dtype=np.uint32)
    return (i % ncol) * 64 + (i // ncol) % 64


def sliced_roofline(kmers_step, h, table_bytes, build_ms, apply_ms, split2_ms, hbm_peak):
    """The sliced build (DYIBF_BUILD_SLICED) does NOT move the random read-modify-write traffic SURVEY 8(d) models (192 B per
    k-mer): its updates are L2 atomics into per-bin vectors and only the interleave streams DRAM.  Two kernels, two bounds:
      interleave: read the shadow (= table size), read + write the table -> 3 x table bytes against the measured HBM peak
      scatter   : h REDs per k-mer against the divergent L2 RED.OR rate tools/red_bench.cu measured (profiles/l2_red_peak.json)"""
    scatter_ms = build_ms - apply_ms - split2_ms
    inter = 3 * table_bytes / (apply_ms / 1e3) / 1e9
    red_rate = kmers_step * h / (scatter_ms / 1e3)
    red_peak, red_src = None, "profiles/l2_red_peak.json absent: no measured RED peak, no scatter fraction"
    try:
        rp = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "l2_red_peak.json")))
        red_peak, red_src = float(rp["g_red_per_s"]) * 1e9, rp.get("source")
    except Exception:
        pass
    return {"bound": "hbm", "kernel": "sliced_interleave_kernel<8>", "achieved": inter, "peak": hbm_peak, "unit": "GB/s",
            "frac": inter / hbm_peak, "kernel_ms": apply_ms, "algorithmic_bytes_per_launch": 3 * table_bytes,
            "traffic": 78998419000.0 * table_bytes / float(1 << 34),
            "traffic_source": "DERIVED, not measured in this run: profiles/r02_sliced_interleave_ncu.md (dram__bytes_read 61.86 GB + "
                              "dram__bytes_write 17.14 GB per 16 GiB pass, round 2; kernel unchanged since), scaled by table size",
            "build_kernels_ms": build_ms, "scatter_ms": scatter_ms, "split2_ms": split2_ms,
            "scatter": {"bound": "l2 atomics", "kernel": "sliced_scatter_kernel<%d>" % h, "achieved": red_rate / 1e9, "unit": "G RED/s",
                        "peak": red_peak / 1e9 if red_peak else None, "frac": red_rate / red_peak if red_peak else None,
                        "peak_source": red_src, "reds_per_launch": kmers_step * h},
            "sector_model_gbs": kmers_step * h * 64 / (build_ms / 1e3) / 1e9,
            "note": "build mode 3 (sliced).  frac = the interleave pass against HBM; the scatter is bound by the L2's divergent RED rate "
                    "(scatter.frac).  sector_model_gbs is SURVEY 8(d)'s 192 B per k-mer random-RMW model over the whole step: it exceeds "
                    "the HBM peak because this build does not move those bytes, so it is an equivalence, not a roofline"}


def _cpu_build(cf, genomes, bin_of_seq):
    """reference build_filter (d_build_filter.cpp:217-272): bins over threads, addKmers per bin.  The reference's threads
    own contiguous bin ranges and do a non-atomic `word |= mask` (d_bloom_filter.h:300): with 1024 bins and 16 threads a
    range is exactly one 64-bin column, so its threads never share a word.  Here a thread owns whole columns for every
    core count, which keeps the host build race-free (= the `-t 1` filter the GPU must reproduce bit for bit)."""
    from concurrent.futures import ThreadPoolExecutor
    cols = sorted(set(int(b) // 64 for b in bin_of_seq))
    threads = max(1, min(cf.cores, len(cols)))
    owner = {c: j % threads for j, c in enumerate(cols)}
    work = [[] for _ in range(threads)]
    for i, b in enumerate(bin_of_seq):
        work[owner[int(b) // 64]].append((i, int(b)))
    if cf.kind == "reference":
        add = lambda i, b: cf.ref.add_kmers(cf.f, genomes[i], b)             # noqa: E731
    else:
        add = lambda i, b: cf.o.add_kmers(cf.words, cf.p, genomes[i], b)     # noqa: E731

    def run(items):
        for i, b in items:
            add(i, b)
    t = time.perf_counter()
    with ThreadPoolExecutor(max_workers=threads) as ex:                      # ctypes releases the GIL
        list(ex.map(run, work))
    return time.perf_counter() - t, threads


def run_reference(args, B):
    """The reference's own addKmers on the host cores: each step adds a bounded sample of bins (whole 64-bin columns)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    from bench_support import synth
    nbits = args.filter_gib * (1 << 33) + 256
    cf = B.CpuFilter(args.bins, H, K, nbits)
    per_bin = args.genome_len - K + 1
    assert args.bins % 64 == 0, "cfg5 wants whole 64-bin columns"
    ncol = args.bins // 64
    probe = synth.genomes(ncol, args.genome_len)
    dt, threads = _cpu_build(cf, probe, _spread_bins(ncol, args.bins))
    rate = ncol * per_bin / dt
    nb = int(min(max(round(rate * 4.0 / per_bin / ncol), 1) * ncol, args.bins))   # ~4 s of CPU work per step
    times = []
    for s in range(args.warmup + args.steps):
        g = synth.genomes(nb, args.genome_len, first_bin=(s + 1) * args.bins)
        dt, threads = _cpu_build(cf, g, _spread_bins(nb, args.bins))
        if s >= args.warmup:
            times.append(dt)
    total = sum(times)
    value = nb * per_bin * args.steps / total
    line = {"impl": "reference", "metric": args.metric, "value": value, "unit": args.unit, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic", "config": _config(args, args.gpus),
            "cpu_baseline": {"value": value, "unit": args.unit, "cores": threads, "kind": cf.kind,
                             "sample": "%d contigs x %d bp per step spread over the %d 64-bin columns, columns over threads "
                                       "(build_filter's bins-over-threads split, race-free)" % (nb, args.genome_len, ncol)},
            "e2e": {"value": value, "unit": args.unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def run_b200(args, B):
    import torch
    import torch.distributed as dist
    import dream_yara_b200 as dy
    from dream_yara_b200._lib import ENC_DNA5, MEM_DEVICE, MEM_HOST, check
    from bench_support import synth

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the IBF build has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    L = dy.lib()
    bins, glen = args.bins, args.genome_len
    nbits = args.filter_gib * (1 << 33) + 256
    W = (bins + 63) // 64
    P = (nbits - 256) // (64 * W)
    kmers_step = bins * (glen - K + 1)
    n_chunks = min(args.warmup + args.steps, MAX_CHUNKS)
    chunk_bytes = bins * glen

    # ---- setup (untimed): distinct contig chunks in pinned host memory and in HBM, two empty filters
    pinned, h_chunks, d_chunks = [], [], []
    for c in range(n_chunks):
        p = C.c_void_p(); check(L.dyibf_host_alloc(C.byref(p), chunk_bytes))
        a = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(bins, glen))
        synth.genomes(bins, glen, out=a, first_bin=(rank * MAX_CHUNKS + c) * bins)
        pinned.append(p); h_chunks.append(a); d_chunks.append(torch.from_numpy(a.reshape(-1)).to(dev))
    h_off = np.arange(bins + 1, dtype=np.uint64) * glen
    h_bin = np.arange(bins, dtype=np.uint32)
    d_off = torch.from_numpy(h_off.view(np.int64)).to(dev)
    d_bin = torch.from_numpy(h_bin.view(np.int32)).to(dev)
    words_dev = torch.zeros(P * W, dtype=torch.int64, device=dev)
    words_host = torch.zeros(P * W, dtype=torch.int64, device=dev)
    f_dev = dy.SeqAnBloomFilter.attach(bins, H, K, nbits, words_dev.data_ptr(), device=local, keep=words_dev)
    f_host = dy.SeqAnBloomFilter.attach(bins, H, K, nbits, words_host.data_ptr(), device=local, keep=words_host)
    stream = torch.cuda.current_stream().cuda_stream

    def device_step(f, c, n_seqs=bins):
        check(L.dyibf_add_sequences(f._h, C.c_void_p(d_chunks[c].data_ptr()), C.c_void_p(d_off.data_ptr()), n_seqs,
                                    C.c_void_p(d_bin.data_ptr()), ENC_DNA5, MEM_DEVICE, C.c_void_p(stream)))

    def host_step(f, c):
        check(L.dyibf_add_sequences(f._h, pinned[c], C.c_void_p(h_off.ctypes.data), bins, C.c_void_p(h_bin.ctypes.data),
                                    ENC_DNA5, MEM_HOST, None))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- CPU baseline + byte-identity check (rank 0, N = 1): the reference's addKmers builds whole 64-bin columns of
    #      chunk 0 on the host; the same bins go through the GPU build into the still-empty filter; columns must be equal
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cf = B.CpuFilter(bins, H, K, nbits)
        per_bin = glen - K + 1
        assert bins % 64 == 0, "cfg5 wants whole 64-bin columns"
        ncol = bins // 64
        t_cpu, threads = _cpu_build(cf, h_chunks[0], _spread_bins(ncol, bins))
        per_col = int(max(1, min(round(15.0 / max(t_cpu, 1e-3)), 8, 64)))
        n_chk = ncol * per_col
        chk_bins = _spread_bins(n_chk, bins)
        if per_col > 1:
            t_cpu += _cpu_build(cf, h_chunks[0][ncol:], chk_bins[ncol:])[0]
        d_chk = torch.from_numpy(chk_bins.view(np.int32)).to(dev)
        for f in (f_dev, f_host):                                             # both tables get the same sample
            check(L.dyibf_add_sequences(f._h, C.c_void_p(d_chunks[0].data_ptr()), C.c_void_p(d_off.data_ptr()), n_chk,
                                        C.c_void_p(d_chk.data_ptr()), ENC_DNA5, MEM_DEVICE, C.c_void_p(stream)))
        torch.cuda.synchronize()
        want = cf.words[:P * W]
        same = True
        for a in range(0, P * W, 1 << 27):                                    # 1 GiB pieces
            got = words_dev[a:a + (1 << 27)].cpu().numpy().view(np.uint64)
            same = same and bool(np.array_equal(got, want[a:a + (1 << 27)]))
        cpu = {"value": n_chk * per_bin / t_cpu, "unit": args.unit, "cores": threads, "kind": cf.kind,
               "sample": "first %d contigs of chunk 0 (x %d bp) spread over the %d 64-bin columns, %.1f s, columns over "
                         "threads (build_filter's bins-over-threads split, race-free)" % (n_chk, glen, ncol, t_cpu),
               "filter_words_equal_gpu": same, "bits_set_in_sample": int(np.bitwise_count(want).sum(dtype=np.int64))}
        if not same:
            raise AssertionError("GPU-built filter differs from the CPU reference build")
        del cf, got, want

    # ---- value: contig bytes resident in HBM, CUDA events on the launching stream
    order = [s % n_chunks for s in range(args.warmup + args.steps)]
    for s in range(args.warmup):
        device_step(f_dev, order[s])
    f_dev.profile_enable(True); f_dev.profile(reset=True)
    sampler = B.ClockSampler(local)
    barrier()
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s in range(args.warmup, args.warmup + args.steps):
        device_step(f_dev, order[s])
    e1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    prof = f_dev.profile(reset=True); f_dev.profile_enable(False)
    t = torch.tensor([e0.elapsed_time(e1), prof["build_ms"] / args.steps, prof["pack_ms"] / args.steps,
                      prof["build_apply_ms"] / args.steps, prof["build_split2_ms"] / args.steps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, build_ms, pack_ms, apply_ms, split2_ms = (float(x) for x in t.tolist())
    value = world * kmers_step * args.steps / (ms_total / 1e3)
    launches = int(prof["build_launches"] + prof["pack_launches"])
    bucket = C.c_int64(0)
    L.dyibf_get_option(f_dev._h, 2, C.byref(bucket))

    # ---- e2e: the public host-buffer call on pinned contig bytes (H2D inside), then ONE download of the finished filter
    e2e = None
    if not args.no_e2e:
        n_data = P * W
        out_words = np.empty(n_data, dtype=np.uint64)
        out_words[::512] = 0                                              # touch the pages before the timed region
        for s in range(args.warmup):
            host_step(f_host, order[s])
        barrier()
        t0 = time.perf_counter()
        for s in range(args.warmup, args.warmup + args.steps):
            host_step(f_host, order[s])
        t_add = time.perf_counter() - t0
        check(L.dyibf_download_words(f_host._h, C.c_void_p(out_words.ctypes.data), 0, n_data))
        barrier()
        dt = torch.tensor([time.perf_counter() - t0, t_add], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": world * kmers_step * args.steps / float(dt[0]), "unit": args.unit,
               "h2d_bytes_per_step": world * (chunk_bytes + (bins + 1) * 8 + bins * 4),
               "d2h_bytes_per_step": world * n_data * 8 // args.steps,
               "add_only_value": world * kmers_step * args.steps / float(dt[1]),
               "timing": "host wall clock around K synchronous dyibf_add_sequences(host) calls + one dyibf_download_words "
                         "of the whole filter (its bytes spread over the K steps), max over ranks"}
        # both paths added the same contigs (the verification adds of chunk 0 are idempotent): same table, and the
        # downloaded words are that table
        if not torch.equal(words_dev, words_host):
            raise AssertionError("host-mode filter != device-mode filter")
        chk = torch.from_numpy(out_words[: 1 << 24].view(np.int64)).to(dev)
        if not torch.equal(chk, words_host[: 1 << 24]):
            raise AssertionError("downloaded words differ from the table in HBM")
        del out_words, chk
        # optional A/B of DYIBF_OPT_BUILD_CHUNK_MIB (re-adding the same contigs leaves the table as it is)
        if os.environ.get("DYIBF_BENCH_BUILD_CHUNKS"):
            by_chunk = {}
            for mib in [int(x) for x in os.environ["DYIBF_BENCH_BUILD_CHUNKS"].split(",")]:
                check(L.dyibf_set_option(f_host._h, 3, mib))
                host_step(f_host, order[0])
                barrier()
                t0 = time.perf_counter()
                for s in range(2):
                    host_step(f_host, order[s])
                barrier()
                by_chunk[str(mib)] = world * kmers_step * 2 / (time.perf_counter() - t0)
            check(L.dyibf_set_option(f_host._h, 3, 256))
            e2e["add_only_by_chunk_mib"] = by_chunk

    # ---- extra (not the headline): the same steps with the bucketed build off = one atomic OR per (k-mer, hash) straight
    #      into the table (DYIBF_OPT_BUILD_BUCKET_MIB = 0); ORs of bits that are already set cost the same traffic
    direct = None
    if bucket.value != 0:
        f_host.set_build_bucket_mib(0)
        device_step(f_host, 0)
        f_host.profile_enable(True); f_host.profile(reset=True)
        barrier()
        e0.record()
        for s in range(2):
            device_step(f_host, order[s])
        e1.record()
        barrier()
        p2 = f_host.profile(reset=True); f_host.profile_enable(False)
        t2 = torch.tensor([e0.elapsed_time(e1), p2["build_ms"] / 2], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t2, op=dist.ReduceOp.MAX)
        direct = {"value": world * kmers_step * 2 / (float(t2[0]) / 1e3), "unit": args.unit, "kernel_ms": float(t2[1]),
                  "sector_gbs": kmers_step * H * 64 / (float(t2[1]) / 1e3) / 1e9,
                  "note": "DYIBF_OPT_BUILD_BUCKET_MIB=0: build_kernel<3> only, 3 random 64-bit RED.OR per k-mer"}
        if not torch.equal(words_dev, words_host):
            raise AssertionError("direct-atomics filter != bucketed filter")

    # ---- extra, N > 1 (SURVEY §8(e) "Build"): ONE filter built by all GPUs — rank r adds only its own bin range to its own
    #      column slice of every block (a self-contained IBF with bins/N bins, same P), then the slices are interleaved into
    #      the replicated layout with NCCL all-gathers; no OR-reduction anywhere
    sharded = None
    if world > 1 and W % world == 0:
        from dream_yara_b200 import dist as D
        first, sb, snbits, Wr, _ = D.bin_shard_params(bins, nbits, rank, world)
        f_host.close()                                                   # frees its scratch; its table becomes the assembly target
        f_sh = dy.SeqAnBloomFilter(sb, H, K, snbits, device=local)
        d_bin_sh = torch.arange(sb, dtype=torch.int32, device=dev)
        def shard_step(c):
            check(L.dyibf_add_sequences(f_sh._h, C.c_void_p(d_chunks[c].data_ptr() + first * glen), C.c_void_p(d_off.data_ptr()), sb,
                                        C.c_void_p(d_bin_sh.data_ptr()), ENC_DNA5, MEM_DEVICE, C.c_void_p(stream)))
        shard_step(order[0])
        shard_view = D._tensor_from_ptr(L.dyibf_device_words(f_sh._h), P * Wr, dev)
        D.assemble_replicated(shard_view, bins, nbits, out=words_host)   # warm-up of the collective
        barrier()
        e0.record()
        for s in range(args.warmup, args.warmup + args.steps):
            shard_step(order[s])
        e1.record()
        barrier()
        e2 = torch.cuda.Event(enable_timing=True)
        D.assemble_replicated(shard_view, bins, nbits, out=words_host)
        e2.record()
        barrier()
        # every rank must hold the same table, and its own slice of it must be its shard
        own = words_host.view(P, W)[:, rank * Wr:(rank + 1) * Wr]
        ok_own = bool(torch.equal(own, shard_view.view(P, Wr)))
        chk = torch.stack([words_host.sum(), words_host[::4097].sum()]).to(torch.int64)
        allc = [torch.empty_like(chk) for _ in range(world)]
        dist.all_gather(allc, chk)
        ok_same = all(bool(torch.equal(c_, allc[0])) for c_ in allc)
        t3 = torch.tensor([e0.elapsed_time(e1), e1.elapsed_time(e2)], dtype=torch.float64, device=dev)
        dist.all_reduce(t3, op=dist.ReduceOp.MAX)
        tot_kmers = bins * (glen - K + 1) * args.steps                   # all ranks together add `bins` contigs per step
        sharded = {"value": tot_kmers / ((float(t3[0]) + float(t3[1])) / 1e3), "unit": args.unit, "scaling": "strong",
                   "build_ms_per_step": float(t3[0]) / args.steps, "assemble_ms": float(t3[1]),
                   "assemble_gbs_per_gpu": (world - 1) * P * Wr * 8 / (float(t3[1]) / 1e3) / 1e9,
                   "bins_per_gpu": sb, "shard_gib": P * Wr * 8 / 2**30, "own_slice_equals_shard": ok_own, "replicas_identical": ok_same,
                   "note": "one %d-bin filter built by %d GPUs: bin-sharded add (%d bins each) for %d steps, then one "
                           "all-gather interleave of the %d GiB table; value counts the assembly" % (bins, world, sb, args.steps, args.filter_gib)}
        if not (ok_own and ok_same):
            raise AssertionError("assembled filter differs between ranks or from the rank's shard")
        f_sh.close()

    # ---- roofline of the build kernels (scatter + apply, or the direct kernel) against the measured HBM peak.
    # h x (32 B sector read + 32 B sector write) per k-mer for random single-word RMW.  The bucketed
    # build replaces that traffic by h x (4 B record written + 4 B read) + one streaming pass over the table per apply
    # sweep, so its `achieved` (same algorithmic definition) may exceed the peak: that is the point of the bucketing.
    peak, peak_src = B.measured_peak()
    alg = kmers_step * H * 64
    achieved = alg / (build_ms / 1e3) / 1e9
    model = kmers_step * H * 8 + 2 * P * W * 8
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "kernel": "split1_kernel<3,u32> + split2_kernel<u32> + bucket_apply_kernel" if bucket.value != 0 else "build_kernel<3>",
                "kernel_ms": build_ms, "split1_ms": build_ms - apply_ms - split2_ms, "split2_ms": split2_ms, "apply_ms": apply_ms,
                "sweeps_per_step": prof["build_apply_launches"] / args.steps, "pack_kernel_ms": pack_ms, "algorithmic_bytes_per_launch": alg, "peak_source": peak_src,
                "bucketed_model_bytes_per_step": model, "bucketed_model_gbs": model / (build_ms / 1e3) / 1e9,
                "note": "algorithmic bytes = the random-RMW sector model (192 B/k-mer); the bucketed build moves "
                        "~%.0f B/k-mer instead (records + one table sweep per step)" % (model / kmers_step)}
    try:
        if f_dev.last_build_mode() == 3:                                   # sliced: the 192 B model is not what this build moves
            roofline = dict(sliced_roofline(kmers_step, H, P * W * 8, build_ms, apply_ms, split2_ms, peak), peak_source=peak_src,
                            pack_kernel_ms=pack_ms)
    except Exception:
        pass
    if rank == 0:
        line = {"metric": args.metric, "value": value, "unit": args.unit, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "u64", "data": "synthetic", "config": _config(args, world), "clocks": clocks,
                "e2e": e2e, "gpu_launches": launches, "roofline": roofline, "cpu_baseline": cpu, "direct_atomics": direct,
                "sharded_build": sharded, "distinct_chunks": n_chunks, "impl": "b200"}
        print(json.dumps(line), flush=True)
    f_dev.close(); f_host.close()                                       # close() is idempotent
    for p_ in pinned:
        L.dyibf_host_free(p_)
    if world > 1:
        dist.destroy_process_group()
