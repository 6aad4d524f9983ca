"""bench.py extras: measurements that ride in the default JSON line next to the headline (cfg2) numbers, so that every
BASELINE.json config, the multi-GPU parity checks and the stages either side of the kernel (filter load / replication,
asynchronous batches, one process driving all GPUs) are measured by the driver's own runs.

Each function takes the bench context (a SimpleNamespace made by bench.run_b200) and returns a dict, or None where it does
not apply.  Nothing here is part of `value`, `e2e` or `roofline` of the headline.  The CPU reference (oracle/_ref, or the
oracle port) is only ever the checker / the timed baseline, outside the GPU timed regions.
"""
import ctypes as C
import os
import shutil
import tempfile
import time
from collections import deque

import numpy as np

K, H = 18, 3


def _vp(a):
    return C.c_void_p(a.ctypes.data)


def _maxr(ctx, vals):
    """max over ranks of a list of floats"""
    t = ctx.torch.tensor(list(vals), dtype=ctx.torch.float64, device=ctx.dev)
    if ctx.world > 1:
        ctx.dist.all_reduce(t, op=ctx.dist.ReduceOp.MAX)
    return [float(x) for x in t.tolist()]


def _lists_from_rows(rows, bins):
    """per-bin ascending row ids (CSR) of selection rows, in numpy"""
    bits = np.unpackbits(np.ascontiguousarray(rows).view(np.uint8), axis=1, bitorder="little")[:, :bins]
    rr, bb = np.nonzero(bits.T)                                       # sorted by bin, then by row
    offs = np.zeros(bins + 1, dtype=np.uint64)
    offs[1:] = np.cumsum(np.bincount(rr, minlength=bins))
    return offs, bb.astype(np.uint32)


# ---------------------------------------------------------------------------------------------------------------------
# asynchronous drop-in batches: dyibf_classify_begin / _end at the reference's batch size (src/mapper.h:109), as a
# prefetching caller (src/file_prefetched.h:137-205) feeds them
# ---------------------------------------------------------------------------------------------------------------------
def _restrict(boff, ids, bins, lo, hi):
    """the per-bin lists (CSR) restricted to the ids in [lo, hi), renumbered from 0: what a classify of those rows alone returns"""
    binno = np.repeat(np.arange(bins, dtype=np.uint32), np.diff(boff.astype(np.int64)))
    m = (ids >= lo) & (ids < hi)
    off = np.zeros(bins + 1, dtype=np.uint64)
    off[1:] = np.cumsum(np.bincount(binno[m], minlength=bins))
    return off, (ids[m] - np.uint32(lo)).astype(np.uint32)


def dropin_async(ctx, batches_pageable, batches_pinned, per_call, rb, group, depth=4):
    L, f, bins = ctx.L, ctx.f, ctx.bins
    from dream_yara_b200._lib import ENC_DNA5, check
    offs_b = np.arange(per_call + 1, dtype=np.uint64) * ctx.RL
    cap = 4 * rb
    outs = [(np.zeros(bins + 1, dtype=np.uint64), np.zeros(cap, dtype=np.uint32)) for _ in range(depth + 1)]
    tot = C.c_uint64(0)

    def run(srcs, n_calls):
        q = deque(); pairs = 0; slot = 0
        def end():
            nonlocal pairs
            t, o = q.popleft()
            rc = L.dyibf_classify_end(t, _vp(o[0]), _vp(o[1]), cap, C.byref(tot))
            check(rc)
            pairs += tot.value
            return o
        last = None
        ctx.barrier()
        t0 = time.perf_counter()
        for i in range(n_calls):
            if len(q) == depth:
                last = end()
            t = C.c_void_p()
            check(L.dyibf_classify_begin(f._h, _vp(srcs[i % len(srcs)]), _vp(offs_b), per_call, ENC_DNA5, group, 1, ctx.thr, C.byref(t)))
            q.append((t, outs[slot])); slot = (slot + 1) % len(outs)
        while q:
            last = end()
        dt = time.perf_counter() - t0
        ctx.barrier()
        return _maxr(ctx, [dt])[0], pairs, last

    n_calls = 5 * len(batches_pageable)
    run(batches_pageable, len(batches_pageable))                      # warm-up: ticket buffers get allocated
    dt, pairs, last = run(batches_pageable, n_calls)
    # the last batch's lists must equal the synchronous call's
    bo = np.zeros(bins + 1, dtype=np.uint64); ids = np.zeros(cap, dtype=np.uint32)
    src = batches_pageable[(n_calls - 1) % len(batches_pageable)]
    check(L.dyibf_classify(f._h, _vp(src), _vp(offs_b), per_call, ENC_DNA5, group, 1, ctx.thr, _vp(bo), _vp(ids), cap, C.byref(tot)))
    same = bool(np.array_equal(bo, last[0]) and np.array_equal(ids[:tot.value], last[1][:tot.value]))
    if not same:
        raise AssertionError("asynchronous classify lists differ from the synchronous call's")
    out = {"value": ctx.units_world * rb * n_calls / dt, "unit": ctx.args.unit, "reads_per_call": rb, "calls": n_calls,
           "in_flight": depth, "ms_per_call": dt / n_calls * 1e3, "read_bin_pairs": int(pairs), "lists_equal_synchronous": same,
           "note": "dyibf_classify_begin/_end, %d batches in flight, pageable host reads in, per-bin lists out, host wall "
                   "clock, max over ranks; what a prefetching caller (file_prefetched.h) sees at the reference's batch size" % depth}
    if batches_pinned:
        dtp, _, _ = run(batches_pinned, n_calls)
        out["pinned_input"] = ctx.units_world * rb * n_calls / dtp
    return out


# ---------------------------------------------------------------------------------------------------------------------
# filter load / save as a measured stage ("Filter loading time", src/d_mapper.h:913-921)
# ---------------------------------------------------------------------------------------------------------------------
def filter_io(ctx):
    if ctx.rank != 0:
        return None
    L, f, torch = ctx.L, ctx.f, ctx.torch
    from dream_yara_b200 import dist as D
    from dream_yara_b200._lib import IoStats, check
    size = 8 + ((ctx.nbits + 63) // 64) * 8
    where = None
    for cand in ("/dev/shm", tempfile.gettempdir()):
        try:
            if shutil.disk_usage(cand).free > size + (4 << 30):
                where = cand
                break
        except OSError:
            pass
    if where is None:
        return {"skipped": "no file system with %d GiB free" % (size >> 30)}
    path = os.path.join(where, "dyibf_bench_%d.filter" % os.getpid())
    out = {"file_bytes": size, "where": where + (" (tmpfs: page-cache-warm by construction)" if where == "/dev/shm" else "")}
    try:
        # measured H2D bandwidth of this box from pinned memory (what a perfect loader could reach)
        n_h2d = 1 << 30
        hp = torch.empty(n_h2d, dtype=torch.uint8).pin_memory()
        dp = torch.empty(n_h2d, dtype=torch.uint8, device=ctx.dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dp.copy_(hp, non_blocking=True); torch.cuda.synchronize()
        e0.record()
        for _ in range(3):
            dp.copy_(hp, non_blocking=True)
        e1.record(); torch.cuda.synchronize()
        h2d = 3 * n_h2d / (e0.elapsed_time(e1) / 1e3) / 1e9
        e0.record()
        for _ in range(3):
            hp.copy_(dp, non_blocking=True)
        e1.record(); torch.cuda.synchronize()
        d2h = 3 * n_h2d / (e0.elapsed_time(e1) / 1e3) / 1e9
        del hp, dp
        st = IoStats()
        ncpu = os.cpu_count() or 8
        io_threads = max(4, min(16, ncpu))
        check(L.dyibf_set_io_threads(io_threads))
        t0 = time.perf_counter()
        check(L.dyibf_save_file(f._h, path.encode()))
        t_save = time.perf_counter() - t0
        check(L.dyibf_last_io_stats(C.byref(st)))
        out.update({"save_s": t_save, "save_gbs": size / t_save / 1e9, "save_pwrite_s": st.file_seconds, "save_threads": st.threads,
                    "host_cores": ncpu, "h2d_pinned_gbs": h2d, "d2h_pinned_gbs": d2h})

        def load(threads, chunk_mib):
            check(L.dyibf_set_io_threads(threads))
            check(L.dyibf_set_io_chunk_mib(chunk_mib))
            h = C.c_void_p()
            t0 = time.perf_counter()
            check(L.dyibf_load_file(path.encode(), ctx.local, C.byref(h)))
            t_load = time.perf_counter() - t0
            check(L.dyibf_last_io_stats(C.byref(st)))
            loaded = D._tensor_from_ptr(L.dyibf_device_words(h), ctx.P * ctx.W, ctx.dev)
            same = bool(torch.equal(loaded, ctx.words))
            del loaded
            L.dyibf_destroy(h)
            if not same:
                raise AssertionError("filter words after save + load differ from the table in HBM")
            return {"threads": st.threads, "chunk_mib": st.chunk_bytes >> 20, "load_s": t_load, "load_gbs": size / t_load / 1e9,
                    "transfer_s": st.seconds, "pread_s_max_thread": st.file_seconds, "copy_wait_s": st.copy_wait_seconds,
                    "submit_s": st.submit_seconds, "starved_s": st.starve_seconds, "pinned_pool_alloc_s": st.pool_seconds}

        loads = [load(t, 4) for t in sorted({1, 4, 8, io_threads})]
        loads += [load(io_threads, 8)]
        check(L.dyibf_set_io_threads(0)); check(L.dyibf_set_io_chunk_mib(0))
        best = max(loads, key=lambda d: d["load_gbs"])
        out.update({"loads": loads, "load_gbs": best["load_gbs"], "load_s": best["load_s"], "load_threads": best["threads"],
                    "load_chunk_mib": best["chunk_mib"],
                    "frac_of_h2d": best["load_gbs"] / h2d, "words_equal_after_roundtrip": True,
                    "note": "dyibf_save_file then dyibf_load_file of the %d GiB cfg2 filter (file -> ring of pinned chunks filled by N "
                            "pread threads -> one thread issuing the copies -> HBM); load_s includes cudaMalloc + zero fill of the "
                            "table; h2d_pinned_gbs = this box's measured pinned host->HBM rate; starved_s = time the copy thread "
                            "waited for the pread threads (page cache -> pinned memcpy rate of the host)" % (size >> 30)})
    finally:
        try:
            os.unlink(path)
        except OSError:
            pass
    return out


# ---------------------------------------------------------------------------------------------------------------------
# multi-GPU parity on the real NCCL run: rank-merged results against the reference's CPU code
# ---------------------------------------------------------------------------------------------------------------------
def multi_gpu_parity(ctx):
    """N > 1.  read-sharded: every rank classifies its own slice of a global read set with its replica of the cfg2 filter,
    the per-rank lists are merged in rank order (dist.merge_bin_lists) and must equal the lists the reference's
    clasifyLoadedReads loop (src/d_mapper.h:338-375, oracle/_ref) produces for the whole set on the same filter words.
    bin-sharded: a %d-bin filter whose word columns are spread over the ranks (every rank builds its own bins), every rank
    classifies ALL reads against its shard, rows concatenate column-wise (dist.all_gather_columns) and lists by bin range;
    both must equal the reference's on the shard-concatenated table."""
    torch, dist, L, dy = ctx.torch, ctx.dist, ctx.L, ctx.dy
    from dream_yara_b200 import dist as D
    from dream_yara_b200._lib import ENC_DNA5, check
    from bench_support import synth
    import bench as B
    world, rank, RL = ctx.world, ctx.rank, ctx.RL
    out = {}
    # ---- read-sharded, on the cfg2 filter every rank holds
    m = max(131072 // world, 1024)
    mine = np.ascontiguousarray(ctx.h_reads[:m])
    offs, ids = ctx.f.classify((mine.reshape(-1), np.arange(m + 1, dtype=np.uint64) * RL), ctx.thr, group=2, derive_rc=True, encoding=ENC_DNA5)
    moffs, mids = D.merge_bin_lists(offs, ids, rank * m, ctx.bins)
    g = torch.empty((world * m, RL), dtype=torch.uint8, device=ctx.dev)
    dist.all_gather_into_tensor(g, torch.from_numpy(mine).to(ctx.dev))
    ok_read = True; t_cpu = 0.0
    if rank == 0:
        cf = ctx.cpu_filter()                                         # the reference's object with this GPU's words (cached for the other checks)
        allr = g.cpu().numpy()
        t0 = time.perf_counter()
        rows = cf.classify(B.full_layout_codes(allr), ctx.thr, False)
        t_cpu = time.perf_counter() - t0
        woffs, wids = _lists_from_rows(rows, ctx.bins)
        ok_read = bool(np.array_equal(woffs, moffs) and np.array_equal(wids, mids))
        out["read_sharded_cpu_s"] = t_cpu
        del rows
    out["reads"] = world * m
    # ---- bin-sharded, small table (the reference needs the whole table in host RAM): Wr = 2 word columns per rank
    sb = 128; tb = sb * world; nbits_t = (1 << 33) + 256
    Wt = tb // 64; Pt = (nbits_t - 256) // (64 * Wt); snbits = Pt * 64 * 2 + 256
    glen = 200_000
    gen = synth.genomes(sb, glen, first_bin=rank * sb)
    fs = dy.SeqAnBloomFilter(sb, H, K, snbits, device=ctx.local)
    fs.addSequences((gen.reshape(-1), np.arange(sb + 1, dtype=np.uint64) * glen), np.arange(sb, dtype=np.uint32))
    nr = max(32768 // world, 512)
    my_reads = synth.reads(gen, 4, rank * nr, nr, 100)
    allq = torch.empty((world * nr, 100), dtype=torch.uint8, device=ctx.dev)
    dist.all_gather_into_tensor(allq, torch.from_numpy(my_reads).to(ctx.dev))
    q = allq.cpu().numpy()
    thr_b = int(L.dyibf_threshold(3, K, 100))
    qb = (q.reshape(-1), np.arange(q.shape[0] + 1, dtype=np.uint64) * 100)
    rows_slice = fs.whichBinsBatch(qb, thr_b, group=2, derive_rc=True, encoding=ENC_DNA5)
    so, si = fs.classify(qb, thr_b, group=2, derive_rc=True, encoding=ENC_DNA5)
    rows_full = D.all_gather_columns(torch.from_numpy(rows_slice.view(np.int64)).to(ctx.dev)).cpu().numpy().view(np.uint64)
    boffs, bids = D.merge_bin_lists_bin_sharded(so, si, tb)
    shard = D._tensor_from_ptr(L.dyibf_device_words(fs._h), Pt * 2, ctx.dev)
    parts = [torch.empty_like(shard) for _ in range(world)]
    dist.all_gather(parts, shard.contiguous())
    ok_bin = True
    if rank == 0:
        cf = B.CpuFilter(tb, H, K, nbits_t)
        table = cf.words[:Pt * Wt].reshape(Pt, Wt)
        for r, part in enumerate(parts):
            table[:, 2 * r:2 * r + 2] = part.cpu().numpy().view(np.uint64).reshape(Pt, 2)
        t0 = time.perf_counter()
        rows = cf.classify(B.full_layout_codes(q), thr_b, False)
        out["bin_sharded_cpu_s"] = time.perf_counter() - t0
        woffs, wids = _lists_from_rows(rows, tb)
        ok_bin = bool(np.array_equal(rows, rows_full) and np.array_equal(woffs, boffs) and np.array_equal(wids, bids))
        out["bin_sharded_pairs"] = int(woffs[-1])
        del cf
    fs.close()
    flags = torch.tensor([int(ok_read), int(ok_bin)], dtype=torch.int64, device=ctx.dev)
    dist.broadcast(flags, src=0)
    out.update({"read_sharded": bool(flags[0].item()), "bin_sharded": bool(flags[1].item()),
                "bin_sharded_config": "%d bins over %d GPUs (128 bins = 2 word columns each), 1 GiB table, %d reads x 100 bp"
                                      % (tb, world, world * nr),
                "checker": "the reference's own clasifyLoadedReads loop (oracle/_ref) on rank 0, outside every timed region"})
    if not (out["read_sharded"] and out["bin_sharded"]):
        raise AssertionError("multi-GPU results differ from the reference: %r" % out)
    return out


# ---------------------------------------------------------------------------------------------------------------------
# cfg3 (BASELINE.json configs[2]): paired-end, e = 0..5 sweep, on the resident cfg2 filter
# ---------------------------------------------------------------------------------------------------------------------
def cfg3(ctx, steps=2):
    torch, L = ctx.torch, ctx.L
    from dream_yara_b200._lib import ENC_DNA5, MEM_DEVICE
    from bench_support import synth
    import bench as B
    n, RL = ctx.n // 2, ctx.RL                                         # pairs per step: the same bytes as a cfg2 step
    if n < 1000:
        return None
    h_pairs = ctx.h_reads.reshape(-1)[: 2 * n * RL].reshape(2, n, RL)    # reuse the pinned read buffer as [m1 | m2]
    synth.pairs(ctx.genomes, 3, ctx.rank * n, n, RL, out=h_pairs)
    d_pairs = torch.from_numpy(h_pairs.reshape(-1)).to(ctx.dev)
    d_rows = ctx.d_rows[: n * ctx.W]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sweep = {}
    f = ctx.f

    def step(thr):
        f.query_device(d_pairs.data_ptr(), ctx.d_off.data_ptr(), 2 * n, thr, d_rows.data_ptr(), group=4, derive_rc=True,
                       encoding=ENC_DNA5, max_seq_len=RL, stream=ctx.stream)
    full_rows = {}
    for e in range(6):
        thr = int(L.dyibf_threshold(e, K, RL))
        res = {"threshold": thr}
        for mode in ("full", "early_exit"):
            f.set_early_exit(mode == "early_exit")
            step(thr)
            f.profile_enable(True); f.profile(reset=True)
            ctx.barrier()
            e0.record()
            for _ in range(steps):
                step(thr)
            e1.record()
            ctx.barrier()
            p = f.profile(reset=True); f.profile_enable(False)
            ms, kms = _maxr(ctx, [e0.elapsed_time(e1), p["query_ms"] / max(p["query_launches"], 1)])
            res[mode] = ctx.units_world * n * steps / (ms / 1e3)
            res[mode + "_kernel_ms"] = kms
            if mode == "full":
                full_rows[e] = d_rows.clone() if e == 3 else None
                res["frac"] = n * B.algorithmic_bytes_per_read(ctx.bins, RL, True) / (kms / 1e3) / 1e9 / ctx.peak
            elif e == 3 and not torch.equal(full_rows[3], d_rows):
                raise AssertionError("cfg3: early-exit rows differ from the full pass")
        sweep["e%d" % e] = res
    f.set_early_exit(False)
    # parity on a prefix of pairs against the reference's loop (N = 1 only: the CPU filter words come from this GPU)
    cpu = None
    if ctx.rank == 0 and ctx.world == 1 and not ctx.args.no_cpu_baseline:
        m = 50_000
        sub = np.ascontiguousarray(h_pairs[:, :m])
        rows_gpu = f.whichBinsBatch((sub.reshape(-1), np.arange(2 * m + 1, dtype=np.uint64) * RL), sweep["e3"]["threshold"], group=4,
                                    derive_rc=True, encoding=ENC_DNA5, max_seq_len=RL)
        cf = ctx.cpu_filter()
        t0 = time.perf_counter()
        rows_cpu = cf.classify(B.full_layout_codes(sub), sweep["e3"]["threshold"], True)
        dt = time.perf_counter() - t0
        same = bool(np.array_equal(rows_cpu, rows_gpu))
        cpu = {"value": m / dt, "unit": "pairs/s", "cores": cf.cores, "kind": cf.kind, "sample": "first %d pairs, e=3" % m, "rows_equal_gpu": same}
        if not same:
            raise AssertionError("cfg3: GPU rows differ from the CPU reference on the sampled pairs")
    del d_pairs
    return {"workload": "cfg3: %d paired-end 2x%d bp pairs per GPU per step on the cfg2 filter, e = 0..5, %d steps per point"
                        % (n, RL, steps), "unit": "pairs/s", "value": sweep["e3"]["full"], "frac": sweep["e3"]["frac"],
            "e_sweep": sweep, "cpu_baseline": cpu,
            "algorithmic_bytes_per_pair": B.algorithmic_bytes_per_read(ctx.bins, RL, True)}


# ---------------------------------------------------------------------------------------------------------------------
# cfg5 (BASELINE.json configs[4]): IBF build, bounded
# ---------------------------------------------------------------------------------------------------------------------
def cfg5(ctx, steps=3):
    """addKmers of 1024 bins x genome_len bases per step into a fresh 16 GiB filter: device-resident rate, host-buffer rate,
    e2e incl. the download of the finished table, and identity of the words with the reference's addKmers on >= 1 Gbp."""
    torch, L, dy = ctx.torch, ctx.L, ctx.dy
    from dream_yara_b200._lib import ENC_DNA5, MEM_DEVICE, MEM_HOST, check
    from bench_support import synth
    from bench_support.build_bench import _cpu_build, _spread_bins
    bins, glen = ctx.bins, ctx.args.genome_len
    kmers_step = bins * (glen - K + 1)
    fb = dy.SeqAnBloomFilter(bins, H, K, ctx.nbits, device=ctx.local)
    chunks = [ctx.genomes, synth.genomes(bins, glen, first_bin=(ctx.rank + 1) * 8 * bins)]
    pinned, d_chunks = [], []
    for a in chunks:
        p = C.c_void_p(); check(L.dyibf_host_alloc(C.byref(p), a.size))
        ha = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=a.shape); ha[:] = a
        pinned.append(p); d_chunks.append(torch.from_numpy(ha.reshape(-1)).to(ctx.dev))
    h_off = np.arange(bins + 1, dtype=np.uint64) * glen
    h_bin = np.arange(bins, dtype=np.uint32)
    d_off = torch.from_numpy(h_off.view(np.int64)).to(ctx.dev)
    d_bin = torch.from_numpy(h_bin.view(np.int32)).to(ctx.dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def dstep(c, n_seqs=bins, binp=None):
        check(L.dyibf_add_sequences(fb._h, C.c_void_p(d_chunks[c].data_ptr()), C.c_void_p(d_off.data_ptr()), n_seqs,
                                    C.c_void_p((binp if binp is not None else d_bin).data_ptr()), ENC_DNA5, MEM_DEVICE, C.c_void_p(ctx.stream)))

    def hstep(c):
        check(L.dyibf_add_sequences(fb._h, pinned[c], _vp(h_off), bins, _vp(h_bin), ENC_DNA5, MEM_HOST, None))
    out = {"workload": "cfg5: addKmers of %d bins x %d bp per step (%d k-mers, k=%d, h=%d) into a %d GiB filter"
                       % (bins, glen, kmers_step, K, H, ctx.args.filter_gib), "unit": "k-mers/s", "steps": steps}
    try:
        # ---- identity with the reference's addKmers (N = 1): >= 1 Gbp, whole 64-bin columns, into the EMPTY filter
        if ctx.rank == 0 and ctx.world == 1 and not ctx.args.no_cpu_baseline:
            import bench as B
            cf = B.CpuFilter(bins, H, K, ctx.nbits)
            n_chk = min(bins, max(bins // 64, int(np.ceil(1.0e9 / glen / (bins // 64))) * (bins // 64)))
            chk_bins = _spread_bins(n_chk, bins)
            t_cpu, threads = _cpu_build(cf, ctx.genomes, chk_bins)
            d_chk = torch.from_numpy(chk_bins.view(np.int32)).to(ctx.dev)
            dstep(0, n_chk, d_chk)
            torch.cuda.synchronize()
            got = np.empty(ctx.P * ctx.W, dtype=np.uint64)
            check(L.dyibf_download_words(fb._h, _vp(got), 0, ctx.P * ctx.W))
            same = bool(np.array_equal(got, cf.words[: ctx.P * ctx.W]))
            out["cpu_baseline"] = {"value": n_chk * (glen - K + 1) / t_cpu, "unit": "k-mers/s", "cores": threads, "kind": cf.kind,
                                   "sample": "%d contigs x %d bp = %.2f Gbp spread over the %d 64-bin columns, %.1f s"
                                             % (n_chk, glen, n_chk * glen / 1e9, bins // 64, t_cpu),
                                   "filter_words_equal_gpu": same}
            del cf, got
            if not same:
                raise AssertionError("cfg5: GPU-built filter differs from the reference's addKmers")
        # ---- value: contig bytes resident in HBM
        dstep(1)
        fb.profile_enable(True); fb.profile(reset=True)
        ctx.barrier()
        e0.record()
        for s in range(steps):
            dstep(s & 1)
        e1.record()
        ctx.barrier()
        p = fb.profile(reset=True); fb.profile_enable(False)
        ms, bms, ams, sms = _maxr(ctx, [e0.elapsed_time(e1), p["build_ms"] / steps, p["build_apply_ms"] / steps, p["build_split2_ms"] / steps])
        value = ctx.world * kmers_step * steps / (ms / 1e3)
        table_bytes = ctx.P * ctx.W * 8
        scatter_ms = bms - ams - sms
        sliced = None
        try:
            mode = fb.last_build_mode()
            if mode == 3:
                from bench_support.build_bench import sliced_roofline
                sliced = sliced_roofline(kmers_step, H, table_bytes, bms, ams, sms, ctx.peak)
        except Exception:
            mode = -1
        if sliced is not None:
            out.update({"value": value, "ms_per_step": ms / steps, "roofline": sliced})
        else:
            alg = kmers_step * H * 64
            out.update({"value": value, "ms_per_step": ms / steps,
                        "roofline": {"bound": "hbm", "achieved": alg / (bms / 1e3) / 1e9, "peak": ctx.peak, "unit": "GB/s",
                                     "frac": alg / (bms / 1e3) / 1e9 / ctx.peak, "kernel_ms": bms, "split1_ms": scatter_ms,
                                     "split2_ms": sms, "apply_ms": ams, "algorithmic_bytes_per_launch": alg, "build_mode": mode,
                                     "note": "192 B per k-mer: the random read-modify-write sector model of SURVEY.md 8(d)"}})
        # ---- host buffers in (pinned), then the finished table out once
        hstep(0)
        ctx.barrier()
        t0 = time.perf_counter()
        for s in range(steps):
            hstep(s & 1)
        t_add = time.perf_counter() - t0
        outw = np.empty(ctx.P * ctx.W, dtype=np.uint64)
        t1 = time.perf_counter()
        check(L.dyibf_download_words(fb._h, _vp(outw), 0, ctx.P * ctx.W))
        t_dl = time.perf_counter() - t1
        ctx.barrier()
        t_add, t_all = _maxr(ctx, [t_add, t_add + t_dl])
        out["e2e"] = {"value": ctx.world * kmers_step * steps / t_all, "unit": "k-mers/s", "add_only_value": ctx.world * kmers_step * steps / t_add,
                      "h2d_bytes_per_step": ctx.world * (bins * glen + (bins + 1) * 8 + bins * 4),
                      "d2h_bytes_per_step": ctx.world * ctx.P * ctx.W * 8 // steps, "download_s": t_dl,
                      "note": "dyibf_add_sequences(host, pinned) x %d then one dyibf_download_words of the %d GiB table" % (steps, ctx.args.filter_gib)}
        del outw
    finally:
        fb.close()
        for p_ in pinned:
            L.dyibf_host_free(p_)
    return out


# ---------------------------------------------------------------------------------------------------------------------
# cfg4 (BASELINE.json configs[3]): bin-sharded, 1024 bins and a 64 GiB shard per GPU
# ---------------------------------------------------------------------------------------------------------------------
def cfg4(ctx, steps=3, shard_gib=64, n_reads=10_000_000, read_len=100):
    """N > 1: an (N x 1024)-bin, (N x 64 GiB) filter bin-sharded over the N GPUs; every GPU sees every read; lists concatenate
    by bin range.  value = reads/s of the device-resident step (pack + query + compaction), e2e = dyibf_classify with pinned
    host reads in and per-bin lists out.  Must be called after the cfg2 filter has been freed."""
    torch, dist, L, dy = ctx.torch, ctx.dist, ctx.L, ctx.dy
    from dream_yara_b200._lib import ENC_DNA5, MEM_DEVICE, check
    from bench_support import synth
    import bench as B
    world, rank = ctx.world, ctx.rank
    bins = 1024
    nbits = shard_gib * (1 << 33) + 256
    W = bins // 64
    thr = int(L.dyibf_threshold(3, K, read_len))
    free_b, _ = torch.cuda.mem_get_info()
    if free_b < (shard_gib << 30) + (12 << 30):
        return {"skipped": "only %.0f GiB free on the GPU" % (free_b / 2**30)}
    n = (n_reads // world) * world
    gen = synth.genomes(bins, ctx.args.genome_len, first_bin=(16 + rank) * bins)
    fs = dy.SeqAnBloomFilter(bins, H, K, nbits, device=ctx.local)
    fs.set_early_exit(False)           # every-k-mer pass: the roofline below counts fixed bytes per read
    try:
        fs.addSequences((gen.reshape(-1), np.arange(bins + 1, dtype=np.uint64) * ctx.args.genome_len), np.arange(bins, dtype=np.uint32))
        # every rank draws n/world reads from its own bins; one NCCL all-gather gives every GPU the whole batch
        per = n // world
        mine = torch.from_numpy(synth.reads(gen, 4, rank * per, per, read_len)).to(ctx.dev)
        d_reads = torch.empty((n, read_len), dtype=torch.uint8, device=ctx.dev)
        dist.all_gather_into_tensor(d_reads, mine)
        del mine
        hp = C.c_void_p(); check(L.dyibf_host_alloc(C.byref(hp), n * read_len))
        h_reads = np.ctypeslib.as_array(C.cast(hp, C.POINTER(C.c_uint8)), shape=(n, read_len))
        h_reads[:] = d_reads.cpu().numpy()
        h_off = np.arange(n + 1, dtype=np.uint64) * read_len
        d_off = torch.from_numpy(h_off.view(np.int64)).to(ctx.dev)
        d_rows = torch.zeros(n * W, dtype=torch.int64, device=ctx.dev)
        cap = 4 * n
        d_boff = torch.zeros(bins + 1, dtype=torch.int64, device=ctx.dev)
        d_ids = torch.zeros(cap, dtype=torch.int32, device=ctx.dev)
        npairs = C.c_uint64(0)

        def dstep():
            fs.query_device(d_reads.data_ptr(), d_off.data_ptr(), n, thr, d_rows.data_ptr(), group=2, derive_rc=True, encoding=ENC_DNA5,
                            max_seq_len=read_len, stream=ctx.stream)
            rc = L.dyibf_compact(fs._h, C.c_void_p(d_rows.data_ptr()), n, C.c_void_p(d_boff.data_ptr()), C.c_void_p(d_ids.data_ptr()),
                                 cap, C.byref(npairs), MEM_DEVICE, C.c_void_p(ctx.stream))
            if rc not in (0, -6):
                check(rc)
        dstep()
        fs.profile_enable(True); fs.profile(reset=True)
        sampler = B.ClockSampler(ctx.local)
        ctx.barrier()
        if rank == 0:
            sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            dstep()
        e1.record()
        ctx.barrier()
        clocks = sampler.stop() if rank == 0 else None
        p = fs.profile(reset=True); fs.profile_enable(False)
        ms, qms = _maxr(ctx, [e0.elapsed_time(e1), p["query_ms"] / max(p["query_launches"], 1)])
        value = n * steps / (ms / 1e3)
        per_read = B.algorithmic_bytes_per_read(bins, read_len)
        # e2e: the host-buffer call, lists out; the bin ranges' lists are concatenated on rank 0 (object gather, untimed)
        h_boff = np.zeros(bins + 1, dtype=np.uint64)
        ip = C.c_void_p(); check(L.dyibf_host_alloc(C.byref(ip), cap * 4))
        tot = C.c_uint64(0)

        def hstep():
            check(L.dyibf_classify(fs._h, hp, _vp(h_off), n, ENC_DNA5, 2, 1, thr, _vp(h_boff), ip, cap, C.byref(tot)))
        hstep()
        ctx.barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            hstep()
        dt = time.perf_counter() - t0
        ctx.barrier()
        dt = _maxr(ctx, [dt])[0]
        h_ids = np.ctypeslib.as_array(C.cast(ip, C.POINTER(C.c_uint32)), shape=(cap,))
        same = int(tot.value) == int(npairs.value) and np.array_equal(h_boff.view(np.int64), d_boff.cpu().numpy()) and \
            np.array_equal(h_ids[: tot.value].view(np.int32), d_ids[: tot.value].cpu().numpy())
        tots = _maxr(ctx, [float(tot.value)])
        pair_sum = torch.tensor([float(tot.value)], dtype=torch.float64, device=ctx.dev)
        dist.all_reduce(pair_sum)
        # rank 0's shard at its REAL shape (1024 bins, 64 GiB, P = 2^29 blocks, query_kernel<16,3,3,8>) against the reference's loop:
        # the shard's words are downloaded into the reference's own filter object on the host (needs that much RAM)
        shard_ref = {"skipped": "rank != 0"}
        if rank == 0:
            try:
                import psutil
                avail = psutil.virtual_memory().available
            except Exception:
                avail = 0
            if avail < (shard_gib << 30) + (32 << 30):
                shard_ref = {"skipped": "host has %.0f GiB available, the reference filter object needs %d GiB" % (avail / 2**30, shard_gib)}
            else:
              try:
                m = 65536
                t0 = time.perf_counter()
                cf = B.CpuFilter(bins, H, K, nbits)
                Pw = (nbits - 256) // (64 * W) * W
                check(L.dyibf_download_words(fs._h, _vp(cf.words), 0, Pw))
                t_dl = time.perf_counter() - t0
                t0 = time.perf_counter()
                rows_ref = cf.classify(B.full_layout_codes(h_reads[:m]), thr, False)
                t_cl = time.perf_counter() - t0
                rows_gpu = d_rows.view(n, W)[:m].cpu().numpy().view(np.uint64)
                wo, wi = _lists_from_rows(rows_ref, bins)
                go, gi = _restrict(h_boff, h_ids[: int(tot.value)], bins, 0, m)
                shard_ref = {"rows_equal_reference": bool(np.array_equal(rows_ref, rows_gpu)),
                             "lists_equal_reference": bool(np.array_equal(wo, go) and np.array_equal(wi, gi)),
                             "reads": m, "pairs": int(wo[-1]), "host_filter_s": t_dl, "cpu_classify_s": t_cl,
                             "what": "rank 0's shard (%d bins, %d GiB) downloaded into the reference's SeqAnBloomFilter on the host; its "
                                     "whichBins loop on the first %d reads vs the GPU rows and the ids < %d of the GPU lists" % (bins, shard_gib, m, m)}
                del cf, rows_ref
              except Exception as ex:
                shard_ref = {"skipped": "checker failed: " + repr(ex)[:200]}
        L.dyibf_host_free(hp); L.dyibf_host_free(ip)
        if not same:
            raise AssertionError("cfg4: host-mode lists != device-mode lists")
        if shard_ref.get("rows_equal_reference") is False or shard_ref.get("lists_equal_reference") is False:
            raise AssertionError("cfg4: rank 0's shard differs from the reference: %r" % shard_ref)
        return {"workload": "cfg4: synthetic %d-bin %d GiB IBF (k=%d, h=%d) bin-sharded over %d GPUs (%d bins, %d GiB per GPU), %d "
                            "single-end %d bp reads per step seen by every GPU, e=3 (threshold %d)"
                            % (bins * world, shard_gib * world, K, H, world, bins, shard_gib, n, read_len, thr),
                "value": value, "unit": "reads/s", "steps": steps, "ms_per_step": ms / steps, "clocks": clocks,
                "e2e": {"value": n * steps / dt, "unit": "reads/s", "h2d_bytes_per_step": world * (n * read_len + (n + 1) * 8),
                        "d2h_bytes_per_step": int(pair_sum.item()) * 4 + world * (bins + 1) * 8,
                        "call": "dyibf_classify per GPU on its shard: pinned host reads in, the shard's per-bin lists out"},
                "roofline": {"bound": "hbm", "achieved": n * per_read / (qms / 1e3) / 1e9, "peak": ctx.peak, "unit": "GB/s",
                             "frac": n * per_read / (qms / 1e3) / 1e9 / ctx.peak, "kernel_ms": qms,
                             "algorithmic_bytes_per_launch": n * per_read, "per": "GPU (each gathers its own 128-byte slice of every block)"},
                "read_bin_pairs_per_step": int(pair_sum.item()), "host_lists_equal_device": True, "shard_vs_reference": shard_ref}
    finally:
        fs.close()


# ---------------------------------------------------------------------------------------------------------------------
# one process, all GPUs: the dyibf_group_* entry points driven by rank 0 while the other ranks sit idle on the host
# ---------------------------------------------------------------------------------------------------------------------
def single_process(ctx, steps=4):
    """N > 1.  What the reference's ONE mapper process would see (src/d_mapper.h:911-922,338-375): one filter object over the
    N GPUs, ONE pinned host batch of N x reads, dyibf_group_classify -> one CSR.  Replication inside the process:
    dyibf_broadcast_filter (NVLink peer copies), timed.  The other ranks must not run GPU work meanwhile: they wait on the
    rendezvous store (host side), not in an NCCL kernel."""
    torch, dist, L, dy = ctx.torch, ctx.dist, ctx.L, ctx.dy
    from dream_yara_b200._lib import ENC_DNA5, check
    from bench_support import synth
    import bench as B
    world, rank, RL, bins = ctx.world, ctx.rank, ctx.RL, ctx.bins
    store = dist.distributed_c10d._get_default_store()
    key = "dyibf_single_process_done"
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    if rank != 0:
        store.wait([key])                                             # host-side wait: no kernel occupies this GPU
        return None
    out = None
    try:
        n = ctx.n * world
        g = dy.FilterGroup(bins, H, K, ctx.nbits, devices=list(range(world)), mode="read")
        g.set_early_exit(False)        # comparable with the headline (every-k-mer pass)
        m0, _, _ = g.member(0)
        goff = np.arange(bins + 1, dtype=np.uint64) * ctx.args.genome_len
        t0 = time.perf_counter()
        m0.addSequences((ctx.genomes.reshape(-1), goff), np.arange(bins, dtype=np.uint32))
        t_build = time.perf_counter() - t0
        t0 = time.perf_counter()
        g.broadcast(0)
        t_bc = time.perf_counter() - t0
        hp = C.c_void_p(); check(L.dyibf_host_alloc(C.byref(hp), n * RL))
        h_reads = np.ctypeslib.as_array(C.cast(hp, C.POINTER(C.c_uint8)), shape=(n, RL))
        h_reads[: ctx.n] = ctx.h_reads                                # rank 0's reads, then fresh ones for the rest
        for r in range(1, world):
            synth.reads(ctx.genomes, 2, r * ctx.n, ctx.n, RL, out=h_reads[r * ctx.n:(r + 1) * ctx.n])
        h_off = np.arange(n + 1, dtype=np.uint64) * RL
        cap = 4 * n
        ip = C.c_void_p(); check(L.dyibf_host_alloc(C.byref(ip), cap * 4))
        h_boff = np.zeros(bins + 1, dtype=np.uint64); tot = C.c_uint64(0)

        def step():
            check(L.dyibf_group_classify(g._g, hp, _vp(h_off), n, ENC_DNA5, 2, 1, ctx.thr, _vp(h_boff), ip, cap, C.byref(tot)))
        step(); step()
        t0 = time.perf_counter()
        for _ in range(steps):
            step()
        dt = time.perf_counter() - t0
        # the first ctx.n rows are rank 0's reads: their lists must be the restriction of the group's lists to ids < ctx.n
        h_ids = np.ctypeslib.as_array(C.cast(ip, C.POINTER(C.c_uint32)), shape=(cap,))
        o0, i0 = ctx.f.classify((ctx.h_reads.reshape(-1), np.arange(ctx.n + 1, dtype=np.uint64) * RL), ctx.thr, group=2, derive_rc=True,
                                encoding=ENC_DNA5)
        same = True
        for b in range(0, bins, 37):
            seg = h_ids[int(h_boff[b]):int(h_boff[b + 1])]
            same = same and np.array_equal(seg[seg < ctx.n], i0[int(o0[b]):int(o0[b + 1])]) and bool(np.all(np.diff(seg.astype(np.int64)) > 0))
        # ... and against the REFERENCE's loop (src/d_mapper.h:338-375, oracle/_ref) on 131 072 rows that straddle device boundaries
        # (device r filters rows [r n/G, (r+1) n/G)): the group's lists restricted to those rows, ALL bins; outside every timed region
        cuts = [ctx.n] if world == 2 else [ctx.n, (world - 1) * ctx.n]
        half = 131072 // (2 * len(cuts))
        ref_equal, t_ref, ref_err = True, time.perf_counter(), None
        try:
            cf = ctx.cpu_filter()                                     # the words of rank 0's table = what member 0 built and broadcast
            for c in cuts:
                lo, hi = c - half, c + half
                rows_ref = cf.classify(B.full_layout_codes(h_reads[lo:hi]), ctx.thr, False)
                wo, wi = _lists_from_rows(rows_ref, bins)
                go, gi = _restrict(h_boff, h_ids[: int(tot.value)], bins, lo, hi)
                ref_equal = ref_equal and bool(np.array_equal(wo, go) and np.array_equal(wi, gi))
        except Exception as ex:                                       # the checker itself failed (memory, ...): say so, keep the line
            ref_equal, ref_err = None, "checker failed: " + repr(ex)[:200]
        t_ref = time.perf_counter() - t_ref
        out = {"value": n * steps / dt, "unit": ctx.args.unit, "reads_per_call": n, "steps": steps, "ms_per_step": dt / steps * 1e3,
               "lists_equal_reference": ref_equal, "reference_check_error": ref_err,
               "reference_check": "rows [c - %d, c + %d) around the device boundaries c = %s of the %d-read batch: the group's per-bin lists "
                                  "restricted to those rows == the reference's clasifyLoadedReads loop on them, all %d bins (%.1f s of CPU, untimed)"
                                  % (half, half, cuts, n, bins, t_ref),
               "h2d_bytes_per_step": n * RL + (n + 1) * 8, "d2h_bytes_per_step": int(tot.value) * 4 + (bins + 1) * 8,
               "host_read_gbs": n * RL / (dt / steps) / 1e9,
               "filter_build_s": t_build, "broadcast_s": t_bc, "broadcast_gbs_per_replica": ctx.P * ctx.W * 8 / t_bc / 1e9,
               "broadcast_gbs_source_egress": (world - 1) * ctx.P * ctx.W * 8 / t_bc / 1e9,
               "lists_consistent_with_rank0": bool(same),
               "call": "dyibf_group_classify: ONE process, ONE pinned host batch of %d reads split over %d GPUs (read-sharded), one "
                       "CSR out; filter replicated by dyibf_broadcast_filter (cudaMemcpyPeerAsync over NVLink)" % (n, world),
               "limiter": "one host thread per GPU issues the copies; the batch is read once from host memory at host_read_gbs"}
        if not same:
            raise AssertionError("single-process group lists are not consistent with rank 0's own lists")
        if ref_equal is False:
            raise AssertionError("single-process group lists differ from the reference's on the rows around the device boundaries")
        # ---- the same entry points on a BIN-SHARDED group (one filter's word columns over the N devices), built by the group itself
        # (dyibf_group_add_sequences routes every contig to the device that owns its bin) and checked against the reference's own
        # build + filtering loop: table words of every shard and all lists
        try:
            sb = 128; tb = sb * world; nbits_t = (1 << 33) + 256; glen_b = 200_000
            Wt = tb // 64; Pt = (nbits_t - 256) // (64 * Wt); Wr = Wt // world
            gen_b = synth.genomes(tb, glen_b, first_bin=5000)
            gb = dy.FilterGroup(tb, H, K, nbits_t, devices=list(range(world)), mode="bin")
            gb.addSequences((gen_b.reshape(-1), np.arange(tb + 1, dtype=np.uint64) * glen_b), np.arange(tb, dtype=np.uint32))
            qn = 32768
            qb_ = synth.reads(gen_b, 4, 0, qn, 100)
            thr_b = int(L.dyibf_threshold(3, K, 100))
            bo, bi = gb.classify((qb_.reshape(-1), np.arange(qn + 1, dtype=np.uint64) * 100), thr_b, group=2, derive_rc=True, encoding=ENC_DNA5)
            cfb = B.CpuFilter(tb, H, K, nbits_t)
            cfb.build(gen_b)
            wo, wi = _lists_from_rows(cfb.classify(B.full_layout_codes(qb_), thr_b, False), tb)
            table = cfb.words[: Pt * Wt].reshape(Pt, Wt)
            words_eq = True
            for i in range(world):
                mi, _, _ = gb.member(i)
                words_eq = words_eq and bool(np.array_equal(mi.download_words()[: Pt * Wr].reshape(Pt, Wr), table[:, i * Wr:(i + 1) * Wr]))
            out["bin_sharded_group"] = {"lists_equal_reference": bool(np.array_equal(wo, bo) and np.array_equal(wi, bi)),
                                        "words_equal_reference": words_eq, "pairs": int(wo[-1]),
                                        "config": "%d bins over %d GPUs in one process (%d word columns each), 1 GiB table built by dyibf_group_add_sequences "
                                                  "from %d x %d bp, %d reads x 100 bp; checker: the reference's addKmers + clasifyLoadedReads loop"
                                                  % (tb, world, Wr, tb, glen_b, qn)}
            gb.close(); del cfb, table
        except Exception as ex:
            out["bin_sharded_group"] = {"lists_equal_reference": None, "words_equal_reference": None, "error": "checker failed: " + repr(ex)[:200]}
        if out["bin_sharded_group"]["lists_equal_reference"] is False or out["bin_sharded_group"]["words_equal_reference"] is False:
            raise AssertionError("bin-sharded in-process group differs from the reference: %r" % out["bin_sharded_group"])
        # ---- A/B: the same batch in a pinned buffer spread over the host's NUMA nodes (DYIBF_HOST_ALLOC_INTERLEAVE): several GPUs
        # on different sockets read it at once
        if world >= 4:
            try:
                os.environ["DYIBF_HOST_ALLOC_INTERLEAVE"] = "1"
                hp2 = C.c_void_p(); check(L.dyibf_host_alloc(C.byref(hp2), n * RL))
                os.environ.pop("DYIBF_HOST_ALLOC_INTERLEAVE", None)
                h2 = np.ctypeslib.as_array(C.cast(hp2, C.POINTER(C.c_uint8)), shape=(n, RL)); h2[:] = h_reads

                def step2():
                    check(L.dyibf_group_classify(g._g, hp2, _vp(h_off), n, ENC_DNA5, 2, 1, ctx.thr, _vp(h_boff), ip, cap, C.byref(tot)))
                step2()
                t0 = time.perf_counter()
                for _ in range(steps):
                    step2()
                dt2 = time.perf_counter() - t0
                out["interleaved_host_buffer"] = {"value": n * steps / dt2, "unit": ctx.args.unit, "host_read_gbs": n * RL / (dt2 / steps) / 1e9,
                                                  "note": "the same calls with the batch in a pinned buffer whose pages are interleaved over the "
                                                          "host's NUMA nodes (DYIBF_HOST_ALLOC_INTERLEAVE=1 at allocation)"}
                L.dyibf_host_free(hp2)
            except Exception as ex:                                    # an A/B, never a reason to lose the line
                out["interleaved_host_buffer"] = {"error": str(ex)[:200]}
            finally:
                os.environ.pop("DYIBF_HOST_ALLOC_INTERLEAVE", None)
        # ---- the same filter built by ALL devices of the group (build_filter's bins-over-threads, d_build_filter.cpp:227-260, with
        # GPUs as the threads): every sequence goes to the device that owns its bin's word column, the replicas then exchange
        # the columns over NVLink.  Must equal the table one GPU built.
        from dream_yara_b200 import dist as D
        g2 = dy.FilterGroup(bins, H, K, ctx.nbits, devices=list(range(world)), mode="read")
        bin_ids = np.arange(bins, dtype=np.uint32)
        t0 = time.perf_counter()
        g2.addSequences((ctx.genomes.reshape(-1), goff), bin_ids)
        t_gb = time.perf_counter() - t0
        t0 = time.perf_counter()
        g2.addSequences((ctx.genomes.reshape(-1), goff), bin_ids)          # second call: scratch allocated, same bits (idempotent)
        t_gb2 = time.perf_counter() - t0
        nd = ctx.P * ctx.W
        eq = True
        for i in sorted({0, world - 1}):
            a_, _, _ = g.member(i); b_, _, _ = g2.member(i)
            dev_i = torch.device("cuda", i)
            eq = eq and bool(torch.equal(D._tensor_from_ptr(a_.device_words(), nd, dev_i), D._tensor_from_ptr(b_.device_words(), nd, dev_i)))
        # ... and the REFERENCE's addKmers (oracle/_ref) for 64 bins, four in every 64-bin word column: those bins' bits of the
        # group-built table (last member: every one of its columns but its own came over NVLink) must be exactly the reference's
        eq_ref, t_rb, sel, last = None, 0.0, np.zeros(4, dtype=np.uint32), world - 1
        try:
            from bench_support.build_bench import _cpu_build, _spread_bins
            sel = _spread_bins(64, bins)
            cfn = B.CpuFilter(bins, H, K, ctx.nbits)
            t_rb, _ = _cpu_build(cfn, ctx.genomes[sel], sel)
            colmask = np.zeros(ctx.W, dtype=np.uint64)
            for b_sel in sel:
                colmask[int(b_sel) // 64] |= np.uint64(1) << np.uint64(int(b_sel) % 64)
            last = world - 1
            dev_l = torch.device("cuda", last)
            tbl = D._tensor_from_ptr(g2.member(last)[0].device_words(), nd, dev_l)
            piece = (1 << 27) // ctx.W * ctx.W                               # ~1 GiB of whole hash blocks per comparison
            eq_ref = True
            with torch.cuda.device(dev_l):
                mask_t = torch.from_numpy(colmask.view(np.int64)).to(dev_l).repeat(piece // ctx.W)
                for a0 in range(0, nd, piece):
                    a1 = min(a0 + piece, nd)
                    ref_t = torch.from_numpy(cfn.words[a0:a1].view(np.int64)).to(dev_l)
                    eq_ref = eq_ref and bool(torch.equal(tbl[a0:a1] & mask_t[: a1 - a0], ref_t))
                    del ref_t
            del cfn
        except Exception as ex:
            eq_ref = None; out["group_build_reference_error"] = "checker failed: " + repr(ex)[:200]
        kmers = bins * (ctx.args.genome_len - K + 1)
        out["group_build"] = {"words_equal_reference": eq_ref,
                              "reference_check": "bins %s...: the reference's addKmers on their %d contigs (%.1f s of CPU) == those bins' bits of "
                                                 "member %d's table, every hash block" % (list(map(int, sel[:4])), len(sel), t_rb, last),
                              "seconds_first_call": t_gb, "seconds": t_gb2, "kmers_per_s": kmers / t_gb2, "unit": "k-mers/s",
                              "single_gpu_build_s": t_build, "words_equal_single_gpu_build": eq,
                              "call": "dyibf_group_add_sequences: %d bins x %d bp from one pageable host buffer, routed by bin column to %d GPUs, "
                                      "columns exchanged over NVLink so that every replica holds the whole filter" % (bins, ctx.args.genome_len, world)}
        g2.close()
        if not eq:
            raise AssertionError("group-built filter differs from the single-GPU build")
        if eq_ref is False:
            raise AssertionError("group-built filter differs from the reference's addKmers on the sampled bins")
        g.close()
        L.dyibf_host_free(hp); L.dyibf_host_free(ip)
    finally:
        store.set(key, "1")
    return out
