"""CPU, world_size 2, gloo: the host-side logic of both multi-GPU shardings (dream_yara_b200/dist.py).
The per-rank filter work is done by the oracle here (no GPU); what is under test is the partitioning of reads and
bins, the replication of the filter words and the order-preserving merges — the merged result must equal the
single-process result bit for bit."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE); sys.path.insert(0, os.path.dirname(HERE))


def _workload(paired):
    from oracle_py import Oracle, full_layout
    o = Oracle()
    rng = np.random.default_rng(42)
    bins, h, k, nbits = 256, 3, 18, (1 << 22) + 256
    p = o.params(bins, h, k, nbits)
    words = o.new_words(p)
    contigs = [rng.integers(0, 4, 1500).astype(np.uint8) for _ in range(bins)]
    for b, c in enumerate(contigs):
        o.add_kmers(words, p, c, b)
    n = 102 if paired else 101
    fwd = []
    for i in range(n * (2 if paired else 1)):
        c = contigs[int(rng.integers(0, bins))]
        st = int(rng.integers(0, 1400)); r = c[st:st + 100].copy()
        if i % 7 == 0:
            r[0] = 4
        fwd.append(r)
    return o, p, words, fwd, n


def _layout(o, seqs):
    from oracle_py import full_layout
    cat = np.concatenate(seqs); off = np.zeros(len(seqs) + 1, dtype=np.uint64); off[1:] = np.cumsum([len(s) for s in seqs])
    return full_layout(cat, off, o)


def _worker(rank, world, port, paired, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from dream_yara_b200 import dist as D
        o, p, words, fwd, n = _workload(paired)
        thr = 20
        # single-process answer
        want_rows = o.classify(words, p, *_layout(o, fwd), paired, thr)
        want_offs, want_ids = o.compact(want_rows, p.bins)

        # ---- filter replication: rank 0 owns the words, everyone else receives them
        data = torch.from_numpy(words[: p.P * p.W].view(np.int64).copy())
        mine = data.clone() if rank == 0 else torch.zeros_like(data)
        D.broadcast_words(mine, src=0, piece=1000)
        assert torch.equal(mine, data)

        # ---- read-sharded
        a, b = D.read_shard_range(n, rank, world)
        local = []
        for lo, hi in D.forward_slice(None, n, paired, a, b):
            local += fwd[lo:hi]
        rows_local = o.classify(mine.numpy().view(np.uint64), p, *_layout(o, local), paired, thr) if b > a else np.zeros((0, p.W), np.uint64)
        rows = D.all_gather_rows(torch.from_numpy(rows_local.view(np.int64)), n).numpy().view(np.uint64)
        assert np.array_equal(rows, want_rows)
        lo_, li_ = o.compact(rows_local, p.bins)
        offs, ids = D.merge_bin_lists(lo_, li_, a, p.bins)
        assert np.array_equal(offs, want_offs) and np.array_equal(ids, want_ids)

        # ---- bin-sharded: own a column slice of every block as a self-contained filter, all reads
        first, sb, snbits, Wr, P = D.bin_shard_params(p.bins, p.nbits, rank, world)
        assert P == p.P and Wr * world == p.W and first == rank * 128 and sb == 128
        shard = D.extract_bin_shard(words, p.bins, p.nbits, rank, world)
        sp = o.params(sb, p.h, p.k, snbits)
        assert sp.P == p.P and sp.W == Wr
        sw = o.new_words(sp); sw[: shard.size] = shard
        rows_slice = o.classify(sw, sp, *_layout(o, fwd), paired, thr)
        rows2 = D.all_gather_columns(torch.from_numpy(rows_slice.view(np.int64))).numpy().view(np.uint64)
        assert np.array_equal(rows2, want_rows)
        so, si = o.compact(rows_slice, sb)
        offs2, ids2 = D.merge_bin_lists_bin_sharded(so, si, p.bins)
        assert np.array_equal(offs2, want_offs) and np.array_equal(ids2, want_ids)

        # ---- sharded build: every rank adds only its own bins to its own column slice, slices are interleaved
        bw = o.new_words(sp)
        rng = np.random.default_rng(42)
        contigs = [rng.integers(0, 4, 1500).astype(np.uint8) for _ in range(p.bins)]       # the contigs of _workload
        mine_bins = np.arange(first, first + sb)
        for b_, lb in zip(mine_bins, D.shard_bin_ids(mine_bins, first)):
            o.add_kmers(bw, sp, contigs[int(b_)], int(lb))
        assert np.array_equal(bw[: shard.size], shard)                                   # = the slice of the one-process build
        full = D.assemble_replicated(torch.from_numpy(bw[: sp.P * sp.W].view(np.int64).copy()), p.bins, p.nbits, piece_bytes=4096)
        assert np.array_equal(full.numpy().view(np.uint64), words[: p.P * p.W])
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("paired", [False, True])
def test_shardings_world2_gloo(paired):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000) + (1 if paired else 0)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, paired, q)) for r in range(2)]
    for p_ in procs:
        p_.start()
    res = [q.get(timeout=300) for _ in procs]
    for p_ in procs:
        p_.join(timeout=60)
    assert all(r[1] == "ok" for r in res), res


def test_shard_arithmetic():
    from dream_yara_b200 import dist as D
    assert [D.read_shard_range(10, r, 4) for r in range(4)] == [(0, 3), (3, 6), (6, 9), (9, 10)]
    assert D.read_shard_range(2, 3, 4) == (2, 2)
    assert D.forward_slice(None, 50, True, 5, 9) == [(5, 9), (55, 59)]
    # BASELINE cfg4: 8192 bins, 512 GiB over 8 GPUs -> 1024 bins, 64 GiB each, same P
    nbits = 512 * (1 << 33) + 256
    first, sb, snb, Wr, P = D.bin_shard_params(8192, nbits, 3, 8)
    assert (first, sb, Wr, P) == (3072, 1024, 16, 1 << 29) and snb == 64 * (1 << 33) + 256
    with pytest.raises(ValueError):
        D.bin_shard_params(130, (1 << 20) + 256, 0, 2)           # W = 3 not divisible by 2
    first, sb, _, _, _ = D.bin_shard_params(1000, (1 << 24) + 256, 1, 2)   # last shard holds the ragged tail
    assert (first, sb) == (512, 488)
