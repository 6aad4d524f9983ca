"""Multi-GPU shardings of the IBF prefilter (one process per GPU, torch.distributed; NCCL over NVLink on B200s,
gloo in the CPU tests).  Neither sharding needs a reduction: the threshold is per bin and per read.

read-sharded (filter fits one GPU's HBM; BASELINE cfg2/cfg3)
    the filter is replicated (one NCCL broadcast of the P*W data words), rank r filters reads
    [r*n/G, (r+1)*n/G) of every batch; rows / per-bin lists concatenate in rank order, which keeps read ids
    ascending inside every bin (the reference's --threads 1 order).
bin-sharded (filter larger than one GPU's HBM; BASELINE cfg4)
    rank r owns words [r*W/G, (r+1)*W/G) of EVERY hash block, stored contiguously as a self-contained IBF with
    bins/G bins and the same number of hash blocks P — hash positions depend only on the k-mer, so the shard is
    queried with the unmodified kernels.  Reads are broadcast, every rank produces its column slice of the rows,
    slices concatenate column-wise.

The reference has no multi-process mode; the single-process semantics being preserved are those of
clasifyLoadedReads (src/d_mapper.h:308-387).
"""
import numpy as np
import torch
import torch.distributed as dist

META_BITS = 256


def read_shard_range(n_reads, rank, world):
    """Contiguous range of read (or pair) ids owned by `rank`: [a, b)."""
    per = (n_reads + world - 1) // world
    a = min(rank * per, n_reads)
    return a, min(a + per, n_reads)


def forward_slice(offsets, n_reads, paired, a, b):
    """Index ranges into the FORWARD read set for reads [a, b): SE -> [(a, b)]; PE [m1 | m2] -> two ranges."""
    if not paired:
        return [(a, b)]
    return [(a, b), (n_reads + a, n_reads + b)]


def bin_shard_params(bins, nbits, rank, world):
    """Parameters of rank's bin shard: (first_bin, shard_bins, shard_nbits, W_shard, P).  Needs world | W."""
    W = (bins + 63) // 64
    if W % world:
        raise ValueError("bin-sharding needs the number of 64-bin words per block (%d) divisible by the world size (%d)" % (W, world))
    Wr = W // world
    P = (nbits - META_BITS) // (64 * W)
    first = rank * Wr * 64
    shard_bins = max(0, min(bins, first + Wr * 64) - first)
    if shard_bins == 0:
        raise ValueError("rank %d would own no bins" % rank)
    return first, shard_bins, P * 64 * Wr + META_BITS, Wr, P


def extract_bin_shard(words, bins, nbits, rank, world):
    """Column slice of the data words ([P*W] -> [P*W/G]) for rank's shard; works on numpy arrays and tensors."""
    _, _, _, Wr, P = bin_shard_params(bins, nbits, rank, world)
    W = Wr * world
    v = words[: P * W].reshape(P, W)[:, rank * Wr:(rank + 1) * Wr]
    return v.contiguous().reshape(-1) if isinstance(v, torch.Tensor) else np.ascontiguousarray(v).reshape(-1)


def broadcast_words(words, src=0, piece=1 << 27):
    """Replicate the filter's data words (int64 tensor) from `src` in 1 GiB pieces."""
    flat = words.view(-1)
    for a in range(0, flat.numel(), piece):
        dist.broadcast(flat[a:a + piece], src=src)


def replicate_filter(filt, src=0, device=None):
    """Read-sharded set-up: every rank ends up with a SeqAnBloomFilter attached to its own copy of the words.

    `filt` is the loaded/built filter on rank `src` (None elsewhere).  Returns (filter, words_tensor)."""
    from .ibf import SeqAnBloomFilter
    rank = dist.get_rank()
    meta = [None]
    if rank == src:
        meta[0] = (filt._noOfBins, filt._noOfHashFunc, filt._kmerSize, filt._noOfBits, filt._noOfHashPos * filt._binIntWidth)
    dist.broadcast_object_list(meta, src=src)
    bins, h, k, nbits, nwords = meta[0]
    device = device if device is not None else torch.device("cuda", torch.cuda.current_device())
    words = torch.empty(nwords, dtype=torch.int64, device=device)
    if rank == src:
        from ._lib import lib
        # device-to-device copy of the source filter's words into the broadcast buffer
        torch.cuda.synchronize()
        words.copy_(_tensor_from_ptr(lib().dyibf_device_words(filt._h), nwords, device))
    broadcast_words(words, src=src)
    torch.cuda.synchronize()
    dev_index = device.index if device.index is not None else torch.cuda.current_device()
    return SeqAnBloomFilter.attach(bins, h, k, nbits, words.data_ptr(), device=dev_index, keep=words), words


def _tensor_from_ptr(ptr, n_int64, device):
    """View foreign device memory as an int64 tensor (no copy) through __cuda_array_interface__."""
    class _Wrap:
        pass
    w = _Wrap()
    w.__cuda_array_interface__ = {"shape": (n_int64,), "typestr": "<i8", "data": (int(ptr), False), "version": 3}
    return torch.as_tensor(w, device=device)


def all_gather_rows(rows_local, n_total, world=None):
    """Read-sharded: concatenate per-rank row blocks ([n_r, W] int64/uint64 tensors) in rank order -> [n_total, W]."""
    world = world or dist.get_world_size()
    W = rows_local.shape[1]
    per = (n_total + world - 1) // world
    pad = torch.zeros((per, W), dtype=rows_local.dtype, device=rows_local.device)
    pad[: rows_local.shape[0]] = rows_local
    out = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(out, pad)
    return torch.cat(out, dim=0)[:n_total]


def all_gather_columns(rows_slice, world=None):
    """Bin-sharded: every rank holds [n, W/G]; result [n, W] (column-wise concatenation in rank order)."""
    world = world or dist.get_world_size()
    out = [torch.empty_like(rows_slice) for _ in range(world)]
    dist.all_gather(out, rows_slice.contiguous())
    return torch.cat(out, dim=1)


def merge_bin_lists(offs_local, ids_local, first_read, bins):
    """Read-sharded per-bin lists: gather every rank's CSR (ids local to its shard), shift ids by the shard's first
    read and concatenate per bin in rank order -> global CSR (numpy).  Ascending order is preserved."""
    world = dist.get_world_size()
    gathered = [None] * world
    dist.all_gather_object(gathered, (np.asarray(offs_local), np.asarray(ids_local), int(first_read)))
    total = sum(int(g[0][bins]) for g in gathered)
    offs = np.zeros(bins + 1, dtype=np.uint64)
    ids = np.zeros(total, dtype=np.uint32)
    pos = 0
    for b in range(bins):
        offs[b] = pos
        for o, i, base in gathered:
            seg = i[int(o[b]):int(o[b + 1])]
            ids[pos:pos + seg.size] = seg + np.uint32(base)
            pos += seg.size
    offs[bins] = pos
    return offs, ids


def merge_bin_lists_bin_sharded(offs_local, ids_local, bins_total):
    """Bin-sharded per-bin lists: rank r's CSR covers its own contiguous bin range; concatenate in rank order."""
    world = dist.get_world_size()
    gathered = [None] * world
    dist.all_gather_object(gathered, (np.asarray(offs_local), np.asarray(ids_local)))
    offs = [np.zeros(1, dtype=np.uint64)]
    ids = []
    base = 0
    for o, i in gathered:
        offs.append(o[1:].astype(np.uint64) + np.uint64(base))
        ids.append(i)
        base += int(o[-1])
    offs = np.concatenate(offs)
    assert offs.size == bins_total + 1, (offs.size, bins_total)
    return offs, (np.concatenate(ids) if ids else np.zeros(0, dtype=np.uint32))


# ---- multi-GPU build (SURVEY §8(e) "Build") ----------------------------------------------------------------------------
# A bitwise OR over partial 16 GiB tables is not an NCCL reduction.  The exact alternative needs none: bins are
# independent columns, so rank r builds only its own bin range into its own column slice of every block — a self-contained
# IBF with bins/G bins and the same P (bin_shard_params) — and the slices are then interleaved into the replicated layout
# with all-gathers (assemble_replicated), or simply stay where they are when the filter is queried bin-sharded.
def shard_bin_ids(bin_ids, first_bin):
    """bin ids of a rank's sequences relative to its shard (bin - first_bin)"""
    return (np.asarray(bin_ids, dtype=np.int64) - int(first_bin)).astype(np.uint32)


def assemble_replicated(shard_words, bins, nbits, out=None, piece_bytes=1 << 30):
    """Every rank holds its column slice [P, W/G] (flat int64 tensor, P*W/G words); returns the full table [P*W] on every
    rank, block p = [slice_0[p] | slice_1[p] | ...] — the reference's layout.  All-gathers of about piece_bytes of the
    full table at a time; no reduction."""
    world, rank = dist.get_world_size(), dist.get_rank()
    _, _, _, Wr, P = bin_shard_params(bins, nbits, rank, world)
    W = Wr * world
    shard = shard_words.view(-1)[: P * Wr].view(P, Wr)
    full = out if out is not None else torch.empty(P * W, dtype=shard.dtype, device=shard.device)
    fv = full.view(-1)[: P * W].view(P, W)
    step = max(1, int(piece_bytes) // (8 * W))
    for a in range(0, P, step):
        b = min(P, a + step)
        mine = shard[a:b].contiguous()
        parts = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(parts, mine)
        for r, part in enumerate(parts):
            fv[a:b, r * Wr:(r + 1) * Wr] = part
    return full
