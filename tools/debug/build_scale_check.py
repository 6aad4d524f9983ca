"""GPU debug: build parity at scale. For several (filter GiB, contig length, mode) compare the GPU-built columns with the C oracle."""
import sys, os, time, ctypes as C
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import dream_yara_b200 as dy
from dream_yara_b200._lib import ENC_DNA5, MEM_DEVICE, MEM_HOST, check
from oracle_py import Oracle
from bench_support import synth

L = dy.lib(); o = Oracle()
bins, h, k = 1024, 3, 18
dev = torch.device("cuda", 0)


def run(gib, glen, nseq, mode, bucket):
    nbits = gib * (1 << 33) + 256
    p = o.params(bins, h, k, nbits)
    g = synth.genomes(nseq, glen)
    w = o.new_words(p)
    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(16) as ex:
        list(ex.map(lambda b: o.add_kmers(w, p, g[b], b), range(nseq)))
    nd = p.P * p.W
    words = torch.zeros(nd, dtype=torch.int64, device=dev)
    f = dy.SeqAnBloomFilter.attach(bins, h, k, nbits, words.data_ptr(), device=0, keep=words)
    f.set_build_bucket_mib(bucket)
    off = np.arange(nseq + 1, dtype=np.uint64) * glen
    bid = np.arange(nseq, dtype=np.uint32)
    if mode == "host":
        f.addSequences((g.reshape(-1), off), bid)
    else:
        dg = torch.from_numpy(g.reshape(-1)).to(dev); do = torch.from_numpy(off.view(np.int64)).to(dev)
        db = torch.from_numpy(bid.view(np.int32)).to(dev)
        check(L.dyibf_add_sequences(f._h, C.c_void_p(dg.data_ptr()), C.c_void_p(do.data_ptr()), nseq, C.c_void_p(db.data_ptr()),
                                    ENC_DNA5, MEM_DEVICE, None))
    torch.cuda.synchronize()
    got = words.cpu().numpy().view(np.uint64)
    want = w[:nd]
    diff = got ^ want
    nz = np.flatnonzero(diff)
    missing = int(np.bitwise_count(want & ~got).sum(dtype=np.int64)); extra = int(np.bitwise_count(got & ~want).sum(dtype=np.int64))
    print("gib=%d glen=%d nseq=%d mode=%s bucket=%d: diff words %d, missing bits %d, extra bits %d, bits want %d; first idx %s"
          % (gib, glen, nseq, mode, bucket, nz.size, missing, extra, int(np.bitwise_count(want).sum(dtype=np.int64)),
             nz[:5].tolist()), flush=True)
    if nz.size:
        i = int(nz[0]); print("   word %d block %d col %d got %016x want %016x" % (i, i // p.W, i % p.W, int(got[i]), int(want[i])))
    f.close(); del words


for cfg in [(1, 100_000, 64, "device", 0), (1, 4_000_000, 16, "device", 0), (1, 4_000_000, 16, "host", 0),
            (1, 4_000_000, 64, "device", -1), (1, 4_000_000, 64, "host", -1), (16, 4_000_000, 64, "device", 0),
            (16, 4_000_000, 64, "device", -1), (16, 4_000_000, 64, "host", -1)]:
    run(*cfg)
