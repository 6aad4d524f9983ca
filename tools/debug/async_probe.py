"""Where does a ticketed 100 000-read batch spend its time?  Host time inside begin / end, per depth, pageable vs pinned."""
import ctypes as C
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import dream_yara_b200 as dy
from dream_yara_b200._lib import ENC_DNA5, check
from bench_support import synth

L = dy.lib()
bins, RL, rb = 1024, 150, 100_000
nbits = 16 * (1 << 33) + 256
gen = synth.genomes(bins, 1_000_000)
f = dy.SeqAnBloomFilter(bins, 3, 18, nbits)
f.addSequences((gen.reshape(-1), np.arange(bins + 1, dtype=np.uint64) * 1_000_000), np.arange(bins, dtype=np.uint32))
thr = 61
nb = 20
reads = synth.reads(gen, 2, 0, rb * nb, RL)
pageable = [np.array(reads[i * rb:(i + 1) * rb]).reshape(-1) for i in range(nb)]
hp = C.c_void_p(); check(L.dyibf_host_alloc(C.byref(hp), reads.size))
hr = np.ctypeslib.as_array(C.cast(hp, C.POINTER(C.c_uint8)), shape=reads.shape); hr[:] = reads
pinned = [hr[i * rb:(i + 1) * rb].reshape(-1) for i in range(nb)]
offs = np.arange(rb + 1, dtype=np.uint64) * RL
op = C.c_void_p(); check(L.dyibf_host_alloc(C.byref(op), offs.nbytes))
offs_p = np.ctypeslib.as_array(C.cast(op, C.POINTER(C.c_uint64)), shape=offs.shape); offs_p[:] = offs
cap = 4 * rb
bo = np.zeros(bins + 1, dtype=np.uint64); ids = np.zeros(cap, dtype=np.uint32); tot = C.c_uint64(0)
vp = lambda a: C.c_void_p(a.ctypes.data)

def run(srcs, off, depth, calls=200):
    q = []; tb = te = 0.0
    t0 = time.perf_counter()
    for i in range(calls):
        if len(q) == depth:
            a = time.perf_counter()
            check(L.dyibf_classify_end(q.pop(0), vp(bo), vp(ids), cap, C.byref(tot)))
            te += time.perf_counter() - a
        t = C.c_void_p()
        a = time.perf_counter()
        check(L.dyibf_classify_begin(f._h, vp(srcs[i % nb]), vp(off), rb, ENC_DNA5, 2, 1, thr, C.byref(t)))
        tb += time.perf_counter() - a
        q.append(t)
    while q:
        check(L.dyibf_classify_end(q.pop(0), vp(bo), vp(ids), cap, C.byref(tot)))
    dt = time.perf_counter() - t0
    return rb * calls / dt / 1e6, dt / calls * 1e3, tb / calls * 1e3, te / calls * 1e3

cases = (("pinned", pinned, offs_p), ("pageable", pageable, offs))
for name, srcs, off in cases:
    for depth in (1, 3):
        run(srcs, off, depth, 40)
        f.profile_enable(True); f.profile(reset=True)
        r = run(srcs, off, depth)
        p = f.profile(reset=True); f.profile_enable(False)
        print("%-26s depth %d: %.1f M reads/s, %.3f ms/call (begin %.3f, end %.3f)" % ((name, depth) + r),
              "kernels ms/launch: pack %.3f query %.3f compact(3 scopes) %.3f" % (p["pack_ms"] / p["pack_launches"], p["query_ms"] / p["query_launches"], 2 * p["compact_ms"] / p["compact_launches"]))
# synchronous dyibf_classify per 100 000-read batch, and one call over all reads
for name, srcs, off in cases:
    t0 = time.perf_counter()
    for i in range(100):
        check(L.dyibf_classify(f._h, vp(srcs[i % nb]), vp(off), rb, ENC_DNA5, 2, 1, thr, vp(bo), vp(ids), cap, C.byref(tot)))
    dt = time.perf_counter() - t0
    print("%-10s synchronous dyibf_classify per batch: %.1f M reads/s" % (name, rb * 100 / dt / 1e6))
offb = np.arange(rb * nb + 1, dtype=np.uint64) * RL
bo2 = np.zeros(bins + 1, dtype=np.uint64); ids2 = np.zeros(4 * rb * nb, dtype=np.uint32)
for name, big in (("pinned", hr.reshape(-1)), ("pageable", reads.reshape(-1))):
    for _ in range(2):
        t0 = time.perf_counter()
        check(L.dyibf_classify(f._h, vp(big), vp(offb), rb * nb, ENC_DNA5, 2, 1, thr, vp(bo2), vp(ids2), ids2.size, C.byref(tot)))
        dt = time.perf_counter() - t0
    print("%-10s one synchronous call of %d reads: %.1f M reads/s" % (name, rb * nb, rb * nb / dt / 1e6))
