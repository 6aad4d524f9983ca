"""cfg5-shaped build step (1024 bins x 4 Mbp into 16 GiB): kernel times of the three build modes, device-resident input."""
import ctypes as C, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import dream_yara_b200 as dy
from dream_yara_b200._lib import ENC_DNA5, MEM_DEVICE, MEM_HOST, check
from bench_support import synth
L = dy.lib()
bins, glen = 1024, 4_000_000
nbits = 16 * (1 << 33) + 256
gen = synth.genomes(bins, glen)
d = torch.from_numpy(gen.reshape(-1)).cuda()
off = np.arange(bins + 1, dtype=np.uint64) * glen
d_off = torch.from_numpy(off.view(np.int64)).cuda()
ids = np.arange(bins, dtype=np.uint32)
d_ids = torch.from_numpy(ids.view(np.int32)).cuda()
hp = C.c_void_p(); check(L.dyibf_host_alloc(C.byref(hp), gen.size))
h = np.ctypeslib.as_array(C.cast(hp, C.POINTER(C.c_uint8)), shape=gen.shape); h[:] = gen
st = torch.cuda.current_stream().cuda_stream
kmers = bins * (glen - 17)
ref = None
for mode, name in ((3, "sliced"), (2, "bucketed")):
    f = dy.SeqAnBloomFilter(bins, 3, 18, nbits)
    f.set_build_mode(mode)
    def dstep():
        check(L.dyibf_add_sequences(f._h, C.c_void_p(d.data_ptr()), C.c_void_p(d_off.data_ptr()), bins, C.c_void_p(d_ids.data_ptr()), ENC_DNA5, MEM_DEVICE, C.c_void_p(st)))
    dstep(); torch.cuda.synchronize()
    f.profile_enable(True); f.profile(reset=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): dstep()
    e1.record(); torch.cuda.synchronize()
    p = f.profile(reset=True); f.profile_enable(False)
    ms = e0.elapsed_time(e1) / 3
    print("%-9s device: %.1f ms/step = %.1f G k-mers/s | pack %.1f build %.1f (apply/interleave %.1f, split2 %.1f) launches %d" % (
        name, ms, kmers / ms / 1e6, p["pack_ms"] / 3, p["build_ms"] / 3, p["build_apply_ms"] / 3, p["build_split2_ms"] / 3, p["build_launches"]), flush=True)
    for label, src in (("pinned", hp), ("pageable", C.c_void_p(gen.ctypes.data))):
        for mib in (256, 64, 1024):
            check(L.dyibf_set_option(f._h, 3, mib))
            f.profile_enable(True); f.profile(reset=True)
            t0 = time.perf_counter()
            for _ in range(2):
                check(L.dyibf_add_sequences(f._h, src, C.c_void_p(off.ctypes.data), bins, C.c_void_p(ids.ctypes.data), ENC_DNA5, MEM_HOST, None))
            dt = (time.perf_counter() - t0) / 2
            p = f.profile(reset=True); f.profile_enable(False)
            print("%-9s host %-8s chunk %4d MiB: %.1f ms/step = %.1f G k-mers/s (mode used %d) | pack %.1f build %.1f (apply %.1f split2 %.1f) launches %d" % (
                name, label, mib, dt * 1e3, kmers / dt / 1e9, f.last_build_mode(), p["pack_ms"] / 2, p["build_ms"] / 2, p["build_apply_ms"] / 2, p["build_split2_ms"] / 2, p["build_launches"]), flush=True)
    check(L.dyibf_set_option(f._h, 3, 256))
    w = dy.dist._tensor_from_ptr(f.device_words(), (nbits - 256) // 64, torch.device("cuda", 0)) if False else None
    from dream_yara_b200 import dist as D
    w = D._tensor_from_ptr(f.device_words(), (nbits - 256) // 64, torch.device("cuda", 0))
    if ref is None: ref = w.clone()
    else: print("tables equal:", bool(torch.equal(ref, w)))
    t0 = time.perf_counter(); out = f.download_words(); dt = time.perf_counter() - t0
    print("download 16 GiB: %.2f s = %.1f GB/s" % (dt, out.nbytes / dt / 1e9))
    del out
    f.close()
