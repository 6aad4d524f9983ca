"""Small build run for ncu: a --gib GiB filter, 1024 bins x --glen bases (device resident), two add calls.
    ncu --set full --import-source on -k regex:'split|apply' -c 6 python tools/build_prof.py --mode 2
    ncu --set full --import-source on -k regex:'sliced_scatter|sliced_interleave' -c 4 python tools/build_prof.py --mode 3 --gib 16 --glen 1000000"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gib", type=int, default=2)
    ap.add_argument("--glen", type=int, default=500_000)
    ap.add_argument("--bucket-mib", type=int, default=4096)
    ap.add_argument("--mode", type=int, default=2, help="DYIBF_BUILD_*: 1 direct, 2 bucketed, 3 sliced")
    a = ap.parse_args()
    import ctypes as C
    import torch
    import dream_yara_b200 as dy
    from dream_yara_b200._lib import ENC_DNA5, MEM_DEVICE, check
    bins, k, h = 1024, 18, 3
    f = dy.SeqAnBloomFilter(bins, h, k, a.gib * (1 << 33) + 256)
    f.set_build_bucket_mib(a.bucket_mib)
    f.set_build_mode(a.mode)
    g = torch.randint(0, 4, (bins * a.glen,), dtype=torch.uint8, device="cuda")
    off = (torch.arange(bins + 1, dtype=torch.int64, device="cuda") * a.glen)
    bid = torch.arange(bins, dtype=torch.int32, device="cuda")
    L = dy.lib()
    f.profile_enable(True)
    for it in range(2):
        check(L.dyibf_add_sequences(f._h, C.c_void_p(g.data_ptr()), C.c_void_p(off.data_ptr()), bins, C.c_void_p(bid.data_ptr()),
                                    ENC_DNA5, MEM_DEVICE, None))
        torch.cuda.synchronize()
        p = f.profile(reset=True)
        print("call %d (mode used %d): build %.2f ms (split2 %.2f, apply / interleave %.2f) for %.2f G records" % (
            it, f.last_build_mode(), p["build_ms"], p["build_split2_ms"], p["build_apply_ms"], bins * (a.glen - k + 1) * h / 1e9))
    f.close()


if __name__ == "__main__":
    main()
