"""CPU: the C-ABI library loads, exports every symbol include/dyibf.h declares, and its host-only entry points
(thresholds, argument validation, error reporting) behave — no compute calls (there is no GPU here)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "dyibf.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dyibf_[a-z_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported_and_bound(dy):
    from dream_yara_b200 import _lib
    names = _declared()
    assert len(names) >= 24
    raw = C.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(raw, n), "libdyibf.so does not export " + n
        assert n in _lib.SYMBOLS, "python binding missing for " + n
    assert sorted(_lib.SYMBOLS) == names


def test_thresholds_match_reference_table(dy, golden):
    L = dy.lib()
    for a, e in enumerate(golden["thr_e"]):
        for b, k in enumerate(golden["thr_k"]):
            for c, rl in enumerate(golden["thr_L"]):
                assert L.dyibf_threshold(int(e), int(k), int(rl)) == golden["thr_table"][a, b, c], (e, k, rl)
    assert [L.dyibf_threshold(e, 18, 150) for e in range(6)] == [133, 115, 79, 61, 25, 7]
    opts = dy.DisOptions(kmerSize=18); opts.setErrorPercent(3)
    assert opts.getThreshold(150) == 61 and opts.getThreshold(100) == 29


def test_no_cpu_fallback_and_error_reporting(dy, tmp_path):
    """Without a GPU every compute entry fails loudly with DYIBF_ERR_CUDA; argument errors are reported first."""
    import torch
    L = dy.lib()
    h = C.c_void_p()
    assert L.dyibf_create(0, 3, 18, 1 << 20, 0, C.byref(h)) == -1 and b"bins" in L.dyibf_last_error()
    assert L.dyibf_create(64, 9, 18, 1 << 20, 0, C.byref(h)) == -1
    assert L.dyibf_create(64, 3, 33, 1 << 20, 0, C.byref(h)) == -1
    assert L.dyibf_create(64, 3, 18, 100, 0, C.byref(h)) == -1
    assert L.dyibf_create(4194304, 3, 18, 1 << 40, 0, C.byref(h)) == -1       # W would overflow _binIntWidth (u16)
    if not torch.cuda.is_available():
        assert L.dyibf_create(64, 3, 18, (1 << 20) + 256, 0, C.byref(h)) == -3
        assert not h.value
        with pytest.raises(dy.DyibfError):
            dy.SeqAnBloomFilter(64, 3, 18, (1 << 20) + 256)
    # file peeking is host-only: header + trailer of an sdsl bit_vector
    nbits = 70000                                                # not a multiple of 64: trailer straddles words
    words = np.zeros((nbits + 63) // 64, dtype=np.uint64)
    from oracle_py import Oracle
    o = Oracle(); p = o.params(130, 2, 16, nbits); o.save(words, p, tmp_path / "t.filter")
    assert dy.readFilterMetadata(tmp_path / "t.filter") == (130, 16)
    b, hh, k, nb = C.c_uint32(), C.c_uint32(), C.c_uint32(), C.c_uint64()
    assert L.dyibf_peek_file(str(tmp_path / "t.filter").encode(), C.byref(b), C.byref(hh), C.byref(k), C.byref(nb)) == 0
    assert (b.value, hh.value, k.value, nb.value) == (130, 2, 16, nbits)
    assert L.dyibf_peek_file(b"/nonexistent.filter", None, None, None, None) == -2
    (tmp_path / "bad.filter").write_bytes(b"\0" * 64)
    assert L.dyibf_peek_file(str(tmp_path / "bad.filter").encode(), None, None, None, None) == -4


def test_host_mirror_helpers(dy):
    cat, off, enc = dy.as_batch(["ACGT", b"NNAC"])
    assert enc == dy.ENC_ASCII and list(off) == [0, 4, 8] and bytes(cat) == b"ACGTNNAC"
    cat, off, enc = dy.as_batch([np.array([0, 1, 4], dtype=np.uint8), np.zeros(0, dtype=np.uint8)])
    assert enc == dy.ENC_DNA5 and list(off) == [0, 3, 3]
    rs = dy.ReadSet(["ACGT", "TTTT"], paired=True)
    assert len(rs) == 4 and rs.getReadsCount() == 2 and rs.getPairsCount() == 1 and rs.lengthSum() == 16


def test_cmake_project_configures_for_sm100a(tmp_path):
    """The CMake build (the reference's build system, src/CMakeLists.txt) configures here, compiles every csrc/*.cu the
    in-tree build compiles, and targets sm_100a only.  Configure step only; the compile is build.py's job."""
    import shutil
    import subprocess
    cmake = shutil.which("cmake")
    if cmake is None or shutil.which("nvcc") is None and not os.path.exists("/usr/local/cuda/bin/nvcc"):
        pytest.skip("cmake / nvcc not on this machine")
    env = dict(os.environ, CUDACXX=shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc")
    r = subprocess.run([cmake, "-S", ROOT, "-B", str(tmp_path), "-G", "Ninja"], env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    ninja = open(tmp_path / "build.ninja").read()
    from dream_yara_b200 import build as b
    for src in b.SOURCES:
        assert os.path.basename(src) in ninja, src + " missing from the CMake build"
    assert "arch=compute_100a,code=[compute_100a,sm_100a]" in ninja or "compute_100a" in ninja
    assert not re.search(r"compute_(?!100a)\d+", ninja)
    for tool in ("dyibf_build_filter", "dyibf_update_filter", "dyibf_filter_reads"):
        assert tool in ninja
