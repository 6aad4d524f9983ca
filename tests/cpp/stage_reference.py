#!/usr/bin/env python
"""Stage the reference's translation units for a build against THIS repository's drop-in headers.

The reference's TUs say `#include "d_bloom_filter.h"` (quoted: resolved next to the including file first) and the
reference headers carry no include guard, so an -I flag cannot shadow them.  This script makes a directory of SYMLINKS to
every file of <reference>/src (nothing is copied, nothing of the reference enters the repository; the directory is
git-ignored and gpurun-ignored), except:

  d_bloom_filter.h, d_kdx_filter.h   -> include/ of this repository (the B200 classes)
  dyibf.h, d_mapper_filter.h         -> include/ of this repository
  d_mapper.h                         the reference's file with the INTEGRATION.md §3.1 patch applied in memory: the body of
                                     clasifyLoadedReads (src/d_mapper.h:308-387) forwards to clasifyLoadedReadsB200
  wrap_<tu>.cpp                      three-line wrappers: #undef INT_WIDTH (glibc 2.39's C23 macro collides with
                                     src/d_misc_options.h:41), then #include "<tu>.cpp" — the UNMODIFIED TU

With --unpatched the same stage is made with the reference's own headers (for the comparison binaries).
"""
import argparse
import os
import re
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))

PATCH_INCLUDE = '#include "d_mapper_filter.h"\n'
PATCH_BODY = """{
    // INTEGRATION.md §3.1: one H2D of the forward reads, revcomp + IBF lookup + compaction on the GPU, one D2H of the lists
    clasifyLoadedReadsB200(mainMapper.reads.seqs, filter, disOptions,
                           IsSameType<typename TMainConfig::TSequencing, PairedEnd>::VALUE, mainMapper.timer);
}
"""


def patch_mapper(text):
    """replace the body of clasifyLoadedReads by the one call of INTEGRATION.md §3.1"""
    m = re.search(r"inline void clasifyLoadedReads\([^)]*\)\s*\n\{", text)
    if not m:
        raise SystemExit("clasifyLoadedReads not found in d_mapper.h")
    depth, i = 0, m.end() - 1
    while True:                                         # matching brace of the function body
        c = text[i]
        if c == "{":
            depth += 1
        elif c == "}":
            depth -= 1
            if depth == 0:
                break
        i += 1
    head = text[:m.end() - 1]
    # the include goes right before the function's comment banner (namespace scope, after all the types it needs)
    cut = head.rfind("// ----------------------------------------------------------------------------\n// Function clasifyLoadedReads()")
    if cut < 0:
        raise SystemExit("clasifyLoadedReads banner not found")
    return head[:cut] + PATCH_INCLUDE + head[cut:] + PATCH_BODY + text[i + 1:]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", default="/root/reference")
    ap.add_argument("--out", default=os.path.join(HERE, "_build", "src"))
    ap.add_argument("--unpatched", action="store_true")
    a = ap.parse_args()
    src = os.path.join(a.ref, "src")
    if not os.path.isdir(src):
        raise SystemExit("no reference tree at %s" % a.ref)
    os.makedirs(a.out, exist_ok=True)
    for name in os.listdir(a.out):
        os.unlink(os.path.join(a.out, name))
    ours = {} if a.unpatched else {n: os.path.join(ROOT, "include", n)
                                   for n in ("d_bloom_filter.h", "d_kdx_filter.h", "dyibf.h", "d_mapper_filter.h")}
    for name in sorted(os.listdir(src)):
        if name in ours or (name == "d_mapper.h" and not a.unpatched):
            continue
        os.symlink(os.path.join(src, name), os.path.join(a.out, name))
    for name, path in ours.items():
        os.symlink(path, os.path.join(a.out, name))
    if not a.unpatched:
        with open(os.path.join(src, "d_mapper.h")) as fh:
            text = fh.read()
        with open(os.path.join(a.out, "d_mapper.h"), "w") as fh:
            fh.write(patch_mapper(text))
    for tu in ("d_build_filter", "d_update_filter", "d_mapper"):
        with open(os.path.join(a.out, "wrap_%s.cpp" % tu), "w") as fh:
            fh.write('#include <climits>\n#include <limits.h>\n#undef INT_WIDTH\n#include "%s.cpp"\n' % tu)
    print("staged", a.out, "(unpatched)" if a.unpatched else "(B200 headers)", file=sys.stderr)


if __name__ == "__main__":
    main()
