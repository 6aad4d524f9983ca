"""Aggregate `ncu --page source --print-source cuda,sass --csv` output by CUDA source line: samples and warp instructions.
    ncu -i x.ncu-rep --page source --print-source cuda,sass --csv --kernel-name regex:NAME --launch-count 1 > x.csv
    python tools/ncu_lines.py x.csv [top]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
hi = next(i for i, r in enumerate(rows) if r and r[0] == 'Line No' and '# Samples' in r)
h = rows[hi]; si = h.index('# Samples'); ex = h.index('Instructions Executed')
fname = ""
lines = []
for r in rows[hi + 1:]:
    if r and r[0] == 'File Path':
        fname = r[1].split('/')[-1]
    if r and r[0].isdigit() and len(r) > ex:
        try:
            lines.append((fname, int(r[0]), int(r[si] or 0), int(r[ex] or 0), r[1].strip()))
        except ValueError:
            pass
tot = sum(l[2] for l in lines); totex = sum(l[3] for l in lines)
print("samples", tot, "warp instructions", totex)
for f, n, s, e, src in sorted(lines, key=lambda l: -l[3])[:top]:
    print("%-18s %4d  samples %5.1f%%  instr %5.1f%%  %s" % (f[:18], n, 100 * s / max(tot, 1), 100 * e / max(totex, 1), src[:100]))
