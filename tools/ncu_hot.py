"""Print the hottest SASS instructions (stall samples) of a kernel from `ncu --page source --csv` output."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
frac = float(sys.argv[2]) if len(sys.argv) > 2 else 0.012
hi = next(i for i, r in enumerate(rows) if 'Address' in r)
hdr = rows[hi]; si = hdr.index('# Samples'); src = hdr.index('Source'); ex = hdr.index('Instructions Executed')
body = []
for r in rows[hi + 1:]:
    if 'Address' in r or (r and r[0] == 'Kernel Name'):
        break                                   # next kernel instance
    if len(r) > max(si, src, ex):
        body.append(r)
tot = sum(int(r[si] or 0) for r in body)
print('total samples', tot, 'instructions', len(body))
for n, r in enumerate(body):
    s = int(r[si] or 0)
    if s > tot * frac:
        print(n, s, round(100 * s / tot, 1), r[ex], r[src][:100])
