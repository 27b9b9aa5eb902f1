"""Bucket an `ncu --page source --csv --print-source sass` export into regions: share of executed warp
instructions and of stall samples, plus the top opcodes of each region.  usage: sass_regions.py file.csv [bucket] [lo hi]"""
import csv, sys

rows = list(csv.reader(open(sys.argv[1])))
h = rows[1]
ia, ie, isamp, it = h.index('Source'), h.index('Instructions Executed'), h.index('# Samples'), h.index('Thread Instructions Executed')
body = [r for r in rows[2:] if len(r) > it and r[ie].isdigit()]
B = int(sys.argv[2]) if len(sys.argv) > 2 else 100
if len(sys.argv) > 4:
    for i in range(int(sys.argv[3]), int(sys.argv[4])):
        r = body[i]
        print(i, r[ie].rjust(9), r[isamp].rjust(5), r[ia][:100])
    sys.exit(0)
tot = sum(int(r[ie]) for r in body)
ts = sum(int(r[isamp]) for r in body)
print('instructions', len(body), 'executed', tot, 'samples', ts)
for b in range(0, len(body), B):
    ch = body[b:b + B]
    e = sum(int(r[ie]) for r in ch)
    s = sum(int(r[isamp]) for r in ch)
    th = sum(int(r[it]) for r in ch)
    ops = {}
    for r in ch:
        p = r[ia].split()
        op = (p[1] if p[0].startswith('@') else p[0]).split('.')[0]
        ops[op] = ops.get(op, 0) + int(r[ie])
    top = sorted(ops.items(), key=lambda x: -x[1])[:6]
    if e > tot * 0.003:
        print(b, f'{100 * e / tot:5.1f}% inst {100 * s / ts:5.1f}% samp lanes {th / max(e, 1):4.1f}',
              ' '.join(f'{k}:{100 * v / tot:.1f}' for k, v in top))
