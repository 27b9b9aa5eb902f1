"""Summarises `ncu -i X.ncu-rep --page source --csv` (SASS view) of ONE kernel: executed warp instructions per opcode
and per straight-line block (runs of equal execution count), with the sampled stall weight of each.
usage: python tools/ncu_src_summary.py src.csv [n_blocks]"""
import collections, csv, sys

rows = list(csv.reader(open(sys.argv[1])))
h = next(i for i, r in enumerate(rows) if "Instructions Executed" in r)
ix = {k: i for i, k in enumerate(rows[h])}
data = []
for r in rows[h + 1:]:
    if len(r) <= ix["Instructions Executed"] or not r[ix["Instructions Executed"]].isdigit():
        continue
    data.append((int(r[ix["Instructions Executed"]]), int(r[ix["# Samples"]]), r[ix["Source"]].strip()))
tot = sum(d[0] for d in data)
smp = sum(d[1] for d in data)
ops, os_ = collections.Counter(), collections.Counter()
for n, s, src in data:
    t = src.split()
    op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
    ops[op] += n
    os_[op] += s
print(f"warp instructions {tot}, samples {smp}, SASS lines {len(data)}")
for op, n in ops.most_common(22):
    print(f"  {op:10s} {n:11d} {n / tot:6.3f}  samples {os_[op] / max(smp, 1):6.3f}")
blocks, start = [], 0
for i in range(1, len(data) + 1):
    if i == len(data) or abs(data[i][0] - data[i - 1][0]) > 0.02 * max(data[i][0], data[i - 1][0], 1):
        blocks.append((start, i - 1))
        start = i
k = int(sys.argv[2]) if len(sys.argv) > 2 else 14
big = sorted(blocks, key=lambda b: -sum(d[0] for d in data[b[0]:b[1] + 1]))[:k]
for a, b in sorted(big):
    n = sum(d[0] for d in data[a:b + 1]); s = sum(d[1] for d in data[a:b + 1])
    kinds = collections.Counter((d[2].split()[1] if d[2].startswith("@") else d[2].split()[0]).split(".")[0] for d in data[a:b + 1])
    print(f"  lines {a:5d}-{b:5d} x{data[a][0]:9d}  inst {n / tot:6.3f}  samples {s / max(smp, 1):6.3f}  " +
          " ".join(f"{k}:{v}" for k, v in kinds.most_common(6)))
