"""Per-kernel SASS opcode histogram of libhydra_gnn_b200.so (cuobjdump -sass): the mnemonics that show which hardware path
a kernel uses (B200_PROFILING.md: UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG/UTMASTG/UBLKCP = TMA / bulk copies,
HMMA = legacy mma.sync, SYNCS = mbarrier ops, LDGSTS = cp.async).
    python tools/sass_hist.py > profiles/r02_sass_opcodes.txt"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "hydra_gnn_b200", "libhydra_gnn_b200.so")
KEYS = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "SYNCS", "HMMA", "LDGSTS", "LDG.E.128", "LDG.E.64",
        "LDS.128", "STS.128", "STG.E.128", "FFMA", "MUFU", "ATOMG", "REDG", "RED.E", "BAR.SYNC"]
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
fn, hist, total = None, {}, collections.Counter()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = m.group(1)
        hist[fn] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if fn and m:
        op = m.group(1)
        total[fn] += 1
        for k in KEYS:
            if op == k or op.startswith(k + ".") or (k in ("UTCHMMA", "HMMA", "MUFU", "SYNCS") and op.startswith(k)):
                hist[fn][k] += 1
print(f"# {os.path.basename(so)}: SASS opcode counts per kernel (static instruction counts, sm_100a)")
print("# " + " ".join(KEYS))
for fn in sorted(hist, key=lambda f: -total[f]):
    name = re.sub(r"\(anonymous namespace\)::|hg::|\(hg::<unnamed>::\w+\)|\(int\)", "", demangle(fn))
    name = re.sub(r"^void ", "", name)[:90]
    cells = " ".join(f"{k}:{hist[fn][k]}" for k in KEYS if hist[fn][k])
    print(f"{name:90s} n={total[fn]:5d}  {cells}")
