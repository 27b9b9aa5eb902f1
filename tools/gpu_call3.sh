#!/bin/bash
# full GPU suite + default bench (with extras) + NT window A/B + ncu of the objects window launch
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -x > $O/gpu_tests.log 2>&1; echo "exit $?" >> $O/gpu_tests.log; tail -4 $O/gpu_tests.log
python bench.py > $O/bench_full.json 2> $O/bench_full.err; echo "bench exit $?"; tail -c 300 $O/bench_full.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref exit $?"
python tools/bench_nt.py --only nt_het_gat --check 0 > $O/nt_win.json 2> $O/nt_win.err
HG_GAT_WINDOW=0 python tools/bench_nt.py --only nt_het_gat --check 0 > $O/nt_v6.json 2> $O/nt_v6.err
python tools/bench_train.py --no-cpu --steps 10 > $O/train.json 2> $O/train.err
python tools/prof_gat.py > $O/prof_gat.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_gat_window -s 10 -c 1 -o $O/gat_window python tools/prof_gat.py > $O/ncu_full.log 2>&1
python - <<'PY'
import json
d=json.load(open("gpurun_out/bench_full.json"))
print("value",round(d["value"]),"ms",d["ms_per_step"],"e2e",round(d["e2e"]["value"]),"step_ms",d.get("step_ms"))
print("roofline",d["roofline"]); print("agg",d.get("roofline_aggregation")); print("extra",json.dumps(d.get("extra"))[:1500]); print("cpu",d.get("cpu_baseline"))
for f in ("nt_win","nt_v6"):
    try:
        m=json.load(open(f"gpurun_out/{f}.json"))["models"]["nt_het_gat"]; print(f, m["ms_per_forward_graph"], m["graphs_per_s"]); [print("   ",k["kernel"],k["launches_per_step"],k["mean_us"],k["gbs"]) for k in m["kernels"][:6]]
    except Exception as e: print(f,"failed",e)
try:
    t=json.load(open("gpurun_out/train.json")); print("train",round(t["value"]),t["ms_per_step"]); [print("   ",k["kernel"],k["launches_per_step"],k["mean_us"]) for k in t["kernels"][:6]]
except Exception as e: print("train failed",e)
PY
