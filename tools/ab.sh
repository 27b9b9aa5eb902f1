#!/bin/bash
# A/B harness: bench the in-tree library ("base") and every build_ab/lib_*.so given as argument names.
for v in "$@"; do
  if [ "$v" = base ]; then unset HYDRA_GNN_B200_SO; else export HYDRA_GNN_B200_SO=/root/repo/hydra_gnn_b200/ab/lib_$v.so; fi
  python bench.py --steps 10 --warmup 5 --no-cpu --no-extra > gpurun_out/bench_$v.json 2> gpurun_out/bench_$v.err || tail -3 gpurun_out/bench_$v.err
  python - "$v" <<'PY'
import json, sys
v = sys.argv[1]
d = json.load(open(f"gpurun_out/bench_{v}.json"))
print(v, round(d["value"]), round(d["ms_per_step"], 4))
for k in d["kernels"]:
    if k["mean_us"] * k["launches_per_step"] > 40:
        print("   ", k["kernel"], k["launches_per_step"], k["mean_us"], k["gbs"])
PY
done
