#!/bin/bash
# A/B of the aggregate-first launches: resident CTAs of the wide form (library variants) and tile rows (environment)
run() {
  python bench.py --steps 10 --warmup 5 --no-cpu --no-extra > gpurun_out/ab_$1.json 2> gpurun_out/ab_$1.err || tail -3 gpurun_out/ab_$1.err
  python - "$1" <<'PY'
import json, sys
v = sys.argv[1]
d = json.load(open(f"gpurun_out/ab_{v}.json"))
print(v, round(d["value"]), round(d["ms_per_step"], 4), " | ".join(f"{k['kernel'].split('[')[1][:28]} {k['mean_us']}" for k in d["kernels"] if "aggregate-first" in k["kernel"]))
PY
}
run base
HYDRA_GNN_B200_SO=/root/repo/hydra_gnn_b200/ab/lib_occ5.so run occ5
HYDRA_GNN_B200_SO=/root/repo/hydra_gnn_b200/ab/lib_occ6.so run occ6
HYDRA_GNN_B200_SO=/root/repo/hydra_gnn_b200/ab/lib_occ6.so HG_GAT_WIDE_TILE=8 run occ6_t8
HG_GAT_AGG_TILE=16 run agg16
HG_GAT_AGG_TILE=8 run agg8
HG_GAT_WIDE_TILE=8 run wide8
