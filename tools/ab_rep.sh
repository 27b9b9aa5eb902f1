#!/bin/bash
# repeated A/B of the step time: in-tree library vs hydra_gnn_b200/ab/lib_$1.so, alternating, $2 rounds
for i in $(seq 1 ${2:-3}); do
  for v in base $1; do
    if [ $v = base ]; then unset HYDRA_GNN_B200_SO; else export HYDRA_GNN_B200_SO=/root/repo/hydra_gnn_b200/ab/lib_$v.so; fi
    python bench.py --steps 20 --warmup 5 --no-cpu --no-extra 2>/dev/null | python -c "
import json,sys; d=json.load(sys.stdin); print('$v', round(d['value']), round(d['ms_per_step'],4), d['step_ms']['median'])"
  done
done
