#!/bin/bash
# timeline of the forward, training A/B of the tile backward, window tests
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_ops.py -m gpu -q -x -k "window or gat" 2>&1 | tail -3
python tools/timeline.py > $O/timeline.txt 2>&1; cat $O/timeline.txt
python tools/bench_train.py --no-cpu --steps 10 > $O/train_base.json 2> $O/train_base.err
HG_GAT_BWD_TILE=1 python tools/bench_train.py --no-cpu --steps 10 > $O/train_tile.json 2> $O/train_tile.err
python - <<'PY'
import json
for f in ("train_base","train_tile"):
    try:
        t=json.loads([l for l in open(f"gpurun_out/{f}.json").read().splitlines() if l.startswith("{")][-1]); print(f,round(t["value"]),t["ms_per_step"],t["our_kernels_share_of_step"]); [print("   ",k["kernel"],k["launches_per_step"],k["mean_us"]) for k in t["kernels"][:8]]
    except Exception as e: print(f,"failed",e, open(f"gpurun_out/{f}.err").read()[-300:])
PY
