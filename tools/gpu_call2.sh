#!/bin/bash
# window kernel: tests, bench, isolated timing, one full capture of the objects-layer launch
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_ops.py -m gpu -q -x -k "window or plan or gat" > $O/t_ops.log 2>&1; echo "exit $?" >> $O/t_ops.log
tail -5 $O/t_ops.log
timeout 1200 python -m pytest tests -m gpu -q -x > $O/gpu_tests.log 2>&1; echo "exit $?" >> $O/gpu_tests.log
tail -5 $O/gpu_tests.log
python bench.py --no-cpu --no-extra > $O/bench_win.json 2> $O/bench_win.err; echo "bench exit $?"
HG_AGG_FIRST_WIDTH=1 python bench.py --no-cpu --no-extra > $O/bench_pf.json 2> $O/bench_pf.err
python tools/prof_gat.py > $O/prof_gat.log 2>&1 && cat $O/prof_gat.log &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_gat_window -s 10 -c 1 -o $O/gat_window \
   python tools/prof_gat.py > $O/ncu_full.log 2>&1
python - <<'PY'
import json
for f in ("bench_win","bench_pf"):
    try:
        d=json.load(open(f"gpurun_out/{f}.json")); print(f, round(d["value"]), d["ms_per_step"], d["roofline"])
    except Exception as e: print(f, "failed", e)
PY
