#!/bin/bash
# Round-2 first GPU call: whole GPU suite, the no-alloc test alone, bench + launch list + full capture of the
# top kernel, first runs of the two opt-in training kernels, sanitizer passes.
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -q -rxX -x > $O/gpu_tests.log 2>&1; echo "pytest exit $?" >> $O/gpu_tests.log
python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/zz_alone.log 2>&1; echo "pytest exit $?" >> $O/zz_alone.log
python bench.py > $O/bench.json 2> $O/bench.err; echo "bench exit $?" >> $O/bench.err
timeout 300 python tools/check_wgrad_tc.py > $O/wgrad_tc.log 2>&1; echo "exit $?" >> $O/wgrad_tc.log
timeout 300 python tools/check_gat_bwd_tile.py > $O/bwd_tile.log 2>&1; echo "exit $?" >> $O/bwd_tile.log
python tools/prof_gat.py > $O/prof_gat.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv \
   python bench.py --steps 2 --warmup 3 --no-cpu --no-graph > $O/ncu_launch.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_gat_fused -s 12 -c 4 -o $O/gat_fused_base \
   python tools/prof_gat.py > $O/ncu_full.log 2>&1
timeout 500 bash tools/sanitize.sh $O > $O/sanitize.log 2>&1
tail -3 $O/gpu_tests.log $O/zz_alone.log $O/bench.err $O/wgrad_tc.log $O/bwd_tile.log
cat $O/bench.json | head -c 3000
