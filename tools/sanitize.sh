#!/bin/bash
# compute-sanitizer passes over the op tests at small sizes: memcheck on the three kernel families,
# racecheck (shared-memory hazards) on the kernels with mbarrier / TMEM pipelines and staged tiles.  Summaries go to
# gpurun_out/sanitize_*.txt; copy the last lines into profiles/ when they change.
#   usage: tools/sanitize.sh [outdir]
OUT=${1:-gpurun_out}
mkdir -p "$OUT"
K_MEM='(test_linear and (1000 or 777 or 129 or 600 or 513)) or gat_fused or seg_reduce or relative_pos or plan_bit_exact'
K_RACE='(test_linear and tf32x3 and (1000 or 777 or 129)) or gat_fused_two_relations or gat_fused_group_mean or seg_reduce_epilogue'
timeout 240 compute-sanitizer --tool memcheck --error-exitcode 7 --print-limit 20 \
    python -m pytest tests/test_gpu_ops.py tests/test_gpu_data.py -m gpu -q -x -k "$K_MEM" > "$OUT/sanitize_memcheck.txt" 2>&1
echo "memcheck exit $?" >> "$OUT/sanitize_memcheck.txt"
timeout 240 compute-sanitizer --tool racecheck --error-exitcode 7 --print-limit 20 \
    python -m pytest tests/test_gpu_ops.py -m gpu -q -x -k "$K_RACE" > "$OUT/sanitize_racecheck.txt" 2>&1
echo "racecheck exit $?" >> "$OUT/sanitize_racecheck.txt"
tail -4 "$OUT/sanitize_memcheck.txt" "$OUT/sanitize_racecheck.txt"
