#!/bin/bash
# Final verification on one B200: whole GPU suite, smoke(), the default bench line, the reference arm.
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -q -x > $O/final_tests.log 2>&1; echo "pytest exit $?" >> $O/final_tests.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/final_smoke.log 2>&1; echo "exit $?" >> $O/final_smoke.log
python bench.py > $O/final_bench.json 2> $O/final_bench.err; echo "bench exit $?" >> $O/final_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/final_bench_ref.json 2> $O/final_bench_ref.err
tail -n 3 $O/final_tests.log $O/final_smoke.log $O/final_bench.err
head -c 2500 $O/final_bench.json; echo; head -c 600 $O/final_bench_ref.json
