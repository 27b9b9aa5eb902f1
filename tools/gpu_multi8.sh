#!/bin/bash
# N-GPU lines for the docs: forward bench, captured training step, 1M-graph job
N=${1:-8}; O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541"
timeout 600 $TR bench.py --gpus $N --no-cpu > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "bench n$N exit $?"
timeout 400 $TR tools/bench_train.py --no-cpu --steps 10 --warmup 3 > $O/train_graph_n$N.json 2> $O/train_graph_n$N.err; echo "train graph exit $?"
timeout 600 $TR tools/bench_1m.py > $O/b1m_n$N.json 2> $O/b1m_n$N.err; echo "1m exit $?"
timeout 200 $TR tools/h2d_probe.py > $O/h2d_n$N.json 2> $O/h2d_n$N.err; cat $O/h2d_n$N.json
python - $N <<'PY'
import json,sys
n=sys.argv[1]
for f in (f"bench_n{n}", f"train_graph_n{n}", f"b1m_n{n}"):
    try:
        d=json.loads([l for l in open(f"gpurun_out/{f}.json").read().splitlines() if l.startswith("{")][-1])
        print(f, round(d["value"]), d.get("ms_per_step", d.get("ms_per_job")), d.get("e2e",{}).get("value"), d.get("ms_all_repeats"), d.get("clocks"))
    except Exception as e: print(f,"failed",e, open(f"gpurun_out/{f}.err").read()[-400:])
PY
