#!/bin/bash
# multi-GPU checks: forward bench with the pipelined logit gather, training step captured with the NCCL all-reduce inside
N=${1:-2}; O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541"
timeout 600 $TR bench.py --gpus $N --no-cpu > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "bench n$N exit $?"; tail -c 600 $O/bench_n$N.err
timeout 300 $TR tools/bench_train.py --no-cpu --steps 10 --warmup 3 > $O/train_eager_n$N.json 2> $O/train_eager_n$N.err; echo "train eager exit $?"
timeout 300 $TR tools/bench_train.py --no-cpu --steps 10 --warmup 3 --graph > $O/train_graph_n$N.json 2> $O/train_graph_n$N.err; echo "train graph exit $?"; tail -c 400 $O/train_graph_n$N.err
python - $N <<'PY'
import json,sys
n=sys.argv[1]
for f in (f"bench_n{n}", f"train_eager_n{n}", f"train_graph_n{n}"):
    try:
        d=json.load(open(f"gpurun_out/{f}.json")); print(f, round(d["value"]), round(d["ms_per_step"],4), d.get("e2e",{}).get("value"))
    except Exception as e: print(f,"failed",e)
PY
