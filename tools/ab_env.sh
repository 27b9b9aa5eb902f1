#!/bin/bash
# A/B of run-time knobs: each argument is "name:ENV=VAL,ENV=VAL" (name "base" alone = no knob)
for spec in "$@"; do
  name=${spec%%:*}; envs=${spec#*:}; [ "$envs" = "$spec" ] && envs=""
  env $(echo $envs | tr ',' ' ') python bench.py --steps 20 --warmup 5 --no-cpu --no-extra 2>/dev/null | python -c "
import json,sys; d=json.load(sys.stdin); print('$name', round(d['value']), round(d['ms_per_step'],4), d['step_ms']['median'], ' '.join(str(k['mean_us']) for k in d['kernels'] if 'M=20277' in k['kernel']))"
done
