#!/bin/bash
# same-box A/B of one environment switch on the v0 / v1 bench lines: $1 = variable, $2.. = values
mkdir -p gpurun_out
var=$1; shift
for rep in 1; do for val in "$@"; do for w in v0 v1; do
  env $var=$val timeout -s KILL 200 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/ab_$w.json 2> gpurun_out/ab_$w.err
  python -c "
import json
l=json.load(open('gpurun_out/ab_$w.json')); print('$var=$val $w', round(l['ms_per_step'],4))"
done; done; done
