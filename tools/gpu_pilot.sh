#!/bin/bash
# short guarded run for a risky kernel change: tensor-core parity tests, then quick traces/benches
mkdir -p gpurun_out
timeout -s KILL 150 python -m pytest tests/test_gpu_tensorcore.py -q -m gpu --timeout 60 -x > gpurun_out/pytest_pilot.log 2>&1; rc=$?; echo "pytest pilot rc=$rc"
tail -5 gpurun_out/pytest_pilot.log
[ $rc -ne 0 ] && exit 0
for w in v0 v1; do
  timeout -s KILL 120 python tools/trace_train.py $w 2>&1 | sed 's/_GLOBAL__N__[0-9a-f_]*net_tc_cu_[0-9a-f]*:://g' > gpurun_out/trace_$w.txt
  timeout -s KILL 200 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/q_bench_$w.json 2> gpurun_out/q_bench_$w.err
  python -c "
import json
l=json.load(open('gpurun_out/q_bench_$w.json')); print('$w', l['value'], l['ms_per_step'], l['e2e']['value'], l['roofline']['frac'])"
done
