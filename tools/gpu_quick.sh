#!/bin/bash
# quick engine iteration: tensor-core parity tests + per-kernel traces at v0/v1 + short bench lines
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tensorcore.py -q -m gpu --timeout 120 -x > gpurun_out/pytest_tc.log 2>&1; echo "pytest tc rc=$?"
tail -3 gpurun_out/pytest_tc.log
for w in v0 v1; do
  python tools/trace_train.py $w 2>&1 | sed 's/_GLOBAL__N__[0-9a-f_]*net_tc_cu_[0-9a-f]*:://g' > gpurun_out/trace_$w.txt
  timeout 300 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/q_bench_$w.json 2> gpurun_out/q_bench_$w.err
  python -c "
import json
l=json.load(open('gpurun_out/q_bench_$w.json')); print('$w', l['value'], l['ms_per_step'], l['e2e']['value'], l['roofline']['frac'])"
done
