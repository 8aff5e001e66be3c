#!/bin/bash
# quick iteration: tensor-core + parity tests, serial trace at v1, headline bench without extras
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu --timeout 300 -x > gpurun_out/pytest_all.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_all.log
for m in ${MATHS:-tf32 bf16}; do
  ALEGPU_TC_SERIAL=1 python tools/trace_train.py ${W:-v1} $m 2>&1 | sed "s/_GLOBAL__N__[0-9a-f_]*net_tc_cu_[0-9a-f]*:://g; s/(anonymous namespace):://g" | cut -c1-200 > gpurun_out/trace_${W:-v1}_$m.txt
  head -30 gpurun_out/trace_${W:-v1}_$m.txt
  timeout 300 python bench.py --workload ${W:-v1} --math $m --no-extras --no-cpu-baseline --min-seconds 1 > gpurun_out/c_bench_$m.json 2> gpurun_out/c_bench_$m.err
  python -c "
import json
l=json.load(open('gpurun_out/c_bench_$m.json')); print('${W:-v1} $m value', round(l['value']), 'ms', round(l['ms_per_step'],3), 'train', round(l['stages_ms']['train'],3), 'pre', round(l['stages_ms']['preprocess'],3), 'frac', round(l['roofline']['frac'],4), 'launches/step', l['gpu_launches']/l['timed_steps'])"
done
tail -n 3 gpurun_out/c_bench_*.err
