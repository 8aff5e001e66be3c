#!/bin/bash
# round 2, first GPU session: tf32 engine + network parity, then the whole GPU suite, short bench lines
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tensorcore.py -q -m gpu --timeout 180 -k "tf32" > gpurun_out/pytest_tf32.log 2>&1; echo "pytest tf32 rc=$?"
tail -40 gpurun_out/pytest_tf32.log
timeout 1200 python -m pytest tests -m gpu -q --timeout 300 --deselect tests/test_gpu_tensorcore.py -x > gpurun_out/pytest_rest.log 2>&1; echo "pytest rest rc=$?"
tail -5 gpurun_out/pytest_rest.log
timeout 600 python -m pytest tests/test_gpu_tensorcore.py -q -m gpu --timeout 180 -k "not tf32" > gpurun_out/pytest_bf16.log 2>&1; echo "pytest bf16 rc=$?"
tail -5 gpurun_out/pytest_bf16.log
for m in tf32 bf16; do
  timeout 300 python bench.py --workload v1 --math $m --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/a_bench_v1_$m.json 2> gpurun_out/a_bench_v1_$m.err; echo "bench v1 $m rc=$?"
  python -c "
import json
l=json.load(open('gpurun_out/a_bench_v1_$m.json')); print('v1 $m', l['value'], l['ms_per_step'], l['e2e']['value'], l['roofline_tensor'])"
done
tail -n 5 gpurun_out/a_bench_*.err
