#!/bin/bash
# One GPU-box session: parity tests, smoke, short bench lines.  Output -> gpurun_out/
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus.txt 2>&1
nproc >> gpurun_out/gpus.txt
timeout 900 python -m pytest tests -m gpu -q -s --timeout 300 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
for w in v0 v1; do
  for m in ${BENCH_MATHS:-bf16}; do
    timeout 600 python bench.py --workload $w --math $m --steps 10 --warmup 3 > gpurun_out/bench_${w}_${m}.json 2> gpurun_out/bench_${w}_${m}.err; echo "bench $w $m rc=$?"
  done
done
tail -n 8 gpurun_out/pytest_gpu.log; grep "grad report" gpurun_out/pytest_gpu.log; cat gpurun_out/smoke.log
for f in gpurun_out/bench_*.json; do echo $f; python -c "
import json,sys
l=json.load(open('$f'))
print({k:l[k] for k in ('value','ms_per_step','dtype','gpu_launches')}, l['e2e']['value'], l.get('cpu_baseline',{}).get('value'), l['roofline']['frac'])
"; done
tail -n 3 gpurun_out/bench_*.err
