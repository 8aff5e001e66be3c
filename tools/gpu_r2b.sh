#!/bin/bash
# round 2: smoke, the default bench line (N=1), the reference arm, and the clock-sampler A/B at v0
mkdir -p gpurun_out
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/b_bench_default.json 2> gpurun_out/b_bench_default.err; echo "bench default rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/b_bench_reference.json 2> gpurun_out/b_bench_reference.err; echo "bench reference rc=$?"
python - <<'PY'
import json
l=json.load(open('gpurun_out/b_bench_default.json'))
print({k:l[k] for k in ('value','ms_per_step','dtype','gpu_launches','repeats','timed_region_s')})
print('e2e', l['e2e']['value'], l['e2e']['ms_per_step'])
print('cpu', l.get('cpu_baseline'))
print('roofline', l['roofline'])
print('hbm', l['roofline_hbm'])
print('clocks', l['clocks'])
for k,v in l['also'].items(): print(k, v)
r=json.load(open('gpurun_out/b_bench_reference.json')); print('reference arm', r['value'], r['config'], r['cpu_baseline']['sample'])
PY
tail -n 5 gpurun_out/*.err
