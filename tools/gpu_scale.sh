#!/bin/bash
# the driver's multi-GPU launch of bench.py on N GPUs of this box:  bash tools/gpu_scale.sh N [extra bench args]
N=$1; shift
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus $N --steps 5 --warmup 3 "$@" > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
echo "rc=$?"
python - <<PY
import json
for ln in open('gpurun_out/scale_n$N.json'):
    if ln.startswith('{'):
        l=json.loads(ln)
        print('N', l['n_gpus'], l['config']['workload'], l['dtype'], 'value', round(l['value']), 'ms', round(l['ms_per_step'],2), 'e2e', round(l['e2e']['value']), 'ms', round(l['e2e']['ms_per_step'],1))
        print('  solo', l['also'].get('single_gpu_same_slice'), 'stages', l['stages_ms'])
PY
tail -n 5 gpurun_out/scale_n$N.err
