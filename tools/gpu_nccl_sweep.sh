#!/bin/bash
# v0 at N GPUs under a few NCCL settings (which algorithm/protocol hides best behind the backward)
N=${1:-8}
mkdir -p gpurun_out
run() {
  tag=$1; shift
  env "$@" timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --workload v0 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/nccl_$tag.json 2> gpurun_out/nccl_$tag.err
  python -c "
import json
l=json.load(open('gpurun_out/nccl_$tag.json')); print('$tag', l['ms_per_step'], l['value'], l['e2e']['ms_per_step'])" 2>&1 | tail -1
}
run default X=1
run nvls NCCL_ALGO=NVLS
run tree NCCL_ALGO=Tree
run ll128 NCCL_PROTO=LL128
run ll NCCL_PROTO=LL
run nooverlap ALEGPU_NO_COMM_OVERLAP=1
