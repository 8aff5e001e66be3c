#!/bin/bash
# multi-GPU bench: $1 = N, $2 = workload
N=${1:-2}; W=${2:-v1}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus_multi.txt
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --workload $W --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${W}_n${N}.json 2> gpurun_out/bench_${W}_n${N}.err
echo "rc=$?"; tail -5 gpurun_out/bench_${W}_n${N}.err; cat gpurun_out/bench_${W}_n${N}.json | cut -c1-600
