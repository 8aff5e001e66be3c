#!/bin/bash
# launch list (per-kernel device time) of one bench command; $1 = workload, $2 = math
W=${1:-v0}; M=${2:-bf16}
mkdir -p gpurun_out
CMD="python bench.py --workload $W --math $M --steps 2 --warmup 1 --no-cpu-baseline --no-graph"
$CMD > gpurun_out/plain_${W}_${M}.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv \
    --log-file gpurun_out/launches_${W}_${M}.csv $CMD > gpurun_out/ncu_${W}_${M}.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/ncu_${W}_${M}.log; wc -l gpurun_out/launches_${W}_${M}.csv
