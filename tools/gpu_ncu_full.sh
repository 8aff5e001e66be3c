#!/bin/bash
# ncu --set full of selected kernels of one bench command.
# $1 workload $2 math $3 kernel regex $4 skip $5 count $6 tag $7 extra ncu flags (e.g. "--cache-control none")
W=${1:-v0}; M=${2:-bf16}; K=${3:-gemm_tc}; S=${4:-0}; C=${5:-3}; TAG=${6:-prof}; X=${7:-}
mkdir -p gpurun_out
CMD="python bench.py --workload $W --math $M --steps 1 --warmup 1 --no-cpu-baseline --no-graph"
$CMD > gpurun_out/plain_full.log 2>&1 &&
ncu --set full --clock-control none --import-source on $X -k regex:$K -s $S -c $C -f -o gpurun_out/$TAG $CMD > gpurun_out/ncu_full.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/ncu_full.log | cut -c1-300
