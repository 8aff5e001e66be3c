#!/bin/bash
# Round evidence for profiles/: launch list of the default bench command + ncu --set full of the
# dominant memory-bound kernel (preprocess) and of the tensor-core GEMMs of one minibatch step.
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/plain_default.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv \
    --log-file gpurun_out/launches_default.csv $CMD > gpurun_out/ncu_default.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/launches_default.csv
ncu --set full --clock-control none --import-source on -k regex:preprocess_rollout_pipe -s 1 -c 1 -f \
    -o gpurun_out/r01_preprocess_v0 $CMD > gpurun_out/ncu_pre_v0.log 2>&1
echo "preprocess v0 rc=$?"
CMD1="python bench.py --workload v1 --steps 1 --warmup 1 --no-cpu-baseline"
$CMD1 > gpurun_out/plain_v1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:preprocess_rollout_pipe -s 1 -c 1 -f \
    -o gpurun_out/r01_preprocess_v1 $CMD1 > gpurun_out/ncu_pre_v1.log 2>&1
echo "preprocess v1 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 11 -c 11 -f \
    -o gpurun_out/r01_gemm_v1 $CMD1 > gpurun_out/ncu_gemm_v1.log 2>&1
echo "gemm v1 rc=$?"
tail -2 gpurun_out/ncu_default.log | cut -c1-200
