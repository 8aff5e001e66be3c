#!/bin/bash
# Round evidence for profiles/: launch list of the default bench command + ncu --set full of the
# dominant memory-bound kernel (preprocess) and of the tensor-core GEMMs of one minibatch step.
# $1 = tag prefix (default r01d)
TAG=${1:-r01e}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/plain_default.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv \
    --log-file gpurun_out/${TAG}_launches_v0_default.csv $CMD > gpurun_out/ncu_default.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/${TAG}_launches_v0_default.csv
ncu --set full --clock-control none --import-source on -k regex:preprocess_rollout_pipe -s 1 -c 1 -f \
    -o gpurun_out/${TAG}_preprocess_v0 $CMD > gpurun_out/ncu_pre_v0.log 2>&1
echo "preprocess v0 rc=$?"
CMD1="python bench.py --workload v1 --steps 1 --warmup 1 --no-cpu-baseline --no-graph"
$CMD1 > gpurun_out/plain_v1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:preprocess_rollout_pipe -s 1 -c 1 -f \
    -o gpurun_out/${TAG}_preprocess_v1 $CMD1 > gpurun_out/ncu_pre_v1.log 2>&1
echo "preprocess v1 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 11 -c 11 -f \
    -o gpurun_out/${TAG}_gemm_v1 $CMD1 > gpurun_out/ncu_gemm_v1.log 2>&1
echo "gemm v1 rc=$?"
CMD0="python bench.py --workload v0 --steps 1 --warmup 1 --no-cpu-baseline --no-graph"
ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 176 -c 11 -f \
    -o gpurun_out/${TAG}_gemm_v0 $CMD0 > gpurun_out/ncu_gemm_v0.log 2>&1
echo "gemm v0 rc=$?"
ncu --set full --clock-control none -k regex:"adam_clip|gae_scan|heads_kernel|obs_to_s2d|grad_sumsq" -s 0 -c 8 -f \
    -o gpurun_out/${TAG}_membound_v1 $CMD1 > gpurun_out/ncu_mem_v1.log 2>&1
echo "membound v1 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"resize_gray_kernel|preprocess_gray_step" -s 20 -c 2 -f \
    -o gpurun_out/${TAG}_livepath python tools/time_live.py > gpurun_out/ncu_live.log 2>&1
echo "live path rc=$?"
tail -2 gpurun_out/ncu_default.log | cut -c1-200
# keep the evidence small: text summaries of every report, then drop the reports (gpurun_out <= 64 MiB)
for r in gpurun_out/${TAG}_*.ncu-rep; do
  python tools/ncu_raw_summary.py $r > ${r%.ncu-rep}_ncu_summary.txt 2>/dev/null
  ncu -i $r --page source --csv > /tmp/src.csv 2>/dev/null && python tools/ncu_src_top.py /tmp/src.csv 12 > ${r%.ncu-rep}_stalls.txt 2>/dev/null
done
python tools/ncu_agg.py gpurun_out/${TAG}_launches_v0_default.csv > gpurun_out/${TAG}_launches_v0_default.txt 2>/dev/null
rm -f gpurun_out/${TAG}_*.ncu-rep
ls -la gpurun_out | head -40
