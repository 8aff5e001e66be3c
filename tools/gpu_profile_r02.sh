#!/bin/bash
# Round-2 evidence for profiles/: ncu launch list of the default bench command (v1, tf32) + ncu --set full
# of the tensor-core kernels of one minibatch step, the preprocess kernel and the other memory-bound
# kernels.  Each ncu pass runs only after the same command has exited 0 without ncu.
TAG=${1:-r02}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-extras --min-seconds 0"
$CMD > gpurun_out/${TAG}_plain_default.json 2> gpurun_out/plain_default.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv \
    --log-file gpurun_out/${TAG}_launches_v1_default.csv $CMD > gpurun_out/ncu_default.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/${TAG}_launches_v1_default.csv
CMD1="python bench.py --workload v1 --steps 1 --warmup 1 --no-cpu-baseline --no-extras --min-seconds 0 --no-graph"
$CMD1 > gpurun_out/plain_v1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:preprocess_rollout_pipe -s 1 -c 1 -f \
    -o gpurun_out/${TAG}_preprocess_v1 $CMD1 > gpurun_out/ncu_pre_v1.log 2>&1
echo "preprocess v1 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"gemm_tc|conv_shift" -s 11 -c 11 -f \
    -o gpurun_out/${TAG}_gemm_v1_tf32 $CMD1 > gpurun_out/ncu_gemm_v1.log 2>&1
echo "gemm v1 tf32 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"gemm_tc|conv_shift" -s 11 -c 11 -f \
    -o gpurun_out/${TAG}_gemm_v1_bf16 $CMD1 --math bf16 > gpurun_out/ncu_gemm_v1b.log 2>&1
echo "gemm v1 bf16 rc=$?"
ncu --set full --clock-control none -k regex:"adam_fused|gae_scan|heads_kernel|obs_to_s2d|heads_wgrad|reduce_" -s 0 -c 10 -f \
    -o gpurun_out/${TAG}_membound_v1 $CMD1 > gpurun_out/ncu_mem_v1.log 2>&1
echo "membound v1 rc=$?"
for r in gpurun_out/${TAG}_*.ncu-rep; do
  python tools/ncu_raw_summary.py $r > ${r%.ncu-rep}_ncu_summary.txt 2>/dev/null
  ncu -i $r --page source --csv > /tmp/src.csv 2>/dev/null && python tools/ncu_src_top.py /tmp/src.csv 12 > ${r%.ncu-rep}_stalls.txt 2>/dev/null
done
python tools/ncu_agg.py gpurun_out/${TAG}_launches_v1_default.csv > gpurun_out/${TAG}_launches_v1_default.txt 2>/dev/null
rm -f gpurun_out/${TAG}_*.ncu-rep
head -30 gpurun_out/${TAG}_launches_v1_default.txt | cut -c1-200
