#!/bin/bash
# One GPU-box session: parity tests, smoke, short bench lines.  Output -> gpurun_out/
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus.txt 2>&1
nproc >> gpurun_out/gpus.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
timeout 600 python bench.py --workload v0 --steps 10 --warmup 3 > gpurun_out/bench_v0.json 2> gpurun_out/bench_v0.err; echo "bench v0 rc=$?"
timeout 600 python bench.py --workload v1 --steps 5 --warmup 3 > gpurun_out/bench_v1.json 2> gpurun_out/bench_v1.err; echo "bench v1 rc=$?"
tail -5 gpurun_out/pytest_gpu.log gpurun_out/smoke.log gpurun_out/bench_v0.json gpurun_out/bench_v0.err gpurun_out/bench_v1.json gpurun_out/bench_v1.err
