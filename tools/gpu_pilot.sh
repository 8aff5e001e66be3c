#!/bin/bash
# short guarded run for a risky kernel change: tensor-core parity tests only, hard 150 s limit
mkdir -p gpurun_out
timeout -s KILL 150 python -m pytest tests/test_gpu_tensorcore.py -q -m gpu --timeout 60 -x > gpurun_out/pytest_pilot.log 2>&1; echo "pytest pilot rc=$?"
tail -30 gpurun_out/pytest_pilot.log
