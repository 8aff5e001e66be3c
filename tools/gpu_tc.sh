#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_tensorcore.py -q -m gpu --timeout 120 > gpurun_out/pytest_tc.log 2>&1; echo "pytest tc rc=$?"
tail -60 gpurun_out/pytest_tc.log
