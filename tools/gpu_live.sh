#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_live_path.py tests/test_abi.py -q -m gpu --timeout 200 > gpurun_out/pytest_live.log 2>&1; echo "pytest live rc=$?"
tail -25 gpurun_out/pytest_live.log
timeout 200 python tools/time_live.py 2>&1 | tee gpurun_out/time_live.json
