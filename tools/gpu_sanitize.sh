#!/bin/bash
# NOTE: compute-sanitizer is CLOSED on this pool (profiles/r02_sanitizer_unavailable.txt); the script is kept for a pool where it runs.
# compute-sanitizer evidence (SURVEY §5 race-detection row): memcheck over the tensor-core engine
# (mbarrier / TMEM / TMA / shift convolutions), the preprocess pipe and one train(); racecheck over the
# shared-memory CUDA-core kernels (preprocess, GAE, heads / loss, optimiser).  Summaries -> gpurun_out/.
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
run() { # tag tool pytest-args...
  local tag=$1 tool=$2; shift 2
  timeout 600 $CS --tool $tool --print-limit 20 --log-file gpurun_out/sanitizer_${tag}.log \
    python -m pytest "$@" -q -m gpu --timeout 500 -x -p no:cacheprovider > gpurun_out/sanitizer_${tag}_pytest.log 2>&1
  echo "$tag ($tool) rc=$? : $(tail -1 gpurun_out/sanitizer_${tag}_pytest.log)"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY" gpurun_out/sanitizer_${tag}.log | tail -2
}
run memcheck_engine memcheck tests/test_gpu_tensorcore.py -k "(engine and 304) or (engine_tf32 and 256-136) or shift or (tf32_vs_oracle and 37) or tma_im2col"
run memcheck_path memcheck tests/test_gpu_parity.py -k "preprocess_rollout or ingest or buffer_get_matches or train_clipped or shuffle_and or vision"
run racecheck_path racecheck tests/test_gpu_parity.py -k "preprocess_matches or preprocess_rollout_equals or (buffer_get_matches and 33) or loss_stage or train_clipped"
