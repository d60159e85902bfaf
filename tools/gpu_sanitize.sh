#!/bin/bash
# compute-sanitizer passes over the region kernels (memcheck + racecheck on the
# shared-memory transposes and the strip exchange).  Usage: bash tools/gpu_sanitize.sh tag
TAG=${1:-r2}
O=gpurun_out
mkdir -p $O
SEL='(frames_match_checker and not 1920) or inverse_batch_sizes or lossless_haar_mode'
for tool in memcheck racecheck; do
  timeout 1500 compute-sanitizer --tool $tool --error-exitcode 9 \
    python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "$SEL" > $O/sanitizer_${tool}_$TAG.log 2>&1
  echo "$tool rc=$?"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed" $O/sanitizer_${tool}_$TAG.log | tail -4
done
