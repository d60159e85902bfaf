#!/bin/bash
# Quick GPU iteration: parity tests, one 4K bench line (device-resident only), one ncu capture.
# Usage: bash tools/gpu_quick.sh tag [kernel-regex] [skip-launches]
TAG=${1:-q}
KRE=${2:-tdlt}
SKIP=${3:-6}
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?"
tail -15 $O/pytest_gpu_$TAG.log
timeout 600 python bench.py --no-e2e --no-cpu > $O/bench_4k_$TAG.json 2> $O/bench_4k_$TAG.err; echo "bench4k rc=$?"
cat $O/bench_4k_$TAG.json; tail -3 $O/bench_4k_$TAG.err
timeout 300 python bench.py --workload 1080p8x8 --no-cpu --no-e2e > $O/bench_1080p8x8_$TAG.json 2> $O/bench_1080p8x8_$TAG.err; echo "bench1080p8x8 rc=$?"
cat $O/bench_1080p8x8_$TAG.json
CMD="python bench.py --steps 2 --warmup 3 --frames 8 --no-e2e --no-cpu"
timeout 300 $CMD > $O/plain_$TAG.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$KRE -s $SKIP -c 2 \
  -f -o $O/prof_$TAG $CMD > $O/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"
