#!/bin/bash
# One GPU-box visit: parity tests, bench lines, ncu launch list and one full capture.
# Usage (from the repo root, under gpurun): bash tools/gpu_round.sh [tag]
TAG=${1:-r1}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi_$TAG.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?"
tail -3 $O/pytest_gpu_$TAG.log
timeout 600 python bench.py > $O/bench_4k_$TAG.json 2> $O/bench_4k_$TAG.err; echo "bench4k rc=$?"
cat $O/bench_4k_$TAG.json; tail -3 $O/bench_4k_$TAG.err
timeout 300 python bench.py --workload 1080p --no-cpu --no-e2e > $O/bench_1080p_$TAG.json 2> $O/bench_1080p_$TAG.err; echo "bench1080p rc=$?"
cat $O/bench_1080p_$TAG.json
timeout 300 python bench.py --workload 1080p8x8 --no-cpu --no-e2e > $O/bench_1080p8x8_$TAG.json 2> $O/bench_1080p8x8_$TAG.err; echo "bench1080p8x8 rc=$?"
cat $O/bench_1080p8x8_$TAG.json
CMD="python bench.py --steps 2 --warmup 3 --frames 8 --no-e2e --no-cpu"
timeout 300 $CMD > $O/plain_$TAG.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:tdlt -c 10 --csv \
  --log-file $O/launches_$TAG.csv $CMD > $O/ncu_launch_$TAG.log 2>&1
echo "ncu launches rc=$?"
timeout 300 $CMD > $O/plain2_$TAG.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tdlt -s 6 -c 2 \
  -f -o $O/prof_$TAG $CMD > $O/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"
ls -la $O
