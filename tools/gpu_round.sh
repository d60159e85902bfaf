#!/bin/bash
# One full GPU-box visit: parity tests, the bench lines the driver will run, the ncu
# launch list of the bench command and one full capture of the two region kernels.
# Usage (from the repo root, under gpurun): bash tools/gpu_round.sh [tag]
TAG=${1:-r1}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi_$TAG.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?"
tail -3 $O/pytest_gpu_$TAG.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke_$TAG.log 2>&1; echo "smoke rc=$?"
timeout 900 python bench.py > $O/bench_4k_$TAG.json 2> $O/bench_4k_$TAG.err; echo "bench4k rc=$?"
cat $O/bench_4k_$TAG.json; tail -3 $O/bench_4k_$TAG.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref_$TAG.json 2> $O/bench_ref_$TAG.err; echo "benchref rc=$?"
cat $O/bench_ref_$TAG.json
timeout 300 python bench.py --workload 1080p --no-cpu --no-e2e > $O/bench_1080p_$TAG.json 2> $O/bench_1080p_$TAG.err; echo "bench1080p rc=$?"
cat $O/bench_1080p_$TAG.json
timeout 300 python bench.py --workload 1080p8x8 --no-cpu --no-e2e > $O/bench_1080p8x8_$TAG.json 2> $O/bench_1080p8x8_$TAG.err; echo "bench1080p8x8 rc=$?"
cat $O/bench_1080p8x8_$TAG.json
timeout 300 python tools/bench_cfl.py > $O/cfl_4k_$TAG.json 2> $O/cfl_4k_$TAG.err; echo "cfl rc=$?"
cat $O/cfl_4k_$TAG.json
timeout 300 python tools/bench_dering.py > $O/dering_4k_$TAG.json 2> $O/dering_4k_$TAG.err; echo "dering rc=$?"
cat $O/dering_4k_$TAG.json
timeout 300 python tools/bench_dering.py --no-cpu --flag-prob 1.0 --skip-prob 0 > $O/dering_4k_all_$TAG.json 2>> $O/dering_4k_$TAG.err
cat $O/dering_4k_all_$TAG.json
timeout 300 python tools/bench_coding.py > $O/coding_4k_$TAG.json 2> $O/coding_4k_$TAG.err; echo "coding rc=$?"
cat $O/coding_4k_$TAG.json
CMD="python bench.py --steps 2 --warmup 3 --no-cpu"
timeout 600 $CMD > $O/plain_$TAG.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
  --log-file $O/launches_$TAG.csv $CMD > $O/ncu_launch_$TAG.log 2>&1
echo "ncu launches rc=$?"
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e"
timeout 600 $CMD > $O/plain2_$TAG.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:region -s 6 -c 2 \
  -f -o $O/prof_$TAG $CMD > $O/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"
ls -la $O
