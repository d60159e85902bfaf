#!/bin/bash
# One full GPU-box visit: parity tests, smoke, the bench lines the driver will run, the other
# workloads, the ncu launch list of the bench command and one full capture of the region kernels.
# Usage (from the repo root, under gpurun): bash tools/gpu_round.sh [tag]
TAG=${1:-r2}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi_$TAG.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?"
tail -3 $O/pytest_gpu_$TAG.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke_$TAG.log 2>&1; echo "smoke rc=$?"
timeout 1200 python bench.py > $O/bench_4k_$TAG.json 2> $O/bench_4k_$TAG.err; echo "bench4k rc=$?"
head -c 1500 $O/bench_4k_$TAG.json; echo; tail -3 $O/bench_4k_$TAG.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref_$TAG.json 2> $O/bench_ref_$TAG.err; echo "benchref rc=$?"
head -c 600 $O/bench_ref_$TAG.json; echo
timeout 300 python bench.py --workload 1080p --no-cpu --no-e2e --no-extras > $O/bench_1080p_$TAG.json 2> $O/bench_1080p_$TAG.err; echo "bench1080p rc=$?"
timeout 300 python bench.py --workload 1080p8x8 --no-cpu --no-e2e --no-extras > $O/bench_1080p8x8_$TAG.json 2> $O/bench_1080p8x8_$TAG.err; echo "bench1080p8x8 rc=$?"
TAG=$TAG python - <<'PY'
import json, os
for n in ("1080p", "1080p8x8"):
    try:
        d = json.load(open("gpurun_out/bench_%s_%s.json" % (n, os.environ["TAG"])))
        print(n, round(d["value"]), d["roofline"]["fwd_ms"], d["roofline"]["inv_ms"], d["roofline"]["frac"])
    except Exception as e:
        print(n, "unreadable", e)
PY
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-extras"
timeout 600 $CMD > $O/plain_$TAG.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
  --log-file $O/launches_$TAG.csv $CMD > $O/ncu_launch_$TAG.log 2>&1
echo "ncu launches rc=$?"
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-extras"
timeout 600 $CMD > $O/plain2_$TAG.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:region_kernel -s 6 -c 2 \
  -f -o $O/prof_$TAG $CMD > $O/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"
ls -la $O | tail -20
