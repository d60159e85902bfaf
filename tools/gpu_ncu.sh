#!/bin/bash
# One ncu --set full capture of the two region kernels on the bench workload.
# Usage: bash tools/gpu_ncu.sh tag [frames] [kernel-regex] [skip]
TAG=${1:-n}
FR=${2:-64}
KRE=${3:-region_kernel}
SKIP=${4:-6}
O=gpurun_out
mkdir -p $O
CMD="python bench.py --steps 2 --warmup 3 --frames $FR --no-e2e --no-cpu"
timeout 300 $CMD > $O/plain_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 $O/plain_$TAG.log; exit 1; }
cat $O/plain_$TAG.log | head -3
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:$KRE -s $SKIP -c 2 \
  -f -o $O/prof_$TAG $CMD > $O/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"
ls -la $O/prof_$TAG.ncu-rep
