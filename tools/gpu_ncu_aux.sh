#!/bin/bash
# ncu --set full capture of one neighbour kernel under its bench tool.
# Usage: bash tools/gpu_ncu_aux.sh tag tool.py kernel-regex [frames]
TAG=${1:-aux}
TOOL=${2:-tools/bench_dering.py}
KRE=${3:-dering_kernel}
FR=${4:-16}
O=gpurun_out
mkdir -p $O
CMD="python $TOOL --frames $FR --steps 2 --no-cpu"
timeout 300 $CMD > $O/plain_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 $O/plain_$TAG.log; exit 1; }
head -c 600 $O/plain_$TAG.log; echo
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$KRE -s ${SKIP:-3} -c ${COUNT:-1} \
  -f -o $O/prof_$TAG $CMD > $O/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"
ls -la $O/prof_$TAG.ncu-rep
