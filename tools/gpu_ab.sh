#!/bin/bash
# A/B visit: parity tests of the product library, then tools/ablate.py over the product
# library and every variant under daala_b200/variants/.  Usage: bash tools/gpu_ab.sh tag [pytest -k expr]
TAG=${1:-ab}
O=gpurun_out
mkdir -p $O
if [ -n "$2" ]; then
  timeout 900 python -m pytest tests -m gpu -x -q -k "$2" > $O/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?"
else
  timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?"
fi
tail -15 $O/pytest_gpu_$TAG.log
timeout 900 python tools/ablate.py 64 quad 1 > $O/ablate_$TAG.txt 2>&1
cat $O/ablate_$TAG.txt
timeout 600 python tools/ablate.py 64 uniform1 1 > $O/ablate_u1_$TAG.txt 2>&1
cat $O/ablate_u1_$TAG.txt
