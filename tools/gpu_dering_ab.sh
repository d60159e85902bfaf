#!/bin/bash
# Deringing A/B: parity tests of the product library, then tools/bench_dering.py for the
# product library and every variant under daala_b200/variants/.  Usage: bash tools/gpu_dering_ab.sh tag
TAG=${1:-dr}
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -k "dering or smooth" > $O/pytest_dering_$TAG.log 2>&1; echo "pytest rc=$?"
tail -5 $O/pytest_dering_$TAG.log
for lib in daala_b200/libdaala_tdlt_cuda.so daala_b200/variants/lib_*.so; do
  [ -f "$lib" ] || continue
  echo "== $lib"
  DAALA_TDLT_LIB=$PWD/$lib timeout 300 python tools/bench_dering.py --no-cpu 2>&1 | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('ms %.4f frac %.3f' % (d['ms_per_step'], d['roofline']['frac']), d.get('parity'))
"
done 2>&1 | tee $O/dering_ab_$TAG.txt
