#!/bin/bash
# Scaling visit: bench.py under torch.distributed.run on N GPUs of one box (the driver's launch
# line), own arm and reference arm.  Usage (under gpurun --gpus N): bash tools/gpu_scale.sh N tag
N=${1:-2}
TAG=${2:-scale}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=index,name,clocks.max.sm --format=csv > $O/smi_${N}gpu_$TAG.txt 2>&1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
  --master-port 29517 bench.py --gpus $N --no-extras > $O/bench_4k_${N}gpu_$TAG.json 2> $O/bench_4k_${N}gpu_$TAG.err
echo "bench rc=$?"; tail -c 1500 $O/bench_4k_${N}gpu_$TAG.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
  --master-port 29518 bench.py --impl reference --gpus $N > $O/bench_ref_${N}gpu_$TAG.json 2> $O/bench_ref_${N}gpu_$TAG.err
echo "ref rc=$?"; tail -c 600 $O/bench_ref_${N}gpu_$TAG.json
