#!/bin/bash
# Multi-GPU bench exactly as the driver launches it: $1 = N
set -u
N=$1
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_$N.log 2>&1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
   bench.py --gpus $N --steps 2 --warmup 3 > gpurun_out/bench_n$N.log 2>&1; echo "rc=$?" >> gpurun_out/bench_n$N.log
[ "${2:-}" = nostrong ] || timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 \
   bench.py --gpus $N --steps 2 --warmup 3 --scaling strong --no-cpu-baseline > gpurun_out/bench_strong_n$N.log 2>&1; echo "rc=$?" >> gpurun_out/bench_strong_n$N.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 \
   bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/bench_ref_n$N.log 2>&1; echo "rc=$?" >> gpurun_out/bench_ref_n$N.log
