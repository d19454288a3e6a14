#!/bin/bash
# Multi-GPU session: $1 = N.  The sharded GPU tests, then the bench exactly as the driver launches it (strong scaling is the
# headline, the replica layout is measured in the same run).
set -u
N=$1
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_$N.log 2>&1
timeout 600 python -m pytest tests/test_gpu_sharded.py tests/test_gpu_cv.py -m gpu -x -q > gpurun_out/pytest_gpu_n$N.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_n$N.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
   bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_n$N.log 2>&1; echo "rc=$?" >> gpurun_out/bench_n$N.log
