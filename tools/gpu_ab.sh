#!/bin/bash
# A/B of kernel variants: $1.. = variant names (libmmd_b200.<variant>.so); "default" = the product library.
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
for v in "$@"; do
  if [ "$v" = default ]; then unset MMD_B200_LIB; else export MMD_B200_LIB=$PWD/multimodal_decoding_b200/libmmd_b200.$v.so; fi
  timeout 600 python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/bench_ab_$v.log 2>&1
done
