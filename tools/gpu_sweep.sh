#!/bin/bash
# Sweep of an engine knob through bench.py: $1 = flag (--cluster-size | --batch-centres), $2.. = values
set -u
mkdir -p gpurun_out
flag=$1; shift
for v in "$@"; do
  timeout 600 python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline $flag $v > gpurun_out/bench_sweep_$v.log 2>&1
done
