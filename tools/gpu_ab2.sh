#!/bin/bash
# A/B session: kernel variants (libmmd_b200.<variant>.so) and engine knobs through bench.py, one full-hemisphere step each.
set -u
mkdir -p gpurun_out
B="--steps 2 --warmup 1 --no-e2e --no-cpu-baseline"
timeout 600 python bench.py $B > gpurun_out/bench_ab_default.log 2>&1
for v in "$@"; do
  MMD_B200_LIB=$PWD/multimodal_decoding_b200/libmmd_b200.$v.so timeout 600 python bench.py $B > gpurun_out/bench_ab_$v.log 2>&1
done
for c in 192 384; do timeout 600 python bench.py $B --cluster-size $c > gpurun_out/bench_ab_cluster$c.log 2>&1; done
for g in 6 12; do timeout 600 python bench.py $B --group-size $g > gpurun_out/bench_ab_group$g.log 2>&1; done
