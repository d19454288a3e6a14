#!/bin/bash
# ncu --set full of the per-centre factor_solve launch (the launch before it in every batch is the shared-core one)
set -u
mkdir -p gpurun_out
SMALL="--centres-per-step 4096 --steps 1 --warmup 1 --no-e2e --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on -k factor_solve_kernel -s 1 -c 1 -f -o gpurun_out/factor_solve_kernel \
     python bench.py $SMALL > gpurun_out/ncu_factor_solve_kernel.log 2>&1
