#!/bin/bash
# Full-batch ncu captures of the per-centre kernels + batch-size sweep.
set -u
mkdir -p gpurun_out
SMALL="--centres-per-step 4096 --steps 1 --warmup 1 --no-e2e --no-cpu-baseline"
for bc in 2048 4440 8880; do
  timeout 600 python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --batch-centres $bc > gpurun_out/bench_bc$bc.log 2>&1
done
for kn in factor_solve_kernel predict_kernel score_kernel; do
  timeout 900 ncu --set full --clock-control none --import-source on -k $kn -s 0 -c 1 -f -o gpurun_out/${kn} \
     python bench.py $SMALL > gpurun_out/ncu_${kn}.log 2>&1
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_nt -s 1 -c 1 -f -o gpurun_out/gemm_syrk \
     python bench.py $SMALL > gpurun_out/ncu_gemm.log 2>&1
ls -la gpurun_out
