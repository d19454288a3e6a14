#!/bin/bash
# One GPU-box session: parity tests, smoke, bench (both arms), ncu launch list, full captures of the per-centre kernels.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/smi.log 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2>&1
timeout 900 python bench.py > gpurun_out/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/bench.log
SMALL="--centres-per-step 12288 --steps 2 --warmup 1 --no-e2e --no-cpu-baseline"
KREG='regex:(factor_solve|predict_kernel|predict_i8|slice_|syrk_i8|col(max|scale)_kernel|score_kernel|gemm_nt|prep_tasks|pack_transpose|knn_kernel|radius_kernel|half_to_double|row_sqnorm|gather_rows)'
timeout 600 python bench.py $SMALL > gpurun_out/bench_small.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KREG" -c 400 --csv \
   --log-file gpurun_out/launches.csv python bench.py $SMALL > gpurun_out/ncu_launches.log 2>&1
for kn in predict_i8_kernel score_kernel; do
  timeout 900 ncu --set full --clock-control none --import-source on -k $kn -s 0 -c 1 -f -o gpurun_out/${kn} \
     python bench.py $SMALL > gpurun_out/ncu_${kn}.log 2>&1
done
# factor_solve: one merged launch per batch (the groups' shared cores + the centres)
timeout 900 ncu --set full --clock-control none --import-source on -k factor_solve_kernel -s 0 -c 1 -f -o gpurun_out/factor_solve_kernel \
     python bench.py $SMALL > gpurun_out/ncu_factor_solve_kernel.log 2>&1
# the cluster SYRK: launch 0 of syrk_i8_kernel is the X^T Y of the upload, launch 1 the first batch's union Grams
timeout 900 ncu --set full --clock-control none --import-source on -k syrk_i8_kernel -s 1 -c 1 -f -o gpurun_out/syrk_i8_kernel \
     python bench.py $SMALL > gpurun_out/ncu_syrk_i8.log 2>&1
ls -la gpurun_out
