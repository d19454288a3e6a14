#!/bin/bash
# Short GPU-box session: parity tests, a 12288-centre bench, the ncu launch list, full captures of the integer prediction kernels.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
SMALL="--centres-per-step 12288 --steps 2 --warmup 1 --no-e2e --no-cpu-baseline"
KREG='regex:(factor_solve|predict_kernel|predict_i8|slice_|syrk_i8|score_kernel|gemm_nt|prep_tasks)'
timeout 600 python bench.py $SMALL > gpurun_out/bench_small.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KREG" -c 400 --csv \
   --log-file gpurun_out/launches.csv python bench.py $SMALL > gpurun_out/ncu_launches.log 2>&1
if [ "${FULL:-0}" = "1" ]; then
  timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2>&1
  timeout 900 python bench.py > gpurun_out/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/bench.log
fi
for kn in ${CAPTURE:-predict_i8_kernel}; do
  timeout 900 ncu --set full --clock-control none --import-source on -k $kn -s 0 -c 1 -f -o gpurun_out/${kn} \
     python bench.py $SMALL > gpurun_out/ncu_${kn}.log 2>&1
done
ls -la gpurun_out
