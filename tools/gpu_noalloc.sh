#!/bin/bash
# GPU-box check of the allocation-free entry point: alone in a fresh process, then at the end of the whole GPU suite, then smoke().
set -u
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_noalloc_alone.log
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tail -n 4 gpurun_out/pytest_noalloc_alone.log gpurun_out/pytest_gpu.log gpurun_out/smoke.log
