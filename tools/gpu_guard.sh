#!/bin/bash
# Guard mode (MMD_B200_GUARD=1: 4 KiB bands around every device buffer, buffers pre-filled with 0xA5) over the small
# all-kernels case and the whole GPU test suite.  compute-sanitizer is closed on the GPU pool; this is the
# out-of-bounds-write / uninitialised-read check the repository can run there.
set -u
mkdir -p gpurun_out
export MMD_B200_GUARD=1
timeout 600 python tests/tools/sanitize_case.py > gpurun_out/guard_case.log 2>&1; echo "rc=$?" >> gpurun_out/guard_case.log
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/guard_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/guard_pytest.log
unset MMD_B200_GUARD
tail -n 6 gpurun_out/guard_case.log gpurun_out/guard_pytest.log
