#!/bin/bash
# Two-GPU check of the sharded call: (TESTS=1: the 2-rank NCCL test, then) a short vertex-sharded bench (no replica pass).
set -u
mkdir -p gpurun_out
if [ "${TESTS:-0}" = "1" ]; then
  timeout 240 python -m pytest tests/test_gpu_sharded.py -m gpu -x -q -k two_ranks > gpurun_out/pytest_gpu_n2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_n2.log
  tail -n 3 gpurun_out/pytest_gpu_n2.log
fi
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 \
   bench.py --gpus 2 --steps 3 --warmup 3 --no-replicas > gpurun_out/bench_n2.log 2> gpurun_out/bench_n2.err; echo "bench rc=$?"
python - <<'PY'
import json
for line in open("gpurun_out/bench_n2.log"):
    if line.startswith("{"):
        d = json.loads(line)
        e = d["e2e"]
        print(d["value"], d["ms_per_step"], e["value"], e["ms_per_step"], e.get("step_ms_rank0"), e.get("call_ms_rank0"), e["host_ms_last_call_rank0"])
PY
