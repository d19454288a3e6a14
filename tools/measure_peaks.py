"""Measure the matmul peaks MEASURED_PEAKS.json lacks (same method: 8192^3 torch.matmul, best of 10,
CUDA events) plus the fp64/fp32 pipe probes in tools/peak_probe.cu.  Writes gpurun_out/peaks_extra.json."""
import json
import os
import subprocess
import sys

import torch


def matmul_tflops(dtype, n=8192, tf32=False, reps=10):
    torch.backends.cuda.matmul.allow_tf32 = tf32
    a = torch.randn(n, n, device="cuda", dtype=dtype)
    b = torch.randn(n, n, device="cuda", dtype=dtype)
    for _ in range(3):
        a @ b
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        a @ b
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return 2 * n ** 3 / best * 1e-9


def main():
    out = {"gpu": torch.cuda.get_device_name(0)}
    out["fp16_tflops"] = matmul_tflops(torch.float16)
    out["bf16_tflops"] = matmul_tflops(torch.bfloat16)
    out["tf32_tflops"] = matmul_tflops(torch.float32, tf32=True)
    out["fp32_tflops"] = matmul_tflops(torch.float32, tf32=False, n=8192, reps=3)
    out["fp64_tflops"] = matmul_tflops(torch.float64, n=8192, reps=3)
    exe = os.path.join(os.path.dirname(os.path.abspath(__file__)), "peak_probe")
    if os.path.exists(exe):
        out["pipes"] = json.loads(subprocess.run([exe], capture_output=True, text=True, check=True).stdout)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open("gpurun_out/peaks_extra.json", "w"), indent=1)
    print(json.dumps(out))


if __name__ == "__main__":
    sys.exit(main())
