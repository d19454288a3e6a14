// Shared device helpers for the searchlight kernels (sm_100a).
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace mmd {

constexpr int EVAL_PAD = 144;     // 140 test + 3 imagery rows, padded to a multiple of 8 (DMMA M granularity)
constexpr int NB = 64;            // panel / tile width of the per-centre factorisation

__host__ __device__ inline int round_up(int x, int m) { return (x + m - 1) / m * m; }

// fp64 tensor-core MMA, D(8x8) += A(8x4, row) * B(4x8, col).
//   lane l holds A[l/4][l%4], B[l%4][l/4], C[l/4][2*(l%4) + {0,1}].
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    uint32_t s = static_cast<uint32_t>(__cvta_generic_to_shared(smem));
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N));
}

__device__ __forceinline__ double h2d(__half h) { return static_cast<double>(__half2float(h)); }

}  // namespace mmd
