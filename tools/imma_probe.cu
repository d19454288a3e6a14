// Micro-probe for the next round's exact-integer route (candidates): what do the warp-level
// (mma.sync) int8 and fp16 tensor instructions reach on a B200?  If the legacy IMMA path is fast enough, the int8
// slice GEMMs of the prediction stage can be written with mma.sync fragments instead of tcgen05 + TMEM.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o imma_probe imma_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

template <int ILP>
__global__ void imma_kernel(int* out, int iters) {
    int c[ILP][4];
    unsigned a0 = 0x01020304u + threadIdx.x, a1 = a0 * 3u, a2 = a0 ^ 0x11111111u, a3 = a0 + 7u;
    unsigned b0 = 0x04030201u - threadIdx.x, b1 = b0 * 5u;
#pragma unroll
    for (int i = 0; i < ILP; ++i) { c[i][0] = i; c[i][1] = -i; c[i][2] = 2 * i; c[i][3] = 1; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i)
            asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+r"(c[i][0]), "+r"(c[i][1]), "+r"(c[i][2]), "+r"(c[i][3])
                         : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
    int s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int ILP>
__global__ void hmma_kernel(float* out, int iters) {
    float c[ILP][4];
    unsigned a0 = 0x3c003c00u, a1 = 0x3c003800u, a2 = 0x38003c00u, a3 = 0x3c003c00u;   // halves near 1
    unsigned b0 = 0x3c003c00u, b1 = 0x38003800u;
#pragma unroll
    for (int i = 0; i < ILP; ++i) { c[i][0] = i; c[i][1] = -i; c[i][2] = 2 * i; c[i][3] = 1; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i)
            asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3])
                         : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
static float time_ms(F launch) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    launch(); cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) {
        cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    return best;
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int sms = p.multiProcessorCount;
    printf("{\"gpu\": \"%s\", \"sms\": %d", p.name, sms);
    int* dout; CK(cudaMalloc(&dout, sizeof(int) * sms * 8 * 1024));
    const int iters = 20000;
    for (int threads : {256, 512, 1024}) {
        int blocks = sms * (2048 / threads);
        float ms = time_ms([&] { imma_kernel<8><<<blocks, threads>>>(dout, iters); });
        printf(", \"imma_s8_m16n8k32_tops_t%d\": %.1f", threads, 2.0 * 16 * 8 * 32 * 8 * iters * (double)blocks * (threads / 32) / ms * 1e-9);
        ms = time_ms([&] { hmma_kernel<8><<<blocks, threads>>>((float*)dout, iters); });
        printf(", \"hmma_f16_m16n8k16_tflops_t%d\": %.1f", threads, 2.0 * 16 * 8 * 16 * 8 * iters * (double)blocks * (threads / 32) / ms * 1e-9);
    }
    CK(cudaGetLastError());
    printf("}\n");
    return 0;
}
