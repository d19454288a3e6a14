// Micro-probes for the roofline denominators MEASURED_PEAKS.json does not hold:
// fp64 FMA pipe (DFMA), fp64 tensor pipe (mma.sync m8n8k4 / m16n8k8 .f64), fp32 FMA pipe.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o peak_probe peak_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

template <int ILP>
__global__ void dfma_kernel(double* out, int iters) {
    double a[ILP], b = 1.0000001, c = 0.9999999;
#pragma unroll
    for (int i = 0; i < ILP; ++i) a[i] = threadIdx.x + i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) a[i] = fma(a[i], b, c);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int ILP>
__global__ void ffma_kernel(float* out, int iters) {
    float a[ILP], b = 1.0000001f, c = 0.9999999f;
#pragma unroll
    for (int i = 0; i < ILP; ++i) a[i] = threadIdx.x + i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) a[i] = fmaf(a[i], b, c);
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int ILP>
__global__ void dmma884_kernel(double* out, int iters) {
    double c[ILP][2];
    double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
#pragma unroll
    for (int i = 0; i < ILP; ++i) { c[i][0] = i; c[i][1] = -i; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int ILP>
__global__ void dmma1688_kernel(double* out, int iters) {
    double c[ILP][4];
    double a0 = 1.0 + threadIdx.x * 1e-9, a1 = a0 + 1e-3, a2 = a0 - 1e-3, a3 = a0 * 0.5;
    double b0 = 1.0 - threadIdx.x * 1e-9, b1 = b0 * 0.25;
#pragma unroll
    for (int i = 0; i < ILP; ++i) { c[i][0] = i; c[i][1] = -i; c[i][2] = 2 * i; c[i][3] = 1; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i)
            asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3])
                         : "d"(a0), "d"(a1), "d"(a2), "d"(a3), "d"(b0), "d"(b1));
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}


// DFMA and DMMA issued by the same warp, interleaved: do the fp64 FMA pipe and the fp64 tensor sub-pipe overlap?
template <int ILP, int FPM>
__global__ void mixed_kernel(double* out, int iters) {
    double c[ILP][2], f[ILP * FPM];
    double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
#pragma unroll
    for (int i = 0; i < ILP; ++i) { c[i][0] = i; c[i][1] = -i; }
#pragma unroll
    for (int i = 0; i < ILP * FPM; ++i) f[i] = threadIdx.x + i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) {
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
#pragma unroll
            for (int j = 0; j < FPM; ++j) f[i * FPM + j] = fma(f[i * FPM + j], b, a);
        }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += c[i][0] + c[i][1];
#pragma unroll
    for (int i = 0; i < ILP * FPM; ++i) s += f[i];
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
    double* dout; CK(cudaMalloc(&dout, sizeof(double) * sms * 8 * 1024));
    const int iters = 20000;
    for (int threads : {256, 512, 1024}) {
        int blocks = sms * (2048 / threads);
        float ms = time_ms([&] { dfma_kernel<8><<<blocks, threads>>>(dout, iters); });
        printf(", \"dfma_tflops_t%d\": %.2f", threads, 2.0 * 8 * iters * (double)blocks * threads / ms * 1e-9);
        ms = time_ms([&] { ffma_kernel<8><<<blocks, threads>>>((float*)dout, iters); });
        printf(", \"ffma_tflops_t%d\": %.2f", threads, 2.0 * 8 * iters * (double)blocks * threads / ms * 1e-9);
        ms = time_ms([&] { dmma884_kernel<8><<<blocks, threads>>>(dout, iters); });
        printf(", \"dmma884_tflops_t%d\": %.2f", threads, 2.0 * 256 * 8 * iters * (double)blocks * (threads / 32) / ms * 1e-9);
        ms = time_ms([&] { dmma1688_kernel<4><<<blocks, threads>>>(dout, iters); });
        printf(", \"dmma1688_tflops_t%d\": %.2f", threads, 2.0 * 1024 * 4 * iters * (double)blocks * (threads / 32) / ms * 1e-9);
    }
    {
        const int threads = 512, blocks = sms * 4;
        float ms = time_ms([&] { mixed_kernel<8, 4><<<blocks, threads>>>(dout, iters); });
        // per warp and iteration: 8 DMMA (256 FMA each) + 8*4 DFMA x 32 lanes
        printf(", \"mixed_dmma_plus_4dfma_tflops\": %.2f", 2.0 * (256 * 8 + 32 * 32) * iters * (double)blocks * (threads / 32) / ms * 1e-9);
        ms = time_ms([&] { mixed_kernel<8, 8><<<blocks, threads>>>(dout, iters); });
        printf(", \"mixed_dmma_plus_8dfma_tflops\": %.2f", 2.0 * (256 * 8 + 64 * 32) * iters * (double)blocks * (threads / 32) / ms * 1e-9);
        ms = time_ms([&] { mixed_kernel<8, 2><<<blocks, threads>>>(dout, iters); });
        printf(", \"mixed_dmma_plus_2dfma_tflops\": %.2f", 2.0 * (256 * 8 + 16 * 32) * iters * (double)blocks * (threads / 32) / ms * 1e-9);
    }
    CK(cudaDeviceSynchronize());
    printf("}\n");
    return 0;
}
