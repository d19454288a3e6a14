// Groundwork for the next round's exact-integer prediction GEMM (candidates): a minimal hand-written
// tcgen05.mma kind::i8 GEMM tile -- int8 operands in shared memory (canonical no-swizzle K-major layout), int32
// accumulators in TMEM, tcgen05.ld back to registers -- checked against the CPU, then timed.
//
//   C (128 x 256, int32) = A (128 x K, int8, K-major) . B (256 x K, int8, K-major)^T,   K = 256
//
// Shared-memory layout (UMMA "interleave", no swizzle): 16-byte K-chunks outermost, rows inside:
//   offset(row r, k-byte c) = (c / 16) * (ROWS * 16) + r * 16 + c % 16
// so a core matrix (8 rows x 16 bytes) is 128 contiguous bytes, SBO (next 8 rows) = 128 B and LBO (next K-chunk) =
// ROWS * 16 B.  One MMA consumes K = 32 bytes = two chunks; the descriptor's start address advances by 2 * LBO.
// Every wait in the kernel is bounded, so a mistake ends in an error code, never in a hung GPU.
//
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tcgen05_i8_probe tcgen05_i8_probe.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("{\"error\": \"%s at line %d\"}\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

constexpr int M = 128, K = 256;
constexpr int A_BYTES = M * K;
constexpr int TMEM_COLS = 256;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint32_t lo = ((addr >> 4) & 0x3fffu) | (((lbo_bytes >> 4) & 0x3fffu) << 16);
    uint32_t hi = ((sbo_bytes >> 4) & 0x3fffu) | (1u << 14);          // version 1 (Blackwell), SWIZZLE_NONE, base offset 0
    return ((uint64_t)hi << 32) | lo;
}

// repeats: how often the whole K loop is issued (1 for the correctness run; accumulates on after the first MMA)
// A_MN: A is given as (K x M) with M contiguous -- the form a gathered row of XtY slices has -- and staged in the
// MN-major canonical layout  offset(m, k) = (m / 16) * (K * 16) + (k / 8) * 128 + (k % 8) * 16 + m % 16
// (SBO = K * 16 between 16-byte m-chunks, LBO = 128 between blocks of 8 k); one MMA (K = 32) spans four k-blocks.
template <int N, bool A_MN>
__global__ void __launch_bounds__(128, 1) i8_tile_kernel(const int8_t* __restrict__ A, const int8_t* __restrict__ B,
                                                         int32_t* __restrict__ C, int repeats, int* __restrict__ status) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* As = smem;
    uint8_t* Bs = smem + A_BYTES;
    constexpr int B_BYTES = N * K;
    (void)B_BYTES;
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_base_sm;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    // global (row-major, K contiguous) -> canonical layout, 16 bytes at a time
    if (A_MN) {
        for (int i = tid; i < K * (M / 16); i += 128) {
            const int k = i / (M / 16), mc = i % (M / 16);
            *reinterpret_cast<int4*>(As + mc * (K * 16) + (k / 8) * 128 + (k % 8) * 16) =
                *reinterpret_cast<const int4*>(A + (size_t)k * M + mc * 16);
        }
    } else {
        for (int i = tid; i < M * (K / 16); i += 128) {
            const int r = i / (K / 16), j = i % (K / 16);
            *reinterpret_cast<int4*>(As + j * (M * 16) + r * 16) = *reinterpret_cast<const int4*>(A + (size_t)r * K + j * 16);
        }
    }
    for (int i = tid; i < N * (K / 16); i += 128) {
        const int r = i / (K / 16), j = i % (K / 16);
        *reinterpret_cast<int4*>(Bs + j * (N * 16) + r * 16) = *reinterpret_cast<const int4*>(B + (size_t)r * K + j * 16);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // generic-proxy writes -> visible to the MMA's async proxy
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_sm)), "r"(TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_sm;

    if (tid == 0) {
        // instruction descriptor: D = S32, A = B = signed 8 bit, both K-major, N >> 3, M >> 4
        const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((A_MN ? 1u : 0u) << 15) | ((uint32_t)(N >> 3) << 17) |
                               ((uint32_t)(M >> 4) << 24);
        const uint32_t a0 = smem_u32(As), b0 = smem_u32(Bs);
        for (int rep = 0; rep < repeats; ++rep) {
#pragma unroll 1
            for (int ks = 0; ks < K / 32; ++ks) {
                const uint64_t adesc = A_MN ? make_desc(a0 + ks * 4 * 128, 128, K * 16) : make_desc(a0 + ks * 2 * (M * 16), M * 16, 128);
                const uint64_t bdesc = make_desc(b0 + ks * 2 * (N * 16), N * 16, 128);
                const uint32_t accumulate = (rep | ks) != 0;
                asm volatile(
                    "{\n\t"
                    ".reg .pred p;\n\t"
                    "setp.ne.b32 p, %4, 0;\n\t"
                    "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t"
                    "}\n" ::"r"(tmem_base), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(0), "r"(0), "r"(0), "r"(0)
                    : "memory");
            }
        }
        // arrives on the barrier once every MMA issued so far has completed (implies fence::before_thread_sync)
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    }
    __syncwarp();

    // bounded wait on phase 0
    uint32_t done = 0;
    for (long spin = 0; spin < (1L << 24) && !done; ++spin) {
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t"
            "}\n" : "=r"(done) : "r"(smem_u32(&bar)), "r"(0) : "memory");
    }
    if (!done && lane == 0) atomicExch(status, 2);                  // the MMAs never completed
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

    if (done && C != nullptr) {
        // warp w owns TMEM lanes 32w .. 32w+31 = rows of the tile; 32 columns per load
        for (int c0 = 0; c0 < N; c0 += 16) {
            uint32_t v[16];
            const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + c0;
            asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
                "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
                : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
                  "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                : "r"(taddr));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            int32_t* crow = C + ((size_t)blockIdx.x * M + warp * 32 + lane) * N + c0;
#pragma unroll
            for (int j = 0; j < 16; ++j) crow[j] = (int32_t)v[j];
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS));
}

template <int N, bool A_MN>
int run_case(const cudaDeviceProp& p, const char* name) {
    const int sms = p.multiProcessorCount;
    constexpr int B_BYTES = N * K;
    std::vector<int8_t> hA(A_BYTES), hB(B_BYTES);      // A: (M x K) K contiguous, or (K x M) M contiguous when A_MN
    srand(7);
    for (auto& v : hA) v = (int8_t)(rand() % 129 - 64);
    for (auto& v : hB) v = (int8_t)(rand() % 129 - 64);
    int8_t *dA, *dB;
    int32_t* dC;
    int* dstatus;
    CK(cudaMalloc(&dA, A_BYTES));
    CK(cudaMalloc(&dB, B_BYTES));
    CK(cudaMalloc(&dC, sizeof(int32_t) * M * N));
    CK(cudaMalloc(&dstatus, sizeof(int)));
    CK(cudaMemcpy(dA, hA.data(), A_BYTES, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, hB.data(), B_BYTES, cudaMemcpyHostToDevice));
    CK(cudaMemset(dC, 0xff, sizeof(int32_t) * M * N));
    CK(cudaMemset(dstatus, 0, sizeof(int)));
    const int smem_bytes = A_BYTES + B_BYTES;
    CK(cudaFuncSetAttribute(i8_tile_kernel<N, A_MN>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));

    i8_tile_kernel<N, A_MN><<<1, 128, smem_bytes>>>(dA, dB, dC, 1, dstatus);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    int status = 0;
    CK(cudaMemcpy(&status, dstatus, sizeof(int), cudaMemcpyDeviceToHost));
    std::vector<int32_t> hC((size_t)M * N);
    CK(cudaMemcpy(hC.data(), dC, sizeof(int32_t) * M * N, cudaMemcpyDeviceToHost));
    long mismatches = 0, first = -1;
    for (int i = 0; i < M; ++i)
        for (int j = 0; j < N; ++j) {
            int32_t want = 0;
            for (int k = 0; k < K; ++k)
                want += (int32_t)(A_MN ? hA[(size_t)k * M + i] : hA[(size_t)i * K + k]) * (int32_t)hB[(size_t)j * K + k];
            if (want != hC[(size_t)i * N + j]) {
                if (first < 0) first = (long)i * N + j;
                ++mismatches;
            }
        }
    printf("{\"case\": \"%s\", \"gpu\": \"%s\", \"tile\": \"%dx%dx%d s8 -> s32\", \"status\": %d, \"mismatches\": %ld, \"first_mismatch\": %ld",
           name, p.name, M, N, K, status, mismatches, first);
    if (status == 0 && mismatches == 0) {
        // throughput: one CTA per SM, the K loop issued `reps` times from resident shared memory
        const int reps = 2000;
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        i8_tile_kernel<N, A_MN><<<sms, 128, smem_bytes>>>(dA, dB, nullptr, reps, dstatus);
        CK(cudaDeviceSynchronize());
        float best = 1e30f;
        for (int r = 0; r < 3; ++r) {
            cudaEventRecord(e0);
            i8_tile_kernel<N, A_MN><<<sms, 128, smem_bytes>>>(dA, dB, nullptr, reps, dstatus);
            cudaEventRecord(e1);
            CK(cudaEventSynchronize(e1));
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (ms < best) best = ms;
        }
        CK(cudaMemcpy(&status, dstatus, sizeof(int), cudaMemcpyDeviceToHost));
        printf(", \"timed_status\": %d, \"tcgen05_i8_tops\": %.1f", status, 2.0 * M * N * K * (double)reps * sms / best * 1e-9);
    }
    printf("}\n");
    cudaFree(dA); cudaFree(dB); cudaFree(dC); cudaFree(dstatus);
    return 0;
}

int main() {
    cudaDeviceProp p;
    CK(cudaGetDeviceProperties(&p, 0));
    if (run_case<256, false>(p, "A K-major, N=256")) return 1;
    if (run_case<144, false>(p, "A K-major, N=144 (one eval block)")) return 1;
    if (run_case<144, true>(p, "A MN-major (gathered rows), N=144")) return 1;
    return 0;
}
