// Groundwork, step 2 (see tcgen05_i8_probe.cu for step 1): one tile of the planned integer prediction GEMM, end to end.
//
//   Y^T tile (128 feature columns x 144 evaluation rows, float64) = A (128 x K) . B (144 x K)^T,   A, B float64
//
// Both operands are cut on the host into S = 6 signed digits behind power-of-two scales (per row of A, per row of B;
// the scheme of tests/tools/int8_slice_probe.py).  The kernel keeps all digit planes resident in shared memory
// (A planes MN-major = the form of gathered XtY rows, B planes K-major), accumulates the slice pairs of equal weight
// w = t + u < S in one TMEM accumulator per weight (three accumulators at a time: 512 TMEM columns hold 3 x 144), and
// recombines in float64, smallest weight first:  y = scaleA * scaleB * sum_w 2^-(12 + 7 w) * S_w.
// The host repeats the same integer arithmetic, so the result must match BIT FOR BIT; the error against the plain
// float64 product is printed next to it.  Every wait is bounded.
//
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tcgen05_i8_sliced_probe tcgen05_i8_sliced_probe.cu
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("{\"error\": \"%s at line %d\"}\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

constexpr int M = 128, N = 144, K = 128, S = 6;
constexpr int A_PLANE = M * K, B_PLANE = N * K;          // bytes per digit plane
constexpr int TMEM_COLS = 512, ACC_STRIDE = 160;         // accumulator w % 3 lives at column (w % 3) * 160

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint32_t lo = ((addr >> 4) & 0x3fffu) | (((lbo_bytes >> 4) & 0x3fffu) << 16);
    uint32_t hi = ((sbo_bytes >> 4) & 0x3fffu) | (1u << 14);
    return ((uint64_t)hi << 32) | lo;
}

// A planes: (S x K x M) int8, M contiguous.  B planes: (S x N x K) int8, K contiguous.  out: (M x N) float64.
// repeats > 1 (timing): the whole tile -- 84 MMAs, two barrier phases, the float64 epilogue when EPILOGUE -- is redone
// from the resident planes; every CTA writes its own copy of the tile.
// EPILOGUE: 0 = none (MMA timing), 1 = four warps, partial sums through global memory, 2 = eight warps (launch with
// 256 threads), each thread keeps one feature row x 72 evaluation columns of running sums in registers
template <int EPILOGUE>
__global__ void __launch_bounds__(256, 1) sliced_tile_kernel(const int8_t* __restrict__ A, const int8_t* __restrict__ B,
                                                             const double* __restrict__ scaleA, const double* __restrict__ scaleB,
                                                             double* __restrict__ out_all, int repeats, int* __restrict__ status) {
    double* out = out_all + (size_t)blockIdx.x * M * N;
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* As = smem;                       // plane t at As + t * A_PLANE, MN-major canonical layout
    uint8_t* Bs = smem + S * A_PLANE;         // plane u at Bs + u * B_PLANE, K-major canonical layout
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_base_sm;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    for (int i = tid; i < S * K * (M / 16); i += blockDim.x) {
        const int t = i / (K * (M / 16)), r = i % (K * (M / 16));
        const int k = r / (M / 16), mc = r % (M / 16);
        *reinterpret_cast<int4*>(As + t * A_PLANE + mc * (K * 16) + (k / 8) * 128 + (k % 8) * 16) =
            *reinterpret_cast<const int4*>(A + (size_t)t * A_PLANE + (size_t)k * M + mc * 16);
    }
    for (int i = tid; i < S * N * (K / 16); i += blockDim.x) {
        const int u = i / (N * (K / 16)), r = i % (N * (K / 16));
        const int n = r / (K / 16), j = r % (K / 16);
        *reinterpret_cast<int4*>(Bs + u * B_PLANE + j * (N * 16) + n * 16) =
            *reinterpret_cast<const int4*>(B + (size_t)u * B_PLANE + (size_t)n * K + j * 16);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
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
    const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
    const int row = (warp & 3) * 32 + lane;                   // feature column of the tile = TMEM lane
    bool ok = true;
    double racc[72];                                          // EPILOGUE == 2: columns (warp / 4) * 72 ...

    // two passes over the weights, smallest terms first: w = 5, 4, 3 and then w = 2, 1, 0
    for (int it = 0; it < 2 * repeats && ok; ++it) {
        const int pass = it & 1;
        const int w_hi = pass == 0 ? S - 1 : S - 4;          // 5 | 2
        if (tid == 0) {
            const uint32_t a0 = smem_u32(As), b0 = smem_u32(Bs);
            for (int w = w_hi; w > w_hi - 3; --w) {
                bool first = true;
                for (int t = 0; t < S; ++t) {
                    const int u = w - t;
                    if (u < 0 || u >= S) continue;
#pragma unroll 1
                    for (int ks = 0; ks < K / 32; ++ks) {
                        const uint64_t adesc = make_desc(a0 + t * A_PLANE + ks * 4 * 128, 128, K * 16);
                        const uint64_t bdesc = make_desc(b0 + u * B_PLANE + ks * 2 * (N * 16), N * 16, 128);
                        const uint32_t accumulate = first ? 0u : 1u;
                        first = false;
                        asm volatile(
                            "{\n\t"
                            ".reg .pred p;\n\t"
                            "setp.ne.b32 p, %4, 0;\n\t"
                            "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t"
                            "}\n" ::"r"(tmem_base + (uint32_t)((w % 3) * ACC_STRIDE)), "l"(adesc), "l"(bdesc), "r"(idesc),
                            "r"(accumulate), "r"(0), "r"(0), "r"(0), "r"(0)
                            : "memory");
                    }
                }
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        }
        __syncwarp();
        uint32_t done = 0;
        for (long spin = 0; spin < (1L << 22) && !done; ++spin) {
            asm volatile(
                "{\n\t"
                ".reg .pred p;\n\t"
                "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                "selp.u32 %0, 1, 0, p;\n\t"
                "}\n" : "=r"(done) : "r"(smem_u32(&bar)), "r"((uint32_t)(it & 1)) : "memory");
        }
        if (!done) {
            if (lane == 0) atomicExch(status, 2 + pass);
            ok = false;
        }
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (ok && EPILOGUE == 2) {
            const int cbase = (warp >> 2) * 72;
#pragma unroll
            for (int ch = 0; ch < 9; ++ch) {
                uint32_t v[3][8];
#pragma unroll
                for (int q = 0; q < 3; ++q) {
                    const int w = w_hi - q;
                    const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)((w % 3) * ACC_STRIDE + cbase + ch * 8);
                    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n"
                                 : "=r"(v[q][0]), "=r"(v[q][1]), "=r"(v[q][2]), "=r"(v[q][3]), "=r"(v[q][4]), "=r"(v[q][5]),
                                   "=r"(v[q][6]), "=r"(v[q][7])
                                 : "r"(taddr));
                }
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    double a = pass == 0 ? 0.0 : racc[ch * 8 + j];
#pragma unroll
                    for (int q = 0; q < 3; ++q) {
                        const double wt = __longlong_as_double((long long)(1023 - (12 + 7 * (w_hi - q))) << 52);
                        a = __dadd_rn(a, __dmul_rn((double)(int32_t)v[q][j], wt));
                    }
                    racc[ch * 8 + j] = a;
                }
            }
            if (pass == 1) {
                const double sa = scaleA[row];
#pragma unroll
                for (int c = 0; c < 72; c += 2) {
                    double2 o;
                    o.x = __dmul_rn(__dmul_rn(racc[c], sa), scaleB[cbase + c]);
                    o.y = __dmul_rn(__dmul_rn(racc[c + 1], sa), scaleB[cbase + c + 1]);
                    *reinterpret_cast<double2*>(out + (size_t)row * N + cbase + c) = o;
                }
            }
        }
        if (ok && EPILOGUE == 1 && warp < 4) {
            const double sa = scaleA[row];
            for (int c0 = 0; c0 < N; c0 += 16) {
                double acc[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) acc[j] = pass == 0 ? 0.0 : out[(size_t)row * N + c0 + j];
                for (int w = w_hi; w > w_hi - 3; --w) {
                    uint32_t v[16];
                    const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)((w % 3) * ACC_STRIDE + c0);
                    asm volatile(
                        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
                        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
                        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                        : "r"(taddr));
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    const double wt = __longlong_as_double((long long)(1023 - (12 + 7 * w)) << 52);      // 2^-(12 + 7 w), exact
#pragma unroll
                    for (int j = 0; j < 16; ++j) acc[j] = __dadd_rn(acc[j], __dmul_rn((double)(int32_t)v[j], wt));
                }
#pragma unroll
                for (int j = 0; j < 16; ++j)
                    out[(size_t)row * N + c0 + j] = pass == 0 ? acc[j] : __dmul_rn(__dmul_rn(acc[j], sa), scaleB[c0 + j]);
            }
        }
        // the next pass overwrites the accumulators: every warp must have read them
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    }
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS));
}

// digits of one vector (stride 1, n entries) behind a power-of-two scale: a = scale * sum_t q_t 2^-(6 + 7 t) + rest
static double cut(const double* a, int n, int8_t* planes, size_t plane_stride, size_t elem_stride) {
    double amax = 0;
    for (int i = 0; i < n; ++i) amax = std::fmax(amax, std::fabs(a[i]));
    const double scale = amax > 0 ? std::exp2(std::ceil(std::log2(amax))) : 1.0;
    for (int i = 0; i < n; ++i) {
        double r = a[i] / scale;
        for (int t = 0; t < S; ++t) {
            r *= t == 0 ? 64.0 : 128.0;
            const double q = std::nearbyint(r);
            planes[t * plane_stride + i * elem_stride] = (int8_t)q;
            r -= q;
        }
    }
    return scale;
}

int main() {
    cudaDeviceProp p;
    CK(cudaGetDeviceProperties(&p, 0));
    std::vector<double> Ar((size_t)M * K), Br((size_t)N * K), sA(M), sB(N);
    srand(11);
    auto rnd = [] { return (rand() / (double)RAND_MAX - 0.5) * 2.0; };
    for (int m = 0; m < M; ++m) {
        const double mag = std::exp2((double)(rand() % 9 - 4));        // rows of different magnitude
        for (int k = 0; k < K; ++k) Ar[(size_t)m * K + k] = mag * rnd() * (1.0 + 3.0 * rnd() * rnd());
    }
    for (int n = 0; n < N; ++n) {
        const double mag = std::exp2((double)(rand() % 7 - 3));
        for (int k = 0; k < K; ++k) Br[(size_t)n * K + k] = mag * rnd();
    }
    // planes: A as (S x K x M) with M contiguous, B as (S x N x K) with K contiguous
    std::vector<int8_t> hA((size_t)S * A_PLANE), hB((size_t)S * B_PLANE);
    for (int m = 0; m < M; ++m) sA[m] = cut(&Ar[(size_t)m * K], K, hA.data() + m, A_PLANE, M);
    for (int n = 0; n < N; ++n) sB[n] = cut(&Br[(size_t)n * K], K, hB.data() + (size_t)n * K, B_PLANE, 1);

    int8_t *dA, *dB;
    double *dsA, *dsB, *dout;
    int* dstatus;
    CK(cudaMalloc(&dA, hA.size()));
    CK(cudaMalloc(&dB, hB.size()));
    CK(cudaMalloc(&dsA, sizeof(double) * M));
    CK(cudaMalloc(&dsB, sizeof(double) * N));
    CK(cudaMalloc(&dout, sizeof(double) * M * N * p.multiProcessorCount));
    CK(cudaMalloc(&dstatus, sizeof(int)));
    CK(cudaMemcpy(dA, hA.data(), hA.size(), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, hB.data(), hB.size(), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dsA, sA.data(), sizeof(double) * M, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dsB, sB.data(), sizeof(double) * N, cudaMemcpyHostToDevice));
    CK(cudaMemset(dout, 0xff, sizeof(double) * M * N));
    CK(cudaMemset(dstatus, 0, sizeof(int)));
    const int smem_bytes = S * (A_PLANE + B_PLANE);
    CK(cudaFuncSetAttribute(sliced_tile_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    CK(cudaFuncSetAttribute(sliced_tile_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    CK(cudaFuncSetAttribute(sliced_tile_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    sliced_tile_kernel<2><<<1, 256, smem_bytes>>>(dA, dB, dsA, dsB, dout, 1, dstatus);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    int status = 0;
    CK(cudaMemcpy(&status, dstatus, sizeof(int), cudaMemcpyDeviceToHost));
    std::vector<double> got((size_t)M * N);
    CK(cudaMemcpy(got.data(), dout, sizeof(double) * M * N, cudaMemcpyDeviceToHost));

    long bit_mismatches = 0;
    double max_rel = 0, ref_max = 0;
    for (int m = 0; m < M; ++m)
        for (int n = 0; n < N; ++n) {
            double exact = 0;
            for (int k = 0; k < K; ++k) exact += Ar[(size_t)m * K + k] * Br[(size_t)n * K + k];
            ref_max = std::fmax(ref_max, std::fabs(exact));
        }
    for (int m = 0; m < M; ++m)
        for (int n = 0; n < N; ++n) {
            double acc = 0, exact = 0;
            for (int w = S - 1; w >= 0; --w) {
                long long sw = 0;
                for (int t = 0; t <= w; ++t) {
                    const int u = w - t;
                    if (t >= S || u >= S) continue;
                    for (int k = 0; k < K; ++k)
                        sw += (long long)hA[(size_t)t * A_PLANE + (size_t)k * M + m] * (long long)hB[(size_t)u * B_PLANE + (size_t)n * K + k];
                }
                acc = acc + (double)sw * std::exp2(-(double)(12 + 7 * w));
            }
            const double want = (acc * sA[m]) * sB[n];
            for (int k = 0; k < K; ++k) exact += Ar[(size_t)m * K + k] * Br[(size_t)n * K + k];
            const double g = got[(size_t)m * N + n];
            if (std::memcmp(&g, &want, sizeof(double)) != 0) ++bit_mismatches;
            max_rel = std::fmax(max_rel, std::fabs(g - exact) / ref_max);
        }
    printf("{\"gpu\": \"%s\", \"tile\": \"%dx%dx%d float64 via %d int8 digits, %d slice-pair GEMMs\", \"status\": %d, "
           "\"bit_mismatches_vs_host_integer_arithmetic\": %ld, \"max_error_vs_float64_product_rel_to_max\": %.3e}\n",
           p.name, M, N, K, S, S * (S + 1) / 2, status, bit_mismatches, max_rel);
    if (status == 0 && bit_mismatches == 0) {
        // one CTA per SM, the tile redone `reps` times: time per tile with and without the float64 epilogue
        const int reps = 400, sms = p.multiProcessorCount;
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        float ms_epi = 1e30f, ms_mma = 1e30f, ms_reg = 1e30f;
        for (int r = 0; r < 4; ++r) {
            float ms;
            cudaEventRecord(e0);
            sliced_tile_kernel<1><<<sms, 128, smem_bytes>>>(dA, dB, dsA, dsB, dout, reps, dstatus);
            cudaEventRecord(e1);
            CK(cudaEventSynchronize(e1));
            cudaEventElapsedTime(&ms, e0, e1);
            if (r > 0 && ms < ms_epi) ms_epi = ms;
            cudaEventRecord(e0);
            sliced_tile_kernel<0><<<sms, 128, smem_bytes>>>(dA, dB, dsA, dsB, dout, reps, dstatus);
            cudaEventRecord(e1);
            CK(cudaEventSynchronize(e1));
            cudaEventElapsedTime(&ms, e0, e1);
            if (r > 0 && ms < ms_mma) ms_mma = ms;
            cudaEventRecord(e0);
            sliced_tile_kernel<2><<<sms, 256, smem_bytes>>>(dA, dB, dsA, dsB, dout, reps, dstatus);
            cudaEventRecord(e1);
            CK(cudaEventSynchronize(e1));
            cudaEventElapsedTime(&ms, e0, e1);
            if (r > 0 && ms < ms_reg) ms_reg = ms;
        }
        CK(cudaMemcpy(&status, dstatus, sizeof(int), cudaMemcpyDeviceToHost));
        const double ops = 2.0 * M * N * K * (S * (S + 1) / 2);
        printf("{\"timed_status\": %d, \"us_per_tile_mma_only\": %.3f, \"us_per_tile_with_float64_epilogue\": %.3f, \"us_per_tile_with_register_epilogue_8_warps\": %.3f, "
               "\"mma_only_tops\": %.1f, \"note\": \"per SM; the production tile has K = 768, i.e. 6x the MMAs per epilogue\"}\n",
               status, ms_mma * 1e3 / reps, ms_epi * 1e3 / reps, ms_reg * 1e3 / reps, ops * reps * sms / ms_mma * 1e-9);
    }
    return 0;
}
