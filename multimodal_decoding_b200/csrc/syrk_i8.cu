// Cluster-union Gram on the 5th-generation tensor cores (tcgen05, kind::i8), EXACT:
//
//   G_U = Xt[U] Xt[U]^T        (u x u, lower triangle; the per-centre normal matrices are sub-matrices of it)
//
// The betas are float16.  Behind one power-of-two scale per vertex row every float16 value is a 41-bit fixed-point number
// (row maximum < 2^17), i.e. exactly six signed 7-bit digits:  x = s * sum_t q_t 2^-(6 + 7 t),  |q_t| <= 64.  The Gram entry is
// then  s_i s_j * sum_{t,u} 2^-(12 + 7 (t + u)) * (Q_t Q_u^T)_ij  with every digit-pair product an exact int32 sum (64^2 * 4096
// < 2^31).  Pairs of equal weight w = t + u share one accumulator in tensor memory; the pairs with w >= 8 are dropped: their
// total is below 2^-42 of s_i s_j N, under the rounding of the float64 DMMA path they replace (and the digits t >= 3 are zero
// for all but the smallest elements of a row).  A row that is NOT exact in six digits flags its cluster, and that cluster's
// Gram is recomputed by the float64 kernel (gemm_gather.cu) -- same device-side pattern as the prediction guard.
//
// slice_union_kernel   one warp per union row: scale, six digit planes (int8, row pitch ldt), exactness flag
// syrk_i8_kernel       persistent CTAs over 128 x 128 lower-triangle tiles; warp 0 = TMA producer (tiled loads, 64B swizzle,
//                      K-major for both operands, 2-stage ring of 64-byte k-blocks), warp 1 = MMA issuer (converged, elected
//                      lane), warps 4-19 = epilogue (tcgen05.ld, float64 recombination, scales, store).  Eight weights need
//                      eight 128-column accumulators, tensor memory holds four: two passes over k (weights 0-3 with digit
//                      planes 0-3, then weights 4-7 with all six planes).
#include <cuda.h>

#include "common.cuh"
#include "kernels.h"

namespace mmd {
namespace {

constexpr int S = 6;                             // digits
constexpr int T = 128;                           // tile edge: TMEM lanes (rows) and accumulator columns
constexpr int KB = 64;                           // bytes (= int8 elements) of k per ring stage: one 64B-swizzle atom row
constexpr int PLANE = T * KB;                    // 8 KB: 128 rows x 64 B
constexpr int STAGE_BYTES = 2 * S * PLANE;       // A planes 0..5 then B planes 0..5
constexpr int STAGES = 2;
constexpr int N_EPI_WARPS = 16, EPI_COLS = T / 4;      // 4 lane quarters x 4 column groups of 32
constexpr int EPI_THREADS = N_EPI_WARPS * 32;
constexpr int THREADS = 128 + EPI_THREADS;
constexpr int OFF_BAR = STAGES * STAGE_BYTES;
constexpr int SMEM_BYTES = OFF_BAR + 128 + 1024;
static_assert(STAGE_BYTES % 1024 == 0 && PLANE % 512 == 0, "swizzle atoms must stay aligned");

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// bounded wait: false after ~2 s (the caller records a status code and leaves)
__device__ __forceinline__ bool mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    unsigned long long t0 = 0;
    for (uint32_t spin = 0;; ++spin) {
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t"
            "}\n"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (done) return true;
        if ((spin & 255u) == 255u) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (t0 == 0) t0 = t;
            else if (t - t0 > 2000000000ull) return false;
        }
    }
}
// K-major operand, 64B swizzle: 8 rows of 64 B per atom (512 B apart)
__device__ __forceinline__ uint64_t desc_k_sw64(uint32_t addr) {
    const uint32_t lo = ((addr >> 4) & 0x3fffu) | (1u << 16);
    const uint32_t hi = ((512u >> 4) & 0x3fffu) | (1u << 14) | (4u << 29);
    return ((uint64_t)hi << 32) | lo;
}
// called by all lanes of the converged issuing warp with warp-uniform arguments; one elected lane issues (see predict_i8.cu)
__device__ __forceinline__ void mma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p, e;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "elect.sync _|e, 0xffffffff;\n\t"
        "@e tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void mma_commit(uint32_t bar) {
    asm volatile(
        "{\n\t"
        ".reg .pred e;\n\t"
        "elect.sync _|e, 0xffffffff;\n\t"
        "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t"
        "}\n" ::"r"(bar)
        : "memory");
}
__device__ __forceinline__ void tma_tile2d(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
        "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}

// digit pairs of pass P for one ring stage (two k-steps of 32): weights 4 P .. 4 P + 3, accumulator (w & 3) * 128
template <int PASS>
__device__ __forceinline__ void issue_pairs(uint32_t tmem_base, uint32_t a_base, uint32_t b_base, uint32_t idesc, uint32_t& started) {
#pragma unroll
    for (int w = 4 * PASS; w < 4 * PASS + 4; ++w) {
        const uint32_t acc = tmem_base + (uint32_t)((w & 3) * T);
#pragma unroll
        for (int t = 0; t < S; ++t) {
            const int u = w - t;
            if (u < 0 || u >= S) continue;
#pragma unroll
            for (int kk = 0; kk < KB / 32; ++kk) {
                mma_i8(acc, desc_k_sw64(a_base + t * PLANE + kk * 32), desc_k_sw64(b_base + u * PLANE + kk * 32), idesc,
                       (started >> (w & 3)) & 1u);
                started |= 1u << (w & 3);
            }
        }
    }
}

__device__ __forceinline__ void cut_digits6(double a, double inv_scale, int8_t q[S], bool& exact) {
    double r = a * inv_scale;
#pragma unroll
    for (int t = 0; t < S; ++t) {
        r *= t == 0 ? 64.0 : 128.0;
        const double d = rint(r);
        q[t] = (int8_t)(int)d;
        r -= d;
    }
    exact = exact && r == 0.0;
}

}  // namespace

// one warp per union row (rows >= u of a cluster are left as they are: their products only reach output rows / columns >= u)
__global__ void __launch_bounds__(256) slice_union_kernel(const __half* __restrict__ Xt, int ldt, const SyrkCluster* __restrict__ clusters,
                                                          int8_t* __restrict__ planes) {
    const SyrkCluster cl = clusters[blockIdx.y];
    const int r = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (r >= cl.u) return;
    const __half* row = Xt + (long)(cl.rows ? cl.rows[r] : r) * ldt;
    float m = 0.f;
    for (int j = lane * 8; j < ldt; j += 256) {
        const int4 v = *reinterpret_cast<const int4*>(row + j);
        const __half2* h = reinterpret_cast<const __half2*>(&v);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float2 f = __half22float2(h[e]);
            m = fmaxf(m, fmaxf(fabsf(f.x), fabsf(f.y)));
        }
    }
    for (int o = 16; o; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    // smallest power of two >= the row maximum (1 for an all-zero row); NaN / infinity never reach here (rejected at upload)
    double scale = 1.0;
    if (m > 0.f) {
        int e;
        const float fr = frexpf(m, &e);
        scale = fr == 0.5f ? ldexp(1.0, e - 1) : ldexp(1.0, e);
    }
    const double inv = 1.0 / scale;
    bool exact = true;
    for (int j = lane * 8; j < ldt; j += 256) {
        const int4 v = *reinterpret_cast<const int4*>(row + j);
        const __half* h = reinterpret_cast<const __half*>(&v);
        union { int8_t b[8]; int2 w; } out[S];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            int8_t q[S];
            cut_digits6((double)__half2float(h[e]), inv, q, exact);
#pragma unroll
            for (int t = 0; t < S; ++t) out[t].b[e] = q[t];
        }
#pragma unroll
        for (int t = 0; t < S; ++t)
            *reinterpret_cast<int2*>(planes + ((long)cl.row_base + (long)t * cl.up + r) * ldt + j) = out[t].w;
    }
    if (lane == 0) cl.scales[r] = scale;
    if (!__all_sync(0xffffffffu, exact) && lane == 0) atomicExch(cl.inexact, 1);
}

__global__ void __launch_bounds__(THREADS, 1)
syrk_i8_kernel(const __grid_constant__ CUtensorMap tm, const SyrkItem* __restrict__ items, int n_items,
               const SyrkCluster* __restrict__ clusters, int nk, int* __restrict__ status) {
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* smem = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
    const uint32_t ring = smem_u32(smem);
    const uint32_t bar0 = smem_u32(smem + OFF_BAR);
    auto full_bar = [&](int s) { return bar0 + 8u * s; };
    auto empty_bar = [&](int s) { return bar0 + 8u * (STAGES + s); };
    const uint32_t tfull_bar = bar0 + 8u * (2 * STAGES), tempty_bar = bar0 + 8u * (2 * STAGES + 1);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + OFF_BAR + 8 * (2 * STAGES + 2));
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(full_bar(s), 1);
            mbar_init(empty_bar(s), 1);
        }
        mbar_init(tfull_bar, 1);
        mbar_init(tempty_bar, N_EPI_WARPS);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ------------------------------------------------ TMA producer ------------------------------------------------
        uint32_t q = 0;
        bool ok = true;
        for (int it = blockIdx.x; it < n_items && ok; it += gridDim.x) {
            const SyrkItem item = items[it];
            const bool diag = item.row_a == item.row_b;
            for (int pass = 0; pass < 2 && ok; ++pass) {
                const int n_planes = pass == 0 ? 4 : S;
                for (int kb = 0; kb < nk && ok; ++kb, ++q) {
                    const int slot = q % STAGES;
                    if (!mbar_wait(empty_bar(slot), ((q / STAGES) & 1u) ^ 1u)) {
                        if (lane == 0) atomicExch(status, 1);
                        ok = false;
                        break;
                    }
                    const uint32_t dst = ring + slot * STAGE_BYTES;
                    if (lane == 0) mbar_expect_tx(full_bar(slot), (uint32_t)((diag ? 1 : 2) * n_planes * PLANE));
                    __syncwarp();
                    if (lane < n_planes) tma_tile2d(dst + lane * PLANE, &tm, full_bar(slot), kb * KB, item.row_a + lane * item.up);
                    else if (!diag && lane >= 8 && lane < 8 + n_planes)
                        tma_tile2d(dst + (S + lane - 8) * PLANE, &tm, full_bar(slot), kb * KB, item.row_b + (lane - 8) * item.up_b);
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------ MMA issuer ------------------------------------------------
        // D = S32, A = B = signed 8 bit, both K-major, M = N = 128
        const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(T >> 3) << 17) | ((uint32_t)(T >> 4) << 24);
        uint32_t q = 0, pc = 0;
        bool ok = true;
        for (int it = blockIdx.x; it < n_items && ok; it += gridDim.x) {
            const SyrkItem item = items[it];
            const bool diag = item.row_a == item.row_b;
            for (int pass = 0; pass < 2 && ok; ++pass, ++pc) {
                if (!mbar_wait(tempty_bar, (pc & 1u) ^ 1u)) {
                    if (lane == 0) atomicExch(status, 2);
                    ok = false;
                    break;
                }
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                uint32_t started = 0;
                for (int kb = 0; kb < nk && ok; ++kb, ++q) {
                    const uint32_t slot = q % STAGES;
                    if (!mbar_wait(full_bar(slot), (q / STAGES) & 1u)) {
                        if (lane == 0) atomicExch(status, 3);
                        ok = false;
                        break;
                    }
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t a_base = ring + slot * STAGE_BYTES, b_base = diag ? a_base : a_base + S * PLANE;
                    if (pass == 0) issue_pairs<0>(tmem_base, a_base, b_base, idesc, started);
                    else issue_pairs<1>(tmem_base, a_base, b_base, idesc, started);
                    mma_commit(empty_bar(slot));
                    __syncwarp();
                }
                if (ok) mma_commit(tfull_bar);
                __syncwarp();
            }
        }
    } else if (warp >= 4) {
        // ------------------------------------------------ epilogue ------------------------------------------------
        const int ew = warp - 4, cg = ew >> 2, lq = warp & 3;
        const int row = lq * 32 + lane;                                // tile row = TMEM lane
        uint32_t pc = 0;
        bool ok = true;
        double racc[EPI_COLS];
        for (int it = blockIdx.x; it < n_items && ok; it += gridDim.x) {
            const SyrkItem item = items[it];
            for (int pass = 0; pass < 2 && ok; ++pass, ++pc) {
                if (!mbar_wait(tfull_bar, pc & 1u)) {
                    if (lane == 0) atomicExch(status, 4);
                    ok = false;
                    break;
                }
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
                for (int ch = 0; ch < EPI_COLS / 8; ++ch) {
                    uint32_t v[4][8];
#pragma unroll
                    for (int a = 0; a < 4; ++a) {
                        const uint32_t taddr = tmem_base + ((uint32_t)(lq * 32) << 16) + (uint32_t)(a * T + cg * EPI_COLS + ch * 8);
                        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n"
                                     : "=r"(v[a][0]), "=r"(v[a][1]), "=r"(v[a][2]), "=r"(v[a][3]), "=r"(v[a][4]), "=r"(v[a][5]),
                                       "=r"(v[a][6]), "=r"(v[a][7])
                                     : "r"(taddr));
                    }
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        // the four weights of a pass span 50 bits: their sum (smallest first) is exact; the two pass sums are
                        // added once, so the result is the truncated digit-pair sum rounded to float64 once
                        double s = 0.0;
#pragma unroll
                        for (int a = 3; a >= 0; --a) {
                            const int w = 4 * pass + a;
                            const double wt = __longlong_as_double((long long)(1023 - (12 + 7 * w)) << 52);      // 2^-(12 + 7 w)
                            s = fma((double)(int32_t)v[a][j], wt, s);
                        }
                        racc[ch * 8 + j] = pass == 0 ? s : racc[ch * 8 + j] + s;
                    }
                }
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive(tempty_bar);                // the accumulators may be overwritten
            }
            if (!ok) break;
            const SyrkCluster cl = clusters[item.cluster];
            const int gr = item.m0 + row;
            if (gr < cl.u) {
                const double sr = cl.scales[gr];
                const int col0 = item.n0 + cg * EPI_COLS;
                const long orow = cl.out_rows ? (long)cl.out_rows[gr] : (long)gr;
                double* out = cl.C + orow * cl.ldc + col0;
                const double* sc = (cl.u_b > 0 ? cl.scales_b : cl.scales) + col0;
                // symmetric product: lower triangle (columns <= row); rectangular product: all u_b columns
                const int n_valid = min(EPI_COLS, (cl.u_b > 0 ? cl.u_b : min(cl.u, gr + 1)) - col0);
#pragma unroll
                for (int c = 0; c < EPI_COLS; ++c)
                    if (c < n_valid) out[c] = racc[c] * sr * sc[c];
            }
        }
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
}

size_t syrk_i8_smem_bytes() { return SMEM_BYTES; }

cudaError_t launch_slice_union(const __half* Xt, int ldt, const SyrkCluster* clusters_dev, int n_clusters, int max_u, int8_t* planes,
                               cudaStream_t s) {
    if (n_clusters <= 0) return cudaSuccess;
    if (ldt % 64 != 0) return cudaErrorInvalidValue;
    dim3 grid((max_u + 7) / 8, n_clusters);
    slice_union_kernel<<<grid, 256, 0, s>>>(Xt, ldt, clusters_dev, planes);
    return cudaGetLastError();
}

cudaError_t launch_syrk_i8(const int8_t* planes, long plane_rows, int ldt, const SyrkItem* items_dev, int n_items,
                           const SyrkCluster* clusters_dev, int n_sms, int* status, cudaStream_t s) {
    if (n_items <= 0) return cudaSuccess;
    if (ldt % KB != 0 || plane_rows > 0x7fffffffL) return cudaErrorInvalidValue;
    alignas(64) CUtensorMap tm;
    cudaError_t e = make_tensor_map_u8(&tm, planes, (uint64_t)ldt, (uint64_t)plane_rows, (uint64_t)ldt, KB, T, 64);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(syrk_i8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return e;
    const int grid = n_items < n_sms ? n_items : n_sms;
    syrk_i8_kernel<<<grid, THREADS, SMEM_BYTES, s>>>(tm, items_dev, n_items, clusters_dev, ldt / KB, status);
    return cudaGetLastError();
}

}  // namespace mmd
