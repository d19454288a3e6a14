// Predictions for the evaluation trials of every centre on the 5th-generation tensor cores (tcgen05, kind::i8):
//
//   Yhat(144 x Dp) = Zt(144 x kp) * XtY[nbr, :](kp x Dp)          (same product as predict.cu; searchlight.py:58-59)
//
// Both float64 operands are cut into S = 6 signed 7-bit digits behind power-of-two scales (one scale per feature
// column of XtY, fixed at upload; one per evaluation row of Zt, per centre):
//     a = scale * sum_t q_t 2^-(6 + 7 t) + rest,   |q_t| <= 64,   |rest| <= scale 2^-42.
// The product is the sum of the digit-pair GEMMs with t + u < S (21 of them); pairs of equal weight w = t + u share one
// int32 accumulator in tensor memory, so every partial sum is exact, and the six accumulators are recombined in
// float64, smallest weight first.  What is dropped is bounded (|error| <= 8 kp 2^-42 sX[d] sZ[e] per element); the
// epilogue hands the score kernel what it needs to turn that bound into a per-row tolerance on the cosine distances,
// and a centre with a comparison inside its tolerance is redone by the float64 kernels (predict.cu), see score.cu.
//
// One persistent CTA per SM walks (centre, 128-feature tile) items:
//   warps 0,2,3 TMA producers, one digit plane of a ring slot each: the neighbour rows of the XtY digit planes arrive by
//               cp.async.bulk.tensor ... tile::gather4 (4 rows x 128 B per instruction, 128B-swizzled, MN-major operand
//               A), the Zt digit planes by tiled TMA (144 x 64 B boxes, 64B-swizzled, K-major operand B); ring of 4
//               slots of 3 + 3 planes x 64 neighbours
//   warp 1      issues tcgen05.mma (M = 128 features, N = 144 evaluation rows, K = 32 per instruction; the warp stays
//               converged and one elected lane issues, descriptors in uniform registers) and commits the slots back
//   warps 4-15  epilogue: tcgen05.ld, float64 recombination, per-feature z-scoring of the three prediction groups
//               (StandardScaler semantics, as predict.cu), stores
// TMEM holds three 144-column accumulators, so a tile takes two passes over its neighbours: weights 5,4,3 (all six
// planes of both operands), then 2,1,0 (planes 0-2).  Every wait is bounded: a protocol error ends in a status code.
#include <cuda.h>

#include "common.cuh"
#include "kernels.h"

namespace mmd {
namespace {

constexpr int S = I8_DIGITS;
constexpr int TM = 128;                          // feature columns per tile = TMEM lanes
constexpr int TN = EVAL_PAD;                     // 144 evaluation rows = accumulator columns
constexpr int KS = 64;                           // neighbours per ring slot
constexpr int HALF = 3;                          // digit planes per ring slot (of each operand)
constexpr int SLOTS = 4;
constexpr int A_PLANE = KS * TM;                 // 8192 B: 64 neighbour rows x 128 feature bytes
constexpr int B_PLANE = TN * KS;                 // 9216 B: 144 evaluation rows x 64 neighbour bytes
constexpr int SLOT_BYTES = HALF * (A_PLANE + B_PLANE);
constexpr int ACC_STRIDE = 160;                  // accumulator of weight w lives at TMEM column (w % 3) * 160
constexpr int TMEM_COLS = 512;
constexpr int N_EPI_WARPS = 12, EPI_COLS = TN / 3;     // 3 column groups x 4 lane quarters
constexpr int EPI_THREADS = N_EPI_WARPS * 32;
constexpr int THREADS = 128 + EPI_THREADS;     // warps 0-3: three producers + the MMA issuer (one warpgroup); warps 4-15: epilogue
constexpr int MAX_KP = 2048;
static_assert(S == 6, "pass structure below is written for six digits");
static_assert(SLOT_BYTES % 1024 == 0 && A_PLANE % 1024 == 0 && B_PLANE % 512 == 0, "swizzle atoms must stay aligned");

// shared memory carve-up (after manual 1024-byte alignment of the dynamic window)
constexpr int OFF_RING = 0;
constexpr int OFF_NBR = OFF_RING + SLOTS * SLOT_BYTES;
constexpr int OFF_RED = OFF_NBR + MAX_KP * 4;
constexpr int OFF_BAR = OFF_RED + 3 * 3 * TM * 8;
constexpr int OFF_END = OFF_BAR + 128;
constexpr int SMEM_BYTES = OFF_END + 1024;

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

// UMMA shared-memory descriptors (sm_100 format: version 1 at bit 46, layout type at bits 61-63)
__device__ __forceinline__ uint64_t desc_a_mn_sw128(uint32_t addr) {   // MN-major, 128B swizzle: 8 k-rows of 128 B per atom
    const uint32_t lo = ((addr >> 4) & 0x3fffu) | (1u << 16);
    const uint32_t hi = ((1024u >> 4) & 0x3fffu) | (1u << 14) | (2u << 29);
    return ((uint64_t)hi << 32) | lo;
}
__device__ __forceinline__ uint64_t desc_b_k_sw64(uint32_t addr) {     // K-major, 64B swizzle: 8 rows of 64 B per atom
    const uint32_t lo = ((addr >> 4) & 0x3fffu) | (1u << 16);
    const uint32_t hi = ((512u >> 4) & 0x3fffu) | (1u << 14) | (4u << 29);
    return ((uint64_t)hi << 32) | lo;
}

// Both are called by ALL lanes of the (converged) issuing warp with warp-uniform arguments; one elected lane issues.
// Keeping the control flow convergent lets ptxas hold the descriptors in uniform registers instead of wrapping every
// instruction in a per-lane broadcast loop.
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

__device__ __forceinline__ void tma_gather4(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int col, int r0, int r1, int r2, int r3) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::"r"(dst),
        "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar), "r"(col), "r"(r0), "r"(r1), "r"(r2), "r"(r3)
        : "memory");
}
__device__ __forceinline__ void tma_tile2d(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
        "l"(reinterpret_cast<uint64_t>(tm)), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}

__device__ __forceinline__ void epi_barrier() { asm volatile("bar.sync 1, %0;" ::"n"(EPI_THREADS) : "memory"); }
__device__ __forceinline__ void prod_barrier() { asm volatile("bar.sync 2, 96;" ::: "memory"); }

}  // namespace

size_t predict_i8_smem_bytes() { return SMEM_BYTES; }

// ---------------------------------------------------------------------------------------------------------------
// digit planes
// ---------------------------------------------------------------------------------------------------------------
// a / scale cut into S digits: exact float64 steps (scaling by powers of two, rint, difference of nearby numbers)
__device__ __forceinline__ void cut_digits(double a, double inv_scale, int8_t q[S]) {
    double r = a * inv_scale;
#pragma unroll
    for (int t = 0; t < S; ++t) {
        r *= t == 0 ? 64.0 : 128.0;
        const double d = rint(r);
        q[t] = (int8_t)(int)d;
        r -= d;
    }
}
// smallest power of two >= amax (1 for amax == 0); amax finite
__device__ __forceinline__ double pow2_ceil(double amax) {
    if (!(amax > 0.0)) return 1.0;
    int e;
    const double m = frexp(amax, &e);            // amax = m 2^e, 0.5 <= m < 1
    return m == 0.5 ? ldexp(1.0, e - 1) : ldexp(1.0, e);
}

// column maxima of |A| (n_rows x n_cols, row-major) -> bits of the maximum per column (non-negative doubles order like integers)
__global__ void colmax_kernel(const double* __restrict__ A, const int* __restrict__ rows, long n_rows, int n_cols,
                              unsigned long long* __restrict__ colmax) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_cols) return;
    double m = 0.0;
    for (long r = blockIdx.y; r < n_rows; r += gridDim.y) m = fmax(m, fabs(A[(rows ? (long)rows[r] : r) * n_cols + c]));
    atomicMax(colmax + c, (unsigned long long)__double_as_longlong(m));
}
__global__ void colscale_kernel(const unsigned long long* __restrict__ colmax, int n_cols, double* __restrict__ scale) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < n_cols) scale[c] = pow2_ceil(__longlong_as_double((long long)colmax[c]));
}
// planes[t][r][c] = digit t of A[r][c] / scale[c]; one thread cuts 16 consecutive columns of a row
__global__ void slice_columns_kernel(const double* __restrict__ A, const int* __restrict__ rows, long n_list, long n_rows,
                                     int n_cols, const double* __restrict__ scale, int8_t* __restrict__ planes) {
    const long chunks_per_row = n_cols / 16;
    const long total = n_list * chunks_per_row;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const long r = rows ? (long)rows[i / chunks_per_row] : i / chunks_per_row;
        const int c0 = (int)(i % chunks_per_row) * 16;
        union { int8_t b[16]; int4 v; } out[S];
#pragma unroll
        for (int j = 0; j < 16; j += 2) {
            const double2 a = *reinterpret_cast<const double2*>(A + r * n_cols + c0 + j);
            int8_t q[S];
            cut_digits(a.x, 1.0 / scale[c0 + j], q);
#pragma unroll
            for (int t = 0; t < S; ++t) out[t].b[j] = q[t];
            cut_digits(a.y, 1.0 / scale[c0 + j + 1], q);
#pragma unroll
            for (int t = 0; t < S; ++t) out[t].b[j + 1] = q[t];
        }
#pragma unroll
        for (int t = 0; t < S; ++t)
            *reinterpret_cast<int4*>(planes + ((long)t * n_rows + r) * n_cols + c0) = out[t].v;
    }
}

cudaError_t launch_slice_xty(const double* XtY, int V, int Dp, int8_t* planes, double* sX, unsigned long long* colmax_ws,
                             const int* rows, int n_rows, cudaStream_t s) {
    if (Dp % 16 != 0) return cudaErrorInvalidValue;
    cudaError_t e = cudaMemsetAsync(colmax_ws, 0, (size_t)Dp * sizeof(unsigned long long), s);
    if (e != cudaSuccess) return e;
    dim3 g((Dp + 255) / 256, 592);
    const long n_list = rows ? n_rows : V;
    colmax_kernel<<<g, 256, 0, s>>>(XtY, rows, n_list, Dp, colmax_ws);
    colscale_kernel<<<(Dp + 255) / 256, 256, 0, s>>>(colmax_ws, Dp, sX);
    slice_columns_kernel<<<148 * 16, 256, 0, s>>>(XtY, rows, n_list, V, Dp, sX, planes);
    return cudaGetLastError();
}

// Zt of every task -> digit planes zplanes[task][u][e][kp_max] and row scales sZ[task][e]; one CTA per task, a warp per row
__global__ void __launch_bounds__(256) slice_zt_kernel(const CentreTask* __restrict__ tasks, int kp_max, const double* __restrict__ work,
                                                       int8_t* __restrict__ zplanes, double* __restrict__ sZ) {
    const int task_id = blockIdx.x;
    const int kp = tasks[task_id].kp;
    const double* Zt = work + (long)task_id * (kp_max + EVAL_PAD) * kp_max + (long)kp * kp;
    int8_t* out = zplanes + (long)task_id * S * TN * kp_max;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int e = warp; e < TN; e += 8) {
        const double* row = Zt + (long)e * kp;
        double m = 0.0;
        for (int j = lane * 2; j < kp; j += 64) {
            const double2 v = *reinterpret_cast<const double2*>(row + j);
            m = fmax(m, fmax(fabs(v.x), fabs(v.y)));
        }
        for (int o = 16; o; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
        const double scale = pow2_ceil(m), inv = 1.0 / scale;
        if (lane == 0) sZ[(long)task_id * TN + e] = scale;
        for (int j = lane * 4; j < kp; j += 128) {
            const double2 v0 = *reinterpret_cast<const double2*>(row + j);
            const double2 v1 = *reinterpret_cast<const double2*>(row + j + 2);
            int8_t q0[S], q1[S], q2[S], q3[S];
            cut_digits(v0.x, inv, q0);
            cut_digits(v0.y, inv, q1);
            cut_digits(v1.x, inv, q2);
            cut_digits(v1.y, inv, q3);
#pragma unroll
            for (int u = 0; u < S; ++u)
                *reinterpret_cast<char4*>(out + ((long)u * TN + e) * kp_max + j) = make_char4(q0[u], q1[u], q2[u], q3[u]);
        }
    }
}

cudaError_t launch_slice_zt(const CentreTask* tasks, int n_tasks, int kp_max, const double* work, int8_t* zplanes, double* sZ,
                            cudaStream_t s) {
    if (n_tasks <= 0) return cudaSuccess;
    slice_zt_kernel<<<n_tasks, 256, 0, s>>>(tasks, kp_max, work, zplanes, sZ);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------
// the GEMM
// ---------------------------------------------------------------------------------------------------------------
// The digit-pair MMAs of one ring step (64 neighbours) for the three weights W_HI, W_HI - 1, W_HI - 2; fully unrolled, so
// every shared-memory offset is an immediate.  base0 holds planes 0-2 of both operands, base1 planes 3-5.
template <int W_HI>
__device__ __forceinline__ void issue_pairs(uint32_t tmem_base, uint32_t base0, uint32_t base1, uint32_t idesc, uint32_t& started) {
#pragma unroll
    for (int w = W_HI; w > W_HI - 3; --w) {
        const uint32_t acc = tmem_base + (uint32_t)((w % 3) * ACC_STRIDE);
#pragma unroll
        for (int t = 0; t <= w; ++t) {
            const int u = w - t;
            if (t >= S || u >= S) continue;
            const uint32_t a_addr = (t < HALF ? base0 : base1) + (t % HALF) * A_PLANE;
            const uint32_t b_addr = (u < HALF ? base0 : base1) + HALF * A_PLANE + (u % HALF) * B_PLANE;
#pragma unroll
            for (int kk = 0; kk < KS / 32; ++kk) {
                mma_i8(acc, desc_a_mn_sw128(a_addr + kk * 32 * TM), desc_b_k_sw64(b_addr + kk * 32), idesc,
                       (started >> (w % 3)) & 1u);
                started |= 1u << (w % 3);
            }
        }
    }
}

__global__ void __launch_bounds__(THREADS, 1)
predict_i8_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, PredictI8Params prm) {
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* smem = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
    const uint32_t ring = smem_u32(smem + OFF_RING);
    int* nbr_sm = reinterpret_cast<int*>(smem + OFF_NBR);
    double* red = reinterpret_cast<double*>(smem + OFF_RED);          // [3 column groups][3 prediction groups][TM]
    const uint32_t bar0 = smem_u32(smem + OFF_BAR);
    auto full_bar = [&](int s) { return bar0 + 8u * s; };
    auto empty_bar = [&](int s) { return bar0 + 8u * (SLOTS + s); };
    const uint32_t tfull_bar = bar0 + 8u * (2 * SLOTS), tempty_bar = bar0 + 8u * (2 * SLOTS + 1);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + OFF_BAR + 8 * (2 * SLOTS + 2));

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n_ft = prm.Dp / TM;
    const long n_items = (long)prm.n_tasks * n_ft;

    if (tid == 0) {
        for (int s = 0; s < SLOTS; ++s) {
            mbar_init(full_bar(s), 1);
            mbar_init(empty_bar(s), 1);
        }
        mbar_init(tfull_bar, 1);
        mbar_init(tempty_bar, N_EPI_WARPS);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;

    if (warp < 4) {
      asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
      if (warp != 1) {
        // ------------------------------------------------ TMA producers ------------------------------------------------
        // three warps (0, 2, 3), one digit plane of the slot each: the issue rate of the per-lane gather4 instructions,
        // not the memory system, bounds a single producer warp (ncu r2 v1: the producer never waits, the MMA warp does)
        const int pw = warp == 0 ? 0 : warp - 1;
        uint32_t q = 0;
        bool ok = true;
        for (long item = blockIdx.x; item < n_items; item += gridDim.x) {
            const int task_id = (int)(item / n_ft), d0 = (int)(item % n_ft) * TM;
            const CentreTask task = prm.tasks[task_id];
            const int k = task.k, kp = task.kp, nk = kp / KS;
            prod_barrier();                                              // the previous item's gathers have read nbr_sm
            for (int i = pw * 32 + lane; i < kp; i += 96) nbr_sm[i] = __ldg(task.nbr + (i < k ? i : 0));
            prod_barrier();
            for (int pass = 0; pass < 2 && ok; ++pass) {
                const int halves = pass == 0 ? 2 : 1;
                for (int ks = 0; ks < nk && ok; ++ks) {
                    for (int half = 0; half < halves; ++half, ++q) {
                        const int slot = q % SLOTS;
                        if (!mbar_wait(empty_bar(slot), ((q / SLOTS) & 1u) ^ 1u)) {
                            if (lane == 0) atomicExch(prm.status, 1);
                            ok = false;
                            break;
                        }
                        const uint32_t dst = ring + slot * SLOT_BYTES;
                        if (pw == 0 && lane == 0) mbar_expect_tx(full_bar(slot), SLOT_BYTES);
                        if (lane < KS / 4) {
                            const int4 r = *reinterpret_cast<const int4*>(nbr_sm + ks * KS + lane * 4);
                            const int off = (half * HALF + pw) * prm.V;
                            tma_gather4(dst + pw * A_PLANE + lane * 512, &tmA, full_bar(slot), d0, r.x + off, r.y + off,
                                        r.z + off, r.w + off);
                        } else if (lane == KS / 4) {
                            tma_tile2d(dst + HALF * A_PLANE + pw * B_PLANE, &tmB, full_bar(slot), ks * KS,
                                       (task_id * S + half * HALF + pw) * TN);
                        }
                    }
                }
            }
        }
      } else if (warp == 1) {
        // ------------------------------------------------ MMA issuer ------------------------------------------------
        // D = S32, A = B = signed 8 bit, A MN-major (gathered rows are contiguous along the features), B K-major
        const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | ((uint32_t)(TN >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);
        uint32_t q = 0, pc = 0;
        bool ok = true;
        for (long item = blockIdx.x; item < n_items && ok; item += gridDim.x) {
            const int task_id = (int)(item / n_ft);
            const int nk = prm.tasks[task_id].kp / KS;
            for (int pass = 0; pass < 2 && ok; ++pass, ++pc) {
                if (!mbar_wait(tempty_bar, (pc & 1u) ^ 1u)) {
                    if (lane == 0) atomicExch(prm.status, 2);
                    ok = false;
                    break;
                }
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const int halves = pass == 0 ? 2 : 1;
                const int w_hi = pass == 0 ? S - 1 : S - 4;                  // weights w_hi, w_hi - 1, w_hi - 2
                uint32_t started = 0;                                        // bit (w % 3): accumulator holds a value
                for (int ks = 0; ks < nk && ok; ++ks) {
                    const uint32_t slot0 = q % SLOTS, slot1 = (q + 1) % SLOTS;
                    if (!mbar_wait(full_bar(slot0), (q / SLOTS) & 1u) ||
                        (halves == 2 && !mbar_wait(full_bar(slot1), ((q + 1) / SLOTS) & 1u))) {
                        if (lane == 0) atomicExch(prm.status, 3);
                        ok = false;
                        break;
                    }
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    {
                        const uint32_t base0 = ring + slot0 * SLOT_BYTES, base1 = ring + slot1 * SLOT_BYTES;
                        if (pass == 0) issue_pairs<S - 1>(tmem_base, base0, base1, idesc, started);
                        else issue_pairs<S - 4>(tmem_base, base0, base0, idesc, started);
                        mma_commit(empty_bar(slot0));
                        if (halves == 2) mma_commit(empty_bar(slot1));
                    }
                    __syncwarp();
                    q += halves;
                }
                if (ok) mma_commit(tfull_bar);
                __syncwarp();
            }
        }
      }
    } else {
        asm volatile("setmaxnreg.inc.sync.aligned.u32 152;");
        // ------------------------------------------------ epilogue ------------------------------------------------
        const int ew = warp - 4, cg = ew >> 2, lq = warp & 3;
        const int f = lq * 32 + lane;                                 // feature of the tile = TMEM lane
        const int n_cap = prm.n_cap, n_test = prm.n_cap + prm.n_img, n_eval = n_test + prm.n_im;
        uint32_t pc = 0;
        bool ok = true;
        double racc[EPI_COLS];
        for (long item = blockIdx.x; item < n_items && ok; item += gridDim.x) {
            const int task_id = (int)(item / n_ft), d0 = (int)(item % n_ft) * TM;
            const double sx = __ldg(prm.sX + d0 + f);
            const double* sz = prm.sZ + (long)task_id * TN + cg * EPI_COLS;
            for (int pass = 0; pass < 2 && ok; ++pass, ++pc) {
                if (!mbar_wait(tfull_bar, pc & 1u)) {
                    if (lane == 0) atomicExch(prm.status, 4);
                    ok = false;
                    break;
                }
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const int w_hi = pass == 0 ? S - 1 : S - 4;
#pragma unroll
                for (int ch = 0; ch < EPI_COLS / 8; ++ch) {
                    uint32_t v[3][8];
#pragma unroll
                    for (int qq = 0; qq < 3; ++qq) {
                        const int w = w_hi - qq;
                        const uint32_t taddr = tmem_base + ((uint32_t)(lq * 32) << 16) + (uint32_t)((w % 3) * ACC_STRIDE + cg * EPI_COLS + ch * 8);
                        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n"
                                     : "=r"(v[qq][0]), "=r"(v[qq][1]), "=r"(v[qq][2]), "=r"(v[qq][3]), "=r"(v[qq][4]), "=r"(v[qq][5]),
                                       "=r"(v[qq][6]), "=r"(v[qq][7])
                                     : "r"(taddr));
                    }
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        double a = pass == 0 ? 0.0 : racc[ch * 8 + j];
#pragma unroll
                        for (int qq = 0; qq < 3; ++qq) {
                            const double wt = __longlong_as_double((long long)(1023 - (12 + 7 * (w_hi - qq))) << 52);   // 2^-(12 + 7 w)
                            a = fma((double)(int32_t)v[qq][j], wt, a);          // the product with a power of two is exact
                        }
                        // last pass: apply the two power-of-two scales (exact products)
                        if (pass == 1) a = __dmul_rn(__dmul_rn(a, sx), __ldg(sz + ch * 8 + j));
                        racc[ch * 8 + j] = a;
                    }
                }
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive(tempty_bar);                // the accumulators may be overwritten
            }
            if (!ok) break;


            double* out = prm.yhat + (long)task_id * YHAT_ROWS * prm.Dp + d0 + f;
            const int e0 = cg * EPI_COLS;
            if (!prm.standardize) {
                double* o = out + (long)e0 * prm.Dp;
#pragma unroll
                for (int c = 0; c < EPI_COLS; ++c, o += prm.Dp) *o = racc[c];
                continue;
            }
            // per-feature z-scoring of the three prediction groups (two exact passes like StandardScaler: mean, then
            // centred sum of squares); a feature's 144 rows sit in three threads (column groups), partial sums meet in `red`
            const int lo[3] = {0, n_cap, n_test}, hi[3] = {n_cap, n_test, n_eval};
            // (the empty asm statements pin the evaluation order row by row: without them ptxas evaluates all 48 rows
            // speculatively for all three groups and spills)
            // Rows are handled in blocks of 8; a block that lies inside one prediction group (all but two or three of
            // the 18 blocks) takes a path without per-row selects.  group_of() is warp-uniform, so are the branches.
            auto group_of = [&](int e) { return (e >= n_cap) + (e >= n_test) + (e >= n_eval); };
            double part[3] = {0.0, 0.0, 0.0};
#pragma unroll
            for (int b = 0; b < EPI_COLS / 8; ++b) {
                const int g_lo = group_of(e0 + b * 8), g_hi = group_of(e0 + b * 8 + 7);
                if (g_lo == g_hi) {
                    double t = 0.0;
#pragma unroll
                    for (int j = 0; j < 8; ++j) t += racc[b * 8 + j];
                    if (g_lo == 0) part[0] += t;
                    else if (g_lo == 1) part[1] += t;
                    else if (g_lo == 2) part[2] += t;
                } else {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int g = group_of(e0 + b * 8 + j);
                        const double v = racc[b * 8 + j];
                        part[0] += g == 0 ? v : 0.0;
                        part[1] += g == 1 ? v : 0.0;
                        part[2] += g == 2 ? v : 0.0;
                    }
                }
            }
#pragma unroll
            for (int g = 0; g < 3; ++g) red[(cg * 3 + g) * TM + f] = part[g];
            epi_barrier();
            double mu[3], rs[3], sd[3];
#pragma unroll
            for (int g = 0; g < 3; ++g) {
                const int n_g = hi[g] - lo[g];
                mu[g] = n_g > 0 ? ((red[g * TM + f] + red[(3 + g) * TM + f]) + red[(6 + g) * TM + f]) / n_g : 0.0;
                part[g] = 0.0;
            }
            epi_barrier();
#pragma unroll
            for (int b = 0; b < EPI_COLS / 8; ++b) {
                const int g_lo = group_of(e0 + b * 8), g_hi = group_of(e0 + b * 8 + 7);
                if (g_lo == g_hi) {
                    const double m = g_lo == 0 ? mu[0] : g_lo == 1 ? mu[1] : mu[2];
                    double t2 = 0.0;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const double t = racc[b * 8 + j] - m;
                        t2 = fma(t, t, t2);
                    }
                    if (g_lo == 0) part[0] += t2;
                    else if (g_lo == 1) part[1] += t2;
                    else if (g_lo == 2) part[2] += t2;
                } else {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int g = group_of(e0 + b * 8 + j);
                        const double t = racc[b * 8 + j] - (g == 0 ? mu[0] : g == 1 ? mu[1] : mu[2]);
                        const double t2 = t * t;
                        part[0] += g == 0 ? t2 : 0.0;
                        part[1] += g == 1 ? t2 : 0.0;
                        part[2] += g == 2 ? t2 : 0.0;
                    }
                }
            }
#pragma unroll
            for (int g = 0; g < 3; ++g) red[(cg * 3 + g) * TM + f] = part[g];
            epi_barrier();
#pragma unroll
            for (int g = 0; g < 3; ++g) {
                const int n_g = hi[g] - lo[g];
                sd[g] = n_g > 0 ? sqrt(((red[g * TM + f] + red[(3 + g) * TM + f]) + red[(6 + g) * TM + f]) / n_g) : 0.0;
                rs[g] = sd[g] == 0.0 ? 1.0 : 1.0 / sd[g];
            }
#pragma unroll
            for (int b = 0; b < EPI_COLS / 8; ++b) {
                const int eb = e0 + b * 8;
                const int g_lo = group_of(eb), g_hi = group_of(eb + 7);
                double* o = out + (long)eb * prm.Dp;
                if (g_lo == g_hi && g_lo < 2) {
                    const double m = g_lo == 0 ? mu[0] : mu[1], r = g_lo == 0 ? rs[0] : rs[1];
#pragma unroll
                    for (int j = 0; j < 8; ++j) o[(long)j * prm.Dp] = (racc[b * 8 + j] - m) * r;
                } else {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int g = group_of(eb + j);
                        const double raw = racc[b * 8 + j];
                        const double z = (raw - (g == 0 ? mu[0] : g == 1 ? mu[1] : mu[2])) * (g == 0 ? rs[0] : g == 1 ? rs[1] : rs[2]);
                        o[(long)j * prm.Dp] = g < 3 ? z : 0.0;                                   // rows n_eval .. 143 are written as zero
                        if (g == 2) out[(long)(EVAL_PAD + eb + j - n_test) * prm.Dp] = raw;       // *_no_std scores
                    }
                }
            }
            if (cg == 0) {
                for (int r = EVAL_PAD + prm.n_im; r < YHAT_ROWS; ++r) out[(long)r * prm.Dp] = 0.0;
                // what the score kernel needs for its tolerances: sum over the features of (sX / sd_g)^2 and of sX^2
                double a[4];
#pragma unroll
                for (int g = 0; g < 3; ++g) {
                    const double t = sd[g] > 0.0 ? sx / sd[g] : 0.0;
                    a[g] = t * t;
                }
                a[3] = sx * sx;
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    double t = a[g];
                    for (int o = 16; o; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
                    if (lane == 0) atomicAdd(prm.amp2 + (long)task_id * 4 + g, t);
                }
            }
            epi_barrier();                                                   // `red` is free for the next item
        }
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS));
}

namespace {
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled() {
    static EncodeTiledFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}
}  // namespace

// 2-D uint8 tensor map: `cols` contiguous bytes per row, `rows` rows of pitch `pitch` bytes, box box_cols x box_rows
cudaError_t make_tensor_map_u8(void* out_map, const void* base, uint64_t cols, uint64_t rows, uint64_t pitch, uint32_t box_cols,
                               uint32_t box_rows, int swizzle_bytes) {
    EncodeTiledFn fn = encode_tiled();
    if (!fn) return cudaErrorNotSupported;
    const cuuint64_t dims[2] = {cols, rows};
    const cuuint64_t strides[1] = {pitch};
    const cuuint32_t box[2] = {box_cols, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUtensorMapSwizzle sw = swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                                  : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                                  : swizzle_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B
                                                        : CU_TENSOR_MAP_SWIZZLE_NONE;
    const CUresult r = fn(reinterpret_cast<CUtensorMap*>(out_map), CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), dims,
                          strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? cudaSuccess : cudaErrorInvalidValue;
}

cudaError_t launch_predict_i8(const PredictI8Params& p, const int8_t* xplanes, const int8_t* zplanes, int n_sms, cudaStream_t s) {
    if (p.n_tasks <= 0) return cudaSuccess;
    if (p.Dp % TM != 0 || p.kp_max % KS != 0 || p.kp_max > MAX_KP || p.n_im > YHAT_ROWS - EVAL_PAD) return cudaErrorInvalidValue;
    if ((long)p.V * S > 0x7fffffffL || (long)p.n_tasks * S * TN > 0x7fffffffL) return cudaErrorInvalidValue;
    alignas(64) CUtensorMap tmA, tmB;
    cudaError_t e = make_tensor_map_u8(&tmA, xplanes, (uint64_t)p.Dp, (uint64_t)p.V * S, (uint64_t)p.Dp, TM, 1, 128);
    if (e != cudaSuccess) return e;
    e = make_tensor_map_u8(&tmB, zplanes, (uint64_t)p.kp_max, (uint64_t)p.n_tasks * S * TN, (uint64_t)p.kp_max, KS, TN, 64);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(predict_i8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return e;
    const long n_items = (long)p.n_tasks * (p.Dp / TM);
    const int grid = (int)(n_items < n_sms ? n_items : n_sms);
    predict_i8_kernel<<<grid, THREADS, SMEM_BYTES, s>>>(tmA, tmB, p);
    return cudaGetLastError();
}

}  // namespace mmd
