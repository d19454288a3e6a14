// Gathered NT GEMM on the fp64 tensor pipe:  C[i][j] = sum_n A[rowsA[i]][n] * B[rowsB[j]][n]
// with fp16 operands (exact in fp64) and float64 accumulation (DMMA m8n8k4).
//
// Two uses on the searchlight path (both replace per-centre float64 BLAS calls in sklearn's Ridge.fit,
// `_ridge.py` _solve_cholesky, reached from searchlight.py:56):
//   * XtY   : A = Xt (V x N_train), B = Yt (D x N_train), once per searchlight call.  The per-centre
//             X_nbh^T Y equals rows idx of this matrix, so it is never recomputed per centre.
//   * SYRK  : A = B = Xt restricted to the union U of a cluster's neighbourhoods; every centre of
//             the cluster then reads its k x k Gram as a sub-matrix of this one.
#include "common.cuh"
#include "kernels.h"

namespace mmd {

namespace {
constexpr int BM = 128, BN = 64, KC = 64, LDS_H = 72, STAGES = 3;   // LDS_H: smem row stride in halves (144 B)
constexpr int A_STAGE = BM * LDS_H, B_STAGE = BN * LDS_H;           // halves per stage
}  // namespace

size_t gemm_gather_smem_bytes() { return (size_t)STAGES * (A_STAGE + B_STAGE) * sizeof(__half); }

__global__ void __launch_bounds__(256, 2) gemm_nt_f16_f64_kernel(const GemmProblem* __restrict__ probs) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __half* As = reinterpret_cast<__half*>(smem_raw);
    __half* Bs = As + STAGES * A_STAGE;

    const GemmProblem p = probs[blockIdx.z];
    if (p.only_if && *p.only_if == 0) return;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    if (m0 >= p.M || n0 >= p.N) return;
    if (p.lower_only && n0 > m0 + BM - 1) return;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp & 3, wn = warp >> 2;

    // each thread copies 4 A chunks and 2 B chunks (16 B each) per stage
    const int crow = tid >> 3, ccol = (tid & 7) * 8;
    const __half* a_src[4];
    const __half* b_src[2];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        int r = m0 + crow + 32 * i;
        r = r < p.M ? r : p.M - 1;
        long row = p.rowsA ? p.rowsA[r] : r;
        a_src[i] = p.A + row * p.lda + ccol;
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        int r = n0 + crow + 32 * i;
        r = r < p.N ? r : p.N - 1;
        long row = p.rowsB ? p.rowsB[r] : r;
        b_src[i] = p.B + row * p.ldb + ccol;
    }
    auto load_stage = [&](int stage, int kc) {
        __half* as = As + stage * A_STAGE + crow * LDS_H + ccol;
        __half* bs = Bs + stage * B_STAGE + crow * LDS_H + ccol;
#pragma unroll
        for (int i = 0; i < 4; ++i) cp_async16(as + 32 * i * LDS_H, a_src[i] + kc * KC);
#pragma unroll
        for (int i = 0; i < 2; ++i) cp_async16(bs + 32 * i * LDS_H, b_src[i] + kc * KC);
    };

    double acc[4][4][2];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    const int nk = p.K / KC;
    for (int s = 0; s < STAGES - 1; ++s) {
        if (s < nk) load_stage(s, s);
        cp_async_commit();
    }
    const int fr = lane >> 2, fk = lane & 3;
    for (int kc = 0; kc < nk; ++kc) {
        cp_async_wait<STAGES - 2>();
        __syncthreads();
        {
            int nxt = kc + STAGES - 1;
            if (nxt < nk) load_stage(nxt % STAGES, nxt);
            cp_async_commit();
        }
        const __half* as = As + (kc % STAGES) * A_STAGE + (wm * 32 + fr) * LDS_H + fk;
        const __half* bs = Bs + (kc % STAGES) * B_STAGE + (wn * 32 + fr) * LDS_H + fk;
#pragma unroll 4
        for (int k4 = 0; k4 < KC / 4; ++k4) {
            double a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = h2d(as[i * 8 * LDS_H + k4 * 4]);
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = h2d(bs[j * 8 * LDS_H + k4 * 4]);
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
        }
    }
    cp_async_wait<0>();

#pragma unroll
    for (int i = 0; i < 4; ++i) {
        int row = m0 + wm * 32 + i * 8 + fr;
        if (row >= p.M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int col = n0 + wn * 32 + j * 8 + 2 * fk;
            const long crow_out = p.scatter_c && p.rowsA ? p.rowsA[row] : row;
            double* c = p.C + crow_out * p.ldc + col;
            if (col + 1 < p.N) {
                *reinterpret_cast<double2*>(c) = make_double2(acc[i][j][0], acc[i][j][1]);
            } else if (col < p.N) {
                c[0] = acc[i][j][0];
            }
        }
    }
}

cudaError_t launch_gemm_gather(const GemmProblem* probs_dev, int n_problems, int max_m, int max_n, cudaStream_t s) {
    // function attributes are per device: set on every launch (several engines on different GPUs may share the process)
    cudaError_t e = cudaFuncSetAttribute(gemm_nt_f16_f64_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)gemm_gather_smem_bytes());
    if (e != cudaSuccess) return e;
    dim3 grid((max_n + BN - 1) / BN, (max_m + BM - 1) / BM, n_problems);
    gemm_nt_f16_f64_kernel<<<grid, 256, gemm_gather_smem_bytes(), s>>>(probs_dev);
    return cudaGetLastError();
}

}  // namespace mmd
