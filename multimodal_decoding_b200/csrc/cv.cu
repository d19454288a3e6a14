// Per-centre alpha grid x K-fold cross-validation (GridSearchCV semantics, ridge_regression_decoding.py:111-117, applied
// to every searchlight neighbourhood; BASELINE.json north_star (3)).
//
// A fold is a contiguous block of the training trials (KFold without shuffling).  Everything a fold needs is derived
// on the device from the panels mmd_upload already built:
//   Xfold_f  = the fold's columns of Xt (zero padded)     -> union Gram of the fold  G_f = Xfold_f Xfold_f^T  (gemm_gather.cu)
//   XtY_f    = Xfold_f Yfold_f^T                          -> training cross-covariance XtY - XtY_f
//   Xe_{f,c} = 144 of the fold's trials as evaluation rows (chunk c), cand_f = the fold's latents as float64
// and the per-centre kernels run unchanged on  A = G[loc,loc] - G_f[loc,loc] + alpha I  (factor_solve.cu reads both
// Grams), raw predictions (predict.cu), and the scorer below.
//
// cv_score_kernel: pairwise_accuracy(latents_val, predictions_val) of eval.py:107-123 restricted to the rows of one
// chunk: cosine distances of the chunk's <= 144 prediction rows to ALL n_val fold latents, strict '<' against the row's
// own latent; the count is row-separable, so chunks add into one counter per (centre, alpha, fold).
#include "common.cuh"
#include "kernels.h"

namespace mmd {
namespace {
constexpr int THREADS = 576;              // 18 warps: 2 row halves x 9 column groups (the layout of score.cu)
constexpr int MT = 5;
constexpr int KC = 32, LDS = KC + 4;           // 32-feature chunks, three stages (as score.cu)
constexpr int PASS_ROWS = 80;
constexpr int STAGES = 3;
constexpr int A_STAGE = PASS_ROWS * LDS, B_STAGE = EVAL_PAD * LDS;
constexpr int CHUNKS = (PASS_ROWS + EVAL_PAD) * (KC / 2);
}  // namespace

size_t cv_score_smem_bytes() { return (size_t)STAGES * (A_STAGE + B_STAGE) * sizeof(double) + (size_t)(2 * 160) * sizeof(double); }

__global__ void __launch_bounds__(THREADS, 1) cv_score_kernel(CvScoreParams prm) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int Dp = prm.Dp;
    double* As = reinterpret_cast<double*>(smem_raw);            // [STAGES][PASS_ROWS][LDS]
    double* Bs = As + STAGES * A_STAGE;                          // [STAGES][EVAL_PAD][LDS]
    double* nrm_sm = Bs + STAGES * B_STAGE;                      // [160]
    double* diag_sm = nrm_sm + 160;                              // [160]
    __shared__ unsigned long long cnt_sm;

    const int task_id = blockIdx.x;
    const CentreTask task = prm.tasks[task_id];
    const double* yhat = prm.yhat + (long)task_id * YHAT_ROWS * Dp;
    const int tid = threadIdx.x, lane = tid & 31, warp = (tid >> 5) % 9, wm = (tid >> 5) / 9;
    const int n_tiles = (prm.n_cand + EVAL_PAD - 1) / EVAL_PAD, own_tile = prm.row0 / EVAL_PAD;

    if (tid == 0) cnt_sm = 0ull;
    for (int i = tid; i < 160; i += THREADS) nrm_sm[i] = diag_sm[i] = 0.0;
    for (int i = tid; i < STAGES * A_STAGE; i += THREADS) As[i] = 0.0;      // rows past 144 are never loaded: keep them finite
    __syncthreads();

    const int fr = lane >> 2, fk = lane & 3;
    const int nk = Dp / KC;
    unsigned int cnt = 0;

    for (int pass = 0; pass < 2; ++pass) {
        const int row0 = pass * PASS_ROWS;
        const int a_rows = min(PASS_ROWS, EVAL_PAD - row0);
        for (int ti = 0; ti < n_tiles; ++ti) {
            // the tile holding the rows' own latents first: its diagonal is every other tile's reference value
            const int tile = ti == 0 ? own_tile : (ti <= own_tile ? ti - 1 : ti);
            const double* cand = prm.cand + (long)tile * EVAL_PAD * Dp;
            auto load_stage = [&](int stage, int kc) {
                const int d0 = kc * KC;
                double* as = As + stage * A_STAGE;
                double* bs = Bs + stage * B_STAGE;
                for (int c = tid; c < CHUNKS; c += THREADS) {
                    const int r = c / (KC / 2), cc = (c % (KC / 2)) * 2;
                    if (r < PASS_ROWS) {
                        if (r < a_rows) cp_async16(as + r * LDS + cc, yhat + (long)(row0 + r) * Dp + d0 + cc);
                    } else {
                        const int rb = r - PASS_ROWS;
                        cp_async16(bs + rb * LDS + cc, cand + (long)rb * Dp + d0 + cc);
                    }
                }
            };
            double acc[MT][2][2];
#pragma unroll
            for (int i = 0; i < MT; ++i)
#pragma unroll
                for (int j = 0; j < 2; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
            double sq = 0.0, sq2 = 0.0;

            for (int s = 0; s < STAGES - 1; ++s) {
                if (s < nk) load_stage(s, s);
                cp_async_commit();
            }
            for (int kc = 0; kc < nk; ++kc) {
                cp_async_wait<STAGES - 2>();
                __syncthreads();
                {
                    const int nxt = kc + STAGES - 1;
                    if (nxt < nk) load_stage(nxt % STAGES, nxt);
                    cp_async_commit();
                }
                const double* a_st = As + (kc % STAGES) * A_STAGE;
                if (ti == 0 && tid < 4 * PASS_ROWS) {      // four threads per row, two independent chains each
                    const double* x = a_st + (tid >> 2) * LDS + (tid & 3) * (KC / 4);
#pragma unroll
                    for (int e = 0; e < KC / 4; e += 2) {
                        const double2 v = *reinterpret_cast<const double2*>(x + e);
                        sq = fma(v.x, v.x, sq);
                        sq2 = fma(v.y, v.y, sq2);
                    }
                }
                const double* as = a_st + (wm * MT * 8 + fr) * LDS + fk;
                const double* bs = Bs + (kc % STAGES) * B_STAGE + (warp * 16 + fr) * LDS + fk;
#pragma unroll
                for (int k4 = 0; k4 < KC / 4; ++k4) {
                    const double b0 = bs[k4 * 4], b1 = bs[8 * LDS + k4 * 4];
#pragma unroll
                    for (int i = 0; i < MT; ++i) {
                        const double a = as[i * 8 * LDS + k4 * 4];
                        dmma884(acc[i][0][0], acc[i][0][1], a, b0);
                        dmma884(acc[i][1][0], acc[i][1][1], a, b1);
                    }
                }
            }
            cp_async_wait<0>();
            if (ti == 0 && tid < 4 * PASS_ROWS) {
                sq += sq2;
                sq += __shfl_xor_sync(0xffffffffu, sq, 1);
                sq += __shfl_xor_sync(0xffffffffu, sq, 2);
                if ((tid & 3) == 0) nrm_sm[row0 + (tid >> 2)] = sq;
            }
            __syncthreads();

#pragma unroll
            for (int i = 0; i < MT; ++i) {
                const int gm = row0 + (wm * MT + i) * 8 + fr;
                const double rn = sqrt(nrm_sm[gm]);
#pragma unroll
                for (int j = 0; j < 2; ++j)
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const int n = warp * 16 + j * 8 + 2 * fk + e;
                        const double dv = 1.0 - acc[i][j][e] / (rn * sqrt(prm.cand_sq[tile * EVAL_PAD + n]));
                        acc[i][j][e] = dv;
                        if (ti == 0 && n == gm) diag_sm[gm] = dv;
                    }
            }
            if (ti == 0) __syncthreads();
#pragma unroll
            for (int i = 0; i < MT; ++i) {
                const int gm = row0 + (wm * MT + i) * 8 + fr;
                if (gm >= prm.n_rows) continue;
                const double ds = diag_sm[gm];
#pragma unroll
                for (int j = 0; j < 2; ++j)
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const int n = tile * EVAL_PAD + warp * 16 + j * 8 + 2 * fk + e;
                        if (n < prm.n_cand) cnt += ds < acc[i][j][e];
                    }
            }
            __syncthreads();                                     // the ring is reused by the next tile
        }
    }
    for (int o = 16; o; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    if (lane == 0 && cnt) atomicAdd(&cnt_sm, (unsigned long long)cnt);
    __syncthreads();
    if (tid == 0) atomicAdd(prm.counts + (long)task.centre * prm.count_stride + prm.count_off, cnt_sm);
}

cudaError_t launch_cv_score(const CvScoreParams& p, cudaStream_t s) {
    if (p.n_tasks <= 0) return cudaSuccess;
    if (p.Dp % KC != 0 || p.row0 % EVAL_PAD != 0 || p.n_rows > EVAL_PAD || p.row0 + p.n_rows > p.n_cand) return cudaErrorInvalidValue;
    const size_t smem = cv_score_smem_bytes();
    cudaError_t e = cudaFuncSetAttribute(cv_score_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    cv_score_kernel<<<p.n_tasks, THREADS, smem, s>>>(p);
    return cudaGetLastError();
}

// dst[r][j] = src[r][col0 + j] for j < n_cols, 0 for n_cols <= j < ld_dst  (fold columns of the vertex-major panels)
__global__ void copy_cols_f16_kernel(const __half* __restrict__ src, long ld_src, int col0, int n_cols, __half* __restrict__ dst,
                                     int ld_dst, long n_rows) {
    const long total = n_rows * ld_dst;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const long r = i / ld_dst;
        const int j = (int)(i % ld_dst);
        dst[i] = j < n_cols ? src[r * ld_src + col0 + j] : __float2half(0.f);
    }
}
cudaError_t launch_copy_cols_f16(const __half* src, long ld_src, int col0, int n_cols, __half* dst, int ld_dst, long n_rows,
                                 cudaStream_t s) {
    copy_cols_f16_kernel<<<148 * 8, 256, 0, s>>>(src, ld_src, col0, n_cols, dst, ld_dst, n_rows);
    return cudaGetLastError();
}

// inout = a - inout
__global__ void rsub_kernel(const double* __restrict__ a, double* __restrict__ inout, long n) {
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) inout[i] = a[i] - inout[i];
}
cudaError_t launch_rsub(const double* a, double* inout, long n, cudaStream_t s) {
    rsub_kernel<<<148 * 8, 256, 0, s>>>(a, inout, n);
    return cudaGetLastError();
}

// cand[r][d] = (double) Yt[d][col0 + r] for r < n_rows (0 below, up to n_rows_pad); sq[r] = sum_d cand[r][d]^2.  One CTA per row.
__global__ void __launch_bounds__(256) fold_latents_kernel(const __half* __restrict__ Yt, long ldy, int col0, int n_rows, int Dp,
                                                           double* __restrict__ cand, double* __restrict__ sq) {
    const int r = blockIdx.x;
    double acc = 0.0;
    for (int d = threadIdx.x; d < Dp; d += blockDim.x) {
        const double v = r < n_rows ? h2d(Yt[(long)d * ldy + col0 + r]) : 0.0;
        cand[(long)r * Dp + d] = v;
        acc += v * v;
    }
    __shared__ double red[8];
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; ++w) t += red[w];
        sq[r] = t;
    }
}
cudaError_t launch_fold_latents(const __half* Yt, long ldy, int col0, int n_rows, int n_rows_pad, int Dp, double* cand, double* sq,
                                cudaStream_t s) {
    fold_latents_kernel<<<n_rows_pad, 256, 0, s>>>(Yt, ldy, col0, n_rows, Dp, cand, sq);
    return cudaGetLastError();
}

// flags[idx[i]] = 0
__global__ void clear_bytes_kernel(unsigned char* __restrict__ flags, const int* __restrict__ idx, int n) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) flags[idx[i]] = 0;
}
cudaError_t launch_clear_bytes(unsigned char* flags, const int* idx, int n, cudaStream_t s) {
    clear_bytes_kernel<<<148, 256, 0, s>>>(flags, idx, n);
    return cudaGetLastError();
}

}  // namespace mmd
