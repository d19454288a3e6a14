// The eight pairwise-accuracy scores of one searchlight centre, fused:
//   per-feature z-scoring of each prediction group (eval.py:118, StandardScaler semantics, ddof = 0,
//   zero-variance features keep scale 1) -> cosine distances to the candidate latents (eval.py:103-104)
//   -> strict '<' comparison counts against each row's own candidate (eval.py:107-113)
//   -> count / (rows*cols - diag) for the metric table of SURVEY.md §3.2 (eval.py:126-181,
//   searchlight.py:81-88).
//
// Row layout of yhat / candidates (set up by the host): [captions | images | imagery | zero pad].
// One CTA per centre.  The (152 x 144) x Dp inner products run on the fp64 tensor pipe; rows
// 144..146 of the product are the un-standardised imagery predictions (the *_no_std scores).
#include "common.cuh"
#include "kernels.h"

namespace mmd {
namespace {
constexpr int THREADS = 288;
constexpr int KC = 16, LDS = KC + 4;
constexpr int PASS_ROWS = 80;             // 10 m-tiles per pass, two passes cover 152 rows (19 tiles + 1 idle)
constexpr int RAW_ROW0 = EVAL_PAD;        // first un-standardised imagery row
constexpr int M_ROWS = EVAL_PAD + 8;
}  // namespace

size_t score_smem_bytes(int Dp) {
    return (size_t)6 * Dp * sizeof(double)                                   // mean / rstd for 3 groups
           + (size_t)2 * (PASS_ROWS + EVAL_PAD) * LDS * sizeof(double)       // double-buffered A and B chunks
           + (size_t)(3 * 160) * sizeof(double);                             // row norms, same/cross diagonals
}

__global__ void __launch_bounds__(THREADS, 1) score_kernel(ScoreParams prm) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int Dp = prm.Dp;
    double* mean_sm = reinterpret_cast<double*>(smem_raw);       // [3][Dp]
    double* rstd_sm = mean_sm + 3 * Dp;                          // [3][Dp]
    double* As = rstd_sm + 3 * Dp;                               // [2][PASS_ROWS][LDS]
    double* Bs = As + 2 * PASS_ROWS * LDS;                       // [2][EVAL_PAD][LDS]
    double* nrm_sm = Bs + 2 * EVAL_PAD * LDS;                    // [160]
    double* dsame_sm = nrm_sm + 160;                             // [160]
    double* dcross_sm = dsame_sm + 160;                          // [160]
    __shared__ int cnt_sm[8];

    const int task_id = blockIdx.x;
    const CentreTask task = prm.tasks[task_id];
    const double* yhat = prm.yhat + (long)task_id * EVAL_PAD * Dp;
    const int n_cap = prm.n_cap, n_img = prm.n_img, n_im = prm.n_im;
    const int n_test = n_cap + n_img, n_eval = n_test + n_im;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    if (tid < 8) cnt_sm[tid] = 0;
    for (int i = tid; i < 160; i += THREADS) nrm_sm[i] = 0.0;

    // ---- per-feature statistics of the three prediction groups ----
    for (int d = tid; d < Dp; d += THREADS) {
        const int lo[3] = {0, n_cap, n_test}, hi[3] = {n_cap, n_test, n_eval};
#pragma unroll
        for (int g = 0; g < 3; ++g) {
            const int n = hi[g] - lo[g];
            double s = 0.0;
            for (int r = lo[g]; r < hi[g]; ++r) s += yhat[(long)r * Dp + d];
            const double mu = n > 0 ? s / n : 0.0;
            double v = 0.0;
            for (int r = lo[g]; r < hi[g]; ++r) {
                const double t = yhat[(long)r * Dp + d] - mu;
                v += t * t;
            }
            const double sd = n > 0 ? sqrt(v / n) : 0.0;
            mean_sm[g * Dp + d] = mu;
            rstd_sm[g * Dp + d] = sd == 0.0 ? 1.0 : 1.0 / sd;
        }
    }
    __syncthreads();

    const int fr = lane >> 2, fk = lane & 3;
    int cnt[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) cnt[q] = 0;
    const int min_ci = n_cap < n_img ? n_cap : n_img;

    for (int pass = 0; pass < 2; ++pass) {
        const int row0 = pass * PASS_ROWS;
        // staging map: 640 double2 of A per chunk (3 slots per thread), one 16-byte load of B per thread
        int a_row[3], a_src[3], a_grp[3], a_col[3];
        double a_sq[3] = {0.0, 0.0, 0.0};
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            const int idx = tid + i * THREADS;
            const int r = idx / 8;
            a_col[i] = (idx % 8) * 2;
            a_row[i] = idx < PASS_ROWS * 8 ? r : -1;
            const int gm = row0 + r;
            int src = -1, grp = -1;
            if (gm < n_eval) {
                src = gm;
                grp = gm < n_cap ? 0 : (gm < n_test ? 1 : 2);
            } else if (gm >= RAW_ROW0 && gm < RAW_ROW0 + n_im) {
                src = n_test + (gm - RAW_ROW0);
            }
            a_src[i] = src;
            a_grp[i] = grp;
        }
        const int b_row = tid / 2, b_col = (tid % 2) * 8;

        double2 a_reg[3];
        uint4 b_reg;
        auto fetch = [&](int kc) {
            const int d0 = kc * KC;
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                double2 v = make_double2(0.0, 0.0);
                if (a_row[i] >= 0 && a_src[i] >= 0) {
                    v = *reinterpret_cast<const double2*>(yhat + (long)a_src[i] * Dp + d0 + a_col[i]);
                    if (a_grp[i] >= 0) {
                        const double* mu = mean_sm + a_grp[i] * Dp + d0 + a_col[i];
                        const double* rs = rstd_sm + a_grp[i] * Dp + d0 + a_col[i];
                        v.x = (v.x - mu[0]) * rs[0];
                        v.y = (v.y - mu[1]) * rs[1];
                    }
                }
                a_reg[i] = v;
            }
            b_reg = *reinterpret_cast<const uint4*>(prm.cand + (long)b_row * Dp + d0 + b_col);
        };
        auto stash = [&](int buf) {
            double* as = As + buf * PASS_ROWS * LDS;
            double* bs = Bs + buf * EVAL_PAD * LDS;
#pragma unroll
            for (int i = 0; i < 3; ++i)
                if (a_row[i] >= 0) {
                    *reinterpret_cast<double2*>(as + a_row[i] * LDS + a_col[i]) = a_reg[i];
                    a_sq[i] += a_reg[i].x * a_reg[i].x + a_reg[i].y * a_reg[i].y;
                }
            const __half* h = reinterpret_cast<const __half*>(&b_reg);
            double* dst = bs + b_row * LDS + b_col;
#pragma unroll
            for (int e = 0; e < 8; e += 2)
                *reinterpret_cast<double2*>(dst + e) = make_double2(h2d(h[e]), h2d(h[e + 1]));
        };

        double acc[10][2][2];
#pragma unroll
        for (int i = 0; i < 10; ++i)
#pragma unroll
            for (int j = 0; j < 2; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

        const int nk = Dp / KC;
        fetch(0);
        stash(0);
        __syncthreads();
        for (int kc = 0; kc < nk; ++kc) {
            if (kc + 1 < nk) fetch(kc + 1);
            const double* as = As + (kc & 1) * PASS_ROWS * LDS + fr * LDS + fk;
            const double* bs = Bs + (kc & 1) * EVAL_PAD * LDS + (warp * 16 + fr) * LDS + fk;
#pragma unroll
            for (int k4 = 0; k4 < KC / 4; ++k4) {
                const double b0 = bs[k4 * 4], b1 = bs[8 * LDS + k4 * 4];
#pragma unroll
                for (int i = 0; i < 10; ++i) {
                    const double a = as[i * 8 * LDS + k4 * 4];
                    dmma884(acc[i][0][0], acc[i][0][1], a, b0);
                    dmma884(acc[i][1][0], acc[i][1][1], a, b1);
                }
            }
            if (kc + 1 < nk) stash((kc + 1) & 1);
            __syncthreads();
        }
#pragma unroll
        for (int i = 0; i < 3; ++i)
            if (a_row[i] >= 0) atomicAdd(&nrm_sm[row0 + a_row[i]], a_sq[i]);
        __syncthreads();

        // ---- distances; first publish each row's own ("same") and cross-modal diagonal entries ----
#pragma unroll
        for (int i = 0; i < 10; ++i) {
            const int gm = row0 + i * 8 + fr;
            const double rn = sqrt(nrm_sm[gm < 160 ? gm : 0]);
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int n = warp * 16 + j * 8 + 2 * fk + e;
                    const double dv = 1.0 - acc[i][j][e] / (rn * sqrt(prm.cand_sq[n]));
                    acc[i][j][e] = dv;
                    if (prm.dist_out && gm < NULL_DIST_ROWS)
                        prm.dist_out[((long)task_id * NULL_DIST_ROWS + gm) * EVAL_PAD + n] = dv;
                    if (gm < n_test) {
                        if (n == gm) dsame_sm[gm] = dv;
                        const int li = gm < n_cap ? gm : gm - n_cap;
                        const int cross = gm < n_cap ? n_cap + li : li;
                        if (li < min_ci && n == cross) dcross_sm[gm] = dv;
                    } else if (gm < n_eval) {
                        if (n == gm) dsame_sm[gm] = dv;
                    } else if (gm >= RAW_ROW0 && gm < RAW_ROW0 + n_im) {
                        if (n == n_test + (gm - RAW_ROW0)) dsame_sm[gm] = dv;
                    }
                }
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 10; ++i) {
            const int gm = row0 + i * 8 + fr;
            if (gm >= M_ROWS) continue;
            const bool is_test = gm < n_test, is_std_im = gm >= n_test && gm < n_eval;
            const bool is_raw_im = gm >= RAW_ROW0 && gm < RAW_ROW0 + n_im;
            if (!(is_test || is_std_im || is_raw_im)) continue;
            const double ds = dsame_sm[gm];
            const bool row_cap = gm < n_cap;
            const int li = row_cap ? gm : gm - n_cap;
            const double dc = (is_test && li < min_ci) ? dcross_sm[gm] : 0.0;
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int n = warp * 16 + j * 8 + 2 * fk + e;
                    const double dv = acc[i][j][e];
                    if (is_test) {
                        if (n >= n_test) continue;
                        const bool col_cap = n < n_cap;
                        if (col_cap == row_cap) cnt[row_cap ? 0 : 1] += ds < dv;
                        else if (li < min_ci) cnt[row_cap ? 2 : 3] += dc < dv;
                    } else {
                        if (n >= n_eval) continue;
                        const int base = is_std_im ? 4 : 6;
                        if (n >= n_test) cnt[base] += ds < dv;
                        cnt[base + 1] += ds < dv;
                    }
                }
        }
        __syncthreads();
    }

#pragma unroll
    for (int q = 0; q < 8; ++q) {
        int c = cnt[q];
        for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
        if (lane == 0 && c) atomicAdd(&cnt_sm[q], c);
    }
    __syncthreads();
    if (tid < 8) {
        const long den_cc = (long)n_cap * n_cap - n_cap, den_ii = (long)n_img * n_img - n_img;
        const long den_x = (long)n_cap * n_img - min_ci;
        const long den_im = (long)n_im * n_im - n_im, den_w = (long)n_im * n_eval - n_im;
        const long den[8] = {den_cc, den_ii, den_x, den_x, den_im, den_w, den_im, den_w};
        prm.scores[(long)task.centre * 8 + tid] = den[tid] > 0 ? (double)cnt_sm[tid] / (double)den[tid] : 0.0;
    }
}

cudaError_t launch_score(const ScoreParams& p, cudaStream_t s) {
    const size_t smem = score_smem_bytes(p.Dp);
    cudaError_t e = cudaFuncSetAttribute(score_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    if (p.n_cap + p.n_img + p.n_im > EVAL_PAD - 1 || p.n_im > 8) return cudaErrorInvalidValue;
    score_kernel<<<p.n_tasks, THREADS, smem, s>>>(p);
    return cudaGetLastError();
}

}  // namespace mmd
