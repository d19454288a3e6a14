// The eight pairwise-accuracy scores of one searchlight centre:
//   cosine distances between the (already z-scored, see predict.cu) predictions and the candidate latents
//   (eval.py:103-104) -> strict '<' comparison counts against each row's own candidate (eval.py:107-113)
//   -> count / (rows*cols - diag) for the metric table of SURVEY.md §3.2 (eval.py:126-181,
//   searchlight.py:81-88).
//
// Row layout of yhat / candidates (set up by the host and predict.cu): [captions | images | imagery | zero pad],
// rows 144..146 of yhat are the un-standardised imagery predictions (the *_no_std scores).
// One CTA per centre.  The (152 x 144) x Dp inner products run on the fp64 tensor pipe in two passes of 80
// rows; both operands are float64 in global memory and go through a 3-stage cp.async ring of 32-feature chunks (A/B of 16 x 4 vs 32 x 3: score stage -3 %), so the loop has
// the same shape as predict.cu's.  Row norms are accumulated from the staged tiles.
//
// Guard for the integer prediction GEMM (predict_i8.cu): that kernel's predictions carry a bounded truncation error
//   |yhat_i8 - yhat| <= eps0 sX[d] sZ[e],   eps0 = 7.2 k 2^-42
// (k terms, each short of the exact product by at most (5.03 + 2) 2^-42 scale units, plus the float64 recombination).
// Propagated through the per-feature z-scoring (|z' - z| <= 2 (|z| + 2) eps0 sX sZmax_g / sd, |z| <= (n_g-1)/sqrt(n_g))
// and the cosine distance (a row perturbed by eta moves the difference of two of its distances by at most
// 4 |eta| / |row|) this gives every prediction row a tolerance; a centre with any counted comparison whose two
// distances are closer than the row's tolerance is appended to `reroute_list` and redone by the float64 kernels.
// Every other centre's comparisons are the ones exact arithmetic on the same operands would produce.
#include "common.cuh"
#include "kernels.h"

namespace mmd {
namespace {
constexpr int THREADS = 576;              // 18 warps: 2 row halves x 9 column groups; one CTA per SM
constexpr int MT = 5;                     // m-tiles per warp and pass (40 rows)
#ifndef MMD_SCORE_KC
#define MMD_SCORE_KC 32
#endif
#ifndef MMD_SCORE_STAGES
#define MMD_SCORE_STAGES 3
#endif
constexpr int KC = MMD_SCORE_KC, LDS = KC + 4;
constexpr int PASS_ROWS = 80;             // 10 m-tiles per pass, two passes cover 152 rows (19 tiles + 1 idle)
constexpr int RAW_ROW0 = EVAL_PAD;        // first un-standardised imagery row
constexpr int M_ROWS = YHAT_ROWS;
constexpr int STAGES = MMD_SCORE_STAGES;
constexpr int A_STAGE = PASS_ROWS * LDS, B_STAGE = EVAL_PAD * LDS;   // doubles per stage
constexpr int CHUNKS = (PASS_ROWS + EVAL_PAD) * (KC / 2);             // 16-byte copies per stage
static_assert(M_ROWS <= 2 * PASS_ROWS && 4 * PASS_ROWS <= THREADS && KC % 8 == 0, "row-norm threads: four per row, warps 0-9");
}  // namespace

size_t score_smem_bytes(int) {
    return (size_t)STAGES * (A_STAGE + B_STAGE) * sizeof(double) + (size_t)(3 * 160) * sizeof(double);
}

__global__ void __launch_bounds__(THREADS, 1) score_kernel(ScoreParams prm) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int Dp = prm.Dp;
    double* As = reinterpret_cast<double*>(smem_raw);            // [STAGES][PASS_ROWS][LDS]
    double* Bs = As + STAGES * A_STAGE;                          // [STAGES][EVAL_PAD][LDS]
    double* nrm_sm = Bs + STAGES * B_STAGE;                      // [160]
    double* dsame_sm = nrm_sm + 160;                             // [160]
    double* dcross_sm = dsame_sm + 160;                          // [160]
    __shared__ int cnt_sm[8];
    __shared__ double tolc_sm[160];                              // guard: 4 |eta_row| (0 without the guard)
    __shared__ int risky_sm;
    __shared__ unsigned long long margin_sm;                     // bits of the smallest |d_own - d_other| over counted comparisons

    int task_id = blockIdx.x;
    if (prm.task_list) {
        if (task_id >= *prm.n_list) return;
        task_id = prm.task_list[task_id];
    }
    const CentreTask task = prm.tasks[task_id];
    const double* yhat = prm.yhat + (long)task_id * YHAT_ROWS * Dp;
    const int n_cap = prm.n_cap, n_img = prm.n_img, n_im = prm.n_im;
    const int n_test = n_cap + n_img, n_eval = n_test + n_im;
    const int tid = threadIdx.x, lane = tid & 31, warp = (tid >> 5) % 9, wm = (tid >> 5) / 9;

    if (tid < 8) cnt_sm[tid] = 0;
    if (tid == 0) {
        risky_sm = 0;
        margin_sm = 0x7ff0000000000000ull;
    }
    for (int i = tid; i < 160; i += THREADS) {
        nrm_sm[i] = 0.0;
        double tc = 0.0;
        if (prm.amp2) {
            const double eps0 = prm.eps_unit * (double)task.k;
            const double* a2 = prm.amp2 + (long)task_id * 4;
            const double* sz = prm.sZ + (long)task_id * EVAL_PAD;
            const int g = i < n_eval ? (i >= n_cap) + (i >= n_test) : -1;
            if (g >= 0) {
                const int lo = g == 0 ? 0 : g == 1 ? n_cap : n_test, hi = g == 0 ? n_cap : g == 1 ? n_test : n_eval;
                double szmax = 0.0;
                for (int r = lo; r < hi; ++r) szmax = fmax(szmax, sz[r]);
                const double n_g = (double)(hi - lo);
                const double zmax = n_g > 1.0 ? (n_g - 1.0) / sqrt(n_g) : 0.0;
                tc = 4.0 * 2.0 * (zmax + 2.0) * eps0 * szmax * sqrt(a2[g]);
            } else if (i >= RAW_ROW0 && i < RAW_ROW0 + n_im) {
                tc = 4.0 * eps0 * sz[n_test + (i - RAW_ROW0)] * sqrt(a2[3]);
            }
        }
        tolc_sm[i] = tc;
    }
    // rows of the second pass beyond YHAT_ROWS are never loaded: keep them finite
    for (int i = tid; i < STAGES * A_STAGE; i += THREADS) As[i] = 0.0;
    __syncthreads();

    const int fr = lane >> 2, fk = lane & 3;
    int cnt[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) cnt[q] = 0;
    double min_margin = __longlong_as_double(0x7ff0000000000000ll);
    bool risky = false;
    const int min_ci = n_cap < n_img ? n_cap : n_img;
    const int nk = Dp / KC;

    for (int pass = 0; pass < 2; ++pass) {
        const int row0 = pass * PASS_ROWS;
        const int a_rows = min(PASS_ROWS, M_ROWS - row0);
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
                    cp_async16(bs + rb * LDS + cc, prm.cand + (long)rb * Dp + d0 + cc);
                }
            }
        };

        double acc[MT][2][2];
#pragma unroll
        for (int i = 0; i < MT; ++i)
#pragma unroll
            for (int j = 0; j < 2; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
        double sq = 0.0, sq2 = 0.0;                       // threads 0..319: a quarter of a row of the squared norm

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
            if (tid < 4 * PASS_ROWS) {                     // four threads per row, two independent chains each: the norm
                const double* x = a_st + (tid >> 2) * LDS + (tid & 3) * (KC / 4);    // work must not hold warps 0-9 back
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
        if (tid < 4 * PASS_ROWS) {                         // warps 0-9, the four quarters of a row sit in adjacent lanes
            sq += sq2;
            sq += __shfl_xor_sync(0xffffffffu, sq, 1);
            sq += __shfl_xor_sync(0xffffffffu, sq, 2);
            if ((tid & 3) == 0) nrm_sm[row0 + (tid >> 2)] = sq;
        }
        __syncthreads();

        // ---- distances; first publish each row's own ("same") and cross-modal diagonal entries ----
#pragma unroll
        for (int i = 0; i < MT; ++i) {
            const int gm = row0 + (wm * MT + i) * 8 + fr;
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
        for (int i = 0; i < MT; ++i) {
            const int gm = row0 + (wm * MT + i) * 8 + fr;
            if (gm >= M_ROWS) continue;
            const bool is_test = gm < n_test, is_std_im = gm >= n_test && gm < n_eval;
            const bool is_raw_im = gm >= RAW_ROW0 && gm < RAW_ROW0 + n_im;
            if (!(is_test || is_std_im || is_raw_im)) continue;
            const double ds = dsame_sm[gm];
            const bool row_cap = gm < n_cap;
            const int li = row_cap ? gm : gm - n_cap;
            const double dc = (is_test && li < min_ci) ? dcross_sm[gm] : 0.0;
            // tolerance of this row's distance differences (0 without the guard); a bound that is not small against
            // the row itself, or not finite, sends the centre to the float64 kernels as well
            const double tol = tolc_sm[gm] / sqrt(nrm_sm[gm]);
            if (tolc_sm[gm] != 0.0 && !(tol < 0.05)) risky = true;
            const int own = is_test ? gm : (is_std_im ? gm : n_test + (gm - RAW_ROW0));        // the row's own candidate
            const int cross = row_cap ? n_cap + li : li;
#pragma unroll
            for (int j = 0; j < 2; ++j)
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int n = warp * 16 + j * 8 + 2 * fk + e;
                    const double dv = acc[i][j][e];
                    if (is_test) {
                        if (n >= n_test) continue;
                        const bool col_cap = n < n_cap;
                        if (col_cap == row_cap) {
                            cnt[row_cap ? 0 : 1] += ds < dv;
                            if (n != own) {
                                const double m = fabs(ds - dv);
                                min_margin = fmin(min_margin, m);
                                risky |= m <= tol;
                            }
                        } else if (li < min_ci) {
                            cnt[row_cap ? 2 : 3] += dc < dv;
                            if (n != cross) {
                                const double m = fabs(dc - dv);
                                min_margin = fmin(min_margin, m);
                                risky |= m <= tol;
                            }
                        }
                    } else {
                        if (n >= n_eval) continue;
                        const int base = is_std_im ? 4 : 6;
                        if (n >= n_test) cnt[base] += ds < dv;
                        cnt[base + 1] += ds < dv;
                        if (n != own) {
                            const double m = fabs(ds - dv);
                            min_margin = fmin(min_margin, m);
                            risky |= m <= tol;
                        }
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
    if (__any_sync(0xffffffffu, risky) && lane == 0) risky_sm = 1;
    for (int o = 16; o; o >>= 1) min_margin = fmin(min_margin, __shfl_xor_sync(0xffffffffu, min_margin, o));
    if (lane == 0) atomicMin(&margin_sm, (unsigned long long)__double_as_longlong(min_margin));     // non-negative doubles order like integers
    __syncthreads();
    if (tid == 0) {
        if (prm.margin_out) prm.margin_out[task.centre] = __longlong_as_double((long long)margin_sm);
        if (prm.amp2 && risky_sm && prm.reroute_list) prm.reroute_list[atomicAdd(prm.reroute_count, 1)] = task_id;
    }
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
    if (p.n_cap + p.n_img + p.n_im > EVAL_PAD - 1 || p.n_im > YHAT_ROWS - EVAL_PAD || p.Dp % KC != 0)
        return cudaErrorInvalidValue;
    if (p.n_tasks <= 0) return cudaSuccess;
    score_kernel<<<p.n_tasks, THREADS, smem, s>>>(p);
    return cudaGetLastError();
}

}  // namespace mmd
