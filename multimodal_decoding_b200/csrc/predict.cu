// Predictions for the evaluation trials of every centre:
//   Yhat(144 x Dp) = Zt(144 x kp) * XtY[nbr, :](kp x Dp)
// which equals sklearn's `estimator.predict(X_test)` / `predict(X_imagery)` (searchlight.py:58-59)
// because X_eval W = X_eval A^{-1} X_nbh^T Y and X_nbh^T Y is a row gather of the global X^T Y.
// float64 on the DMMA pipe; both operands staged with cp.async (rows of XtY are gathered by
// neighbour id, 512 contiguous bytes each).
#include "common.cuh"
#include "kernels.h"

namespace mmd {
namespace {
constexpr int THREADS = 256;
constexpr int BN = 64;               // feature columns per CTA
constexpr int KC = 16;               // neighbours per pipeline stage
constexpr int LDA = KC + 4;          // 20 doubles: conflict-free A fragments
constexpr int LDB = BN + 4;          // 68 doubles: conflict-free B fragments
constexpr int STAGES = 3;
constexpr int A_STAGE = EVAL_PAD * LDA, B_STAGE = KC * LDB;   // doubles
constexpr int LDT = BN + 2;          // epilogue tile stride (even: rows stay 16-byte aligned for the double2 stores)
static_assert(STAGES * (A_STAGE + B_STAGE) >= EVAL_PAD * LDT + 4 * BN, "the epilogue tile aliases the pipeline buffers");
static_assert(THREADS == 4 * BN, "epilogue thread map: 4 row quarters x 64 columns");
}  // namespace

size_t predict_smem_bytes() { return (size_t)STAGES * (A_STAGE + B_STAGE) * sizeof(double) + 2048 * sizeof(int); }

__global__ void __launch_bounds__(THREADS, 2) predict_kernel(PredictParams prm) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* As = reinterpret_cast<double*>(smem_raw);
    double* Bs = As + STAGES * A_STAGE;
    int* nbr_sm = reinterpret_cast<int*>(Bs + STAGES * B_STAGE);

    int task_id = blockIdx.y;
    if (prm.task_list) {
        if (task_id >= *prm.n_list) return;
        task_id = prm.task_list[task_id];
    }
    const CentreTask task = prm.tasks[task_id];
    const int k = task.k, kp = task.kp, Dp = prm.Dp;
    const int d0 = blockIdx.x * BN;
    const double* Zt = prm.work + (long)task_id * (prm.kp_max + EVAL_PAD) * prm.kp_max + (long)kp * kp;   // 144 x kp
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp & 1, wn = warp >> 1;           // 2 x 4 warps: 72 rows x 16 columns each

    for (int i = tid; i < kp; i += THREADS) nbr_sm[i] = task.nbr[i < k ? i : 0];
    __syncthreads();

    auto load_stage = [&](int stage, int kc) {
        const int k0 = kc * KC;
        double* as = As + stage * A_STAGE;
        double* bs = Bs + stage * B_STAGE;
        // A: 144 rows x 16 doubles = 1152 chunks of 16 B
        for (int c = tid; c < EVAL_PAD * (KC / 2); c += THREADS) {
            const int r = c / (KC / 2), cc = (c % (KC / 2)) * 2;
            cp_async16(as + r * LDA + cc, Zt + (long)r * kp + k0 + cc);
        }
        // B: 16 gathered rows x 64 doubles = 512 chunks
        for (int c = tid; c < KC * (BN / 2); c += THREADS) {
            const int r = c / (BN / 2), cc = (c % (BN / 2)) * 2;
            cp_async16(bs + r * LDB + cc, prm.XtY + (long)nbr_sm[k0 + r] * Dp + d0 + cc);
        }
    };

    double acc[9][2][2];
#pragma unroll
    for (int i = 0; i < 9; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    const int nk = (k + KC - 1) / KC;                  // columns k.. of Zt are zero (padding): skip their chunks
    for (int s = 0; s < STAGES - 1; ++s) {
        if (s < nk) load_stage(s, s);
        cp_async_commit();
    }
    const int fr = lane >> 2, fk = lane & 3;
    for (int kc = 0; kc < nk; ++kc) {
        cp_async_wait<STAGES - 2>();
        __syncthreads();
        {
            const int nxt = kc + STAGES - 1;
            if (nxt < nk) load_stage(nxt % STAGES, nxt);
            cp_async_commit();
        }
        const double* as = As + (kc % STAGES) * A_STAGE + (wm * 72 + fr) * LDA + fk;
        const double* bs = Bs + (kc % STAGES) * B_STAGE + fk * LDB + wn * 16 + fr;
#pragma unroll
        for (int k4 = 0; k4 < KC / 4; ++k4) {
            double b0 = bs[k4 * 4 * LDB], b1 = bs[k4 * 4 * LDB + 8];
#pragma unroll
            for (int i = 0; i < 9; ++i) {
                const double a = as[i * 8 * LDA + k4 * 4];
                dmma884(acc[i][0][0], acc[i][0][1], a, b0);
                dmma884(acc[i][1][0], acc[i][1][1], a, b1);
            }
        }
    }
    cp_async_wait<0>();

    double* out = prm.yhat + (long)task_id * YHAT_ROWS * Dp;
    const int n_cap = prm.n_cap, n_test = prm.n_cap + prm.n_img, n_eval = n_test + prm.n_im;
    if (prm.standardize) {
        // Per-feature z-scoring of the three prediction groups before the tile leaves the SM: the CTA holds all 144
        // rows of its 64 columns.  The tile is parked in the (now idle) pipeline buffers so the accumulator registers
        // die here; then two exact passes per group (mean, centred sum of squares) like StandardScaler.
        __syncthreads();
        double* T = As;                                    // [EVAL_PAD][LDT]
        double* red = T + EVAL_PAD * LDT;                  // [4][BN]
#pragma unroll
        for (int i = 0; i < 9; ++i)
#pragma unroll
            for (int j = 0; j < 2; ++j)
                *reinterpret_cast<double2*>(T + (wm * 72 + i * 8 + fr) * LDT + wn * 16 + j * 8 + 2 * fk) =
                    make_double2(acc[i][j][0], acc[i][j][1]);
        __syncthreads();
        const int c = tid & (BN - 1), q = tid >> 6;        // column of the tile, quarter of the rows
        const int lo[3] = {0, n_cap, n_test}, hi[3] = {n_cap, n_test, n_eval};
#pragma unroll
        for (int g = 0; g < 3; ++g) {
            const int n_g = hi[g] - lo[g];
            double s = 0.0;
            for (int r = lo[g] + q; r < hi[g]; r += 4) s += T[r * LDT + c];
            red[q * BN + c] = s;
            __syncthreads();
            const double mu = n_g > 0 ? (((red[c] + red[BN + c]) + red[2 * BN + c]) + red[3 * BN + c]) / n_g : 0.0;
            __syncthreads();
            double v = 0.0;
            for (int r = lo[g] + q; r < hi[g]; r += 4) {
                const double t = T[r * LDT + c] - mu;
                v += t * t;
            }
            red[q * BN + c] = v;
            __syncthreads();
            const double sd = n_g > 0 ? sqrt((((red[c] + red[BN + c]) + red[2 * BN + c]) + red[3 * BN + c]) / n_g) : 0.0;
            const double rs = sd == 0.0 ? 1.0 : 1.0 / sd;
            __syncthreads();
            for (int r = lo[g] + q; r < hi[g]; r += 4) {
                const double raw = T[r * LDT + c];
                if (g == 2) out[(long)(EVAL_PAD + r - n_test) * Dp + d0 + c] = raw;   // *_no_std scores use the raw rows
                out[(long)r * Dp + d0 + c] = (raw - mu) * rs;
            }
        }
        // zero rows: n_eval .. EVAL_PAD-1 and EVAL_PAD + n_im .. YHAT_ROWS-1
        for (int r = n_eval + q; r < YHAT_ROWS; r += 4)
            if (r < EVAL_PAD || r >= EVAL_PAD + prm.n_im) out[(long)r * Dp + d0 + c] = 0.0;
        return;
    }
#pragma unroll
    for (int i = 0; i < 9; ++i) {
        const int row = wm * 72 + i * 8 + fr;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int col = d0 + wn * 16 + j * 8 + 2 * fk;
            *reinterpret_cast<double2*>(out + (long)row * Dp + col) = make_double2(acc[i][j][0], acc[i][j][1]);
        }
    }
}

cudaError_t launch_predict(const PredictParams& p, cudaStream_t s) {
    // function attributes are per device: set on every launch (several engines on different GPUs may share the process)
    cudaError_t e = cudaFuncSetAttribute(predict_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)predict_smem_bytes());
    if (e != cudaSuccess) return e;
    if (p.n_tasks <= 0) return cudaSuccess;
    if (p.kp_max > 2048 || p.n_im > YHAT_ROWS - EVAL_PAD) return cudaErrorInvalidValue;
    dim3 grid(p.Dp / BN, p.n_tasks);
    predict_kernel<<<grid, THREADS, predict_smem_bytes(), s>>>(p);
    return cudaGetLastError();
}

}  // namespace mmd
