// Dense float64 building blocks of the whole-brain ridge grid search (ridge_regression_decoding.py:111-135):
// what sklearn's Ridge delegates to BLAS/LAPACK (gemm, potrf, potrs) as this project's own kernels on the fp64 tensor pipe.
//
//   gemm_nt_f64_kernel   C (+/-)= A (M x K) * B (N x K)^T, all row-major float64; 128 x 64 tiles, DMMA m8n8k4, 3-stage cp.async
//   blocked Cholesky     right-looking, 64-wide panels: chol_diag_kernel (factor_solve.cu: the 64 x 64 block, its inverse and the
//                        inverse's transpose), panel = panel * Linv^T and trailing -= panel * panel^T both through the GEMM
//   triangular solves    on TRANSPOSED right-hand sides (d x n, one right-hand side per row) so that every product is NT:
//                        forward  Bt[:, i] = Bt[:, i] Linv_ii^T ; Bt[:, >i] -= Bt[:, i] L[>i, i]^T
//                        backward Bt[:, i] = Bt[:, i] Linv_ii   ; Bt[:, <i] -= Bt[:, i] Lt[<i, i]^T      (Lt = L^T)
// All operands are zero padded to multiples of 64 (the padding of a matrix to be factored carries an identity).
#include "common.cuh"
#include "kernels.h"

namespace mmd {
namespace {
constexpr int BM = 128, BN = 64, KC = 16, LDS = KC + 4, STAGES = 3;
constexpr int A_STAGE = BM * LDS, B_STAGE = BN * LDS;      // doubles
}  // namespace

size_t dense_gemm_smem_bytes() { return (size_t)STAGES * (A_STAGE + B_STAGE) * sizeof(double); }

__global__ void __launch_bounds__(256, 2) gemm_nt_f64_kernel(DenseGemm p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* As = reinterpret_cast<double*>(smem_raw);
    double* Bs = As + STAGES * A_STAGE;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    if (m0 >= p.M || n0 >= p.N) return;
    if (p.lower_only && n0 > m0 + BM - 1) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp & 3, wn = warp >> 2;

    // 16-byte chunks: A tile 128 rows x 8 chunks (4 per thread), B tile 64 rows x 8 chunks (2 per thread)
    const int crow = tid >> 3, ccol = (tid & 7) * 2;
    const double* a_src[4];
    const double* b_src[2];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        int r = m0 + crow + 32 * i;
        r = r < p.M ? r : p.M - 1;
        a_src[i] = p.A + (long)r * p.lda + ccol;
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        int r = n0 + crow + 32 * i;
        r = r < p.N ? r : p.N - 1;
        b_src[i] = p.B + (long)r * p.ldb + ccol;
    }
    auto load_stage = [&](int stage, int kc) {
        double* as = As + stage * A_STAGE + crow * LDS + ccol;
        double* bs = Bs + stage * B_STAGE + crow * LDS + ccol;
#pragma unroll
        for (int i = 0; i < 4; ++i) cp_async16(as + 32 * i * LDS, a_src[i] + kc * KC);
#pragma unroll
        for (int i = 0; i < 2; ++i) cp_async16(bs + 32 * i * LDS, b_src[i] + kc * KC);
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
            const int nxt = kc + STAGES - 1;
            if (nxt < nk) load_stage(nxt % STAGES, nxt);
            cp_async_commit();
        }
        const double* as = As + (kc % STAGES) * A_STAGE + (wm * 32 + fr) * LDS + fk;
        const double* bs = Bs + (kc % STAGES) * B_STAGE + (wn * 32 + fr) * LDS + fk;
#pragma unroll
        for (int k4 = 0; k4 < KC / 4; ++k4) {
            double a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = as[i * 8 * LDS + k4 * 4];
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = bs[j * 8 * LDS + k4 * 4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
        }
    }
    cp_async_wait<0>();
    // in-place use (C aliases A for K <= one tile row set): every read of this CTA's operands has completed
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int row = m0 + wm * 32 + i * 8 + fr;
        if (row >= p.M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int col = n0 + wn * 32 + j * 8 + 2 * fk;
            double* c = p.C + (long)row * p.ldc + col;
            if (col + 1 < p.N) {
                double2 v = make_double2(acc[i][j][0], acc[i][j][1]);
                if (p.mode == 1) {
                    const double2 o = *reinterpret_cast<const double2*>(c);
                    v = make_double2(o.x - v.x, o.y - v.y);
                }
                *reinterpret_cast<double2*>(c) = v;
            } else if (col < p.N) {
                c[0] = p.mode == 1 ? c[0] - acc[i][j][0] : acc[i][j][0];
            }
        }
    }
}

cudaError_t launch_dense_gemm(const DenseGemm& p, cudaStream_t s) {
    if (p.M <= 0 || p.N <= 0) return cudaSuccess;
    if (p.K % KC != 0 || (p.lda & 1) || (p.ldb & 1) || (p.ldc & 1)) return cudaErrorInvalidValue;
    cudaError_t e = cudaFuncSetAttribute(gemm_nt_f64_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dense_gemm_smem_bytes());
    if (e != cudaSuccess) return e;
    dim3 grid((p.N + BN - 1) / BN, (p.M + BM - 1) / BM);
    gemm_nt_f64_kernel<<<grid, 256, dense_gemm_smem_bytes(), s>>>(p);
    return cudaGetLastError();
}

// out (rows_out x ld_out) = in[row_idx[r]][col_idx[c]] (an index of -1, or r/c beyond the lists, gives 0);
// diag_add is added where r == c < n_rows (alpha I), and the padding diagonal r == c >= n_rows gets pad_diag
__global__ void gather_block_kernel(const double* __restrict__ in, long ld_in, const int* __restrict__ row_idx, int n_rows,
                                    const int* __restrict__ col_idx, int n_cols, double* __restrict__ out, int rows_out, int ld_out,
                                    double diag_add, double pad_diag) {
    const long total = (long)rows_out * ld_out;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int r = (int)(i / ld_out), c = (int)(i % ld_out);
        double v = 0.0;
        if (r < n_rows && c < n_cols) {
            const int rr = row_idx ? row_idx[r] : r, cc = col_idx ? col_idx[c] : c;
            if (rr >= 0 && cc >= 0) v = in[(long)rr * ld_in + cc];
            if (r == c) v += diag_add;
        } else if (r == c) {
            v = pad_diag;
        }
        out[i] = v;
    }
}
cudaError_t launch_gather_block(const double* in, long ld_in, const int* row_idx, int n_rows, const int* col_idx, int n_cols,
                                double* out, int rows_out, int ld_out, double diag_add, double pad_diag, cudaStream_t s) {
    gather_block_kernel<<<148 * 8, 256, 0, s>>>(in, ld_in, row_idx, n_rows, col_idx, n_cols, out, rows_out, ld_out, diag_add, pad_diag);
    return cudaGetLastError();
}

// out (n_cols_in x ld_out, zero padded) = in (n_rows_in x n_cols_in, ld_in)^T ; 32 x 32 tiles
__global__ void transpose_f64_kernel(const double* __restrict__ in, long ld_in, int n_rows_in, int n_cols_in, double* __restrict__ out,
                                     long ld_out, int rows_out) {
    __shared__ double tile[32][33];
    const int r0 = blockIdx.y * 32, c0 = blockIdx.x * 32;       // tile of `in`
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int r = r0 + j, c = c0 + threadIdx.x;
        tile[j][threadIdx.x] = (r < n_rows_in && c < n_cols_in) ? in[(long)r * ld_in + c] : 0.0;
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int orow = c0 + j, ocol = r0 + threadIdx.x;       // out[c][r]
        if (orow < rows_out && ocol < ld_out) out[(long)orow * ld_out + ocol] = tile[threadIdx.x][j];
    }
}
// `out` has rows_out rows of ld_out doubles; everything outside the transposed block is written as zero
cudaError_t launch_transpose_f64(const double* in, long ld_in, int n_rows_in, int n_cols_in, double* out, long ld_out, int rows_out,
                                 cudaStream_t s) {
    dim3 grid((unsigned)((rows_out + 31) / 32), (unsigned)((ld_out + 31) / 32));
    transpose_f64_kernel<<<grid, dim3(32, 8), 0, s>>>(in, ld_in, n_rows_in, n_cols_in, out, ld_out, rows_out);
    return cudaGetLastError();
}

// float32 / float64 host layouts -> float64 panels: out (rows_out x ld_out) = in (n_rows x n_cols, ld_in), zero padded
template <typename T>
__global__ void widen_pad_kernel(const T* __restrict__ in, long ld_in, int n_rows, int n_cols, double* __restrict__ out, int rows_out,
                                 long ld_out) {
    const long total = (long)rows_out * ld_out;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const long r = i / ld_out, c = i % ld_out;
        out[i] = (r < n_rows && c < n_cols) ? (double)in[r * ld_in + c] : 0.0;
    }
}
cudaError_t launch_widen_pad(const void* in, int is_f32, long ld_in, int n_rows, int n_cols, double* out, int rows_out, long ld_out,
                             cudaStream_t s) {
    if (is_f32) widen_pad_kernel<float><<<148 * 8, 256, 0, s>>>(reinterpret_cast<const float*>(in), ld_in, n_rows, n_cols, out, rows_out, ld_out);
    else widen_pad_kernel<double><<<148 * 8, 256, 0, s>>>(reinterpret_cast<const double*>(in), ld_in, n_rows, n_cols, out, rows_out, ld_out);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------------------------
// blocked Cholesky and triangular solves (host drivers: sequences of launches on `s`)
// ------------------------------------------------------------------------------------------------------------------
// A: (np x np, ld np) symmetric positive definite, lower triangle read and overwritten by L (np a multiple of 64).
// linv / linvT: (np/64) x 64 x 64 inverses of the diagonal blocks (and their transposes).  flag: set on a non-positive pivot.
cudaError_t dense_potrf(double* A, int np, double* linv, double* linvT, int* flag, cudaStream_t s) {
    if (np % NB != 0) return cudaErrorInvalidValue;
    const int nb = np / NB;
    cudaError_t e;
    for (int j = 0; j < nb; ++j) {
        double* Ajj = A + (long)j * NB * np + j * NB;
        if ((e = launch_chol_diag(Ajj, np, linv + (long)j * NB * NB, linvT + (long)j * NB * NB, flag, s)) != cudaSuccess) return e;
        const int m = np - (j + 1) * NB;
        if (m <= 0) break;
        double* panel = A + (long)(j + 1) * NB * np + j * NB;
        DenseGemm t{};                       // panel = panel * Linv_jj^T (in place: a CTA owns its 128 rows of the 64 columns)
        t.A = panel; t.lda = np; t.B = linv + (long)j * NB * NB; t.ldb = NB; t.C = panel; t.ldc = np;
        t.M = m; t.N = NB; t.K = NB; t.mode = 0; t.lower_only = 0;
        if ((e = launch_dense_gemm(t, s)) != cudaSuccess) return e;
        DenseGemm u{};                       // trailing -= panel * panel^T, lower triangle
        u.A = panel; u.lda = np; u.B = panel; u.ldb = np; u.C = A + (long)(j + 1) * NB * np + (j + 1) * NB; u.ldc = np;
        u.M = m; u.N = m; u.K = NB; u.mode = 1; u.lower_only = 1;
        if ((e = launch_dense_gemm(u, s)) != cudaSuccess) return e;
    }
    return cudaSuccess;
}

// Bt (d x np, ld np): every row is one right-hand side; overwritten by the solution of (L L^T) x = b.
// L: the factor from dense_potrf; Lt: its transpose (np x np); both with ld np.
cudaError_t dense_potrs_t(const double* L, const double* Lt, const double* linv, const double* linvT, int np, double* Bt, int d,
                          cudaStream_t s) {
    const int nb = np / NB;
    cudaError_t e;
    for (int i = 0; i < nb; ++i) {            // forward: L z = b
        double* Bi = Bt + (long)i * NB;
        DenseGemm t{};
        t.A = Bi; t.lda = np; t.B = linv + (long)i * NB * NB; t.ldb = NB; t.C = Bi; t.ldc = np;
        t.M = d; t.N = NB; t.K = NB;
        if ((e = launch_dense_gemm(t, s)) != cudaSuccess) return e;
        const int m = np - (i + 1) * NB;
        if (m <= 0) break;
        DenseGemm u{};                        // Bt[:, >i] -= Bt[:, i] * L[>i, i]^T
        u.A = Bi; u.lda = np; u.B = L + (long)(i + 1) * NB * np + i * NB; u.ldb = np; u.C = Bi + NB; u.ldc = np;
        u.M = d; u.N = m; u.K = NB; u.mode = 1;
        if ((e = launch_dense_gemm(u, s)) != cudaSuccess) return e;
    }
    for (int i = nb - 1; i >= 0; --i) {       // backward: L^T x = z
        double* Bi = Bt + (long)i * NB;
        DenseGemm t{};
        t.A = Bi; t.lda = np; t.B = linvT + (long)i * NB * NB; t.ldb = NB; t.C = Bi; t.ldc = np;
        t.M = d; t.N = NB; t.K = NB;
        if ((e = launch_dense_gemm(t, s)) != cudaSuccess) return e;
        if (i == 0) break;
        DenseGemm u{};                        // Bt[:, <i] -= Bt[:, i] * Lt[<i, i]^T   (Lt[c][i*64+k] = L[i*64+k][c])
        u.A = Bi; u.lda = np; u.B = Lt + i * NB; u.ldb = np; u.C = Bt; u.ldc = np;
        u.M = d; u.N = i * NB; u.K = NB; u.mode = 1;
        if ((e = launch_dense_gemm(u, s)) != cudaSuccess) return e;
    }
    return cudaSuccess;
}

}  // namespace mmd
