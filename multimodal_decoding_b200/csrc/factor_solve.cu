// Per-centre ridge solve in float64 (the part sklearn delegates to LAPACK dposv, `_ridge.py`
// _solve_cholesky, reached from searchlight.py:56-59), restated for one CTA per centre:
//
//   A  = G[loc, loc] + alpha I            (k x k, gathered from the cluster's union Gram)
//   A  = L L^T                            blocked left-looking Cholesky, 64-wide panels
//   Zt = E A^{-1}                         E = the centre's 143 evaluation trials (rows), k columns
//
// Only the 143 evaluation rows are ever multiplied by the ridge weights
// (predictions = X_eval W = (X_eval A^{-1}) (X_nbh^T Y)), so the solve runs on 144 right-hand
// sides instead of the D columns of X^T Y.  E is appended below A as extra rows: the panel
// update/TRSM that produces L then yields the forward solve W = E L^{-T} for free, and the
// backward solve Zt = W L^{-1} is row-independent (one warp per 16 rows, no block barriers).
//
// All inner products run on the fp64 tensor pipe (DMMA m8n8k4), operands straight from
// global/L2 (each warp owns a 16-row strip; the 64 panel rows it multiplies against are
// shared through L1).  Diagonal 64x64 blocks are factored and inverted in shared memory and
// stored as L_jj^{-1}, so every triangular solve is a product.
#include "common.cuh"
#include "kernels.h"

namespace mmd {
namespace {

constexpr int THREADS = 288;          // 9 warps x 16 rows = 144 = EVAL_PAD
constexpr int NWARPS = THREADS / 32;
constexpr int LDD = NB + 1;           // shared-memory stride of the diagonal block (doubles)

// acc(16 x 64) += A(16 x K) * B(64 x K)^T ;  A, B row-major with leading dimension ld.
__device__ __forceinline__ void strip_nt(const double* a_rows, const double* b_rows,
                                         long lda, long ldb, int K, double (&acc)[2][8][2], int lane) {
    const int fr = lane >> 2, fk = lane & 3;
    const double* a0 = a_rows + fr * lda + fk;
    const double* a1 = a0 + 8 * lda;
    const double* b = b_rows + fr * ldb + fk;
    double an0 = a0[0], an1 = a1[0], bn[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) bn[j] = b[j * 8 * ldb];
    for (int k0 = 0; k0 < K; k0 += 4) {
        double ac0 = an0, ac1 = an1, bc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) bc[j] = bn[j];
        if (k0 + 4 < K) {
            an0 = a0[k0 + 4];
            an1 = a1[k0 + 4];
#pragma unroll
            for (int j = 0; j < 8; ++j) bn[j] = b[j * 8 * ldb + k0 + 4];
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            dmma884(acc[0][j][0], acc[0][j][1], ac0, bc[j]);
            dmma884(acc[1][j][0], acc[1][j][1], ac1, bc[j]);
        }
    }
}

// acc(16 x 64) += A(16 x K) * B(K x 64) ;  A row-major (lda), B row-major (ldb), 64 columns used.
__device__ __forceinline__ void strip_nn(const double* a_rows, const double* b_rows,
                                         long lda, long ldb, int K, double (&acc)[2][8][2], int lane) {
    const int fr = lane >> 2, fk = lane & 3;
    const double* a0 = a_rows + fr * lda + fk;
    const double* a1 = a0 + 8 * lda;
    const double* b = b_rows + fk * ldb + fr;
    double an0 = a0[0], an1 = a1[0], bn[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) bn[j] = b[j * 8];
    for (int k0 = 0; k0 < K; k0 += 4) {
        double ac0 = an0, ac1 = an1, bc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) bc[j] = bn[j];
        if (k0 + 4 < K) {
            an0 = a0[k0 + 4];
            an1 = a1[k0 + 4];
#pragma unroll
            for (int j = 0; j < 8; ++j) bn[j] = b[(long)(k0 + 4) * ldb + j * 8];
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            dmma884(acc[0][j][0], acc[0][j][1], ac0, bc[j]);
            dmma884(acc[1][j][0], acc[1][j][1], ac1, bc[j]);
        }
    }
}

__device__ __forceinline__ void zero_acc(double (&acc)[2][8][2]) {
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
}

// dst(16 x 64, ld) = sign_src * src(16 x 64, ld) + sign_acc * acc     (dst may alias src)
__device__ __forceinline__ void store_strip(double* dst, long ld, const double (&acc)[2][8][2], int lane,
                                            bool add_dst, double sign_acc) {
    const int fr = lane >> 2, fk = lane & 3;
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            double2* p = reinterpret_cast<double2*>(dst + (long)(i * 8 + fr) * ld + j * 8 + 2 * fk);
            double2 v = add_dst ? *p : make_double2(0.0, 0.0);
            v.x += sign_acc * acc[i][j][0];
            v.y += sign_acc * acc[i][j][1];
            *p = v;
        }
}

__global__ void __launch_bounds__(THREADS, 1) factor_solve_kernel(SolveParams prm) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* Dsm = reinterpret_cast<double*>(smem_raw);          // NB x LDD   diagonal block / L_jj
    double* Ism = Dsm + NB * LDD;                               // NB x LDD   L_jj^{-1}
    int* loc_sm = reinterpret_cast<int*>(Ism + NB * LDD);       // kp
    __shared__ int bad_pivot;

    const int task_id = blockIdx.x;
    if (task_id >= prm.n_tasks) return;
    const CentreTask task = prm.tasks[task_id];
    const int k = prm.k, kp = prm.kp, nblk = kp / NB;
    const long ld = kp;
    double* M = prm.work + (long)task_id * (kp + EVAL_PAD) * ld;   // rows 0..kp-1: A/L ; rows kp..: E/W/Zt
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    if (tid == 0) bad_pivot = 0;
    for (int i = tid; i < kp; i += THREADS) loc_sm[i] = i < k ? task.loc[i] : -1;
    __syncthreads();

    // ---- gather A = G[loc, loc] + alpha I (lower block triangle; diagonal blocks filled symmetrically) ----
    for (int i = warp; i < kp; i += NWARPS) {
        const int li = loc_sm[i];
        const int ncol = (i / NB + 1) * NB;
        double* row = M + (long)i * ld;
        for (int j = lane; j < ncol; j += 32) {
            const int lj = loc_sm[j];
            double v;
            if (li >= 0 && lj >= 0) {
                const int hi = li > lj ? li : lj, lo = li > lj ? lj : li;
                v = task.gram[(long)hi * task.gram_ld + lo];
                if (i == j) v += prm.alpha;
            } else {
                v = (i == j) ? 1.0 : 0.0;
            }
            row[j] = v;
        }
    }
    // ---- E rows: M[kp + e][i] = Xe[nbr[i]][e] ----
    for (int i = warp; i < kp; i += NWARPS) {
        const __half* src = i < k ? prm.Xe + (long)task.nbr[i] * EVAL_PAD : nullptr;
        for (int e = lane; e < EVAL_PAD; e += 32) M[(long)(kp + e) * ld + i] = src ? h2d(src[e]) : 0.0;
    }
    __syncthreads();

    const int n_estrips = EVAL_PAD / 16;
    double acc[2][8][2];

    for (int j = 0; j < nblk; ++j) {
        const int K = j * NB;
        const double* panel_rows = M + (long)(j * NB) * ld;        // rows j*64.., columns 0..K are L_j,0:K
        // ---- phase 1: P = M[:, panel] - M[:, 0:K] * L_j,0:K^T for every strip at or below the diagonal block
        const int first = j * (NB / 16), n_astrips = kp / 16 - first;
        if (K > 0) {
            for (int t = warp; t < n_astrips + n_estrips; t += NWARPS) {
                const int s = first + t;                              // E strips follow the A strips contiguously
                double* rows = M + (long)s * 16 * ld;
                zero_acc(acc);
                strip_nt(rows, panel_rows, ld, ld, K, acc, lane);
                store_strip(rows + K, ld, acc, lane, true, -1.0);
            }
            __syncthreads();
        }
        // ---- phase 2: factor the diagonal block in shared memory, invert it, store L_jj^{-1} in its place
        for (int e = tid; e < NB * NB; e += THREADS) {
            const int r = e / NB, c = e % NB;
            Dsm[r * LDD + c] = M[(long)(j * NB + r) * ld + K + c];
        }
        __syncthreads();
        for (int c = 0; c < NB; ++c) {
            double d = Dsm[c * LDD + c];
            if (!(d > 0.0)) {
                if (tid == 0) bad_pivot = 1;
                d = 1.0;
            }
            const double piv = sqrt(d), inv = 1.0 / piv;
            __syncthreads();
            if (tid < NB) {
                if (tid > c) Dsm[tid * LDD + c] *= inv;
                else if (tid == c) Dsm[c * LDD + c] = piv;
            }
            __syncthreads();
            for (int r = c + 1 + warp; r < NB; r += NWARPS) {
                const double lrc = Dsm[r * LDD + c];
                for (int cc = c + 1 + lane; cc <= r; cc += 32) Dsm[r * LDD + cc] -= lrc * Dsm[cc * LDD + c];
            }
            __syncthreads();
        }
        // inverse of the lower-triangular block: column c by forward substitution, 4 threads per column
        if (tid < 4 * NB) {
            const int c = tid >> 2, q = tid & 3;
            for (int r = 0; r < NB; ++r) {
                double s = 0.0;
                for (int m = q; m < r; m += 4) s += Dsm[r * LDD + m] * Ism[m * LDD + c];
                s += __shfl_xor_sync(0xffffffffu, s, 1);
                s += __shfl_xor_sync(0xffffffffu, s, 2);
                if (q == 0) {
                    double x;
                    if (r < c) x = 0.0;
                    else if (r == c) x = 1.0 / Dsm[r * LDD + r];
                    else x = -s / Dsm[r * LDD + r];
                    Ism[r * LDD + c] = x;
                }
                __syncwarp();
            }
        }
        __syncthreads();
        for (int e = tid; e < NB * NB; e += THREADS) {
            const int r = e / NB, c = e % NB;
            M[(long)(j * NB + r) * ld + K + c] = Ism[r * LDD + c];
        }
        __syncthreads();
        // ---- phase 3: rows below the block (and the E rows): L_rj = P_rj * L_jj^{-T}
        const double* inv_blk = panel_rows + K;
        const int first3 = (j + 1) * (NB / 16), n3 = kp / 16 - first3 + n_estrips;
        for (int t = warp; t < n3; t += NWARPS) {
            double* rows = M + (long)(first3 + t) * 16 * ld + K;
            zero_acc(acc);
            strip_nt(rows, inv_blk, ld, ld, NB, acc, lane);
            __syncwarp();
            store_strip(rows, ld, acc, lane, false, 1.0);
        }
        __syncthreads();
    }

    // ---- backward solve, one warp per 16 evaluation rows: Zt_i = (W_i - sum_{p>i} Zt_p L_pi) L_ii^{-1}
    if (warp < n_estrips) {
        double* zrows = M + (long)(kp + warp * 16) * ld;
        for (int i = nblk - 1; i >= 0; --i) {
            const int c0 = i * NB, Kb = kp - (c0 + NB);
            if (Kb > 0) {
                zero_acc(acc);
                strip_nn(zrows + c0 + NB, M + (long)(c0 + NB) * ld + c0, ld, ld, Kb, acc, lane);
                store_strip(zrows + c0, ld, acc, lane, true, -1.0);
                __syncwarp();
            }
            zero_acc(acc);
            strip_nn(zrows + c0, M + (long)c0 * ld + c0, ld, ld, NB, acc, lane);
            __syncwarp();
            store_strip(zrows + c0, ld, acc, lane, false, 1.0);
            __syncwarp();
        }
    }
    if (tid == 0) prm.flags[task.centre] = bad_pivot;
}

}  // namespace

size_t factor_solve_smem_bytes() { return 0; }

cudaError_t launch_factor_solve(const SolveParams& p, cudaStream_t s) {
    const size_t smem = 2 * NB * LDD * sizeof(double) + (size_t)p.kp * sizeof(int);
    cudaError_t e = cudaFuncSetAttribute(factor_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    factor_solve_kernel<<<p.n_tasks, THREADS, smem, s>>>(p);
    return cudaGetLastError();
}

}  // namespace mmd
