// Per-centre ridge solve in float64 (the part sklearn delegates to LAPACK dposv, `_ridge.py`
// _solve_cholesky, reached from searchlight.py:56-59), one CTA per centre, two CTAs per SM:
//
//   A  = G[loc, loc] + alpha I            (k x k, read straight out of the cluster's union Gram)
//   A  = L L^T                            blocked left-looking Cholesky, 64-wide panels
//   Zt = E A^{-1}                         E = the centre's 143 evaluation trials (rows), k columns
//
// Only the 143 evaluation rows are ever multiplied by the ridge weights
// (predictions = X_eval W = (X_eval A^{-1}) (X_nbh^T Y)), so the solve runs on 144 right-hand
// sides instead of the D columns of X^T Y.  E is appended below A as extra 16-row strips: the
// panel update that produces L also yields the forward solve W = E L^{-T}; the backward solve
// Zt = W L^{-1} runs afterwards, one warp per strip.
//
// Work decomposition.  A "strip" is 16 rows x 64 columns of the current panel, owned by one
// warp: acc(16x64) = init - strip(16xK) * panel(64xK)^T on the fp64 tensor pipe (DMMA m8n8k4).
// The panel operand (shared by all nine warps) is staged in shared memory with a 3-stage cp.async
// pipeline, the strip operand streams from global/L2 with 16-byte loads one chunk ahead.  The
// initial value of a strip is gathered from the union Gram (or the evaluation trials) at that
// moment, so A itself is never materialised.  The triangular solve against the 64x64 diagonal
// block is a product with its inverse (kept in shared memory), applied to the strip while it is
// still on chip; every strip of L / W is written to global memory exactly once.
#include "common.cuh"
#include "kernels.h"

namespace mmd {
namespace {

constexpr int THREADS = 288;          // 9 warps x 16 rows = 144 = EVAL_PAD
constexpr int NWARPS = THREADS / 32;
constexpr int KC = 16;                // k-values per pipeline chunk
constexpr int STAGES = 3;
constexpr int LDB_NT = 24;            // [n][k] chunk rows: 16 doubles + 8 pad (conflict-free 16-byte fragment loads)
constexpr int LDB_NN = 66;            // [k][n] chunk rows: 64 doubles + 2 pad
constexpr int STAGE_DOUBLES = 1632;   // >= 64*24, >= 16*66; 3 stages = 4896 doubles = 9 warps x 8 x 68 staging
constexpr int LDD = 68;               // stride of the 64x64 diagonal block and of the per-warp staging rows
constexpr int PIPE_DOUBLES = STAGES * STAGE_DOUBLES;
static_assert(PIPE_DOUBLES >= NB * LDD, "diagonal block aliases the pipeline buffers");
static_assert(PIPE_DOUBLES >= NWARPS * 8 * LDD, "strip staging aliases the pipeline buffers");

typedef double Acc[2][8][2];          // C fragments of a 16 x 64 strip: [row half][8-column block][2]

// acc += strip(16 x 16*nchunks) * B^T   (NN = false: B rows are the 64 panel rows, contiguous along k)
// acc += strip(16 x 16*nchunks) * B     (NN = true : B is (16*nchunks) x 64, row-major)
// Every thread of the CTA must call this (cp.async staging and barriers are CTA-wide); `active`
// says whether this warp owns a strip.  The k index inside a group of 8 is permuted the same way for
// both operands (lane fk takes k = 8t + 2fk + p), which lets the strip operand use 16-byte loads.
template <bool NN>
__device__ __forceinline__ void strip_gemm(Acc& acc, const double* __restrict__ a_strip, long lda,
                                           const double* __restrict__ b_src, long ldb, int nchunks,
                                           double* pipe, bool active, int tid, int lane) {
    const int fr = lane >> 2, fk = lane & 3;
    auto load_b = [&](int stage, int c) {
        double* dst = pipe + stage * STAGE_DOUBLES;
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const int idx = tid + q * THREADS;
            if (idx < 512) {
                if (!NN) {
                    const int n = idx >> 3, kk = (idx & 7) * 2;
                    cp_async16(dst + n * LDB_NT + kk, b_src + (long)n * ldb + c * KC + kk);
                } else {
                    const int kk = idx >> 5, n = (idx & 31) * 2;
                    cp_async16(dst + kk * LDB_NN + n, b_src + (long)(c * KC + kk) * ldb + n);
                }
            }
        }
    };
    __syncthreads();                                   // previous users of the pipeline buffers are done
    for (int s = 0; s < STAGES - 1; ++s) {
        if (s < nchunks) load_b(s, s);
        cp_async_commit();
    }
    const double* a0p = a_strip + (long)fr * lda + 2 * fk;
    const double* a1p = a0p + 8 * lda;
    double2 cur0, cur1, nxt0, nxt1;                   // strip operand, half a chunk (8 k-values) ahead
    if (active && nchunks > 0) {
        cur0 = *reinterpret_cast<const double2*>(a0p);
        cur1 = *reinterpret_cast<const double2*>(a1p);
    }
    for (int c = 0; c < nchunks; ++c) {
        cp_async_wait<STAGES - 2>();
        __syncthreads();
        {
            const int n = c + STAGES - 1;
            if (n < nchunks) load_b(n % STAGES, n);
            cp_async_commit();
        }
        if (active) {
            const double* bs = pipe + (c % STAGES) * STAGE_DOUBLES;
#pragma unroll
            for (int t = 0; t < 2; ++t) {
                const int o = c * KC + 8 * (t + 1);
                if (t == 0 || c + 1 < nchunks) {
                    nxt0 = *reinterpret_cast<const double2*>(a0p + o);
                    nxt1 = *reinterpret_cast<const double2*>(a1p + o);
                }
#pragma unroll
                for (int nb = 0; nb < 8; ++nb) {
                    double bx, by;
                    if (!NN) {
                        const double2 b = *reinterpret_cast<const double2*>(bs + (nb * 8 + fr) * LDB_NT + 8 * t + 2 * fk);
                        bx = b.x;
                        by = b.y;
                    } else {
                        bx = bs[(8 * t + 2 * fk) * LDB_NN + nb * 8 + fr];
                        by = bs[(8 * t + 2 * fk + 1) * LDB_NN + nb * 8 + fr];
                    }
                    dmma884(acc[0][nb][0], acc[0][nb][1], cur0.x, bx);
                    dmma884(acc[1][nb][0], acc[1][nb][1], cur1.x, bx);
                    dmma884(acc[0][nb][0], acc[0][nb][1], cur0.y, by);
                    dmma884(acc[1][nb][0], acc[1][nb][1], cur1.y, by);
                }
                cur0 = nxt0;
                cur1 = nxt1;
            }
        }
    }
    cp_async_wait<0>();
    __syncthreads();                                   // all warps are done with the staged chunks (callers reuse the buffers)
}

// acc <- (-acc) * T^T (NN = false) or (-acc) * T (NN = true) for a lower-triangular 64x64 block T held in
// shared memory (stride LDD).  The strip is staged through a warp-private shared buffer, 8 rows at a
// time, to turn C fragments into A fragments; zero blocks of T are skipped.
template <bool NN>
__device__ __forceinline__ void strip_times_tri(Acc& acc, const double* __restrict__ T, double* stage, int lane) {
    const int fr = lane >> 2, fk = lane & 3;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        __syncwarp();
#pragma unroll
        for (int nb = 0; nb < 8; ++nb) {
            *reinterpret_cast<double2*>(stage + fr * LDD + nb * 8 + 2 * fk) = make_double2(-acc[i][nb][0], -acc[i][nb][1]);
            acc[i][nb][0] = acc[i][nb][1] = 0.0;
        }
        __syncwarp();
#pragma unroll 1
        for (int kb = 0; kb < 8; ++kb) {                 // rolled on purpose: bounds the live fragment registers
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int ks = 2 * kb + h;
                const double a = stage[fr * LDD + 4 * ks + fk];
#pragma unroll
                for (int nb = 0; nb < 8; ++nb) {
                    // NT: out[r][n] = sum_c P[r][c] T[n][c], T[n][c] = 0 for c > n  -> k-step ks feeds blocks nb >= ks/2
                    // NN: out[r][n] = sum_c P[r][c] T[c][n], T[c][n] = 0 for c < n  -> k-step ks feeds blocks nb <= ks/2
                    if (!NN) {
                        if (nb >= kb) dmma884(acc[i][nb][0], acc[i][nb][1], a, T[(nb * 8 + fr) * LDD + 4 * ks + fk]);
                    } else {
                        if (nb <= kb) dmma884(acc[i][nb][0], acc[i][nb][1], a, T[(4 * ks + fk) * LDD + nb * 8 + fr]);
                    }
                }
            }
        }
    }
}

// Negated initial value of a strip of the current panel: -(G[loc, loc] + alpha I) for rows of A,
// -E for the evaluation rows (u >= n_a_strips).
__device__ __forceinline__ void strip_init(Acc& acc, const CentreTask& task, const SolveParams& prm, const int* loc_sm,
                                           const int* nbr_sm, bool is_eval, int row0, int col0, int lane) {
    const int fr = lane >> 2, fk = lane & 3;
    const int k = prm.k;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int r = row0 + 8 * i + fr;                 // row of A, or evaluation trial index
        const int li = is_eval ? 0 : loc_sm[r];
#pragma unroll
        for (int nb = 0; nb < 8; ++nb)
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const int c = col0 + nb * 8 + 2 * fk + q;
                double v;
                if (is_eval) {
                    v = c < k ? h2d(prm.Xe[(long)nbr_sm[c] * EVAL_PAD + r]) : 0.0;
                } else {
                    const int lj = loc_sm[c];
                    if (li >= 0 && lj >= 0) {
                        const int hi = li > lj ? li : lj, lo = li > lj ? lj : li;
                        v = task.gram[(long)hi * task.gram_ld + lo];
                        if (r == c) v += prm.alpha;
                    } else {
                        v = (r == c) ? 1.0 : 0.0;        // padding rows/columns: identity
                    }
                }
                acc[i][nb][q] = -v;
            }
    }
}

__device__ __forceinline__ void strip_store(const Acc& acc, double* dst, long ld, int lane) {
    const int fr = lane >> 2, fk = lane & 3;
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int nb = 0; nb < 8; ++nb)
            *reinterpret_cast<double2*>(dst + (long)(8 * i + fr) * ld + nb * 8 + 2 * fk) =
                make_double2(acc[i][nb][0], acc[i][nb][1]);
}

// Cholesky factor of the 64x64 block in D (lower triangle used; overwritten by L_jj) and its inverse in I
// (full block, zeros above the diagonal).  All threads of the CTA take part.
__device__ __forceinline__ void factor_invert_block(double* D, double* I, int* bad_pivot, int tid, int lane, int warp) {
    for (int c = 0; c < NB; ++c) {
        double d = D[c * LDD + c];
        if (!(d > 0.0)) {
            if (tid == 0) *bad_pivot = 1;
            d = 1.0;
        }
        const double piv = sqrt(d), inv = 1.0 / piv;
        __syncthreads();
        if (tid < NB) {
            if (tid > c) D[tid * LDD + c] *= inv;
            else if (tid == c) D[c * LDD + c] = piv;
        }
        __syncthreads();
        for (int r = c + 1 + warp; r < NB; r += NWARPS) {
            const double lrc = D[r * LDD + c];
            for (int cc = c + 1 + lane; cc <= r; cc += 32) D[r * LDD + cc] -= lrc * D[cc * LDD + c];
        }
        __syncthreads();
    }
    // inverse of the lower-triangular block: column c by forward substitution, 4 threads per column
    if (tid < 4 * NB) {
        const int c = tid >> 2, q = tid & 3;
        for (int r = 0; r < NB; ++r) {
            double s = 0.0;
            for (int m = q; m < r; m += 4) s += D[r * LDD + m] * I[m * LDD + c];
            s += __shfl_xor_sync(0xffffffffu, s, 1);
            s += __shfl_xor_sync(0xffffffffu, s, 2);
            if (q == 0) {
                double x;
                if (r < c) x = 0.0;
                else if (r == c) x = 1.0 / D[r * LDD + r];
                else x = -s / D[r * LDD + r];
                I[r * LDD + c] = x;
            }
            __syncwarp();
        }
    }
    __syncthreads();
}

__global__ void __launch_bounds__(THREADS, 2) factor_solve_kernel(SolveParams prm) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* pipe = reinterpret_cast<double*>(smem_raw);         // cp.async stages | diagonal block | strip staging
    double* Ism = pipe + PIPE_DOUBLES;                          // NB x LDD: L_jj^{-1}
    int* loc_sm = reinterpret_cast<int*>(Ism + NB * LDD);       // kp
    int* nbr_sm = loc_sm + prm.kp;                              // kp
    __shared__ int bad_pivot;

    const int task_id = blockIdx.x;
    if (task_id >= prm.n_tasks) return;
    const CentreTask task = prm.tasks[task_id];
    const int k = prm.k, kp = prm.kp, nblk = kp / NB;
    const long ld = kp;
    double* M = prm.work + (long)task_id * (kp + EVAL_PAD) * ld;   // rows 0..kp-1: L ; rows kp..: W then Zt
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    double* stage = pipe + warp * 8 * LDD;
    double* Dsm = pipe;

    if (tid == 0) bad_pivot = 0;
    for (int i = tid; i < kp; i += THREADS) {
        loc_sm[i] = i < k ? task.loc[i] : -1;
        nbr_sm[i] = i < k ? task.nbr[i] : 0;
    }
    __syncthreads();

    Acc acc;
    constexpr int N_ESTRIPS = EVAL_PAD / 16;

    // ---------------- factorisation + forward solve, panel by panel ----------------
    for (int j = 0; j < nblk; ++j) {
        const int K = j * NB;
        const int n_a = kp / 16 - 4 * j;                        // strips of A from the diagonal block down
        const int n_strips = n_a + N_ESTRIPS;
        const double* panel_rows = M + (long)K * ld;            // L[K..K+63][0..K)
        for (int u0 = 0; u0 < n_strips; u0 += NWARPS) {
            const int u = u0 + warp;
            const bool active = u < n_strips;
            const bool is_eval = u >= n_a;
            const int mrow = is_eval ? kp + 16 * (u - n_a) : K + 16 * u;        // first row of the strip in M
            if (active) strip_init(acc, task, prm, loc_sm, nbr_sm, is_eval, is_eval ? 16 * (u - n_a) : mrow, K, lane);
            strip_gemm<false>(acc, M + (long)mrow * ld, ld, panel_rows, ld, K / KC, pipe, active, tid, lane);
            if (u0 == 0) {
                __syncthreads();                                 // pipeline buffers free -> diagonal block
                if (warp < 4) {
                    const int fr = lane >> 2, fk = lane & 3;
#pragma unroll
                    for (int i = 0; i < 2; ++i)
#pragma unroll
                        for (int nb = 0; nb < 8; ++nb)
                            *reinterpret_cast<double2*>(Dsm + (16 * warp + 8 * i + fr) * LDD + nb * 8 + 2 * fk) =
                                make_double2(-acc[i][nb][0], -acc[i][nb][1]);
                }
                __syncthreads();
                factor_invert_block(Dsm, Ism, &bad_pivot, tid, lane, warp);
                for (int e = tid; e < NB * NB; e += THREADS) {
                    const int r = e / NB, c = e % NB;
                    M[(long)(K + r) * ld + K + c] = Ism[r * LDD + c];          // diagonal blocks hold L_jj^{-1}
                }
                __syncthreads();                                 // Dsm no longer needed: staging may reuse it
                if (warp >= 4 && active) {
                    strip_times_tri<false>(acc, Ism, stage, lane);
                    strip_store(acc, M + (long)mrow * ld + K, ld, lane);
                }
            } else if (active) {
                strip_times_tri<false>(acc, Ism, stage, lane);
                strip_store(acc, M + (long)mrow * ld + K, ld, lane);
            }
        }
        __syncthreads();                                         // panel j complete (global writes visible CTA-wide)
    }

    // ---------------- backward solve: Zt_i = (W_i - sum_{p>i} Zt_p L_pi) L_ii^{-1}, one warp per 16 rows ----------------
    {
        double* zrows = M + (long)(kp + 16 * warp) * ld;
        const int fr = lane >> 2, fk = lane & 3;
        for (int i = nblk - 1; i >= 0; --i) {
            const int c0 = i * NB, Kb = kp - (c0 + NB);
#pragma unroll
            for (int h = 0; h < 2; ++h)
#pragma unroll
                for (int nb = 0; nb < 8; ++nb) {
                    const double2 w = *reinterpret_cast<const double2*>(zrows + (long)(8 * h + fr) * ld + c0 + nb * 8 + 2 * fk);
                    acc[h][nb][0] = -w.x;
                    acc[h][nb][1] = -w.y;
                }
            strip_gemm<true>(acc, zrows + c0 + NB, ld, M + (long)(c0 + NB) * ld + c0, ld, Kb / KC, pipe, true, tid, lane);
            __syncthreads();                                     // pipeline idle, previous Ism readers done
            for (int e = tid; e < NB * NB; e += THREADS) {
                const int r = e / NB, c = e % NB;
                Ism[r * LDD + c] = M[(long)(c0 + r) * ld + c0 + c];
            }
            __syncthreads();
            strip_times_tri<true>(acc, Ism, stage, lane);
            strip_store(acc, zrows + c0, ld, lane);
            __syncthreads();                                     // Zt block visible before it is read as an operand
        }
    }
    if (tid == 0) prm.flags[task.centre] = bad_pivot;
}

}  // namespace

size_t factor_solve_smem_bytes() { return (size_t)(PIPE_DOUBLES + NB * LDD) * sizeof(double); }

cudaError_t launch_factor_solve(const SolveParams& p, cudaStream_t s) {
    const size_t smem = factor_solve_smem_bytes() + 2 * (size_t)p.kp * sizeof(int);
    static size_t configured = 0;
    if (smem > configured) {
        cudaError_t e = cudaFuncSetAttribute(factor_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        configured = smem;
    }
    factor_solve_kernel<<<p.n_tasks, THREADS, smem, s>>>(p);
    return cudaGetLastError();
}

}  // namespace mmd
