// Per-centre ridge solve in float64 (the part sklearn delegates to LAPACK dposv, `_ridge.py`
// _solve_cholesky, reached from searchlight.py:56-59), one 4-warp CTA per centre, three CTAs per SM:
//
//   A  = G[loc, loc] + alpha I            (k x k, read straight out of the cluster's union Gram)
//   A  = L L^T                            blocked left-looking Cholesky, 64-wide panels
//   Zt = E A^{-1}                         E = the centre's 143 evaluation trials (rows), k columns
//
// Only the 143 evaluation rows are ever multiplied by the ridge weights
// (predictions = X_eval W = (X_eval A^{-1}) (X_nbh^T Y)), so the solve runs on 144 right-hand
// sides instead of the D columns of X^T Y.  E is appended below A as nine extra 16-row strips: the
// panel update that produces L also yields the forward solve W = E L^{-T}; the backward solve
// Zt = W L^{-1} runs afterwards on the same strips.
//
// Work decomposition.  A "strip" is 16 rows x 64 columns of the current panel:
//   acc(16x64) = init - strip(16xK) * panel(64xK)^T      on the fp64 tensor pipe (DMMA m8n8k4).
// A round handles four strips, one per warp (one warp per SM sub-partition, so the four schedulers
// carry the same load); both operands of the round go through a 4-stage cp.async ring in shared
// memory.  The number of strips of a panel is always 4q + 1: the last one is processed by all four
// warps together, 16 columns each ("split" round), so no warp idles.  The initial value of a strip
// is gathered from the union Gram (or the evaluation trials) while the ring fills, so A itself is
// never materialised.  The triangular solve against the 64x64 diagonal block is a product with its
// inverse (kept in shared memory), applied to the strip while it is still on chip; every strip of
// L / W is written to global memory exactly once.
#include "common.cuh"
#include "kernels.h"

namespace mmd {
namespace {

constexpr int THREADS = 128;
constexpr int NWARPS = THREADS / 32;
constexpr int KC = 8;                 // k-values per pipeline chunk (one 64-byte row segment)
constexpr int STAGES = 4;
#ifndef MMD_FS_PAIRED
#define MMD_FS_PAIRED 0
#endif
constexpr bool PAIRED = MMD_FS_PAIRED != 0;   // two pipeline chunks per CTA barrier in strip_gemm (measured slower: 4.90 vs 4.53 s / 2 hemispheres)
constexpr int A_ROWS = NWARPS * 16;   // strip rows staged per chunk (one 16-row strip per warp)
constexpr int A_STAGE = A_ROWS * KC;  // [row][k], 64-byte rows: conflict-free 16-byte fragment loads
constexpr int LDB_NN = 66;            // [k][n] chunk rows: 64 doubles + 2 pad
constexpr int LDD = 68;               // stride of the 64x64 diagonal block and of the staging rows
constexpr int STAGE_DOUBLES = NB * LDD / STAGES;       // 1088 >= A_STAGE + max(64*KC, KC*LDB_NN)
constexpr int PIPE_DOUBLES = STAGES * STAGE_DOUBLES;   // = NB * LDD: the diagonal block aliases the ring
static_assert(STAGE_DOUBLES >= A_STAGE + KC * LDB_NN && STAGE_DOUBLES >= A_STAGE + NB * KC, "stage size");
static_assert(PIPE_DOUBLES >= NWARPS * 8 * LDD && PIPE_DOUBLES >= 16 * LDD, "strip staging aliases the ring");
static_assert(A_ROWS * (KC / 2) == 2 * THREADS && NB * (KC / 2) == 2 * THREADS, "loader thread map");
static_assert(EVAL_PAD % 16 == 0 && (EVAL_PAD / 16) % NWARPS == 1, "strip count of a panel must be 4q + 1");

// C fragments of a 16-row strip: [row half][8-column block][2]; NBW = 8 (whole strip) or 2 (16 columns)
template <int NBW>
struct Frag {
    double v[2][NBW][2];
};

// acc += strip(16 x 8*nchunks) * B^T   (NN = false: B rows are the 64 panel rows, contiguous along k)
// acc += strip(16 x 8*nchunks) * B     (NN = true : B is (8*nchunks) x 64, row-major)
// NBW = 8: warp w owns staged strip w and all 64 columns.  NBW = 2: all warps share staged strip 0 and
// warp w owns columns 16w .. 16w+15.  `a_row(r)` gives the global address of staged row r at k = 0
// (nullptr: row not staged).  `init()` runs after the first loads are in flight, so the gather of the
// strip's initial value overlaps the pipeline fill.  Every thread of the CTA must call this (staging
// and barriers are CTA-wide).  The k index inside a chunk is permuted the same way for both operands
// (lane fk takes k = 2fk + p), which lets the fragments be 16-byte loads.
template <bool NN, int NBW, typename RowPtr, typename Init>
__device__ __forceinline__ void strip_gemm(Frag<NBW>& acc, RowPtr a_row, const double* __restrict__ b_src, long ldb,
                                           int nchunks, double* pipe, Init init, int tid, int lane, int warp) {
    const int fr = lane >> 2, fk = lane & 3;
    const int lrow = tid >> 2, lpart = (tid & 3) * 2;
    const double* a_src0 = a_row(lrow);
    const double* a_src1 = a_row(lrow + THREADS / 4);
    if (a_src0) a_src0 += lpart;
    if (a_src1) a_src1 += lpart;
    const int a_off0 = lrow * KC + lpart, a_off1 = a_off0 + (THREADS / 4) * KC;
    const double *b_ptr0, *b_ptr1;
    int b_off0, b_off1;
    long b_step;
    if (!NN) {
        b_ptr0 = b_src + (long)lrow * ldb + lpart;
        b_ptr1 = b_ptr0 + (long)(THREADS / 4) * ldb;
        b_off0 = A_STAGE + lrow * KC + lpart;
        b_off1 = b_off0 + (THREADS / 4) * KC;
        b_step = KC;
    } else {
        const int kk = tid >> 5, n = (tid & 31) * 2;
        b_ptr0 = b_src + (long)kk * ldb + n;
        b_ptr1 = b_ptr0 + (long)(KC / 2) * ldb;
        b_off0 = A_STAGE + kk * LDB_NN + n;
        b_off1 = b_off0 + (KC / 2) * LDB_NN;
        b_step = (long)KC * ldb;
    }
    auto load = [&](int stage, int c) {
        double* dst = pipe + stage * STAGE_DOUBLES;
        if (a_src0) cp_async16(dst + a_off0, a_src0 + c * KC);
        if (a_src1) cp_async16(dst + a_off1, a_src1 + c * KC);
        cp_async16(dst + b_off0, b_ptr0 + c * b_step);
        cp_async16(dst + b_off1, b_ptr1 + c * b_step);
    };
    __syncthreads();                                   // previous users of the pipeline buffers are done
    const int a_slot = NBW == 8 ? warp : 0, nb0 = NBW == 8 ? 0 : warp * NBW;
    auto compute = [&](int c) {
        const double* st = pipe + (c % STAGES) * STAGE_DOUBLES;
        const double2 a0 = *reinterpret_cast<const double2*>(st + (a_slot * 16 + fr) * KC + 2 * fk);
        const double2 a1 = *reinterpret_cast<const double2*>(st + (a_slot * 16 + 8 + fr) * KC + 2 * fk);
        const double* bs = st + A_STAGE;
#pragma unroll
        for (int q = 0; q < NBW; q += 2) {
            double bx[2], by[2];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int nb = nb0 + q + e;
                if (!NN) {
                    const double2 b = *reinterpret_cast<const double2*>(bs + (nb * 8 + fr) * KC + 2 * fk);
                    bx[e] = b.x;
                    by[e] = b.y;
                } else {
                    bx[e] = bs[(2 * fk) * LDB_NN + nb * 8 + fr];
                    by[e] = bs[(2 * fk + 1) * LDB_NN + nb * 8 + fr];
                }
            }
            // four independent accumulators between dependent DMMAs
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                dmma884(acc.v[0][q + e][0], acc.v[0][q + e][1], a0.x, bx[e]);
                dmma884(acc.v[1][q + e][0], acc.v[1][q + e][1], a1.x, bx[e]);
            }
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                dmma884(acc.v[0][q + e][0], acc.v[0][q + e][1], a0.y, by[e]);
                dmma884(acc.v[1][q + e][0], acc.v[1][q + e][1], a1.y, by[e]);
            }
        }
    };
    if (PAIRED) {
        // two chunks per barrier (nchunks is even: K is a multiple of 64): half the barriers and wait_groups of
        // the one-chunk schedule; the next pair is in flight while the current one is consumed
        if (nchunks > 0) {
            load(0, 0);
            load(1, 1);
        }
        cp_async_commit();
        init();
        for (int c = 0; c < nchunks; c += 2) {
            cp_async_wait<0>();
            __syncthreads();
            if (c + 2 < nchunks) {
                load((c + 2) % STAGES, c + 2);
                load((c + 3) % STAGES, c + 3);
            }
            cp_async_commit();
            compute(c);
            compute(c + 1);
        }
    } else {
        for (int s = 0; s < STAGES - 1; ++s) {
            if (s < nchunks) load(s, s);
            cp_async_commit();
        }
        init();
        for (int c = 0; c < nchunks; ++c) {
            cp_async_wait<STAGES - 2>();
            __syncthreads();
            {
                const int n = c + STAGES - 1;
                if (n < nchunks) load(n % STAGES, n);
                cp_async_commit();
            }
            compute(c);
        }
    }
    cp_async_wait<0>();
    __syncthreads();                                   // all warps are done with the staged chunks (callers reuse the buffers)
}

// acc <- (-acc) * T^T (NN = false) or (-acc) * T (NN = true) for a lower-triangular 64x64 block T held in
// shared memory (stride LDD).  The strip is staged through a warp-private shared buffer, 8 rows at a
// time, to turn C fragments into A fragments; zero blocks of T are skipped.
template <bool NN>
__device__ __forceinline__ void strip_times_tri(Frag<8>& acc, const double* __restrict__ T, double* stage, int lane) {
    const int fr = lane >> 2, fk = lane & 3;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        __syncwarp();
#pragma unroll
        for (int nb = 0; nb < 8; ++nb) {
            *reinterpret_cast<double2*>(stage + fr * LDD + nb * 8 + 2 * fk) = make_double2(-acc.v[i][nb][0], -acc.v[i][nb][1]);
            acc.v[i][nb][0] = acc.v[i][nb][1] = 0.0;
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
                        if (nb >= kb) dmma884(acc.v[i][nb][0], acc.v[i][nb][1], a, T[(nb * 8 + fr) * LDD + 4 * ks + fk]);
                    } else {
                        if (nb <= kb) dmma884(acc.v[i][nb][0], acc.v[i][nb][1], a, T[(4 * ks + fk) * LDD + nb * 8 + fr]);
                    }
                }
            }
        }
    }
}

// Same product for the strip of a split round: every warp holds 16 of the 64 columns.  The raw strip is
// assembled in shared memory (S, 16 x LDD, aliases the idle pipeline ring), then warp w computes output
// columns 16w .. 16w+15.
template <bool NN>
__device__ __forceinline__ void split_times_tri(Frag<2>& acc, const double* __restrict__ T, double* S, int lane, int warp) {
    const int fr = lane >> 2, fk = lane & 3;
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            *reinterpret_cast<double2*>(S + (8 * i + fr) * LDD + (2 * warp + e) * 8 + 2 * fk) =
                make_double2(-acc.v[i][e][0], -acc.v[i][e][1]);
            acc.v[i][e][0] = acc.v[i][e][1] = 0.0;
        }
    __syncthreads();
#pragma unroll 4
    for (int ks = 0; ks < 16; ++ks) {
        const int kb = ks >> 1;
        const double a0 = S[fr * LDD + 4 * ks + fk], a1 = S[(8 + fr) * LDD + 4 * ks + fk];
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const int nb = 2 * warp + e;
            if (!NN ? nb >= kb : nb <= kb) {
                const double b = !NN ? T[(nb * 8 + fr) * LDD + 4 * ks + fk] : T[(4 * ks + fk) * LDD + nb * 8 + fr];
                dmma884(acc.v[0][e][0], acc.v[0][e][1], a0, b);
                dmma884(acc.v[1][e][0], acc.v[1][e][1], a1, b);
            }
        }
    }
}

// Negated initial value of NBW column blocks of a strip of the current panel, starting at column col0:
// -(G[loc, loc] + alpha I) for rows of A, -E for the evaluation rows.  Two passes: every load is issued
// first (unconditionally, padding indices clamped to a valid address), the fix-ups follow, so the gather
// costs one memory round trip instead of one per element.
template <int NBW>
__device__ __forceinline__ void strip_init(Frag<NBW>& acc, const CentreTask& task, const SolveParams& prm, bool is_eval,
                                           int row0, int col0, int lane) {
    const int fr = lane >> 2, fk = lane & 3;
    const int k = task.k;
    int lj[NBW][2];
#pragma unroll
    for (int nb = 0; nb < NBW; ++nb)
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const int c = col0 + nb * 8 + 2 * fk + q;
            lj[nb][q] = c < k ? __ldg((is_eval ? task.nbr : task.loc) + c) : -1;
        }
    if (is_eval) {
        const unsigned short* xe = reinterpret_cast<const unsigned short*>(prm.Xe);
        unsigned short raw[2][NBW][2];
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int nb = 0; nb < NBW; ++nb)
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                    const int l = lj[nb][q];
                    raw[i][nb][q] = __ldg(xe + (long)(l >= 0 ? l : 0) * EVAL_PAD + row0 + 8 * i + fr);
                }
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int nb = 0; nb < NBW; ++nb)
#pragma unroll
                for (int q = 0; q < 2; ++q)
                    acc.v[i][nb][q] = lj[nb][q] >= 0 ? -h2d(__ushort_as_half(raw[i][nb][q])) : 0.0;
    } else {
        int li[2];
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int r = row0 + 8 * i + fr;
            li[i] = r < k ? __ldg(task.loc + r) : -1;
        }
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int nb = 0; nb < NBW; ++nb)
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                    const int a = li[i] >= 0 ? li[i] : 0, b = lj[nb][q] >= 0 ? lj[nb][q] : 0;
                    const int hi = a > b ? a : b, lo = a > b ? b : a;
                    acc.v[i][nb][q] = __ldg(task.gram + (long)hi * task.gram_ld + lo);
                }
        if (prm.gram_sub_off != 0) {                       // cross-validation: leave out the fold's trials
            const double* gsub = task.gram + prm.gram_sub_off;
#pragma unroll
            for (int i = 0; i < 2; ++i)
#pragma unroll
                for (int nb = 0; nb < NBW; ++nb)
#pragma unroll
                    for (int q = 0; q < 2; ++q) {
                        const int a = li[i] >= 0 ? li[i] : 0, b = lj[nb][q] >= 0 ? lj[nb][q] : 0;
                        const int hi = a > b ? a : b, lo = a > b ? b : a;
                        acc.v[i][nb][q] -= __ldg(gsub + (long)hi * task.gram_ld + lo);
                    }
        }
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int nb = 0; nb < NBW; ++nb)
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                    const bool diag = row0 + 8 * i + fr == col0 + nb * 8 + 2 * fk + q;
                    const bool real = li[i] >= 0 && lj[nb][q] >= 0;
                    const double v = real ? acc.v[i][nb][q] + (diag ? prm.alpha : 0.0) : (diag ? 1.0 : 0.0);   // padding: identity
                    acc.v[i][nb][q] = -v;
                }
    }
}

template <int NBW>
__device__ __forceinline__ void strip_store(const Frag<NBW>& acc, double* dst, long ld, int lane) {
    const int fr = lane >> 2, fk = lane & 3;
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int nb = 0; nb < NBW; ++nb)
            *reinterpret_cast<double2*>(dst + (long)(8 * i + fr) * ld + nb * 8 + 2 * fk) =
                make_double2(acc.v[i][nb][0], acc.v[i][nb][1]);
}

// acc = -src
template <int NBW>
__device__ __forceinline__ void strip_load_neg(Frag<NBW>& acc, const double* src, long ld, int lane) {
    const int fr = lane >> 2, fk = lane & 3;
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int nb = 0; nb < NBW; ++nb) {
            const double2 v = *reinterpret_cast<const double2*>(src + (long)(8 * i + fr) * ld + nb * 8 + 2 * fk);
            acc.v[i][nb][0] = -v.x;
            acc.v[i][nb][1] = -v.y;
        }
}

// Cholesky factor of the 64x64 block in D (overwritten by L_jj, lower block triangle) and its inverse in I
// (lower block triangle; the 8x8 diagonal blocks are written in full, zeros above the diagonal).  All threads
// of the CTA take part.  The block is processed as an 8x8 grid of 8x8 tiles, right-looking:
//   (A) warp 0 factors and inverts the diagonal tile in registers (one row per lane, shuffles only),
//   (B) the tiles below it are multiplied by that inverse, (C) the trailing tiles get their rank-8 update,
// (B) and (C) on the DMMA pipe with fragments read straight from shared memory; (D) builds the inverse of
// the whole block by block forward substitution, block columns dealt to the warps.
__device__ __forceinline__ void factor_invert_block(double* D, double* I, int* bad_pivot, int tid, int lane, int warp) {
    const int fr = lane >> 2, fk = lane & 3;
    for (int kk = 0; kk < 8; ++kk) {
        if (warp == 0) {
            // ---- (A) diagonal tile: lane r < 8 owns row r ----
            __syncwarp();
            double a[8], invd[8];
            const int r = lane & 7;
            const double* src = D + (8 * kk + r) * LDD + 8 * kk;
#pragma unroll
            for (int c = 0; c < 8; ++c) a[c] = src[c];
            bool bad = false;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                double dcc = __shfl_sync(0xffffffffu, a[c], c);
                if (!(dcc > 0.0)) {
                    bad = true;
                    dcc = 1.0;
                }
                const double inv = rsqrt(dcc);
                invd[c] = inv;
                const double l = (r == c) ? dcc * inv : a[c] * inv;
                a[c] = l;
#pragma unroll
                for (int cc = c + 1; cc < 8; ++cc) {
                    const double lcc = __shfl_sync(0xffffffffu, l, cc);
                    a[cc] = fma(-l, lcc, a[cc]);
                }
            }
            if (bad && lane == 0) *bad_pivot = 1;
            // inverse, column `r` per lane: x[q] = -(sum_{m<q} L[q][m] x[m]) / L[q][q], x[m] = 0 for m < r
            double x[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                double sum = 0.0;
#pragma unroll
                for (int m = 0; m < q; ++m) sum = fma(__shfl_sync(0xffffffffu, a[m], q), x[m], sum);
                x[q] = (q == r) ? invd[q] : (q > r ? -sum * invd[q] : 0.0);
            }
            if (lane < 8) {
                double* dl = D + (8 * kk + r) * LDD + 8 * kk;
#pragma unroll
                for (int c = 0; c < 8; ++c) dl[c] = c <= r ? a[c] : 0.0;
#pragma unroll
                for (int q = 0; q < 8; ++q) I[(8 * kk + q) * LDD + 8 * kk + r] = x[q];
            }
        }
        __syncthreads();
        // ---- (B) L[i,kk] = S[i,kk] * Linv_kk^T for the tiles below the diagonal ----
        for (int i = kk + 1 + warp; i < 8; i += NWARPS) {
            double c0 = 0.0, c1 = 0.0;
#pragma unroll
            for (int h = 0; h < 2; ++h)
                dmma884(c0, c1, D[(8 * i + fr) * LDD + 8 * kk + 4 * h + fk], I[(8 * kk + fr) * LDD + 8 * kk + 4 * h + fk]);
            __syncwarp();
            *reinterpret_cast<double2*>(D + (8 * i + fr) * LDD + 8 * kk + 2 * fk) = make_double2(c0, c1);
        }
        __syncthreads();
        // ---- (C) S[i,j] -= L[i,kk] L[j,kk]^T, kk < j <= i; warp 0 takes the next diagonal tile only ----
        {
            const int n = 7 - kk;                         // tiles (i,j) with kk < j <= i < 8, enumerated row by row
            int t = 0;
            for (int i = kk + 1; i < 8; ++i)
                for (int j = kk + 1; j <= i; ++j, ++t) {
                    const int owner = t == 0 ? 0 : 1 + (t - 1) % (NWARPS - 1);
                    if (owner != warp) continue;
                    double2 c = *reinterpret_cast<const double2*>(D + (8 * i + fr) * LDD + 8 * j + 2 * fk);
#pragma unroll
                    for (int h = 0; h < 2; ++h)
                        dmma884(c.x, c.y, -D[(8 * i + fr) * LDD + 8 * kk + 4 * h + fk], D[(8 * j + fr) * LDD + 8 * kk + 4 * h + fk]);
                    *reinterpret_cast<double2*>(D + (8 * i + fr) * LDD + 8 * j + 2 * fk) = c;
                }
            (void)n;
        }
        // no CTA barrier here: the next (A) touches only tile (kk+1,kk+1), which warp 0 itself just updated
    }
    __syncthreads();
    // ---- (D) inverse of the whole block: Linv[i,c] = -Linv_ii * sum_{p=c}^{i-1} L[i,p] Linv[p,c], warp c = block column ----
    for (int c = warp; c < 8; c += NWARPS) {

        for (int i = c + 1; i < 8; ++i) {
            double t0 = 0.0, t1 = 0.0;
            for (int p = c; p < i; ++p)
#pragma unroll
                for (int h = 0; h < 2; ++h)
                    dmma884(t0, t1, D[(8 * i + fr) * LDD + 8 * p + 4 * h + fk], I[(8 * p + 4 * h + fk) * LDD + 8 * c + fr]);
            double* blk = I + (8 * i) * LDD + 8 * c;
            *reinterpret_cast<double2*>(blk + fr * LDD + 2 * fk) = make_double2(t0, t1);
            __syncwarp();
            double r0 = 0.0, r1 = 0.0;
#pragma unroll
            for (int h = 0; h < 2; ++h)
                dmma884(r0, r1, -I[(8 * i + fr) * LDD + 8 * i + 4 * h + fk], blk[(4 * h + fk) * LDD + fr]);
            __syncwarp();
            *reinterpret_cast<double2*>(blk + fr * LDD + 2 * fk) = make_double2(r0, r1);
            __syncwarp();
        }
    }
    __syncthreads();
}

// 64x64 block (row stride ld, 16-byte aligned rows) -> shared memory with stride LDD, as one cp.async group.  The
// next strip_gemm drains it (its final wait_group 0 + barrier), so the copy overlaps that product; the block is
// only read afterwards (strip_times_tri / split_times_tri).
__device__ __forceinline__ void load_block_async(double* dst, const double* __restrict__ src, long ld, int tid) {
    for (int e = tid; e < NB * NB / 2; e += THREADS) {
        const int r = e / (NB / 2), c = (e % (NB / 2)) * 2;
        cp_async16(dst + r * LDD + c, src + (long)r * ld + c);
    }
    cp_async_commit();
}

__global__ void __launch_bounds__(THREADS, 3) factor_solve_kernel(SolveParams prm) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* pipe = reinterpret_cast<double*>(smem_raw);         // cp.async ring | diagonal block | strip staging
    double* Ism = pipe + PIPE_DOUBLES;                          // NB x LDD: L_jj^{-1}
    __shared__ int bad_pivot;

    __shared__ int ticket_sm;
    int slot = blockIdx.x;
    if (prm.n_core > 0) {                                       // merged launch: tickets in the order the CTAs start running
        if (threadIdx.x == 0) ticket_sm = atomicAdd(prm.ticket, 1);
        __syncthreads();
        slot = ticket_sm;
    }
    const bool is_core = prm.core_only || slot < prm.n_core;
    const int task_id = prm.n_core > 0 && !is_core ? slot - prm.n_core : slot;
    if (!is_core && task_id >= prm.n_tasks) return;
    const CentreTask task = prm.n_core > 0 && is_core ? prm.core_tasks[task_id] : prm.tasks[task_id];
    const int kp_stride = prm.n_core > 0 && is_core ? prm.core_kp_max : prm.kp_max;
    double* work = prm.n_core > 0 && is_core ? prm.core_work : prm.work;
    int* flags = prm.n_core > 0 && is_core ? prm.core_flags : prm.flags;
    const int kp = task.kp, nblk = kp / NB;
    const long ld = kp;
    double* M = work + (long)task_id * (kp_stride + EVAL_PAD) * kp_stride;   // rows 0..kp-1: L ; rows kp..: W then Zt
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    double* stage = pipe + warp * 8 * LDD;
    double* Dsm = pipe;
    constexpr int N_ESTRIPS = EVAL_PAD / 16;

    if (tid == 0) bad_pivot = 0;
    if (task.core_done && task.s_p > 0) {                      // merged launch: the group's core must be complete
        if (tid == 0) {
            int done = 0;
            while (!done) {
                asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(done) : "l"(task.core_done) : "memory");
                if (!done) __nanosleep(500);
            }
        }
        __syncthreads();
    }

    // Prefix sharing: the first jc panels of L (and of W) belong to the group's core task.
    const int s_p = task.s_p, jc = s_p / NB;
    const double* core = task.core_work;                        // rows 0..s_p-1: L11, rows s_p..: W1; stride s_p
    const long ldc = s_p;
    // The group's W1 = E L11^{-T} (rows s_p.. of the core block) is read where it is needed -- as the leading K-range of the
    // evaluation strips below and as the initial value of the backward solve -- and never copied into this centre's block.
    const double* coreW = core + (long)s_p * ldc;

    // ---------------- factorisation + forward solve, panel by panel ----------------
    for (int j = 0; j < nblk; ++j) {
        const int K = j * NB;
        if (j < jc) {
            // shared panel: L21[:, K..K+63] for this centre's own rows s_p..kp-1 only; the panel rows of L11 and the
            // inverse of its diagonal block come from the core task
            load_block_async(Ism, core + (long)K * ldc + K, ldc, tid);   // previous readers of Ism passed a barrier
            const double* panel_rows = core + (long)K * ldc;
            for (int u0 = 0; u0 < (kp - s_p) / 16; u0 += NWARPS) {
                const int mrow = s_p + 16 * (u0 + warp);
                Frag<8> acc;
                auto a_row = [&](int r) -> const double* { return M + (long)(s_p + 16 * u0 + r) * ld; };
                auto init = [&]() { strip_init<8>(acc, task, prm, false, mrow, K, lane); };
                strip_gemm<false, 8>(acc, a_row, panel_rows, ldc, K / KC, pipe, init, tid, lane, warp);
                strip_times_tri<false>(acc, Ism, stage, lane);
                strip_store<8>(acc, M + (long)mrow * ld + K, ld, lane);
            }
            __syncthreads();
            continue;
        }
        const int n_a = kp / 16 - 4 * j;                        // strips of A from the diagonal block down
        const int n_strips = n_a + N_ESTRIPS;                   // = 4q + 1
        const double* panel_rows = M + (long)K * ld;            // L[K..K+63][0..K)
        auto strip_row = [&](int u) { return u >= n_a ? kp + 16 * (u - n_a) : K + 16 * u; };   // first row in M

        auto nop0 = [&]() {};
        for (int u0 = 0; u0 + NWARPS <= n_strips; u0 += NWARPS) {
            const int u = u0 + warp;
            const bool is_eval = u >= n_a;
            const int mrow = strip_row(u);
            Frag<8> acc;
            auto init = [&]() { strip_init<8>(acc, task, prm, is_eval, is_eval ? mrow - kp : mrow, K, lane); };
            if (jc > 0) {
                // columns [0, s_p): the evaluation rows come from the group's W1, the rows of A from this centre's L21
                auto a_row1 = [&](int r) -> const double* {
                    const int us = u0 + (r >> 4);
                    return us >= n_a ? coreW + (long)(16 * (us - n_a) + (r & 15)) * ldc : M + (long)(strip_row(us) + (r & 15)) * ld;
                };
                strip_gemm<false, 8>(acc, a_row1, panel_rows, ld, s_p / KC, pipe, init, tid, lane, warp);
                if (K > s_p) {
                    auto a_row2 = [&](int r) -> const double* { return M + (long)(strip_row(u0 + (r >> 4)) + (r & 15)) * ld + s_p; };
                    strip_gemm<false, 8>(acc, a_row2, panel_rows + s_p, ld, (K - s_p) / KC, pipe, nop0, tid, lane, warp);
                }
            } else {
                auto a_row = [&](int r) -> const double* { return M + (long)(strip_row(u0 + (r >> 4)) + (r & 15)) * ld; };
                strip_gemm<false, 8>(acc, a_row, panel_rows, ld, K / KC, pipe, init, tid, lane, warp);
            }
            if (u0 == 0) {
                // the four strips of the diagonal block: factor it, keep its inverse
                const int fr = lane >> 2, fk = lane & 3;
#pragma unroll
                for (int i = 0; i < 2; ++i)
#pragma unroll
                    for (int nb = 0; nb < 8; ++nb)
                        *reinterpret_cast<double2*>(Dsm + (16 * warp + 8 * i + fr) * LDD + nb * 8 + 2 * fk) =
                            make_double2(-acc.v[i][nb][0], -acc.v[i][nb][1]);
                __syncthreads();
                factor_invert_block(Dsm, Ism, &bad_pivot, tid, lane, warp);
                for (int e = tid; e < NB * NB; e += THREADS) {
                    const int r = e / NB, c = e % NB;
                    M[(long)(K + r) * ld + K + c] = c <= (r | 7) ? Ism[r * LDD + c] : 0.0;   // diagonal blocks hold L_jj^{-1}
                }
            } else {
                strip_times_tri<false>(acc, Ism, stage, lane);
                strip_store<8>(acc, M + (long)mrow * ld + K, ld, lane);
            }
        }
        {   // split round: the last evaluation strip, 16 columns per warp
            const int mrow = strip_row(n_strips - 1);
            Frag<2> acc;
            auto init = [&]() { strip_init<2>(acc, task, prm, true, mrow - kp, K + 16 * warp, lane); };
            if (jc > 0) {
                auto a_row1 = [&](int r) -> const double* { return r < 16 ? coreW + (long)(mrow - kp + r) * ldc : nullptr; };
                strip_gemm<false, 2>(acc, a_row1, panel_rows, ld, s_p / KC, pipe, init, tid, lane, warp);
                if (K > s_p) {
                    auto a_row2 = [&](int r) -> const double* { return r < 16 ? M + (long)(mrow + r) * ld + s_p : nullptr; };
                    strip_gemm<false, 2>(acc, a_row2, panel_rows + s_p, ld, (K - s_p) / KC, pipe, nop0, tid, lane, warp);
                }
            } else {
                auto a_row = [&](int r) -> const double* { return r < 16 ? M + (long)(mrow + r) * ld : nullptr; };
                strip_gemm<false, 2>(acc, a_row, panel_rows, ld, K / KC, pipe, init, tid, lane, warp);
            }
            split_times_tri<false>(acc, Ism, pipe, lane, warp);
            strip_store<2>(acc, M + (long)mrow * ld + K + 16 * warp, ld, lane);
        }
        __syncthreads();                                         // panel j complete (global writes visible CTA-wide)
    }

    if (is_core) {
        if (tid == 0 && bad_pivot) flags[task.centre] = 1;         // the caller zeroes the flags
        if (prm.n_core > 0) {                                      // publish: every thread's writes, then the flag
            __threadfence();
            __syncthreads();
            if (tid == 0) asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(prm.core_done + task_id), "r"(1) : "memory");
        }
        return;
    }

    // ---------------- backward solve: Zt_i = (W_i - sum_{p>i} Zt_p L_pi) L_ii^{-1} ----------------
    // Rows of L below a shared panel (i < jc) come from two places: c0+64 .. s_p-1 from the core task, s_p .. kp-1 from
    // this centre's own rows; the product is accumulated in two pipeline runs.
    for (int i = nblk - 1; i >= 0; --i) {
        const int c0 = i * NB, Kb = kp - (c0 + NB);
        const bool shared = i < jc;
        const int Kb1 = shared ? s_p - (c0 + NB) : 0;            // k-extent served by the core task
        const double* lpanel = M + (long)(c0 + NB + Kb1) * ld + c0;            // own rows, row-major (k x n)
        const double* cpanel = core + (long)(c0 + NB) * ldc + c0;              // core rows (used when Kb1 > 0)
        const double* dsrc = shared ? core + (long)c0 * ldc + c0 : M + (long)c0 * ld + c0;
        const long dld = shared ? ldc : ld;
        load_block_async(Ism, dsrc, dld, tid);                   // previous readers of Ism passed a barrier
        auto nop = [&]() {};
        for (int u0 = 0; u0 + NWARPS <= N_ESTRIPS; u0 += NWARPS) {
            double* zrows = M + (long)(kp + 16 * (u0 + warp)) * ld;
            Frag<8> acc;
            auto init = [&]() {
                if (shared) strip_load_neg<8>(acc, coreW + (long)(16 * (u0 + warp)) * ldc + c0, ldc, lane);
                else strip_load_neg<8>(acc, zrows + c0, ld, lane);
            };
            if (Kb1 > 0) {
                auto a_row1 = [&](int r) -> const double* { return M + (long)(kp + 16 * u0 + r) * ld + c0 + NB; };
                strip_gemm<true, 8>(acc, a_row1, cpanel, ldc, Kb1 / KC, pipe, init, tid, lane, warp);
                auto a_row2 = [&](int r) -> const double* { return M + (long)(kp + 16 * u0 + r) * ld + c0 + NB + Kb1; };
                strip_gemm<true, 8>(acc, a_row2, lpanel, ld, (Kb - Kb1) / KC, pipe, nop, tid, lane, warp);
            } else {
                auto a_row = [&](int r) -> const double* { return M + (long)(kp + 16 * u0 + r) * ld + c0 + NB; };
                strip_gemm<true, 8>(acc, a_row, lpanel, ld, Kb / KC, pipe, init, tid, lane, warp);
            }
            strip_times_tri<true>(acc, Ism, stage, lane);
            strip_store<8>(acc, zrows + c0, ld, lane);
        }
        {
            double* zrows = M + (long)(kp + 16 * (N_ESTRIPS - 1)) * ld;
            Frag<2> acc;
            auto init = [&]() {
                if (shared) strip_load_neg<2>(acc, coreW + (long)(16 * (N_ESTRIPS - 1)) * ldc + c0 + 16 * warp, ldc, lane);
                else strip_load_neg<2>(acc, zrows + c0 + 16 * warp, ld, lane);
            };
            if (Kb1 > 0) {
                auto a_row1 = [&](int r) -> const double* { return r < 16 ? zrows + (long)r * ld + c0 + NB : nullptr; };
                strip_gemm<true, 2>(acc, a_row1, cpanel, ldc, Kb1 / KC, pipe, init, tid, lane, warp);
                auto a_row2 = [&](int r) -> const double* { return r < 16 ? zrows + (long)r * ld + c0 + NB + Kb1 : nullptr; };
                strip_gemm<true, 2>(acc, a_row2, lpanel, ld, (Kb - Kb1) / KC, pipe, nop, tid, lane, warp);
            } else {
                auto a_row = [&](int r) -> const double* { return r < 16 ? zrows + (long)r * ld + c0 + NB : nullptr; };
                strip_gemm<true, 2>(acc, a_row, lpanel, ld, Kb / KC, pipe, init, tid, lane, warp);
            }
            split_times_tri<true>(acc, Ism, pipe, lane, warp);
            strip_store<2>(acc, zrows + c0 + 16 * warp, ld, lane);
        }
        __syncthreads();                                         // Zt block visible before it is read as an operand
    }
    if (tid == 0 && task.core_flag && *task.core_flag) bad_pivot = 1;
    if (tid == 0 && bad_pivot) flags[task.centre] = 1;             // zeroed by the caller; sticky over the launches of a call
}

// One 64 x 64 diagonal block of the dense blocked Cholesky (dense.cu): same in-shared-memory factorisation as above.
__global__ void __launch_bounds__(THREADS) chol_diag_kernel(double* A, long lda, double* linv, double* linvT, int* flag) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* Dsm = reinterpret_cast<double*>(smem_raw);
    double* Ism = Dsm + NB * LDD;
    __shared__ int bad_pivot;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) bad_pivot = 0;
    for (int e = tid; e < NB * NB; e += THREADS) {
        const int r = e / NB, c = e % NB;
        Dsm[r * LDD + c] = A[(long)r * lda + c];
        Ism[r * LDD + c] = 0.0;
    }
    __syncthreads();
    factor_invert_block(Dsm, Ism, &bad_pivot, tid, lane, warp);
    for (int e = tid; e < NB * NB; e += THREADS) {
        const int r = e / NB, c = e % NB;
        A[(long)r * lda + c] = c <= r ? Dsm[r * LDD + c] : 0.0;
        const double v = c <= r ? Ism[r * LDD + c] : 0.0;
        linv[r * NB + c] = v;
        linvT[c * NB + r] = v;
    }
    if (tid == 0 && bad_pivot) *flag = 1;
}

}  // namespace

cudaError_t launch_chol_diag(double* A, long lda, double* linv, double* linvT, int* flag, cudaStream_t s) {
    const size_t smem = (size_t)2 * NB * LDD * sizeof(double);
    cudaError_t e = cudaFuncSetAttribute(chol_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    chol_diag_kernel<<<1, THREADS, smem, s>>>(A, lda, linv, linvT, flag);
    return cudaGetLastError();
}

size_t factor_solve_smem_bytes() { return (size_t)(PIPE_DOUBLES + NB * LDD) * sizeof(double); }

cudaError_t launch_factor_solve(const SolveParams& p, cudaStream_t s) {
    const size_t smem = factor_solve_smem_bytes();
    // function attributes are per device: set on every launch (several engines on different GPUs may share the process)
    cudaError_t e = cudaFuncSetAttribute(factor_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(factor_solve_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
    if (e != cudaSuccess) return e;
    if (p.kp_max % NB != 0) return cudaErrorInvalidValue;
    factor_solve_kernel<<<p.n_tasks + (p.n_core > 0 ? p.n_core : 0), THREADS, smem, s>>>(p);
    return cudaGetLastError();
}

}  // namespace mmd
