// Internal launch interface between the C-ABI (api.cu) and the kernels.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace mmd {

// ---- pack.cu -----------------------------------------------------------------------------
void launch_pack_transpose(const __half* X, long ldx, const int* ids_dev, int n_ids, int V, __half* out,
                           int ld_out, unsigned char* bad_row, int* any_bad, cudaStream_t s);
__global__ void row_sqnorm_kernel(const __half* Y, long ldy, const int* rows, int n_rows, int D, double* sq);
__global__ void half_to_double_kernel(const __half* in, double* out, long n);

// ---- gemm_gather.cu ----------------------------------------------------------------------
struct GemmProblem {
    const __half* A;
    long lda;
    const int* rowsA;   // optional gather list (device), nullptr = identity
    int M;
    const __half* B;
    long ldb;
    const int* rowsB;
    int N;
    int K;              // multiple of 64; rows are zero padded up to it
    double* C;
    long ldc;           // even
    int lower_only;     // skip tiles strictly above the diagonal (symmetric rank-k update)
    int scatter_c;      // 1: row i of the product is written to C row rowsA[i] (a row subset of a full-size result)
    const int* only_if; // optional (device): the problem is skipped unless *only_if != 0 (float64 redo of a cluster whose
                        // betas were not exact in the integer SYRK's six digits, syrk_i8.cu)
};
size_t gemm_gather_smem_bytes();
cudaError_t launch_gemm_gather(const GemmProblem* probs_dev, int n_problems, int max_m, int max_n, cudaStream_t s);

// ---- syrk_i8.cu ---------------------------------------------------------------------------
// exact integer cluster Gram on tcgen05 (see the file header); planes: per cluster six digit planes of `up` rows x ldt bytes
struct SyrkCluster {
    long row_base;        // first row of the cluster's plane 0 in the batch's plane tensor
    int u, up;            // union size, rounded up to 128 (row stride between the planes)
    const int* rows;      // the union's vertex ids (device)
    double* scales;       // (up) power-of-two scale of every union row, written by the slicing kernel
    int* inexact;         // set when a row is not exact in six digits: the float64 kernel redoes this cluster
    double* C;            // the cluster's Gram (u x ldc, lower triangle written)
    long ldc;
    // rectangular product A B^T (the global X^T Y of mmd_upload) instead of a symmetric one: u_b > 0 columns with their own
    // scales; all of them are written; out_rows (optional) scatters output row r to C row out_rows[r]
    int u_b;
    const double* scales_b;
    const int* out_rows;
};
struct SyrkItem {
    int row_a, row_b;     // plane-0 tensor rows of the tile's rows (m0) and columns (n0)
    int up, cluster;
    int m0, n0;           // symmetric product: n0 <= m0 (lower-triangle tiles)
    int up_b;             // plane stride of operand B (= up for the symmetric product)
};
size_t syrk_i8_smem_bytes();
// rows of `Xt` (row pitch ldt halfs) listed by each cluster's `rows` (nullptr: rows 0..u-1) -> six digit planes and scales
cudaError_t launch_slice_union(const __half* Xt, int ldt, const SyrkCluster* clusters_dev, int n_clusters, int max_u, int8_t* planes,
                               cudaStream_t s);
cudaError_t launch_syrk_i8(const int8_t* planes, long plane_rows, int ldt, const SyrkItem* items_dev, int n_items,
                           const SyrkCluster* clusters_dev, int n_sms, int* status, cudaStream_t s);

// ---- factor_solve.cu ---------------------------------------------------------------------
// One searchlight centre = one CTA.  All pointers are device pointers.
struct CentreTask {
    const double* gram;   // the cluster's union Gram (lower triangle valid), row-major
    long gram_ld;
    const int* loc;       // k ascending positions of this centre's neighbours inside the union
    const int* nbr;       // k vertex ids in the same (ascending) order
    int centre;           // output slot (index into scores / flags)
    int k;                // this centre's neighbour count (ragged radius-mode neighbourhoods: searchlight.py:164-171)
    int kp;               // k rounded up to a multiple of 64: order of this centre's padded normal matrix
    // Prefix sharing: the first s_p entries of loc / nbr (a multiple of 64, 0 = none) are common to a group of
    // nearby centres; the leading s_p columns of L and of W = E L^{-T} were computed once for the group by a
    // "core" task (k = kp = s_p) and are read from core_work (rows 0..s_p-1: L11 with inverted diagonal blocks,
    // rows s_p..s_p+143: W1; row stride s_p).
    int s_p;
    const double* core_work;
    const int* core_flag;     // non-positive pivot met while factoring the shared core
    const int* core_done;     // merged launch: set (release) by the group's core task when core_work is complete
};
struct SolveParams {
    const CentreTask* tasks;
    int n_tasks;
    int kp_max;           // largest task.kp of the launch; the per-task stride of `work` is (kp_max + EVAL_PAD) * kp_max
    double alpha;
    const __half* Xe;     // (V x EVAL_PAD) evaluation trials per vertex
    double* work;         // per task: (task.kp + EVAL_PAD) x task.kp doubles at the start of its stride
    int* flags;           // per task slot: 1 if a non-positive pivot was met
    int core_only;        // 1: tasks are shared cores: factorisation + forward solve only (no backward solve)
    // Merged launch (n_core > 0): ONE grid runs the groups' shared cores and the centres.  A CTA draws a ticket; tickets
    // 0..n_core-1 are core tasks (core_tasks[t], written to core_work / core_flags, stride (core_kp_max + EVAL_PAD) *
    // core_kp_max), the rest are the centres.  The first CTAs to run draw the lowest tickets, so a core never waits for a
    // slot behind a centre that spins on it; a centre waits for its group's core_done flag.  The centre work fills the
    // half-empty second wave a separate core launch leaves.
    int n_core;
    const CentreTask* core_tasks;
    double* core_work;
    int core_kp_max;
    int* core_flags;
    int* core_done;       // (n_core) zeroed by the caller
    int* ticket;          // one int, zeroed by the caller
    long gram_sub_off;    // 0, or the offset (in doubles) from task.gram to a second Gram of the same layout that is subtracted:
                          // A = G[loc,loc] - G_f[loc,loc] + alpha I, the normal matrix without the trials of fold f (cv.cu)
};
size_t factor_solve_smem_bytes();
cudaError_t launch_factor_solve(const SolveParams& p, cudaStream_t s);

// ---- predict.cu --------------------------------------------------------------------------
struct PredictParams {
    const CentreTask* tasks;
    int n_tasks;
    int kp_max;           // stride of `work` as in SolveParams
    const double* work;   // as above; Zt = rows task.kp.. of each task's block
    const double* XtY;    // (V x Dp)
    int Dp;               // feature count rounded up to 64
    double* yhat;         // per task: YHAT_ROWS x Dp (layout below)
    int n_cap, n_img, n_im;
    int standardize;      // 1: rows are z-scored per prediction group (the scoring input); 0: raw predictions
    // optional indirection: blockIdx.y = b works on task task_list[b] if b < *n_list and leaves otherwise
    const int* task_list;
    const int* n_list;
};
// Rows of a task's yhat block, written by predict and read by score:
//   0 .. n_eval-1            predictions of [caption test | image test | imagery] trials; with `standardize` each
//                            group is z-scored per feature (StandardScaler semantics: ddof 0, scale 1 if the
//                            variance is 0; eval.py:118 through calc_all_pairwise_accuracy_scores)
//   n_eval .. EVAL_PAD-1     zero
//   EVAL_PAD .. +n_im-1      the un-standardised imagery predictions (the *_no_std scores, searchlight.py:84-87)
//   .. YHAT_ROWS-1           zero
constexpr int YHAT_ROWS = 152;
size_t predict_smem_bytes();
cudaError_t launch_predict(const PredictParams& p, cudaStream_t s);

// ---- predict_i8.cu -----------------------------------------------------------------------
// The same product on tcgen05 tensor cores: both operands cut into I8_DIGITS signed 7-bit digits behind power-of-two
// scales, digit-pair GEMMs (kind::i8, int32 accumulators in tensor memory), float64 recombination (see the file header).
constexpr int I8_DIGITS = 6;
struct PredictI8Params {
    const CentreTask* tasks;
    int n_tasks;
    int kp_max;           // row pitch of the Zt digit planes (bytes) = largest task.kp of the launch, a multiple of 64
    int V;                // rows per XtY digit plane
    int Dp;               // feature count rounded up to 128
    const double* sX;     // (Dp) power-of-two scale of every XtY column
    const double* sZ;     // (n_tasks x EVAL_PAD) power-of-two scale of every Zt row
    double* yhat;         // as PredictParams::yhat
    double* amp2;         // (n_tasks x 4), zeroed by the caller: sum over the features of (sX / sd_g)^2 for the three
                          // prediction groups and of sX^2 -- the score kernel's tolerance input
    int n_cap, n_img, n_im;
    int standardize;
    int* status;          // set non-zero when a bounded wait ran out (protocol error; 0 in every correct run)
};
size_t predict_i8_smem_bytes();
// planes: (I8_DIGITS x V x Dp) int8, sX: (Dp), colmax_ws: (Dp) scratch
// rows (optional, device): only these n_rows rows of XtY exist (mmd_upload_active); maxima and planes cover them alone
cudaError_t launch_slice_xty(const double* XtY, int V, int Dp, int8_t* planes, double* sX, unsigned long long* colmax_ws,
                             const int* rows, int n_rows, cudaStream_t s);
// zplanes: (n_tasks x I8_DIGITS x EVAL_PAD x kp_max) int8, sZ: (n_tasks x EVAL_PAD); Zt is read from `work` as in PredictParams
cudaError_t launch_slice_zt(const CentreTask* tasks, int n_tasks, int kp_max, const double* work, int8_t* zplanes, double* sZ,
                            cudaStream_t s);
cudaError_t launch_predict_i8(const PredictI8Params& p, const int8_t* xplanes, const int8_t* zplanes, int n_sms, cudaStream_t s);
cudaError_t make_tensor_map_u8(void* out_map, const void* base, uint64_t cols, uint64_t rows, uint64_t pitch, uint32_t box_cols,
                               uint32_t box_rows, int swizzle_bytes);

// ---- score.cu ----------------------------------------------------------------------------
struct ScoreParams {
    const CentreTask* tasks;
    int n_tasks;
    const double* yhat;        // per task YHAT_ROWS x Dp, standardised by predict (layout at PredictParams)
    int Dp;
    const double* cand;        // (EVAL_PAD x Dp) candidate latents as float64, rows ordered like yhat rows
    const double* cand_sq;     // squared norms of the candidate rows
    int n_cap, n_img, n_im;
    double* scores;            // (n_centres x 8), written at row task.centre
    double* dist_out;          // optional: per task NULL_DIST_ROWS x EVAL_PAD cosine distances (null-distribution input)
    // optional indirection: CTA b works on task task_list[b] if b < *n_list and leaves otherwise (device pointers)
    const int* task_list;
    const int* n_list;
    // optional guard for predictions that came from predict_i8.cu (all nullptr / 0 for float64 predictions):
    const double* amp2;        // PredictI8Params::amp2
    const double* sZ;          // PredictI8Params::sZ
    double eps_unit;           // per-term truncation bound of the integer GEMM in scale units (7.2 * 2^-42)
    int* reroute_list;         // tasks with a comparison inside its tolerance are appended here ...
    int* reroute_count;        // ... at position atomicAdd(reroute_count, 1)
    double* margin_out;        // optional (n_centres): smallest |d_own - d_other| over the counted comparisons
};
// rows 0..n_eval-1: standardised predictions (captions, images, imagery); rows EVAL_PAD..: raw imagery predictions
constexpr int NULL_DIST_ROWS = 160;
size_t score_smem_bytes(int Dp);
cudaError_t launch_score(const ScoreParams& p, cudaStream_t s);

// ---- null_score.cu ----------------------------------------------------------------------
struct NullParams {
    const CentreTask* tasks;
    int n_tasks;
    const double* dist;        // as written by the score kernel
    const int* perm_test;      // (n_perm x n_test) candidate column playing the role of test row g's own latent
    const int* perm_imagery;   // (n_perm x n_im)   imagery candidate (0..n_im-1) playing the role of imagery row i's latent
    const int* perm_imagery_centre;   // optional (n_centres x n_perm x n_im), indexed by task.centre: one draw per centre and
                                      // permutation like searchlight.py:65; overrides perm_imagery
    int n_perm, n_cap, n_img, n_im;
    double* null_out;          // (n_centres x n_perm x 6), written at row task.centre
};
cudaError_t launch_null_score(const NullParams& p, cudaStream_t s);

// ---- pairwise.cu ------------------------------------------------------------------------
cudaError_t launch_standardize_columns(double* A, int n, int D, long ld, cudaStream_t s);
cudaError_t launch_cosine_dist(const double* P, int np, const double* C, int nc, int D, long ld, double* psq, double* csq,
                               double* dist, cudaStream_t s);
cudaError_t launch_pairwise_count(const double* dist, int r, int c, unsigned long long* count, cudaStream_t s);

// ---- cv.cu ------------------------------------------------------------------------------
struct CvScoreParams {
    const CentreTask* tasks;
    int n_tasks;
    const double* yhat;        // per task YHAT_ROWS x Dp raw predictions; rows 0..n_rows-1 = the chunk's validation trials
    int Dp;
    const double* cand;        // (n_cand rounded up to EVAL_PAD) x Dp: the fold's latents as float64, zero rows below
    const double* cand_sq;     // their squared norms
    int n_cand;                // validation trials of the fold
    int row0;                  // first validation trial of the chunk (multiple of EVAL_PAD)
    int n_rows;                // trials of the chunk (<= EVAL_PAD)
    unsigned long long* counts;   // counts[task.centre * count_stride + count_off] += #{(i, j): d[i][own(i)] < d[i][j]}
    long count_stride;
    int count_off;
};
size_t cv_score_smem_bytes();
cudaError_t launch_cv_score(const CvScoreParams& p, cudaStream_t s);
cudaError_t launch_copy_cols_f16(const __half* src, long ld_src, int col0, int n_cols, __half* dst, int ld_dst, long n_rows,
                                 cudaStream_t s);
cudaError_t launch_clear_bytes(unsigned char* flags, const int* idx, int n, cudaStream_t s);
cudaError_t launch_rsub(const double* a, double* inout, long n, cudaStream_t s);
cudaError_t launch_fold_latents(const __half* Yt, long ldy, int col0, int n_rows, int n_rows_pad, int Dp, double* cand, double* sq,
                                cudaStream_t s);

// ---- dense.cu (+ chol_diag_kernel in factor_solve.cu) ------------------------------------------------------------
// float64 NT product on the DMMA pipe: C = A B^T (mode 0) or C -= A B^T (mode 1); K a multiple of 16, even leading dimensions
struct DenseGemm {
    const double* A;
    long lda;
    const double* B;
    long ldb;
    double* C;
    long ldc;
    int M, N, K;
    int mode;
    int lower_only;     // skip tiles strictly above the diagonal
};
cudaError_t launch_dense_gemm(const DenseGemm& p, cudaStream_t s);
cudaError_t launch_gather_block(const double* in, long ld_in, const int* row_idx, int n_rows, const int* col_idx, int n_cols,
                                double* out, int rows_out, int ld_out, double diag_add, double pad_diag, cudaStream_t s);
cudaError_t launch_transpose_f64(const double* in, long ld_in, int n_rows_in, int n_cols_in, double* out, long ld_out, int rows_out,
                                 cudaStream_t s);
cudaError_t launch_widen_pad(const void* in, int is_f32, long ld_in, int n_rows, int n_cols, double* out, int rows_out, long ld_out,
                             cudaStream_t s);
// Cholesky factor of one 64 x 64 block (lower triangle of A read, L written back with zeros above the diagonal), its inverse
// and the inverse's transpose (64 x 64 row-major each); *flag is set on a non-positive pivot
cudaError_t launch_chol_diag(double* A, long lda, double* linv, double* linvT, int* flag, cudaStream_t s);
cudaError_t dense_potrf(double* A, int np, double* linv, double* linvT, int* flag, cudaStream_t s);
cudaError_t dense_potrs_t(const double* L, const double* Lt, const double* linv, const double* linvT, int np, double* Bt, int d,
                          cudaStream_t s);

// ---- prep.cu -----------------------------------------------------------------------------
// task t reads its neighbour list adjacency[task_row_off[t] .. + task_k[t]) and writes kp_max-strided rows
// task_core_off / task_core_n (optional): the task's shared core, an ascending list inside `adjacency`
// (task_core_n[t] entries at adjacency[task_core_off[t]..], a subset of the task's own list); the output is
// ordered [core | remaining neighbours], both parts ascending.
cudaError_t launch_prep_tasks(const int* adjacency, const long* task_row_off, const int* task_k, int k_max, int kp_max,
                              const long* task_union_off, const int* task_union_n, const int* union_vertices,
                              const long* task_core_off, const int* task_core_n,
                              int n_vertices, const unsigned char* vertex_bad, int n_tasks, int* nbr_sorted, int* loc,
                              int* bad_index_flag, cudaStream_t s);
__global__ void gather_rows_f16_kernel(const __half* Y, long ldy, const int* rows, int n_rows, int D, __half* out,
                                       int n_out, int ld_out);

// ---- knn.cu ------------------------------------------------------------------------------
// coords: (V x 3) float64 device; queries q_begin..q_end-1 are rows of coords; outputs are indexed from q_begin.
cudaError_t launch_knn(const double* coords, int V, int k, int q_begin, int q_end, int64_t* idx_out,
                       double* dist_out, int* overflow_flag, cudaStream_t s);

// ---- stats.cu ----------------------------------------------------------------------------
// data: (n_subjects x V) float64; t_out: (V).  relative: the "hat" limit is sigma * var (per-vertex call), else sigma.
cudaError_t launch_ttest(const double* data, int n_subjects, int V, double popmean, double sigma, int relative,
                         int constant_is_nan, double* t_out, cudaStream_t s);
// values: (n_maps x V) float64; edges: (E x 2) int32; label_ws / count_ws: (n_maps x V) int32 workspaces; out: (n_maps x V)
cudaError_t launch_tfce(const double* values, int n_maps, int V, const int* edges, int E, double h, double e, double dh,
                        int* label_ws, int* count_ws, float* out, cudaStream_t s);
cudaError_t launch_cluster_labels(const double* values, int V, const int* edges, int E, double threshold, int* label_ws,
                                  int* count_ws, int* labels_out, cudaStream_t s);

// ---- radius.cu ---------------------------------------------------------------------------
cudaError_t launch_radius_count(const double* coords, int V, double radius, int q_begin, int q_end, int* counts,
                                cudaStream_t s);
cudaError_t launch_radius_fill(const double* coords, int V, double radius, int q_begin, int q_end, const long* offsets,
                               int* indices, cudaStream_t s);

}  // namespace mmd
