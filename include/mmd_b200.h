/* C ABI of the B200 searchlight ridge-decoding engine (libmmd_b200.so).
 *
 * The reference (mitjanikolaus/multimodal_decoding) is pure Python and has no FFI layer; its
 * boundary for this path is a set of Python functions.  Each entry point below names the reference
 * function whose work it takes over; the Python mirror of those functions
 * (multimodal_decoding_b200/analyses/decoding/searchlight/searchlight.py) binds them with ctypes,
 * see INTEGRATION.md.  Paths are relative to the reference root.
 *
 * Conventions
 *   - plain pointers and sizes only; every data pointer may be a HOST pointer or a DEVICE pointer on
 *     the handle's GPU (detected with cudaPointerGetAttributes); host buffers are copied by the
 *     library, device buffers are borrowed for the duration of the call;
 *   - every function returns 0 on success; on failure a non-zero code is returned and
 *     mmd_last_error() describes it (CUDA error string or argument error);
 *   - a handle owns one CUDA stream and its workspaces; handles are independent, one handle must not
 *     be used from two threads at once;
 *   - there is no CPU fallback: without a usable CUDA device mmd_create fails.
 */
#ifndef MMD_B200_H
#define MMD_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mmd_handle mmd_handle;

#define MMD_N_SCORES 8          /* order: see MMD_SCORE_NAMES in mmd_score_name() */
#define MMD_EVAL_PAD 144        /* test + imagery trials per centre must be <= 143 */

/* Number of stage timers reported by mmd_get_stats (pack, xty, prep, syrk, factor_solve, predict, score, knn, null, reroute). */
#define MMD_N_STAGES 10

int mmd_create(int device, mmd_handle** out);
void mmd_destroy(mmd_handle* h);
const char* mmd_last_error(const mmd_handle* h);
const char* mmd_version(void);

/* The CUDA stream (cudaStream_t) every kernel of this handle is launched on, so that a caller can bracket calls
 * with its own CUDA events (bench.py does; the reference only has wall-clock timing, searchlight.py:250-257). */
void* mmd_stream(mmd_handle* h);

/* Guard mode (set MMD_B200_GUARD=1 in the environment before the library is loaded): every device buffer of the
 * library carries a 4 KiB band on either side and is pre-filled with 0xA5 bytes.  mmd_check_guards synchronises the
 * handle's stream and returns the number of buffers whose bands were overwritten since the last call (0 = clean,
 * always 0 with guard mode off; details on stderr and in mmd_last_error).  Test infrastructure for the GPU pool,
 * where compute-sanitizer is not available; the reference has no counterpart. */
int mmd_check_guards(mmd_handle* h);
/* Proves the detector: writes one byte behind and one byte in front of a scratch buffer and checks that both are
 * reported.  0 = detector works, -1 = guard mode is off, other = detector failed. */
int mmd_guard_selftest(mmd_handle* h);

/* Name of score column i (0..7), e.g. "pairwise_acc_captions"; eval.py:9-44 and searchlight.py:87. */
const char* mmd_score_name(int i);

/* Replaces the k-NN branch of get_adjacency_matrix (analyses/decoding/searchlight/searchlight.py:152-178,
 * i.e. sklearn NearestNeighbors().fit(coords).kneighbors(coords, k)).
 *   coords      (n_vertices x 3) float64
 *   queries     rows q_begin .. q_end-1 of coords
 *   idx_out     ((q_end-q_begin) x k) int64, ascending distance, the query itself first
 *   dist_out    ((q_end-q_begin) x k) float64 Euclidean distances */
int mmd_knn(mmd_handle* h, const double* coords, int n_vertices, int k, int q_begin, int q_end,
            int64_t* idx_out, double* dist_out);

/* Replaces the radius branch of get_adjacency_matrix (searchlight.py:164-171, sklearn
 * NearestNeighbors(radius=r).radius_neighbors_graph(coords) turned into one index array per vertex: ascending
 * vertex id, the vertex itself included, squared distance <= r*r).  Two passes over queries q_begin .. q_end-1:
 *   count pass  (indices_out == NULL): counts_out[q - q_begin] = number of neighbours
 *   fill pass   (indices_out != NULL): offsets = exclusive prefix sum of the counts ((q_end-q_begin) + 1 entries,
 *               offsets[0] = 0); neighbours of query q are written to indices_out[offsets[q-q_begin] ..)
 * coords (n_vertices x 3) float64 HOST or DEVICE; offsets, counts_out, indices_out HOST. */
int mmd_radius_neighbors(mmd_handle* h, const double* coords, int n_vertices, double radius, int q_begin, int q_end,
                         const int64_t* offsets, int32_t* counts_out, int32_t* indices_out);

/* Stages the arrays custom_search_light receives (searchlight.py:117-129; built at :217-248) on the GPU.
 *   X            (n_total x n_vertices) float16 row-major, trials x vertices
 *   Y            (n_total x n_dims)     float16 row-major, trials x embedding dims
 *   train_ids / test_ids / imagery_ids  row indices into X and Y (searchlight.py:229-232)
 *   test_is_caption[n_test]             1 where TEST_STIM_TYPES == "caption" (data.py:182-187), else 0
 * Builds the vertex-major panels and X_train^T Y_train once (used by every centre).  Returns once X and Y have been
 * read; the kernels stay in flight on the handle's stream.  NaN / infinity in X or in the training rows of Y is
 * recorded here and reported by the searchlight calls (below).
 * Ordering contract for DEVICE pointers: X / Y are read by kernels on the handle's own non-blocking stream (mmd_stream),
 * which is not ordered behind any other stream.  The caller must make sure the work that produced them has completed
 * (stream or event synchronisation) before calling; the Python Engine synchronises torch's current stream for tensors. */
int mmd_upload(mmd_handle* h, const void* X, int n_total, int n_vertices, const void* Y, int n_dims,
               const int32_t* train_ids, int n_train, const int32_t* test_ids, int n_test,
               const uint8_t* test_is_caption, const int32_t* imagery_ids, int n_imagery);

/* mmd_upload for a rank that scores only part of the surface (vertex-sharded searchlight, north_star (5); the reference's
 * joblib chunks of searchlight.py:130-148): `active_vertices` (n_active ascending vertex ids, HOST; NULL / 0 = all) lists
 * the vertices the rank's neighbourhoods use.  X^T Y rows and the digit planes of the integer prediction GEMM are built
 * for those vertices only (the per-upload work shrinks with the shard); a later searchlight call whose neighbourhoods
 * leave the set fails with a message naming mmd_upload_active. */
/* Host helper for the list above: mask[v] = 1 for every vertex id that occurs in `indices` (n entries; int64 when is_i64,
 * else int32), 0 elsewhere; returns 2 when an id is outside [0, n_vertices).  Runs on host threads. */
int mmd_mark_vertices(const void* indices, int is_i64, int64_t n, int n_vertices, uint8_t* mask);
int mmd_upload_active(mmd_handle* h, const void* X, int n_total, int n_vertices, const void* Y, int n_dims,
                      const int32_t* train_ids, int n_train, const int32_t* test_ids, int n_test,
                      const uint8_t* test_is_caption, const int32_t* imagery_ids, int n_imagery,
                      const int32_t* active_vertices, int n_active);

/* Replaces the loop of custom_group_iter_search_light + train_and_test (searchlight.py:38-114) for
 * `n_centres` centres with a fixed neighbourhood size k: ridge fit with `alpha`
 * (Ridge(alpha, fit_intercept=False), searchlight.py:240), prediction of the test and imagery trials and
 * the 8 pairwise-accuracy scores (eval.py:126-181).
 *   adjacency    (n_centres x k) int32 vertex ids (columns of X), HOST pointer
 *   scores_out   (n_centres x 8) float64, HOST or DEVICE
 *   flags_out    (n_centres) int32, 1 where the normal matrix was not positive definite; may be NULL
 * Fails (message starting "Input X contains NaN or infinity" / "Input y ...") when a listed neighbourhood uses a
 * vertex whose betas are not finite or the training latents are not finite: sklearn's Ridge.fit / predict raise
 * ValueError on such inputs (check_array; searchlight.py:56-59), and the Python mirror raises ValueError too. */
int mmd_searchlight(mmd_handle* h, const int32_t* adjacency, int n_centres, int k, double alpha,
                    double* scores_out, int32_t* flags_out);

/* mmd_searchlight for the int64 index matrix sklearn's kneighbors returns (searchlight.py:173), without a copy on
 * the caller's side. */
int mmd_searchlight_i64(mmd_handle* h, const int64_t* adjacency, int n_centres, int k, double alpha,
                        double* scores_out, int32_t* flags_out);

/* The same loop for ragged neighbourhoods, i.e. the radius branch of get_adjacency_matrix
 * (searchlight.py:164-171: one index array of its own length per centre).
 *   indices      concatenated vertex ids of all neighbourhoods, int32, HOST pointer
 *   offsets      (n_centres + 1) int64, HOST pointer: centre i owns indices[offsets[i] .. offsets[i+1]); every
 *                neighbourhood needs 1 .. min(2000, n_train) entries
 *   anchors      (n_centres) int32 or NULL: the vertex each centre sits on; only used to group nearby centres so that
 *                they share one Gram computation (NULL: the first entry of each list is used) */
int mmd_searchlight_ragged(mmd_handle* h, const int32_t* indices, const int64_t* offsets, const int32_t* anchors,
                           int n_centres, double alpha, double* scores_out, int32_t* flags_out);

/* mmd_searchlight plus the null-distribution branch of train_and_test (searchlight.py:61-79): for each of
 * `n_perm` (1..1700 per call; the reference's default is 100) target permutations the fixed predictions are re-scored
 * against permuted targets.
 *   perm_test     (n_perm x n_test) int32: entry [p][g] is the test trial (position in test_ids order) whose latent
 *                 takes the place of trial g's latent, i.e. `indices` of y_test[indices] (data.py:540-561)
 *   perm_imagery  (n_perm x n_imagery) int32: the same for the imagery latents (the reference draws one of the two
 *                 3-derangements per permutation, searchlight.py:65)
 *   null_out      (n_centres x n_perm x 6) float64, HOST or DEVICE; per permutation: captions, images, imagery,
 *                 imagery_whole_test_set, imagery_no_std, imagery_whole_test_set_no_std (the reference's dict order) */
#define MMD_N_NULL_SCORES 6
int mmd_searchlight_null(mmd_handle* h, const int32_t* adjacency, int n_centres, int k, double alpha,
                         const int32_t* perm_test, const int32_t* perm_imagery, int n_perm, double* scores_out,
                         int32_t* flags_out, double* null_out);
/* The same with one imagery permutation per CENTRE and permutation, as the reference draws them (searchlight.py:65 runs inside
 * the per-centre train_and_test): perm_imagery_centre is (n_centres x n_perm x n_imagery) int32, HOST, indexed by the centre's
 * row in `adjacency`; NULL = the shared table perm_imagery of mmd_searchlight_null. */
int mmd_searchlight_null_pc(mmd_handle* h, const int32_t* adjacency, int n_centres, int k, double alpha,
                            const int32_t* perm_test, const int32_t* perm_imagery, const int32_t* perm_imagery_centre, int n_perm,
                            double* scores_out, int32_t* flags_out, double* null_out);
/* ... and for ragged (radius-mode) neighbourhoods, arguments as mmd_searchlight_ragged (searchlight.py:61-79 with :164-171). */
int mmd_searchlight_null_ragged(mmd_handle* h, const int32_t* indices, const int64_t* offsets, const int32_t* anchors, int n_centres,
                                double alpha, const int32_t* perm_test, const int32_t* perm_imagery,
                                const int32_t* perm_imagery_centre, int n_perm, double* scores_out, int32_t* flags_out,
                                double* null_out);

/* Alpha grid x K-fold cross-validation per neighbourhood: what
 *   GridSearchCV(Ridge(fit_intercept=False), param_grid=dict(alpha=alphas), scoring=make_scorer(pairwise_accuracy), cv=n_folds)
 * (ridge_regression_decoding.py:111-117) computes in cv_results_["split<f>_test_score"] when it is fitted on the training
 * trials of one centre, X[train_ids][:, adjacency[i]] -> latents[train_ids] (BASELINE.json north_star (3)).
 *   alphas        (n_alphas) float64 > 0, HOST
 *   fold_bounds   (n_folds + 1) int32, HOST: fold f holds the training trials at positions fold_bounds[f] ..
 *                 fold_bounds[f+1]-1 of train_ids (KFold without shuffling, sklearn/model_selection/_split.py: the first
 *                 n_train % n_folds folds have one trial more)
 *   fold_scores_out (n_centres x n_alphas x n_folds) float64, HOST: pairwise_accuracy(latents_val, predictions_val)
 *                 (eval.py:116-123, cosine, no standardisation) of the ridge fitted without the fold
 * The caller averages over the folds, takes the first best alpha (GridSearchCV's rank rule) and refits with
 * mmd_searchlight; the Python mirror does exactly that (custom_search_light with a GridSearchCV estimator).
 * Per cluster of centres the Gram of every fold's trials is computed once and subtracted inside the per-centre gather
 * (A = G - G_f + alpha I); the validation trials run through the solver in chunks of 144 right-hand sides. */
int mmd_searchlight_cv(mmd_handle* h, const int32_t* adjacency, int n_centres, int k, const double* alphas, int n_alphas,
                       const int32_t* fold_bounds, int n_folds, double* fold_scores_out, int32_t* flags_out);

/* Whole-brain ridge decoder with an alpha grid and K-fold cross-validation: the computation behind
 *   GridSearchCV(Ridge(fit_intercept=False), param_grid=dict(alpha=...), scoring=make_scorer(pairwise_accuracy), cv=5, refit=True)
 *   .fit(train_fmri_betas, train_latents)                                       ridge_regression_decoding.py:111-123
 * on this library's own float64 kernels (csrc/dense.cu: DMMA GEMM, blocked Cholesky, triangular solves; no BLAS / LAPACK).
 *   mmd_ridge_load     X (n x v) and Y (n x d), HOST or DEVICE, float32 (is_f32 = 1) or float64, row-major.  Forms X X^T (n x n)
 *                      when v > n -- sklearn's kernel form of the cholesky solver -- else X^T X and X^T Y.
 *   mmd_ridge_cv       fold_scores_out (n_alphas x n_folds) = cv_results_["split<f>_test_score"][a]: ridge fitted without fold f
 *                      (contiguous folds, fold_bounds as in mmd_searchlight_cv), pairwise_accuracy(Y_val, X_val W) (eval.py:116-123)
 *   mmd_ridge_fit      refit on all trials; coef_out (d x v) float64 HOST = sklearn's coef_ (may be NULL); the weights stay on the
 *                      device for mmd_ridge_predict
 *   mmd_ridge_predict  out (m x d) float64 HOST = X_new W for X_new (m x v) */
int mmd_ridge_load(mmd_handle* h, const void* X, int x_is_f32, int n, int v, const void* Y, int y_is_f32, int d);
int mmd_ridge_cv(mmd_handle* h, const double* alphas, int n_alphas, const int32_t* fold_bounds, int n_folds, double* fold_scores_out);
int mmd_ridge_fit(mmd_handle* h, double alpha, double* coef_out);
int mmd_ridge_predict(mmd_handle* h, const void* X, int x_is_f32, int m, double* out);

/* eval.py primitives for arbitrary shapes, HOST float64 row-major buffers.
 *   mmd_cosine_distances          get_distance_matrix(predictions, originals, "cosine")   eval.py:103-104
 *   mmd_dist_mat_to_pairwise_acc  dist_mat_to_pairwise_acc(dist_mat)                      eval.py:107-113
 *   mmd_pairwise_accuracy         pairwise_accuracy(latents, predictions, "cosine",
 *                                 standardize_predictions, standardize_latents)          eval.py:116-123 */
int mmd_cosine_distances(mmd_handle* h, const double* predictions, int n_pred, const double* originals, int n_orig,
                         int n_dims, double* dist_out);
int mmd_dist_mat_to_pairwise_acc(mmd_handle* h, const double* dist_mat, int n_rows, int n_cols, double* acc_out);
int mmd_pairwise_accuracy(mmd_handle* h, const double* latents, int n_lat, const double* predictions, int n_pred,
                          int n_dims, int standardize_predictions, int standardize_latents, double* acc_out);

/* mmd_pairwise_accuracy for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  This entry point has no counterpart in the reference.
 * Same arithmetic and kernels as mmd_pairwise_accuracy, but
 *   - `latents` (n_lat x n_dims), `predictions` (n_pred x n_dims) and `acc_out` (one double) are DEVICE memory of GPU
 *     `device`, float64 row-major; nothing is staged, the inputs are not modified;
 *   - the scratch (standardised copies, row norms, distance matrix, counter) is a file-scope `static __device__` array of
 *     MMD_STATIC_SCRATCH_BYTES (4.5 MiB) inside the library: every GPU that loads the library holds it until the process
 *     exits.  Limits: n_lat, n_pred <= MMD_STATIC_MAX_ROWS, n_dims <= MMD_STATIC_MAX_DIMS; larger problems are rejected
 *     with an error (use mmd_pairwise_accuracy);
 *   - the call makes `device` current, calls no cudaMalloc / cudaFree / allocating library routine, and does NOT
 *     synchronise: the kernels are enqueued on `stream` (a cudaStream_t; NULL = the default stream) and `acc_out` is
 *     valid once that stream has passed them;
 *   - all calls on a GPU share the one array (it keeps the previous call's contents; what must start at zero is cleared
 *     on the stream): calls that overlap on the GPU -- different streams, threads or graph launches in flight -- race
 *     and must be serialised by the caller;
 *   - the first call on a device resolves the array's address (cudaGetSymbolAddress) and loads the kernels: make it
 *     outside any stream capture.
 * No handle: on failure a non-zero code is returned and, if `err` is not NULL, a message is written to err[0..err_len). */
#define MMD_STATIC_MAX_ROWS 256
#define MMD_STATIC_MAX_DIMS 1024
#define MMD_STATIC_SCRATCH_BYTES 4722944
int mmd_pairwise_accuracy_static(int device, void* stream, const double* latents, int n_lat, const double* predictions,
                                 int n_pred, int n_dims, int standardize_predictions, int standardize_latents,
                                 double* acc_out, char* err, int err_len);

/* Group-level statistics on the score maps the searchlight writes (HOST buffers).
 *   mmd_ttest_1samp     calc_image_t_values(data, popmean, sigma=sigma): per vertex the one-sample t-value of
 *                       `data[:, v] - popmean` over subjects, NaN where all subjects agree
 *                       (analyses/decoding/searchlight/searchlight_permutation_testing.py:213-258);
 *                       data (n_subjects x n_vertices) float64, t_out (n_vertices) float64; flags bit 0: the variance
 *                       floor is sigma * var ("relative", the reference's default) instead of sigma; bit 1: NaN where
 *                       all subjects agree (calc_t_value) -- without it the plain ttest_1samp_no_p quotient
 *   mmd_tfce            calc_tfce_values for one hemisphere and many maps (analyses/cluster_analysis.py:248-312, cluster
 *                       extent = number of vertices): values (n_maps x n_vertices) float64, edges (n_edges x 2) int32
 *                       vertex pairs of the mesh, dh <= 0 = "auto" (max / 100); tfce_out (n_maps x n_vertices) float32
 *   mmd_cluster_labels  calc_clusters at one threshold (analyses/cluster_analysis.py:15-75): labels_out[v] = smallest
 *                       vertex id of v's cluster, -1 for vertices outside clusters */
int mmd_ttest_1samp(mmd_handle* h, const double* data, int n_subjects, int n_vertices, double popmean, double sigma,
                    int flags, double* t_out);
int mmd_tfce(mmd_handle* h, const double* values, int n_maps, int n_vertices, const int32_t* edges, int n_edges,
             double h_exponent, double e_exponent, double dh, float* tfce_out);
int mmd_cluster_labels(mmd_handle* h, const double* values, int n_vertices, const int32_t* edges, int n_edges,
                       double threshold, int32_t* labels_out);

/* Test/diagnostic hook: predictions of one centre, (143 x n_dims) float64 in the internal row order
 * [caption test trials, image test trials, imagery trials]; HOST pointer. */
int mmd_predict_centre(mmd_handle* h, const int32_t* neighbours, int k, double alpha, double* yhat_out);

/* Cumulative GPU time per stage in milliseconds (CUDA events on the handle's stream) and the number of
 * kernel launches since the last mmd_reset_stats. */
int mmd_get_stats(mmd_handle* h, float stage_ms[MMD_N_STAGES], int64_t* n_launches, int64_t* n_clusters,
                  double* syrk_flops);
void mmd_reset_stats(mmd_handle* h);
const char* mmd_stage_name(int i);

/* More counters since the last mmd_reset_stats: [0] batches (every batch launches each per-centre kernel once),
 * [1] shared-core factorisations, [2] centres that used a shared core, [3] sum of their shared prefix lengths,
 * [4] host microseconds spent grouping centres (clusters, unions, shared cores), [5] host microseconds to hand the
 * adjacency to the device, [6] centres whose predictions were redone by the float64 kernels because a comparison sat
 * inside the error bound of the integer prediction GEMM (mmd_set_predict_mode), [7] searchlight calls that reused the
 * host plan of the previous call (same adjacency: clusters, unions and shared cores are kept, the adjacency is not copied
 * to the device again);
 * solve_flops = float64 flops executed by the factorisation / triangular-solve stage. */
#define MMD_N_COUNTERS 8
int mmd_get_counters(mmd_handle* h, int64_t counters[MMD_N_COUNTERS], double* solve_flops);

/* Prefix sharing: up to `group_size` neighbouring centres factor the part of their normal matrices that belongs to
 * the neighbours they all share only once (1 = off; default 8).  The scores do not depend on it. */
int mmd_set_group_size(mmd_handle* h, int group_size);

/* How estimator.predict (searchlight.py:58-59) is executed.
 *   1 (default)  tcgen05 tensor cores: both float64 operands are cut into six signed 7-bit digits, the digit-pair
 *                products run as kind::i8 GEMMs with exact int32 accumulation in tensor memory and are recombined in
 *                float64; the scoring kernel bounds the effect of the dropped digits on every comparison and hands the
 *                centres it cannot vouch for to the float64 kernels (counter [6] above), so the scores are the ones
 *                float64 arithmetic gives;
 *   0            float64 tensor pipe (DMMA) for every centre.
 * The environment variable MMD_B200_PREDICT=f64 selects 0 at mmd_create. */
int mmd_set_predict_mode(mmd_handle* h, int mode);
/* Cluster-union Gram: 1 (default) = exact integer arithmetic on tcgen05 (six 7-bit digits per float16 beta, int32 accumulators
 * in tensor memory; a cluster whose betas are not exact in six digits is redone by the float64 kernel), 0 = float64 DMMA kernel.
 * Environment override at mmd_create: MMD_B200_SYRK=f64. */
int mmd_set_syrk_mode(mmd_handle* h, int mode);

/* Diagnostics: after the next mmd_searchlight* call mmd_get_margins returns, per centre, the smallest absolute
 * difference between a prediction's distance to its own candidate and to any other candidate it is compared with
 * (eval.py:107-113) -- how far the centre is from a flipped comparison.  margins_out: (n_centres) float64 HOST. */
int mmd_enable_margins(mmd_handle* h, int on);
int mmd_get_margins(mmd_handle* h, double* margins_out, int n_centres);

/* Tuning knobs (0 keeps the default): centres per cluster, max centres per batch. */
int mmd_configure(mmd_handle* h, int cluster_size, int batch_centres, int host_threads);

#ifdef __cplusplus
}
#endif
#endif /* MMD_B200_H */
