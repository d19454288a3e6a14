// Null-distribution scores of a searchlight centre (searchlight.py:61-79): the predictions stay fixed and
// the *targets* are permuted (data.py:540-561 for the test trials, a 3-derangement for the imagery
// trials), so every permuted score is a re-indexing of the centre's cosine-distance blocks that
// score.cu already computed:
//     d'(i, j) = d(i, perm(j)),   count = #{(i, j) : d(i, perm(i)) < d(i, perm(j))}.
// Six values per permutation, in the reference's dict order: captions, images (no cross-modal scores:
// comp_cross_decoding_scores=False), imagery, imagery_whole_test_set, and the two *_no_std imagery scores.
#include "common.cuh"
#include "kernels.h"

namespace mmd {

namespace {
constexpr int NULL_THREADS = 512;
constexpr int LDR = NULL_DIST_ROWS + 1;                      // row index runs fastest in the transposed copy
}  // namespace

// The centre's distance block is copied to shared memory transposed (Dt[col][row]): consecutive threads work on
// consecutive rows of one permutation, so the gathers d(row, perm(j)) of a warp hit consecutive addresses.
__global__ void __launch_bounds__(NULL_THREADS) null_score_kernel(NullParams prm) {
    extern __shared__ __align__(16) unsigned char null_smem[];
    double* Dt = reinterpret_cast<double*>(null_smem);       // [EVAL_PAD][LDR]
    int* cnt_sm = reinterpret_cast<int*>(Dt + EVAL_PAD * LDR);   // [n_perm][6]
    const int task_id = blockIdx.x;
    const CentreTask task = prm.tasks[task_id];
    const double* D = prm.dist + (long)task_id * NULL_DIST_ROWS * EVAL_PAD;
    for (int i = threadIdx.x; i < NULL_DIST_ROWS * EVAL_PAD; i += blockDim.x)
        Dt[(i % EVAL_PAD) * LDR + i / EVAL_PAD] = D[i];
    const int n_cap = prm.n_cap, n_test = prm.n_cap + prm.n_img, n_im = prm.n_im, n_eval = n_test + n_im;
    const int n_perm = prm.n_perm;
    for (int i = threadIdx.x; i < n_perm * 6; i += blockDim.x) cnt_sm[i] = 0;
    __syncthreads();

    const int rows_per_perm = n_test + 2 * n_im;             // test rows, standardised imagery rows, raw imagery rows
    for (int w = threadIdx.x; w < n_perm * rows_per_perm; w += blockDim.x) {
        const int p = w / rows_per_perm, rr = w % rows_per_perm;
        const int* pt = prm.perm_test + (long)p * n_test;
        const int* pi = prm.perm_imagery_centre ? prm.perm_imagery_centre + ((long)task.centre * n_perm + p) * n_im
                                                : prm.perm_imagery + (long)p * n_im;
        if (rr < n_test) {
            const bool cap = rr < n_cap;
            const double* row = Dt + rr;                     // row[x * LDR] = d(rr, x)
            const double diag = row[pt[rr] * LDR];
            const int j0 = cap ? 0 : n_cap, j1 = cap ? n_cap : n_test;
            int c = 0;
            for (int j = j0; j < j1; ++j) c += diag < row[pt[j] * LDR];
            if (c) atomicAdd(&cnt_sm[p * 6 + (cap ? 0 : 1)], c);
        } else {
            const int q = rr - n_test;
            const bool raw = q >= n_im;
            const int i = raw ? q - n_im : q;
            const double* row = Dt + (raw ? EVAL_PAD + i : n_test + i);
            const double diag = row[(n_test + pi[i]) * LDR];
            int c_im = 0, c_all = 0;
            for (int j = 0; j < n_im; ++j) c_im += diag < row[(n_test + pi[j]) * LDR];
            for (int j = 0; j < n_test; ++j) c_all += diag < row[pt[j] * LDR];
            c_all += c_im;
            if (c_im) atomicAdd(&cnt_sm[p * 6 + (raw ? 4 : 2)], c_im);
            if (c_all) atomicAdd(&cnt_sm[p * 6 + (raw ? 5 : 3)], c_all);
        }
    }
    __syncthreads();
    const long n_img = prm.n_img;
    const double den[6] = {(double)((long)n_cap * n_cap - n_cap), (double)(n_img * n_img - n_img),
                           (double)((long)n_im * n_im - n_im),    (double)((long)n_im * n_eval - n_im),
                           (double)((long)n_im * n_im - n_im),    (double)((long)n_im * n_eval - n_im)};
    double* out = prm.null_out + (long)task.centre * n_perm * 6;
    for (int i = threadIdx.x; i < n_perm * 6; i += blockDim.x) out[i] = den[i % 6] > 0.0 ? cnt_sm[i] / den[i % 6] : 0.0;
}

cudaError_t launch_null_score(const NullParams& p, cudaStream_t s) {
    const size_t smem = (size_t)EVAL_PAD * LDR * sizeof(double) + (size_t)p.n_perm * 6 * sizeof(int);
    if (smem > 227 * 1024) return cudaErrorInvalidValue;      // up to ~1,900 permutations per call
    cudaError_t e = cudaFuncSetAttribute(null_score_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    null_score_kernel<<<p.n_tasks, NULL_THREADS, smem, s>>>(p);
    return cudaGetLastError();
}

}  // namespace mmd
