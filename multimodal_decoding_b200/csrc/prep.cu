// Per-task index preparation: order a centre's neighbour list by vertex id and locate every
// neighbour inside its cluster's (sorted) union list.  With both lists ascending, the centre's
// k x k Gram is the sub-matrix G_U[loc, loc] and (row >= col) always lands in the stored lower triangle.
// The ridge solution does not depend on the order of the neighbourhood columns
// (searchlight.py:110 only fixes *which* columns enter the fit).
#include "common.cuh"
#include "kernels.h"

namespace mmd {

__global__ void __launch_bounds__(256) prep_tasks_kernel(const int* __restrict__ adjacency,
                                                         const long* __restrict__ task_row_off,
                                                         const int* __restrict__ task_k, int kp,
                                                         const long* __restrict__ task_union_off,
                                                         const int* __restrict__ task_union_n,
                                                         const int* __restrict__ union_vertices,
                                                         const long* __restrict__ task_core_off,
                                                         const int* __restrict__ task_core_n, int n_vertices,
                                                         const unsigned char* __restrict__ vertex_bad,
                                                         int* __restrict__ nbr_sorted, int* __restrict__ loc,
                                                         int* __restrict__ bad_index_flag) {
    extern __shared__ int sm[];                 // next power of two >= the largest k of the launch
    const int t = blockIdx.x, tid = threadIdx.x;
    const int k = task_k[t];
    int np2 = 1;
    while (np2 < k) np2 <<= 1;
    const int* row = adjacency + task_row_off[t];
    for (int i = tid; i < np2; i += blockDim.x) {
        int v = i < k ? row[i] : 0x7fffffff;
        if (i < k && (v < 0 || v >= n_vertices)) {
            atomicOr(bad_index_flag, 2);        // bit 1 = vertex id out of range
            v = 0;
        } else if (i < k && vertex_bad[v]) {
            // bit 2 = a neighbour's betas hold NaN or infinity; bit 3 = the vertex is outside the active set of mmd_upload_active
            atomicOr(bad_index_flag, vertex_bad[v] == 2 ? 8 : 4);
        }
        sm[i] = v;
    }
    __syncthreads();
    for (int size = 2; size <= np2; size <<= 1)
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int p = tid; p < np2 / 2; p += blockDim.x) {
                const int lo = 2 * p - (p & (stride - 1)), hi = lo + stride;
                const bool up = (lo & size) == 0;
                const int a = sm[lo], b = sm[hi];
                if ((a > b) == up) {
                    sm[lo] = b;
                    sm[hi] = a;
                }
            }
            __syncthreads();
        }
    const int* uni = union_vertices + task_union_off[t];
    const int un = task_union_n[t];
    const int n_core = task_core_n ? task_core_n[t] : 0;
    const int* core = n_core ? adjacency + task_core_off[t] : nullptr;
    for (int i = tid; i < kp; i += blockDim.x) nbr_sorted[(long)t * kp + i] = -1, loc[(long)t * kp + i] = -1;
    __syncthreads();
    for (int i = tid; i < k; i += blockDim.x) {
        int v = -1, pos = -1;
        int dst = i;
        {
            v = sm[i];
            if (n_core) {                        // [core | rest]: both lists ascending, core is a subset of the row
                int lo = 0, hi = n_core;
                while (lo < hi) {
                    const int mid = (lo + hi) >> 1;
                    if (core[mid] < v) lo = mid + 1;
                    else hi = mid;
                }
                dst = (lo < n_core && core[lo] == v) ? lo : n_core + (i - lo);
            }
            int lo = 0, hi = un;                 // lower_bound
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (uni[mid] < v) lo = mid + 1;
                else hi = mid;
            }
            pos = lo;
            if (lo >= un || uni[lo] != v) atomicOr(bad_index_flag, 1);    // bit 0 = not in the union
        }
        nbr_sorted[(long)t * kp + dst] = v;
        loc[(long)t * kp + dst] = pos;
    }
}

cudaError_t launch_prep_tasks(const int* adjacency, const long* task_row_off, const int* task_k, int k_max, int kp_max,
                              const long* task_union_off, const int* task_union_n, const int* union_vertices,
                              const long* task_core_off, const int* task_core_n,
                              int n_vertices, const unsigned char* vertex_bad, int n_tasks, int* nbr_sorted, int* loc,
                              int* bad_index_flag, cudaStream_t s) {
    int np2 = 1;
    while (np2 < k_max) np2 <<= 1;
    prep_tasks_kernel<<<n_tasks, 256, np2 * sizeof(int), s>>>(adjacency, task_row_off, task_k, kp_max, task_union_off,
                                                             task_union_n, union_vertices, task_core_off, task_core_n,
                                                             n_vertices, vertex_bad, nbr_sorted, loc, bad_index_flag);
    return cudaGetLastError();
}

// out[j][d] = Y[rows[j]][d] for j < n_rows, d < D; zero elsewhere (out is n_out x ld_out)
__global__ void gather_rows_f16_kernel(const __half* __restrict__ Y, long ldy, const int* __restrict__ rows,
                                       int n_rows, int D, __half* __restrict__ out, int n_out, int ld_out) {
    const int j = blockIdx.x;
    for (int d = threadIdx.x; d < ld_out; d += blockDim.x) {
        __half v = __float2half(0.f);
        if (j < n_rows && d < D) v = Y[(long)rows[j] * ldy + d];
        out[(long)j * ld_out + d] = v;
    }
}

}  // namespace mmd
