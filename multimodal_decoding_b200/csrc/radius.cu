// Radius neighbourhoods (searchlight.py:164-166, the `radius` branch of get_adjacency_matrix:
// sklearn NearestNeighbors(radius=r).fit(coords).radius_neighbors_graph(coords), then one index array
// per row, ascending vertex id, the vertex itself included).
//
// The reference densifies a V x V matrix (`.toarray()`), which only fits low mesh resolutions; here the
// same index arrays come out in CSR form from two brute-force passes in float64: a count pass, a host
// prefix sum, and a fill pass.  A point is a neighbour when its squared distance, evaluated as
// ((dx*dx + dy*dy) + dz*dz) without FMA contraction like sklearn's kd-tree `rdist`, is <= r*r
// (`_dist_to_rdist`).  One thread per query vertex, candidate points staged through shared memory in
// tiles (every thread of the CTA reads the same point: a shared-memory broadcast); a thread meets the
// candidates in ascending vertex order, so its output list is already sorted.
#include "common.cuh"
#include "kernels.h"

namespace mmd {
namespace {
constexpr int THREADS = 128;
constexpr int TILE = 1024;
}  // namespace

template <bool FILL>
__global__ void __launch_bounds__(THREADS) radius_kernel(const double* __restrict__ coords, int V, double r2,
                                                         int q_begin, int q_end, const long* __restrict__ offsets,
                                                         int* __restrict__ counts, int* __restrict__ indices) {
    __shared__ double pts[TILE * 3];
    const int q = q_begin + blockIdx.x * THREADS + threadIdx.x;
    const bool live = q < q_end;
    const double qx = live ? coords[3 * q] : 0.0, qy = live ? coords[3 * q + 1] : 0.0, qz = live ? coords[3 * q + 2] : 0.0;
    int* out = FILL && live ? indices + offsets[q - q_begin] : nullptr;
    int n = 0;
    for (int p0 = 0; p0 < V; p0 += TILE) {
        const int np = min(TILE, V - p0);
        __syncthreads();
        for (int i = threadIdx.x; i < np * 3; i += THREADS) pts[i] = coords[(long)p0 * 3 + i];
        __syncthreads();
        if (!live) continue;
#pragma unroll 4
        for (int p = 0; p < np; ++p) {
            const double dx = __dsub_rn(qx, pts[3 * p]), dy = __dsub_rn(qy, pts[3 * p + 1]), dz = __dsub_rn(qz, pts[3 * p + 2]);
            const double d2 = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
            if (d2 <= r2) {
                if (FILL) out[n] = p0 + p;
                ++n;
            }
        }
    }
    if (!FILL && live) counts[q - q_begin] = n;
}

cudaError_t launch_radius_count(const double* coords, int V, double radius, int q_begin, int q_end, int* counts,
                                cudaStream_t s) {
    if (q_end <= q_begin) return cudaSuccess;
    const int blocks = (q_end - q_begin + THREADS - 1) / THREADS;
    radius_kernel<false><<<blocks, THREADS, 0, s>>>(coords, V, radius * radius, q_begin, q_end, nullptr, counts, nullptr);
    return cudaGetLastError();
}

cudaError_t launch_radius_fill(const double* coords, int V, double radius, int q_begin, int q_end, const long* offsets,
                               int* indices, cudaStream_t s) {
    if (q_end <= q_begin) return cudaSuccess;
    const int blocks = (q_end - q_begin + THREADS - 1) / THREADS;
    radius_kernel<true><<<blocks, THREADS, 0, s>>>(coords, V, radius * radius, q_begin, q_end, offsets, nullptr, indices);
    return cudaGetLastError();
}

}  // namespace mmd
