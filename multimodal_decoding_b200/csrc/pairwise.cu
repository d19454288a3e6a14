// Generic scoring primitives behind the reference's eval.py API (eval.py:103-123), float64:
//   column standardisation (sklearn StandardScaler: population std, zero-variance columns keep scale 1),
//   cosine distance matrix (scipy cdist(..., "cosine"): 1 - u.v / (|u| |v|), rows = predictions),
//   strict-'<' pairwise-accuracy count against each row's diagonal entry (dist_mat_to_pairwise_acc).
// The per-centre searchlight scores do not go through here (score.cu fuses all of it per centre); these
// kernels serve arbitrary shapes: the whole-brain grid search scores 800 x 800 folds with them.
#include "common.cuh"
#include "kernels.h"

namespace mmd {

// one thread per column; rows are strided by ld (coalesced across the threads of a warp)
__global__ void __launch_bounds__(128) standardize_columns_kernel(double* __restrict__ A, int n, int D, long ld) {
    const int d = blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= D || n <= 0) return;
    double s = 0.0;
    for (int r = 0; r < n; ++r) s += A[(long)r * ld + d];
    const double mu = s / n;
    double v = 0.0;
    for (int r = 0; r < n; ++r) {
        const double t = A[(long)r * ld + d] - mu;
        v += t * t;
    }
    const double sd = sqrt(v / n);
    const double inv = sd == 0.0 ? 1.0 : 1.0 / sd;
    for (int r = 0; r < n; ++r) A[(long)r * ld + d] = (A[(long)r * ld + d] - mu) * inv;
}

// one warp per row
__global__ void __launch_bounds__(256) row_sqnorm_f64_kernel(const double* __restrict__ A, int n, int D, long ld,
                                                             double* __restrict__ out) {
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= n) return;
    double s = 0.0;
    for (int d = lane; d < D; d += 32) {
        const double x = A[(long)row * ld + d];
        s = fma(x, x, s);
    }
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) out[row] = s;
}

// dist[i][j] = 1 - P_i . C_j / (|P_i| |C_j|); 64 x 64 output tile per CTA, 4 x 4 per thread, DFMA
__global__ void __launch_bounds__(256) cosine_dist_kernel(const double* __restrict__ P, int np, const double* __restrict__ C,
                                                          int nc, int D, long ld, const double* __restrict__ psq,
                                                          const double* __restrict__ csq, double* __restrict__ dist) {
    __shared__ double ps[16][65], cs[16][65];
    const int i0 = blockIdx.y * 64, j0 = blockIdx.x * 64;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    double acc[4][4] = {};
    for (int d0 = 0; d0 < D; d0 += 16) {
        for (int e = threadIdx.x; e < 64 * 16; e += 256) {
            const int r = e >> 4, kk = e & 15;
            ps[kk][r] = (i0 + r < np && d0 + kk < D) ? P[(long)(i0 + r) * ld + d0 + kk] : 0.0;
            cs[kk][r] = (j0 + r < nc && d0 + kk < D) ? C[(long)(j0 + r) * ld + d0 + kk] : 0.0;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < 16; ++kk) {
            double a[4], b[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                a[u] = ps[kk][ty * 4 + u];
                b[u] = cs[kk][tx * 4 + u];
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int v = 0; v < 4; ++v) acc[u][v] = fma(a[u], b[v], acc[u][v]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int v = 0; v < 4; ++v) {
            const int i = i0 + ty * 4 + u, j = j0 + tx * 4 + v;
            if (i < np && j < nc) dist[(long)i * nc + j] = 1.0 - acc[u][v] / (sqrt(psq[i]) * sqrt(csq[j]));
        }
}

// count[0] += #{(i, j): dist[i][i] < dist[i][j]}, rows i < min(r, c) carry a diagonal entry (eval.py:107-113)
__global__ void __launch_bounds__(256) pairwise_count_kernel(const double* __restrict__ dist, int r, int c,
                                                             unsigned long long* __restrict__ count) {
    const int i = blockIdx.x;
    if (i >= r) return;
    const double diag = dist[(long)i * c + i];
    int local = 0;
    for (int j = threadIdx.x; j < c; j += blockDim.x) local += diag < dist[(long)i * c + j];
    for (int o = 16; o; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(count, (unsigned long long)local);
}

cudaError_t launch_standardize_columns(double* A, int n, int D, long ld, cudaStream_t s) {
    standardize_columns_kernel<<<(D + 127) / 128, 128, 0, s>>>(A, n, D, ld);
    return cudaGetLastError();
}
cudaError_t launch_cosine_dist(const double* P, int np, const double* C, int nc, int D, long ld, double* psq, double* csq,
                               double* dist, cudaStream_t s) {
    row_sqnorm_f64_kernel<<<(np + 7) / 8, 256, 0, s>>>(P, np, D, ld, psq);
    row_sqnorm_f64_kernel<<<(nc + 7) / 8, 256, 0, s>>>(C, nc, D, ld, csq);
    dim3 grid((nc + 63) / 64, (np + 63) / 64);
    cosine_dist_kernel<<<grid, 256, 0, s>>>(P, np, C, nc, D, ld, psq, csq, dist);
    return cudaGetLastError();
}
cudaError_t launch_pairwise_count(const double* dist, int r, int c, unsigned long long* count, cudaStream_t s) {
    cudaError_t e = cudaMemsetAsync(count, 0, sizeof(unsigned long long), s);
    if (e != cudaSuccess) return e;
    const int rows = r < c ? r : c;
    if (rows > 0) pairwise_count_kernel<<<rows, 256, 0, s>>>(dist, rows, c, count);
    return cudaGetLastError();
}

}  // namespace mmd
