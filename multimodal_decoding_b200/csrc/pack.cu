// Layout kernels: build the vertex-major fp16 panels the per-centre kernels read.
//
// The reference keeps X as (trials x vertices) row-major and gathers k strided columns per centre
// (searchlight.py:110), 2 bytes at a time.  Here every vertex owns one contiguous row of trials
//   Xt[v] = [ train trials (zero padded to ld) ]      and      Xe[v] = [ test trials | imagery trials | 0 ]
// so a neighbour is one contiguous, 16-byte aligned row.
#include "common.cuh"
#include "kernels.h"

namespace mmd {

// out[v][n] = X[ids[n]][v]  (n < n_ids), 0 for n_ids <= n < ld_out.   64 x 64 tiles through shared memory.
// A NaN or infinite input marks its output row (`bad_row[v] = 1`) and/or sets `*any_bad`: sklearn's Ridge refuses
// such inputs (`check_array`, ValueError) and the searchlight reports it for the centres that use the vertex.
__global__ void __launch_bounds__(256) pack_transpose_kernel(const __half* __restrict__ X, long ldx,
                                                             const int* __restrict__ ids, int n_ids, int V,
                                                             __half* __restrict__ out, int ld_out,
                                                             unsigned char* __restrict__ bad_row,
                                                             int* __restrict__ any_bad) {
    __shared__ __half tile[64][66];
    const int v0 = blockIdx.x * 64, n0 = blockIdx.y * 64;
    const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;   // 64 x 4
    for (int r = ty; r < 64; r += 4) {
        int n = n0 + r, v = v0 + tx;
        __half val = __float2half(0.f);
        if (n < n_ids && v < V) {
            val = X[(long)ids[n] * ldx + v];
            if ((__half_as_ushort(val) & 0x7c00u) == 0x7c00u) {      // exponent all ones: NaN or infinity
                if (bad_row) bad_row[v] = 1;
                if (any_bad) *any_bad = 1;
            }
        }
        tile[r][tx] = val;
    }
    __syncthreads();
    for (int r = ty; r < 64; r += 4) {
        int v = v0 + r, n = n0 + tx;
        if (v < V && n < ld_out) out[(long)v * ld_out + n] = tile[tx][r];
    }
}

void launch_pack_transpose(const __half* X, long ldx, const int* ids_dev, int n_ids, int V, __half* out,
                           int ld_out, unsigned char* bad_row, int* any_bad, cudaStream_t s) {
    dim3 grid((V + 63) / 64, (ld_out + 63) / 64);
    pack_transpose_kernel<<<grid, 256, 0, s>>>(X, ldx, ids_dev, n_ids, V, out, ld_out, bad_row, any_bad);
}

// sq[j] = sum_d Y[rows[j]][d]^2 in float64 (candidate norms for the cosine distances)
__global__ void row_sqnorm_kernel(const __half* __restrict__ Y, long ldy, const int* __restrict__ rows, int n_rows,
                                  int D, double* __restrict__ sq) {
    int j = blockIdx.x;
    if (j >= n_rows) return;
    const __half* y = Y + (long)rows[j] * ldy;
    double acc = 0;
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
        double v = h2d(y[d]);
        acc += v * v;
    }
    __shared__ double red[32];
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
        sq[j] = t;
    }
}

// out[i] = (double)in[i]: the candidate latents as float64 operands of the scoring GEMM (exact widening)
__global__ void half_to_double_kernel(const __half* __restrict__ in, double* __restrict__ out, long n) {
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) out[i] = h2d(in[i]);
}

}  // namespace mmd
