// Group-level statistics on the searchlight score maps:
//   * one-sample t-maps over subjects -- `calc_image_t_values` / `ttest_1samp_no_p`
//     (analyses/decoding/searchlight/searchlight_permutation_testing.py:213-258),
//   * threshold-free cluster enhancement on the mesh graph -- `calc_tfce_values` / `calc_clusters`
//     (analyses/cluster_analysis.py:15-75, 248-312), for many maps at once (the 10,000-permutation null
//     distribution of searchlight_permutation_testing.py:442-556 is one launch per chunk of maps).
//
// TFCE: one CTA per map.  For every threshold t = dh, 2dh, ... the supra-threshold sub-graph is labelled by
// hooking (atomicMin on root labels) and pointer jumping until nothing changes; component sizes come from one
// atomicAdd per active vertex; every vertex then adds its term to a float32 accumulator in the reference's
// order (ascending thresholds) and rounding (float64 sum, rounded to float32), so h = 2 / e = 1 results are
// bit-identical to numpy's.  Arithmetic that numpy does without FMA is written with explicit _rn intrinsics.
#include "common.cuh"
#include "kernels.h"

namespace mmd {
namespace {

constexpr int TFCE_THREADS = 1024;

// numpy's add.reduce over a strided vector of n doubles (pairwise summation: plain loop below 8 elements, eight
// partial sums up to 128; searchlight studies have 6 subjects)
__device__ double numpy_sum(const double* __restrict__ x, long stride, int n) {
    if (n < 8) {
        double r = x[0];
        for (int i = 1; i < n; ++i) r = __dadd_rn(r, x[i * stride]);
        return r;
    }
    double r[8];
    for (int j = 0; j < 8; ++j) r[j] = x[j * stride];
    int i = 8;
    for (; i < n - (n % 8); i += 8)
        for (int j = 0; j < 8; ++j) r[j] = __dadd_rn(r[j], x[(i + j) * stride]);
    double res = __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])),
                           __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
    for (; i < n; ++i) res = __dadd_rn(res, x[i * stride]);
    return res;
}

constexpr int MAX_SUBJECTS = 128;

__global__ void __launch_bounds__(128) ttest_kernel(const double* __restrict__ data, int n, int V, double popmean,
                                                    double sigma, int relative, int constant_is_nan,
                                                    double* __restrict__ t_out) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= V) return;
    double y[MAX_SUBJECTS];
    bool all_equal = true;
    const double first = data[v];
    for (int s = 0; s < n; ++s) {
        const double x = data[(long)s * V + v];
        all_equal = all_equal && (x == first);
        y[s] = __dsub_rn(x, popmean);
    }
    if (all_equal && constant_is_nan) {                // calc_t_value: undefined when all subjects agree
        t_out[v] = __longlong_as_double(0x7ff8000000000000LL);
        return;
    }
    const double mean = numpy_sum(y, 1, n) / n;
    double d2[MAX_SUBJECTS];
    for (int s = 0; s < n; ++s) {
        const double d = __dsub_rn(y[s], mean);
        d2[s] = __dmul_rn(d, d);
    }
    double var = numpy_sum(d2, 1, n) / (n - 1);
    if (sigma > 0.0) var = __dadd_rn(var, relative ? __dmul_rn(sigma, var) : sigma);   // per-vertex call: max(var) = var
    t_out[v] = mean / sqrt(var / n);
}

__device__ __forceinline__ double block_max(double v, double* red) {
    for (int o = 16; o; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double r = red[0];
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) r = fmax(r, red[w]);
    return r;
}

// Labels of the sub-graph {value >= t}: label[v] = smallest vertex id of v's component (v itself when isolated or
// inactive).  All threads of the CTA take part.
__device__ void label_components(const double* __restrict__ val, int V, const int2* __restrict__ edges, int E, double t,
                                 int* __restrict__ label, int* changed_sm) {
    const int tid = threadIdx.x;
    for (int v = tid; v < V; v += TFCE_THREADS) label[v] = v;
    __syncthreads();
    for (;;) {
        if (tid == 0) *changed_sm = 0;
        __syncthreads();
        bool changed = false;
        for (int i = tid; i < E; i += TFCE_THREADS) {
            const int2 ed = edges[i];
            if (!(val[ed.x] >= t) || !(val[ed.y] >= t)) continue;
            const int ra = label[ed.x], rb = label[ed.y];
            if (ra != rb) {
                atomicMin(&label[ra > rb ? ra : rb], ra > rb ? rb : ra);
                changed = true;
            }
        }
        if (changed) *changed_sm = 1;
        __syncthreads();
        const int any = *changed_sm;
        // pointer jumping: every vertex points at its current root
        for (int v = tid; v < V; v += TFCE_THREADS) {
            int l = label[v];
            while (true) {
                const int p = label[l];
                if (p == l) break;
                l = p;
            }
            label[v] = l;
        }
        __syncthreads();
        if (!any) break;
    }
}

struct TfceParams {
    const double* values;   // (n_maps x V)
    int n_maps, V;
    const int2* edges;
    int E;
    double h, e, dh;        // dh <= 0: max / 100 ("auto")
    int* label;             // workspace (n_maps x V)
    int* count;             // workspace (n_maps x V)
    float* out;             // (n_maps x V)
};

__global__ void __launch_bounds__(TFCE_THREADS) tfce_kernel(TfceParams prm) {
    __shared__ double red[TFCE_THREADS / 32];
    __shared__ int changed_sm;
    const int map = blockIdx.x, tid = threadIdx.x, V = prm.V;
    const double* val = prm.values + (long)map * V;
    int* label = prm.label + (long)map * V;
    int* count = prm.count + (long)map * V;
    float* out = prm.out + (long)map * V;

    double mx = -INFINITY;                               // np.nanmax: NaNs ignored
    bool any_value = false;
    for (int v = tid; v < V; v += TFCE_THREADS) {
        const double x = val[v];
        out[v] = 0.f;
        if (x == x) {
            mx = fmax(mx, x);
            any_value = true;
        }
    }
    const double max_score = block_max(mx, red);
    const bool have = __syncthreads_or(any_value);
    if (!have || isinf(max_score) || !(max_score > 0.0)) return;

    const double step = prm.dh > 0.0 ? prm.dh : max_score / 100.0;
    // np.arange(step, max + step, step): ceil((stop - start) / step) values start + i * step
    const double stop = __dadd_rn(max_score, step);
    const long n_thr = (long)ceil(__dsub_rn(stop, step) / step);
    const bool square = prm.h == 2.0, linear = prm.e == 1.0;

    for (long i = 0; i < n_thr; ++i) {
        const double t = __dadd_rn(step, __dmul_rn((double)i, step));
        for (int v = tid; v < V; v += TFCE_THREADS) count[v] = 0;
        label_components(val, V, prm.edges, prm.E, t, label, &changed_sm);
        for (int v = tid; v < V; v += TFCE_THREADS)
            if (val[v] >= t) atomicAdd(&count[label[v]], 1);
        __syncthreads();
        const double th = square ? __dmul_rn(t, t) : pow(t, prm.h);
        for (int v = tid; v < V; v += TFCE_THREADS) {
            const double x = val[v];
            if (!(x >= t)) continue;
            const int size = count[label[v]];
            double term;
            if (size >= 2) {                              // a cluster has at least one edge
                const double ext = linear ? (double)size : pow((double)size, prm.e);
                term = __dmul_rn(__dmul_rn(ext, th), step);
            } else if (x > t) {                           // supra-threshold vertex outside clusters (strict)
                term = __dmul_rn(th, step);
            } else {
                continue;
            }
            out[v] = (float)__dadd_rn((double)out[v], term);
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(TFCE_THREADS) cluster_labels_kernel(const double* __restrict__ val, int V,
                                                                      const int2* __restrict__ edges, int E, double t,
                                                                      int* __restrict__ label, int* __restrict__ count,
                                                                      int* __restrict__ out) {
    __shared__ int changed_sm;
    const int tid = threadIdx.x;
    for (int v = tid; v < V; v += TFCE_THREADS) count[v] = 0;
    label_components(val, V, edges, E, t, label, &changed_sm);
    for (int v = tid; v < V; v += TFCE_THREADS)
        if (val[v] >= t) atomicAdd(&count[label[v]], 1);
    __syncthreads();
    for (int v = tid; v < V; v += TFCE_THREADS) out[v] = (val[v] >= t && count[label[v]] >= 2) ? label[v] : -1;
}

}  // namespace

cudaError_t launch_ttest(const double* data, int n_subjects, int V, double popmean, double sigma, int relative,
                         int constant_is_nan, double* t_out, cudaStream_t s) {
    if (n_subjects < 2 || n_subjects > MAX_SUBJECTS) return cudaErrorInvalidValue;
    ttest_kernel<<<(V + 127) / 128, 128, 0, s>>>(data, n_subjects, V, popmean, sigma, relative, constant_is_nan, t_out);
    return cudaGetLastError();
}

cudaError_t launch_tfce(const double* values, int n_maps, int V, const int* edges, int E, double h, double e, double dh,
                        int* label_ws, int* count_ws, float* out, cudaStream_t s) {
    TfceParams p{values, n_maps, V, reinterpret_cast<const int2*>(edges), E, h, e, dh, label_ws, count_ws, out};
    tfce_kernel<<<n_maps, TFCE_THREADS, 0, s>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_cluster_labels(const double* values, int V, const int* edges, int E, double threshold, int* label_ws,
                                  int* count_ws, int* labels_out, cudaStream_t s) {
    cluster_labels_kernel<<<1, TFCE_THREADS, 0, s>>>(values, V, reinterpret_cast<const int2*>(edges), E, threshold,
                                                    label_ws, count_ws, labels_out);
    return cudaGetLastError();
}

}  // namespace mmd
