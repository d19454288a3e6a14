// mmd_pairwise_accuracy_static: pairwise_accuracy (eval.py:116-123) for callers that cannot allocate -- code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc.  No counterpart in the reference.
// Same kernels as mmd_pairwise_accuracy (pairwise.cu); what differs is where the scratch lives: the
// standardised copies of both inputs, the squared row norms, the distance matrix and the counter sit in ONE file-scope
// device array of fixed size instead of the handle's growable buffers.
//
//   - the array is part of the library image: every GPU that loads libmmd_b200.so holds its 4.5 MiB until the process exits;
//   - in host code `g_static_scratch` is only a stand-in for the symbol, not a device address: the address is resolved once
//     per device, with that device current, by cudaGetSymbolAddress (static_scratch_ptr below) and that pointer is what the
//     kernels get;
//   - every call shares the array and finds the previous call's contents in it.  All regions are fully rewritten before they
//     are read except the counter, which launch_pairwise_count clears on the stream; calls that overlap on the GPU (two
//     streams, two host threads, two graph launches in flight) RACE on it -- serialise them;
//   - no cudaMalloc / cudaFree / allocating library call on this path, no synchronisation: the work is enqueued on the
//     caller's stream and `acc_out` (device memory) is valid once that stream has reached the end of it.
#include <cstdarg>
#include <cstdio>
#include <mutex>

#include "../../include/mmd_b200.h"
#include "common.cuh"
#include "kernels.h"

namespace mmd {
namespace {

constexpr int MAX_ROWS = MMD_STATIC_MAX_ROWS, MAX_DIMS = MMD_STATIC_MAX_DIMS;
constexpr size_t OFF_PRED = 0;                                                      // [MAX_ROWS][MAX_DIMS] float64
constexpr size_t OFF_LAT = OFF_PRED + (size_t)MAX_ROWS * MAX_DIMS * sizeof(double);  // [MAX_ROWS][MAX_DIMS]
constexpr size_t OFF_DIST = OFF_LAT + (size_t)MAX_ROWS * MAX_DIMS * sizeof(double);  // [MAX_ROWS][MAX_ROWS]
constexpr size_t OFF_SQ = OFF_DIST + (size_t)MAX_ROWS * MAX_ROWS * sizeof(double);   // [2][MAX_ROWS]
constexpr size_t OFF_COUNT = OFF_SQ + (size_t)2 * MAX_ROWS * sizeof(double);         // one unsigned long long
constexpr size_t SCRATCH_BYTES = OFF_COUNT + 256;
static_assert(SCRATCH_BYTES == MMD_STATIC_SCRATCH_BYTES, "include/mmd_b200.h states the size of the static scratch");
static_assert(OFF_LAT % 256 == 0 && OFF_DIST % 256 == 0 && OFF_SQ % 256 == 0 && OFF_COUNT % 256 == 0, "regions stay aligned");

static __device__ __align__(256) unsigned char g_static_scratch[SCRATCH_BYTES];

constexpr int MAX_DEVICES = 64;
std::mutex g_mutex;
unsigned char* g_scratch_ptr[MAX_DEVICES] = {};

// the device address of g_static_scratch on `device` (which must be current); resolved once per device
cudaError_t static_scratch_ptr(int device, unsigned char** out) {
    std::lock_guard<std::mutex> lock(g_mutex);
    if (!g_scratch_ptr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess) return e;
        g_scratch_ptr[device] = static_cast<unsigned char*>(p);
    }
    *out = g_scratch_ptr[device];
    return cudaSuccess;
}

__global__ void static_acc_finish_kernel(const unsigned long long* __restrict__ count, double denom, double* __restrict__ acc_out) {
    if (threadIdx.x == 0 && blockIdx.x == 0) *acc_out = (double)*count / denom;
}

int report(char* err, int err_len, int code, const char* fmt, ...) {
    if (err && err_len > 0) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, (size_t)err_len, fmt, ap);
        va_end(ap);
    }
    return code;
}

bool on_device(const void* p, int device) {
    cudaPointerAttributes a{};
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return (a.type == cudaMemoryTypeDevice && a.device == device) || a.type == cudaMemoryTypeManaged;
}

}  // namespace
}  // namespace mmd

using namespace mmd;

#define NA_CU(call)                                                                                          \
    do {                                                                                                     \
        const cudaError_t e_ = (call);                                                                       \
        if (e_ != cudaSuccess)                                                                               \
            return report(err, err_len, 2, "mmd_pairwise_accuracy_static: %s: %s", #call, cudaGetErrorString(e_)); \
    } while (0)

extern "C" int mmd_pairwise_accuracy_static(int device, void* stream, const double* latents, int n_lat, const double* predictions,
                                            int n_pred, int n_dims, int standardize_predictions, int standardize_latents,
                                            double* acc_out, char* err, int err_len) {
    if (err && err_len > 0) err[0] = 0;
    if (!latents || !predictions || !acc_out) return report(err, err_len, 1, "mmd_pairwise_accuracy_static: null pointer");
    if (n_pred < 1 || n_lat < 1 || n_dims < 1) return report(err, err_len, 1, "mmd_pairwise_accuracy_static: empty input");
    if (n_pred > MAX_ROWS || n_lat > MAX_ROWS || n_dims > MAX_DIMS)
        return report(err, err_len, 1,
                      "mmd_pairwise_accuracy_static: %d predictions x %d latents x %d dims exceed the static scratch (at most %d rows "
                      "each and %d dims; use mmd_pairwise_accuracy for larger problems)",
                      n_pred, n_lat, n_dims, MAX_ROWS, MAX_DIMS);
    if (n_pred > n_lat && n_lat != 1)
        return report(err, err_len, 1,
                      "mmd_pairwise_accuracy_static: %d predictions but %d candidate latents: the diagonal (eval.py:108) cannot be "
                      "broadcast against the distance matrix", n_pred, n_lat);
    if (device < 0 || device >= MAX_DEVICES) return report(err, err_len, 1, "mmd_pairwise_accuracy_static: device %d out of range", device);
    int current = -1;
    NA_CU(cudaGetDevice(&current));
    if (current != device) NA_CU(cudaSetDevice(device));
    if (!on_device(latents, device) || !on_device(predictions, device) || !on_device(acc_out, device))
        return report(err, err_len, 1, "mmd_pairwise_accuracy_static: latents, predictions and acc_out must be device memory of GPU %d "
                                       "(this entry point stages nothing)", device);
    unsigned char* base = nullptr;
    NA_CU(static_scratch_ptr(device, &base));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    double* dist = reinterpret_cast<double*>(base + OFF_DIST);
    double* psq = reinterpret_cast<double*>(base + OFF_SQ);
    double* csq = psq + MAX_ROWS;
    unsigned long long* count = reinterpret_cast<unsigned long long*>(base + OFF_COUNT);

    // standardisation works in place: on a copy in the scratch, never on the caller's rows
    const double* pred = predictions;
    const double* lat = latents;
    if (standardize_predictions) {
        double* p = reinterpret_cast<double*>(base + OFF_PRED);
        NA_CU(cudaMemcpyAsync(p, predictions, (size_t)n_pred * n_dims * sizeof(double), cudaMemcpyDeviceToDevice, s));
        NA_CU(launch_standardize_columns(p, n_pred, n_dims, n_dims, s));
        pred = p;
    }
    if (standardize_latents) {
        double* l = reinterpret_cast<double*>(base + OFF_LAT);
        NA_CU(cudaMemcpyAsync(l, latents, (size_t)n_lat * n_dims * sizeof(double), cudaMemcpyDeviceToDevice, s));
        NA_CU(launch_standardize_columns(l, n_lat, n_dims, n_dims, s));
        lat = l;
    }
    NA_CU(launch_cosine_dist(pred, n_pred, lat, n_lat, n_dims, n_dims, psq, csq, dist, s));
    NA_CU(launch_pairwise_count(dist, n_pred, n_lat, count, s));          // clears the counter on the stream first
    const long rows = n_pred < n_lat ? n_pred : n_lat;
    static_acc_finish_kernel<<<1, 32, 0, s>>>(count, (double)((long)n_pred * n_lat - rows), acc_out);
    NA_CU(cudaGetLastError());
    return 0;
}
