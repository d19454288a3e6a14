// Stand-alone check of csrc/predict_i8.cu (the tcgen05 / TMA prediction GEMM) against the host, stage by stage, so that one
// GPU run localises a mistake:
//   1. digit planes of XtY and Zt (launch_slice_xty / launch_slice_zt)        -- bit-exact against the host's cut
//   2. what the two TMA forms leave in shared memory (tile::gather4 with 128B swizzle, tiled load with 64B swizzle)
//   3. predict_i8_kernel, raw predictions                                      -- bit-exact against the host's integer
//      arithmetic, and its distance from the plain float64 product
//   4. predict_i8_kernel with z-scoring + the tolerance sums                   -- against the host to 1e-12
//   5. timing at the production shape (kp = 768, Dp = 1024), one CTA per SM
// Every device-side wait is bounded.  Build (from the repo root):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Imultimodal_decoding_b200/csrc \
//        -o tools/predict_i8_probe tools/predict_i8_probe.cu multimodal_decoding_b200/csrc/predict_i8.o
#include <cuda.h>
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>

#include "common.cuh"
#include "kernels.h"

using namespace mmd;

#define CK(x) do { cudaError_t e__ = (x); if (e__ != cudaSuccess) { printf("{\"error\": \"%s at line %d\"}\n", cudaGetErrorString(e__), __LINE__); return 1; } } while (0)

constexpr int S = I8_DIGITS, TN = EVAL_PAD, KS = 64, TM = 128;
constexpr int A_PLANE = KS * TM, B_PLANE = TN * KS, DUMP_BYTES = 3 * (A_PLANE + B_PLANE);

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// One warp: the loads of one ring slot (planes 0-2 of both operands, first 64 neighbours), then the slot is copied out.
__global__ void tma_dump_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const int* nbr, int V,
                                int d0, int b_row0, uint32_t expect_bytes, uint8_t* out, int* status) {
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* smem = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
    __shared__ __align__(8) uint64_t bar;
    const int lane = threadIdx.x;
    for (int i = lane; i < DUMP_BYTES / 4; i += 32) reinterpret_cast<uint32_t*>(smem)[i] = 0xEEEEEEEEu;
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    const uint32_t dst = smem_u32(smem), b = smem_u32(&bar);
    if (lane == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(expect_bytes) : "memory");
    if (lane < 16) {
        const int4 r = *reinterpret_cast<const int4*>(nbr + lane * 4);
        for (int p = 0; p < 3; ++p) {
            const int off = p * V;
            asm volatile(
                "cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes"
                " [%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::"r"(dst + p * A_PLANE + lane * 512),
                "l"(reinterpret_cast<uint64_t>(&tmA)), "r"(b), "r"(d0), "r"(r.x + off), "r"(r.y + off), "r"(r.z + off), "r"(r.w + off)
                : "memory");
        }
    } else if (lane < 19) {
        const int p = lane - 16;
        asm volatile(
            "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
            " [%0], [%1, {%3, %4}], [%2];" ::"r"(dst + 3 * A_PLANE + p * B_PLANE),
            "l"(reinterpret_cast<uint64_t>(&tmB)), "r"(b), "r"(0), "r"(b_row0 + p * TN)
            : "memory");
    }
    uint32_t done = 0;
    for (long spin = 0; spin < (1L << 20) && !done; ++spin)
        asm volatile(
            "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
            : "=r"(done) : "r"(b), "r"(0) : "memory");
    if (!done && lane == 0) *status = 1;
    __syncwarp();
    for (int i = lane; i < DUMP_BYTES / 4; i += 32) reinterpret_cast<uint32_t*>(out)[i] = reinterpret_cast<uint32_t*>(smem)[i];
}

static double pow2_ceil_h(double amax) {
    if (!(amax > 0.0)) return 1.0;
    int e;
    const double m = std::frexp(amax, &e);
    return m == 0.5 ? std::ldexp(1.0, e - 1) : std::ldexp(1.0, e);
}
static void cut_h(double a, double inv, int8_t q[S]) {
    double r = a * inv;
    for (int t = 0; t < S; ++t) {
        r *= t == 0 ? 64.0 : 128.0;
        const double d = std::nearbyint(r);
        q[t] = (int8_t)(int)d;
        r -= d;
    }
}

struct Case {
    int V, Dp, n_tasks, kp_max;
    std::vector<int> k, kp;
    std::vector<double> XtY, work;       // work: per task (kp_max + 144) * kp_max doubles, Zt at kp*kp with stride kp
    std::vector<int> nbr;                // n_tasks x kp_max
};

static Case make_case(int V, int Dp, int n_tasks, int kp_max, bool ragged, int window, unsigned seed) {
    Case c;
    c.V = V; c.Dp = Dp; c.n_tasks = n_tasks; c.kp_max = kp_max;
    std::mt19937_64 rng(seed);
    std::uniform_real_distribution<double> U(-1.0, 1.0);
    c.XtY.resize((size_t)V * Dp);
    std::vector<double> colmag(Dp);
    for (int d = 0; d < Dp; ++d) colmag[d] = std::exp2((double)((int)(rng() % 9) - 4)) * 700.0;
    for (size_t i = 0; i < c.XtY.size(); ++i) c.XtY[i] = colmag[i % Dp] * U(rng) * (0.2 + std::fabs(U(rng)));
    for (int v = 0; v < V; ++v) c.XtY[(size_t)v * Dp + (Dp - 1)] = 0.0;            // an all-zero (padding) feature
    const size_t stride = (size_t)(kp_max + EVAL_PAD) * kp_max;
    c.work.assign(stride * n_tasks, 0.0);
    c.nbr.assign((size_t)n_tasks * kp_max, 0);
    for (int t = 0; t < n_tasks; ++t) {
        int kp = kp_max, k = kp_max;
        if (ragged && t % 3 == 1) { kp = std::max(64, kp_max - 64); k = kp - 28; }
        else if (ragged && t % 3 == 0) k = kp_max - 2;
        c.k.push_back(k);
        c.kp.push_back(kp);
        double* Zt = c.work.data() + stride * t + (size_t)kp * kp;
        for (int e = 0; e < EVAL_PAD - 1; ++e) {
            const double mag = std::exp2((double)((int)(rng() % 7) - 12));
            for (int j = 0; j < k; ++j) Zt[(size_t)e * kp + j] = mag * U(rng);
        }
        // sorted distinct neighbours from a window that drifts with the task
        const int w = std::min(window, V);
        const int start = (int)(((long)t * 37) % (V - w + 1));
        std::vector<int> pool(w);
        for (int i = 0; i < w; ++i) pool[i] = start + i;
        std::shuffle(pool.begin(), pool.end(), rng);
        std::sort(pool.begin(), pool.begin() + k);
        for (int j = 0; j < k; ++j) c.nbr[(size_t)t * kp_max + j] = pool[j];
    }
    return c;
}

int main(int argc, char** argv) {
    const bool timing = !(argc > 1 && !strcmp(argv[1], "--no-timing"));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;

    // ---------------------------------------------------------------- small case: stages 1-4
    Case c = make_case(3001, 256, 5, 192, true, 900, 11);
    const int V = c.V, Dp = c.Dp, nt = c.n_tasks, kpm = c.kp_max;
    const size_t stride = (size_t)(kpm + EVAL_PAD) * kpm;
    double *dXtY, *dwork, *dsX, *dsZ, *dyhat, *damp;
    int8_t *dxp, *dzp;
    unsigned long long* dcolmax;
    int *dnbr, *dstatus;
    CentreTask* dtasks;
    CK(cudaMalloc(&dXtY, c.XtY.size() * 8));
    CK(cudaMalloc(&dwork, c.work.size() * 8));
    CK(cudaMalloc(&dsX, Dp * 8));
    CK(cudaMalloc(&dsZ, (size_t)nt * TN * 8));
    CK(cudaMalloc(&dyhat, (size_t)nt * YHAT_ROWS * Dp * 8));
    CK(cudaMalloc(&damp, (size_t)nt * 4 * 8));
    CK(cudaMalloc(&dxp, (size_t)S * V * Dp));
    CK(cudaMalloc(&dzp, (size_t)nt * S * TN * kpm));
    CK(cudaMalloc(&dcolmax, Dp * 8));
    CK(cudaMalloc(&dnbr, c.nbr.size() * 4));
    CK(cudaMalloc(&dstatus, 4));
    CK(cudaMalloc(&dtasks, nt * sizeof(CentreTask)));
    CK(cudaMemcpy(dXtY, c.XtY.data(), c.XtY.size() * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dwork, c.work.data(), c.work.size() * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dnbr, c.nbr.data(), c.nbr.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemset(dzp, 0x55, (size_t)nt * S * TN * kpm));
    CK(cudaMemset(dstatus, 0, 4));
    std::vector<CentreTask> tasks(nt);
    for (int t = 0; t < nt; ++t) {
        memset(&tasks[t], 0, sizeof(CentreTask));
        tasks[t].nbr = dnbr + (size_t)t * kpm;
        tasks[t].centre = t;
        tasks[t].k = c.k[t];
        tasks[t].kp = c.kp[t];
    }
    CK(cudaMemcpy(dtasks, tasks.data(), nt * sizeof(CentreTask), cudaMemcpyHostToDevice));

    CK(launch_slice_xty(dXtY, V, Dp, dxp, dsX, dcolmax, 0));
    CK(launch_slice_zt(dtasks, nt, kpm, dwork, dzp, dsZ, 0));
    CK(cudaDeviceSynchronize());
    std::vector<int8_t> xp((size_t)S * V * Dp), zp((size_t)nt * S * TN * kpm);
    std::vector<double> sX(Dp), sZ((size_t)nt * TN);
    CK(cudaMemcpy(xp.data(), dxp, xp.size(), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(zp.data(), dzp, zp.size(), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(sX.data(), dsX, Dp * 8, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(sZ.data(), dsZ, sZ.size() * 8, cudaMemcpyDeviceToHost));
    long bad_x = 0, bad_z = 0, bad_s = 0;
    for (int d = 0; d < Dp; ++d) {
        double m = 0;
        for (int v = 0; v < V; ++v) m = std::fmax(m, std::fabs(c.XtY[(size_t)v * Dp + d]));
        if (pow2_ceil_h(m) != sX[d]) ++bad_s;
    }
    for (int v = 0; v < V; ++v)
        for (int d = 0; d < Dp; ++d) {
            int8_t q[S];
            cut_h(c.XtY[(size_t)v * Dp + d], 1.0 / sX[d], q);
            for (int t = 0; t < S; ++t) bad_x += xp[((size_t)t * V + v) * Dp + d] != q[t];
        }
    for (int t = 0; t < nt; ++t) {
        const int kp = c.kp[t];
        const double* Zt = c.work.data() + stride * t + (size_t)kp * kp;
        for (int e = 0; e < TN; ++e) {
            double m = 0;
            for (int j = 0; j < kp; ++j) m = std::fmax(m, std::fabs(Zt[(size_t)e * kp + j]));
            if (pow2_ceil_h(m) != sZ[(size_t)t * TN + e]) ++bad_s;
            for (int j = 0; j < kp; ++j) {
                int8_t q[S];
                cut_h(Zt[(size_t)e * kp + j], 1.0 / sZ[(size_t)t * TN + e], q);
                for (int u = 0; u < S; ++u) bad_z += zp[(((size_t)t * S + u) * TN + e) * kpm + j] != q[u];
            }
        }
    }
    printf("{\"stage\": \"digit planes\", \"gpu\": \"%s\", \"scale_mismatches\": %ld, \"xty_digit_mismatches\": %ld, \"zt_digit_mismatches\": %ld}\n",
           prop.name, bad_s, bad_x, bad_z);

    // ---------------------------------------------------------------- stage 2: TMA layouts
    {
        alignas(64) CUtensorMap tmA, tmA4, tmB;
        CK(make_tensor_map_u8(&tmA, dxp, Dp, (uint64_t)V * S, Dp, TM, 1, 128));
        CK(make_tensor_map_u8(&tmA4, dxp, Dp, (uint64_t)V * S, Dp, TM, 4, 128));
        CK(make_tensor_map_u8(&tmB, dzp, kpm, (uint64_t)nt * S * TN, kpm, KS, TN, 64));
        uint8_t* ddump;
        CK(cudaMalloc(&ddump, DUMP_BYTES));
        CK(cudaFuncSetAttribute(tma_dump_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DUMP_BYTES + 1024));
        const int task = 2, d0 = 128;
        for (int variant = 0; variant < 1; ++variant) {      // (a 128 x 4 box is not what tile::gather4 takes)
            CK(cudaMemset(dstatus, 0, 4));
            CK(cudaMemset(ddump, 0, DUMP_BYTES));
            tma_dump_kernel<<<1, 32, DUMP_BYTES + 1024>>>(variant == 0 ? tmA : tmA4, tmB, dnbr + (size_t)task * kpm, V, d0, task * S * TN,
                                                          DUMP_BYTES, ddump, dstatus);
            cudaError_t e = cudaDeviceSynchronize();
            int st = 0;
            if (e == cudaSuccess) cudaMemcpy(&st, dstatus, 4, cudaMemcpyDeviceToHost);
            std::vector<uint8_t> dump(DUMP_BYTES);
            if (e == cudaSuccess) cudaMemcpy(dump.data(), ddump, DUMP_BYTES, cudaMemcpyDeviceToHost);
            long badA = 0, badB = 0, untouched = 0;
            for (int p = 0; p < 3; ++p)
                for (int r = 0; r < KS; ++r)
                    for (int j = 0; j < 8; ++j)
                        for (int b = 0; b < 16; ++b) {
                            const uint8_t got = dump[p * A_PLANE + r * 128 + ((j ^ (r & 7)) << 4) + b];
                            const int row = c.nbr[(size_t)task * kpm + r];
                            const uint8_t want = (uint8_t)xp[((size_t)p * V + row) * Dp + d0 + j * 16 + b];
                            badA += got != want;
                            untouched += got == 0xEE && want != 0xEE;
                        }
            for (int p = 0; p < 3; ++p)
                for (int r = 0; r < TN; ++r)
                    for (int j = 0; j < 4; ++j)
                        for (int b = 0; b < 16; ++b) {
                            const uint8_t got = dump[3 * A_PLANE + p * B_PLANE + r * 64 + ((j ^ ((r >> 1) & 3)) << 4) + b];
                            const uint8_t want = (uint8_t)zp[(((size_t)task * S + p) * TN + r) * kpm + j * 16 + b];
                            badB += got != want;
                        }
            printf("{\"stage\": \"tma layout\", \"gather4_box_rows\": %d, \"cuda\": \"%s\", \"wait_timed_out\": %d, \"a_sw128_mismatches\": %ld, "
                   "\"a_bytes_untouched\": %ld, \"b_sw64_mismatches\": %ld}\n",
                   variant == 0 ? 1 : 4, cudaGetErrorString(e), st, badA, untouched, badB);
            if (e != cudaSuccess) return 1;
        }
        cudaFree(ddump);
    }

    // ---------------------------------------------------------------- stages 3-4: the GEMM
    PredictI8Params prm{};
    prm.tasks = dtasks; prm.n_tasks = nt; prm.kp_max = kpm; prm.V = V; prm.Dp = Dp; prm.sX = dsX; prm.sZ = dsZ; prm.yhat = dyhat;
    prm.amp2 = damp; prm.n_cap = 70; prm.n_img = 70; prm.n_im = 3; prm.standardize = 0; prm.status = dstatus;
    // host: the integer arithmetic of the kernel, and the plain float64 product
    std::vector<double> want((size_t)nt * TN * Dp), exact((size_t)nt * TN * Dp);
    for (int t = 0; t < nt; ++t) {
        const int kp = c.kp[t], k = c.k[t];
        const double* Zt = c.work.data() + stride * t + (size_t)kp * kp;
        for (int e = 0; e < TN; ++e)
            for (int d = 0; d < Dp; ++d) {
                long long sw[S] = {0};
                double ex = 0;
                for (int j = 0; j < kp; ++j) {
                    const int row = c.nbr[(size_t)t * kpm + (j < k ? j : 0)];
                    ex += Zt[(size_t)e * kp + j] * c.XtY[(size_t)row * Dp + d];
                    for (int a = 0; a < S; ++a)
                        for (int b = 0; a + b < S; ++b)
                            sw[a + b] += (long long)xp[((size_t)a * V + row) * Dp + d] * (long long)zp[(((size_t)t * S + b) * TN + e) * kpm + j];
                }
                double acc = 0;
                for (int w = S - 1; w >= 0; --w) acc = acc + (double)sw[w] * std::exp2(-(double)(12 + 7 * w));
                want[((size_t)t * TN + e) * Dp + d] = (acc * sX[d]) * sZ[(size_t)t * TN + e];
                exact[((size_t)t * TN + e) * Dp + d] = ex;
            }
    }
    CK(cudaMemset(dyhat, 0xff, (size_t)nt * YHAT_ROWS * Dp * 8));
    CK(cudaMemset(damp, 0, (size_t)nt * 4 * 8));
    CK(launch_predict_i8(prm, dxp, dzp, sms, 0));
    {
        cudaError_t e = cudaDeviceSynchronize();
        int st = 0;
        if (e == cudaSuccess) cudaMemcpy(&st, dstatus, 4, cudaMemcpyDeviceToHost);
        std::vector<double> got((size_t)nt * YHAT_ROWS * Dp);
        if (e == cudaSuccess) cudaMemcpy(got.data(), dyhat, got.size() * 8, cudaMemcpyDeviceToHost);
        long mism = 0, first = -1;
        double max_rel = 0, ref = 0;
        for (size_t i = 0; i < exact.size(); ++i) ref = std::fmax(ref, std::fabs(exact[i]));
        for (int t = 0; t < nt; ++t)
            for (int e2 = 0; e2 < TN; ++e2)
                for (int d = 0; d < Dp; ++d) {
                    const double g = got[((size_t)t * YHAT_ROWS + e2) * Dp + d], w = want[((size_t)t * TN + e2) * Dp + d];
                    if (memcmp(&g, &w, 8) != 0) {
                        if (first < 0) first = ((long)t * TN + e2) * Dp + d;
                        ++mism;
                    }
                    max_rel = std::fmax(max_rel, std::fabs(g - exact[((size_t)t * TN + e2) * Dp + d]) / ref);
                }
        printf("{\"stage\": \"gemm raw\", \"cuda\": \"%s\", \"status\": %d, \"bit_mismatches_vs_host_integer_arithmetic\": %ld, \"first\": %ld, "
               "\"max_error_vs_float64_product_rel_to_max\": %.3e}\n", cudaGetErrorString(e), st, mism, first, max_rel);
        if (e != cudaSuccess) return 1;
        if (mism) {
            // where do they differ: print a few
            int shown = 0;
            for (int t = 0; t < nt && shown < 6; ++t)
                for (int e2 = 0; e2 < TN && shown < 6; ++e2)
                    for (int d = 0; d < Dp && shown < 6; ++d) {
                        const double g = got[((size_t)t * YHAT_ROWS + e2) * Dp + d], w = want[((size_t)t * TN + e2) * Dp + d];
                        if (memcmp(&g, &w, 8) != 0) {
                            printf("  task %d row %d feature %d: got %.17g want %.17g\n", t, e2, d, g, w);
                            ++shown;
                        }
                    }
        }
    }
    prm.standardize = 1;
    CK(cudaMemset(dyhat, 0xff, (size_t)nt * YHAT_ROWS * Dp * 8));
    CK(cudaMemset(damp, 0, (size_t)nt * 4 * 8));
    CK(cudaMemset(dstatus, 0, 4));
    CK(launch_predict_i8(prm, dxp, dzp, sms, 0));
    {
        cudaError_t e = cudaDeviceSynchronize();
        int st = 0;
        if (e == cudaSuccess) cudaMemcpy(&st, dstatus, 4, cudaMemcpyDeviceToHost);
        std::vector<double> got((size_t)nt * YHAT_ROWS * Dp), amp((size_t)nt * 4);
        if (e == cudaSuccess) {
            cudaMemcpy(got.data(), dyhat, got.size() * 8, cudaMemcpyDeviceToHost);
            cudaMemcpy(amp.data(), damp, amp.size() * 8, cudaMemcpyDeviceToHost);
        }
        double max_err = 0, amp_err = 0;
        long nonfinite = 0;
        const int lo[3] = {0, 70, 140}, hi[3] = {70, 140, 143};
        for (int t = 0; t < nt; ++t) {
            double a_want[4] = {0, 0, 0, 0};
            for (int d = 0; d < Dp; ++d) {
                for (int g = 0; g < 3; ++g) {
                    double mu = 0, var = 0;
                    for (int r = lo[g]; r < hi[g]; ++r) mu += want[((size_t)t * TN + r) * Dp + d];
                    mu /= hi[g] - lo[g];
                    for (int r = lo[g]; r < hi[g]; ++r) { const double q = want[((size_t)t * TN + r) * Dp + d] - mu; var += q * q; }
                    const double sd = std::sqrt(var / (hi[g] - lo[g])), rs = sd == 0 ? 1.0 : 1.0 / sd;
                    if (sd > 0) a_want[g] += (sX[d] / sd) * (sX[d] / sd);
                    for (int r = lo[g]; r < hi[g]; ++r) {
                        const double z = (want[((size_t)t * TN + r) * Dp + d] - mu) * rs;
                        const double gz = got[((size_t)t * YHAT_ROWS + r) * Dp + d];
                        if (!std::isfinite(gz)) ++nonfinite;
                        max_err = std::fmax(max_err, std::fabs(gz - z));
                        if (g == 2) max_err = std::fmax(max_err, std::fabs(got[((size_t)t * YHAT_ROWS + EVAL_PAD + r - 140) * Dp + d] - want[((size_t)t * TN + r) * Dp + d]) / (std::fabs(want[((size_t)t * TN + r) * Dp + d]) + 1e-300));
                    }
                }
                a_want[3] += sX[d] * sX[d];
                for (int r = 143; r < YHAT_ROWS; ++r)
                    if (r < EVAL_PAD || r >= EVAL_PAD + 3) max_err = std::fmax(max_err, std::fabs(got[((size_t)t * YHAT_ROWS + r) * Dp + d]));
            }
            for (int g = 0; g < 4; ++g) amp_err = std::fmax(amp_err, std::fabs(amp[(size_t)t * 4 + g] - a_want[g]) / a_want[g]);
        }
        printf("{\"stage\": \"gemm z-scored\", \"cuda\": \"%s\", \"status\": %d, \"max_abs_error\": %.3e, \"non_finite\": %ld, \"tolerance_sums_rel_error\": %.3e}\n",
               cudaGetErrorString(e), st, max_err, nonfinite, amp_err);
        if (e != cudaSuccess) return 1;
    }
    cudaFree(dXtY); cudaFree(dwork); cudaFree(dsX); cudaFree(dsZ); cudaFree(dyhat); cudaFree(damp); cudaFree(dxp); cudaFree(dzp);
    cudaFree(dnbr); cudaFree(dtasks);
    if (!timing) return 0;

    // ---------------------------------------------------------------- stage 5: timing at the production shape
    {
        const int Vb = 163842, Db = 1024, kb = 768, ntb = 4 * sms * 2;
        Case b;
        b.V = Vb; b.Dp = Db; b.n_tasks = ntb; b.kp_max = kb;
        std::mt19937_64 rng(5);
        const size_t strideb = (size_t)(kb + EVAL_PAD) * kb;
        double *XtYb, *workb, *sXb, *sZb, *yhatb, *ampb;
        int8_t *xpb, *zpb;
        int* nbrb;
        CentreTask* tasksb;
        CK(cudaMalloc(&XtYb, (size_t)Vb * Db * 8));
        CK(cudaMalloc(&workb, strideb * ntb * 8));
        CK(cudaMalloc(&sXb, Db * 8));
        CK(cudaMalloc(&sZb, (size_t)ntb * TN * 8));
        CK(cudaMalloc(&yhatb, (size_t)ntb * YHAT_ROWS * Db * 8));
        CK(cudaMalloc(&ampb, (size_t)ntb * 4 * 8));
        CK(cudaMalloc(&xpb, (size_t)S * Vb * Db));
        CK(cudaMalloc(&zpb, (size_t)ntb * S * TN * kb));
        CK(cudaMalloc(&nbrb, (size_t)ntb * kb * 4));
        CK(cudaMalloc(&tasksb, ntb * sizeof(CentreTask)));
        // contents do not matter for the timing: pseudo-random doubles from a small host block, tiled
        std::vector<double> blk(1 << 20);
        std::uniform_real_distribution<double> U(-1.0, 1.0);
        for (auto& v : blk) v = U(rng) * 900.0;
        for (size_t off = 0; off < (size_t)Vb * Db; off += blk.size())
            CK(cudaMemcpy(XtYb + off, blk.data(), std::min(blk.size(), (size_t)Vb * Db - off) * 8, cudaMemcpyHostToDevice));
        for (auto& v : blk) v *= 1e-6;
        for (size_t off = 0; off < strideb * ntb; off += blk.size())
            CK(cudaMemcpy(workb + off, blk.data(), std::min(blk.size(), strideb * ntb - off) * 8, cudaMemcpyHostToDevice));
        std::vector<int> nb((size_t)ntb * kb);
        std::vector<CentreTask> tk(ntb);
        for (int t = 0; t < ntb; ++t) {
            const int w = 2500, start = (int)(((long)t * 20) % (Vb - w));       // neighbouring centres share most rows
            std::vector<int> pool(w);
            for (int i = 0; i < w; ++i) pool[i] = start + i;
            std::shuffle(pool.begin(), pool.end(), rng);
            std::sort(pool.begin(), pool.begin() + kb);
            for (int j = 0; j < kb; ++j) nb[(size_t)t * kb + j] = pool[j];
            memset(&tk[t], 0, sizeof(CentreTask));
            tk[t].nbr = nbrb + (size_t)t * kb;
            tk[t].centre = t;
            tk[t].k = 750;
            tk[t].kp = kb;
        }
        CK(cudaMemcpy(nbrb, nb.data(), nb.size() * 4, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(tasksb, tk.data(), ntb * sizeof(CentreTask), cudaMemcpyHostToDevice));
        unsigned long long* cm;
        CK(cudaMalloc(&cm, Db * 8));
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        float ms_sx = 0, ms_sz = 0, ms_gemm = 1e30f;
        cudaEventRecord(e0);
        CK(launch_slice_xty(XtYb, Vb, Db, xpb, sXb, cm, 0));
        cudaEventRecord(e1);
        CK(cudaEventSynchronize(e1));
        cudaEventElapsedTime(&ms_sx, e0, e1);
        cudaEventRecord(e0);
        CK(launch_slice_zt(tasksb, ntb, kb, workb, zpb, sZb, 0));
        cudaEventRecord(e1);
        CK(cudaEventSynchronize(e1));
        cudaEventElapsedTime(&ms_sz, e0, e1);
        PredictI8Params pb{};
        pb.tasks = tasksb; pb.n_tasks = ntb; pb.kp_max = kb; pb.V = Vb; pb.Dp = Db; pb.sX = sXb; pb.sZ = sZb; pb.yhat = yhatb; pb.amp2 = ampb;
        pb.n_cap = 70; pb.n_img = 70; pb.n_im = 3; pb.standardize = 1; pb.status = dstatus;
        CK(cudaMemset(dstatus, 0, 4));
        for (int r = 0; r < 4; ++r) {
            CK(cudaMemset(ampb, 0, (size_t)ntb * 4 * 8));
            cudaEventRecord(e0);
            CK(launch_predict_i8(pb, xpb, zpb, sms, 0));
            cudaEventRecord(e1);
            CK(cudaEventSynchronize(e1));
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (r > 0) ms_gemm = std::min(ms_gemm, ms);
        }
        int st = 0;
        CK(cudaMemcpy(&st, dstatus, 4, cudaMemcpyDeviceToHost));
        const double ops = 2.0 * TN * kb * Db * 21.0 * ntb;
        printf("{\"stage\": \"timing\", \"status\": %d, \"centres\": %d, \"slice_xty_ms\": %.3f, \"slice_zt_us_per_centre\": %.3f, "
               "\"predict_i8_us_per_centre\": %.3f, \"int8_tops\": %.1f, \"note\": \"kp = 768, Dp = 1024, V = 163842, one CTA per SM\"}\n",
               st, ntb, ms_sx, ms_sz * 1e3 / ntb, ms_gemm * 1e3 / ntb, ops / ms_gemm * 1e-9);
    }
    return 0;
}
