// C ABI (include/mmd_b200.h) and host orchestration of the searchlight engine.
//
// Host-side work that has no counterpart in the reference: centres are grouped into spatial clusters so
// that the float64 Gram of a cluster's neighbourhood *union* is computed once (symmetric rank-k update on
// the DMMA pipe) and every member centre reads its k x k normal matrix as a sub-matrix of it.  For the
// reference's k-NN adjacency (searchlight.py:173) a 128-centre cluster has a union of ~2k vertices, which
// cuts the Gram work per centre from 2*N*k^2 to ~2*N*u^2/(2*128) (reported as executed flops).
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <future>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/mmd_b200.h"
#include "common.cuh"
#include "kernels.h"

using namespace mmd;

namespace {

enum Stage { ST_PACK = 0, ST_XTY, ST_PREP, ST_SYRK, ST_SOLVE, ST_PREDICT, ST_SCORE, ST_KNN, ST_NULL, ST_REROUTE };
const char* kStageNames[MMD_N_STAGES] = {"pack", "xty", "prep", "syrk", "factor_solve", "predict", "score", "knn", "null", "reroute"};
const char* kScoreNames[MMD_N_SCORES] = {
    "pairwise_acc_captions",
    "pairwise_acc_images",
    "pairwise_acc_cross_captions_to_images",
    "pairwise_acc_cross_images_to_captions",
    "pairwise_acc_imagery",
    "pairwise_acc_imagery_whole_test_set",
    "pairwise_acc_imagery_no_std",
    "pairwise_acc_imagery_whole_test_set_no_std",
};

// Guard mode (environment MMD_B200_GUARD=1; compute-sanitizer is not available on the GPU pool): every device
// buffer gets a 4 KiB band on either side and is filled, bands included, with 0xA5 bytes.  A kernel that writes
// outside its buffer breaks a band (found when the buffer is released or by mmd_check_guards); a kernel that reads
// memory nobody wrote sees 0xA5A5... instead of the zeros a fresh cudaMalloc tends to return.
constexpr size_t GUARD_BYTES = 4096;
std::atomic<int> g_guard_failures(0);
bool guard_mode() {
    static const bool on = [] {
        const char* e = getenv("MMD_B200_GUARD");
        return e && *e && *e != '0';
    }();
    return on;
}

struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    void* base = nullptr;       // == p unless guard mode
    cudaError_t ensure(size_t need) {
        if (need <= bytes) return cudaSuccess;
        release();
        const size_t pad = guard_mode() ? GUARD_BYTES : 0;
        cudaError_t e = cudaMalloc(&base, need + 2 * pad);
        if (e != cudaSuccess) {
            base = nullptr;
            return e;
        }
        if (pad && ((e = cudaMemset(base, 0xA5, need + 2 * pad)) != cudaSuccess ||
                    (e = cudaDeviceSynchronize()) != cudaSuccess))      // the handle's stream does not wait for the null stream
            return e;
        p = static_cast<char*>(base) + pad;
        bytes = need;
        return cudaSuccess;
    }
    // 0 = bands intact (or guard mode off); the device must be idle on this buffer
    int check() const {
        if (!p || !guard_mode()) return 0;
        std::vector<unsigned char> host(2 * GUARD_BYTES);
        if (cudaMemcpy(host.data(), base, GUARD_BYTES, cudaMemcpyDeviceToHost) != cudaSuccess ||
            cudaMemcpy(host.data() + GUARD_BYTES, static_cast<char*>(p) + bytes, GUARD_BYTES, cudaMemcpyDeviceToHost) !=
                cudaSuccess)
            return 4;
        int bad = 0;
        for (size_t i = 0; i < 2 * GUARD_BYTES; ++i)
            if (host[i] != 0xA5) bad |= i < GUARD_BYTES ? 1 : 2;
        if (bad) {
            fprintf(stderr, "mmd_b200 guard: write %s a device buffer of %zu bytes\n",
                    bad == 3 ? "before and behind" : bad == 1 ? "before" : "behind", bytes);
            ++g_guard_failures;
        }
        return bad;
    }
    void release() {
        if (p) {
            check();
            cudaFree(base);
        }
        p = base = nullptr;
        bytes = 0;
    }
    template <typename T>
    T* as() const { return reinterpret_cast<T*>(p); }
};

struct PinnedBuf {
    void* p = nullptr;
    size_t bytes = 0;
    cudaError_t ensure(size_t need) {
        if (need <= bytes) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr;
        bytes = 0;
        cudaError_t e = cudaHostAlloc(&p, need, cudaHostAllocDefault);
        if (e == cudaSuccess) bytes = need;
        return e;
    }
    void release() {
        if (p) cudaFreeHost(p);
        p = nullptr;
        bytes = 0;
    }
};

struct TimedSpan {
    cudaEvent_t a, b;
    int stage;
};

}  // namespace

// Cross-validation state of one mmd_searchlight_cv call (cv.cu): per-fold panels derived from the uploaded ones.
struct CvState {
    int n_folds = 0;
    std::vector<int> lo, hi, ld;                    // fold f = training positions [lo, hi); ld = its padded length
    std::vector<double> alphas;
    std::vector<DevBuf> xfold, xty_tr, cand, cand_sq;      // per fold
    std::vector<std::vector<DevBuf>> xe;                   // per fold and chunk of EVAL_PAD validation trials
    DevBuf yfold, pool, counts, probs;
    size_t pool_stride = 0;                         // doubles between the fold Gram pools of a batch
    void release() {
        for (auto* v : {&xfold, &xty_tr, &cand, &cand_sq})
            for (DevBuf& b : *v) b.release();
        for (auto& v : xe)
            for (DevBuf& b : v) b.release();
        yfold.release();
        pool.release();
        counts.release();
        probs.release();
        xfold.clear(); xty_tr.clear(); cand.clear(); cand_sq.clear(); xe.clear();
    }
};

// Whole-brain ridge grid search state (mmd_ridge_*; dense.cu): padded float64 panels of one (X, Y) problem.
struct RidgeState {
    int n = 0, v = 0, d = 0, np = 0, vp = 0, dp = 0, mp_fit = 0;
    bool dual = false, loaded = false, fitted = false;
    DevBuf X, Xt, Y, Yt, G, bt, A, Lt, linv, linvT, rhs, kval, pred_t, pred, yval, idx_tr, idx_val, flag, coef_t, stage, xft, yft,
        dist, sq, count;
    std::vector<DevBuf*> all() {
        return {&X, &Xt, &Y, &Yt, &G, &bt, &A, &Lt, &linv, &linvT, &rhs, &kval, &pred_t, &pred, &yval, &idx_tr, &idx_val, &flag,
                &coef_t, &stage, &xft, &yft, &dist, &sq, &count};
    }
};

struct mmd_handle {
    int device = 0;
    cudaStream_t stream = nullptr;
    std::string err;
    // uploaded problem
    int V = 0, n_train = 0, ldt = 0, n_test = 0, n_cap = 0, n_img = 0, n_im = 0, D = 0, Dp = 0;
    bool uploaded = false;
    DevBuf Xt, Xe, Yt, cand, cand64, cand_sq, XtY;
    // per-call / per-batch workspaces
    DevBuf adj, union_vertices, core_work, group_flags, nbr_sorted, loc, probs, gram_pool,
        work, yhat, scores, flags, misc_flag, tmp;
    // null-distribution scoring (optional) and the generic eval.py primitives
    DevBuf vertex_bad, y_bad;               // per vertex: 1 = betas hold NaN/inf, 2 = outside the active set; one int: training latents hold NaN/inf
    DevBuf active;                          // mmd_upload_active: the vertices whose XtY rows / digit planes exist
    int n_active = 0;                       // 0 = all vertices
    DevBuf up_x, up_y, up_train, up_eval;   // mmd_upload's staging copies: alive until the packing kernels have run
    cudaEvent_t up_copied = nullptr;        // recorded behind the last host-to-device copy of mmd_upload
    PinnedBuf batch_host[2];        // per-batch task arrays, packed: pinned and double-buffered so that the one copy per
    cudaEvent_t batch_copied[2] = {nullptr, nullptr};      // batch is truly asynchronous (a pageable source may make
    int batch_half = 0;             // cudaMemcpyAsync wait for the stream first)
    DevBuf batch_dev;
    PinnedBuf adj_host;             // int32 staging copy of the caller's adjacency: async H2D overlaps the host grouping
    DevBuf st_in, st_edges, st_label, st_count, st_out;     // group-level statistics
    DevBuf perm_im_centre;          // optional per-centre imagery permutations of the current mmd_searchlight_null_pc call
    bool use_perm_im_centre = false;
    DevBuf dist, null_out, perm_test, perm_im, pa_pred, pa_lat, pa_dist, pa_sq, pa_count;
    // integer prediction GEMM (predict_i8.cu): digit planes of XtY (per upload) and of every centre's Zt (per batch),
    // their scales, the score kernel's tolerance input, the list of centres handed back to the float64 kernels
    DevBuf xplanes, sX, colmax, zplanes, sZ, amp2, i8_status, reroute_list, reroute_counts, margins;
    CvState* cv = nullptr;          // non-null while mmd_searchlight_cv runs
    RidgeState* ridge = nullptr;    // mmd_ridge_load
    struct PlanCache* plan_cache = nullptr;     // clusters / unions / groups of the last adjacency (run_searchlight)
    int64_t plan_hits = 0;
    int predict_mode = 1;           // 1: tcgen05 int8 digit GEMM + float64 reroute of guarded centres; 0: float64 DMMA kernel
    int syrk_mode = 1;              // 1: exact integer cluster Gram on tcgen05 (syrk_i8.cu); 0: float64 DMMA kernel
    DevBuf syrk_planes, syrk_scales, syrk_flags, syrk_status;
    DevBuf core_done;               // merged core + centre launch of factor_solve: per-core completion flags + the ticket
    size_t syrk_plane_rows = 0;     // rows of the current chunk's plane tensor (largest batch)
    int n_sms = 148;
    int call_batches = 0;           // batches of the current searchlight call (index into reroute_counts)
    bool want_margins = false;
    int64_t n_rerouted = 0;
    int n_perm = 0;
    std::vector<int> int_of_test;   // position in test_ids order -> internal evaluation row (captions first)
    // knobs
    int cluster_size = 256, batch_centres = 6144, host_threads = 0, group_size = 8;
    // stats
    float stage_ms[MMD_N_STAGES] = {0};
    int64_t launches = 0, n_clusters = 0, n_batches = 0, n_core_tasks = 0, n_grouped = 0, sum_sp = 0, plan_us = 0,
            adj_h2d_us = 0;
    double syrk_flops = 0.0, solve_flops = 0.0;
    std::vector<cudaEvent_t> event_pool;
    size_t events_used = 0;
    std::vector<TimedSpan> spans;
};

namespace {

int fail(mmd_handle* h, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (h) h->err = buf;
    return 1;
}

#define CU(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t e__ = (call);                                                                        \
        if (e__ != cudaSuccess) return fail(h, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), \
                                            __FILE__, __LINE__);                                         \
    } while (0)

bool is_device_ptr(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

cudaEvent_t next_event(mmd_handle* h) {
    if (h->events_used == h->event_pool.size()) {
        cudaEvent_t e;
        cudaEventCreate(&e);
        h->event_pool.push_back(e);
    }
    return h->event_pool[h->events_used++];
}
void span_begin(mmd_handle* h, int stage) {
    TimedSpan s{next_event(h), next_event(h), stage};
    cudaEventRecord(s.a, h->stream);
    h->spans.push_back(s);
}
void span_end(mmd_handle* h, int n_launches = 1) {
    cudaEventRecord(h->spans.back().b, h->stream);
    h->launches += n_launches;
}
// call after the stream is synchronised
void resolve_spans(mmd_handle* h) {
    for (auto& s : h->spans) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, s.a, s.b) == cudaSuccess) h->stage_ms[s.stage] += ms;
    }
    h->spans.clear();
    h->events_used = 0;
    cudaGetLastError();
}

template <typename T>
int to_device(mmd_handle* h, DevBuf& buf, const T* src, size_t n) {
    CU(buf.ensure(std::max<size_t>(n, 1) * sizeof(T)));
    if (n) CU(cudaMemcpyAsync(buf.p, src, n * sizeof(T), cudaMemcpyDefault, h->stream));
    return 0;
}

// ------------------------------------------------------------------------------------------
// host plan: spatial clusters of centres and the sorted union of their neighbourhoods
// ------------------------------------------------------------------------------------------
struct Plan {
    std::vector<int> cl_off;      // cluster c owns members[cl_off[c] .. cl_off[c+1])
    std::vector<int> members;     // centre (adjacency row) indices
    std::vector<long> un_off;     // union of cluster c = un_vert[un_off[c] .. un_off[c+1])
    std::vector<int> un_vert;     // ascending vertex ids
    // prefix-sharing groups: nearby centres of one cluster whose neighbourhoods share a core of s_p vertices
    std::vector<int> cl_grp_off;  // cluster c owns groups [cl_grp_off[c], cl_grp_off[c+1])
    std::vector<int> grp_off;     // group g owns members[grp_off[g] .. grp_off[g+1]) (groups never straddle clusters)
    std::vector<int> grp_sp;      // shared prefix length of group g: multiple of 64, 0 = none
    std::vector<long> core_off;   // core of group g = core_vert[core_off[g] .. + grp_sp[g]), ascending vertex ids
    std::vector<int> core_vert;
};

// Neighbourhood lists of one call: rectangular (n x k, the k-NN branch of get_adjacency_matrix,
// searchlight.py:173) or ragged CSR (radius branch, searchlight.py:164-166).
struct Rows {
    const int32_t* data;
    const int64_t* off;        // nullptr = rectangular
    int k;                     // rectangular row length
    const int32_t* anchors;    // optional: the vertex a centre sits on (used only to group centres spatially)
    int n;
    long row_off(int i) const { return off ? (long)off[i] : (long)i * k; }
    const int32_t* row(int i) const { return data + row_off(i); }
    int len(int i) const { return off ? (int)(off[i + 1] - off[i]) : k; }
    int anchor(int i) const { return anchors ? anchors[i] : row(i)[0]; }
    long total() const { return off ? (long)off[n] : (long)n * k; }
};

constexpr int MIN_SHARED_PREFIX = 128;

// Spatial clusters of the centres of one call (cheap, sequential): cluster c owns members[cl_off[c] .. cl_off[c+1]).
struct Clustering {
    std::vector<int> centre_of;   // vertex -> the centre sitting on it (-1: none)
    std::vector<int> cl_off, members;
};

}  // namespace
// The host plan depends only on the adjacency (and the grouping knobs): the same adjacency serves every training mode and
// feature model of a subject-hemisphere (searchlight.py:186-236 loops over them), so the plan of the last adjacency is kept
// and reused when a call presents the same one (two independent 64-bit hashes over the staged int32 copy + its shape).
struct PlanCache {
    bool valid = false;
    uint64_t h1 = 0, h2 = 0;
    long total = 0;
    int n = 0, k = 0, V = 0, cluster_size = 0, group_size = 0, batch_cap = 0, ragged = 0;
    bool adj_on_device = false;
    Clustering cl;
    std::vector<int> chunk_start;
    std::vector<Plan> plans;
    // the unions and shared cores of every chunk are page-locked where they lie, so that their copies to the device are
    // asynchronous (a pageable source makes cudaMemcpyAsync wait for the stream first: the GPU would idle at every chunk)
    std::vector<void*> registered;
    void pin(void* p, size_t bytes) {
        if (p && bytes && cudaHostRegister(p, bytes, cudaHostRegisterDefault) == cudaSuccess) registered.push_back(p);
        else cudaGetLastError();
    }
    void unpin_all() {
        for (void* p : registered) cudaHostUnregister(p);
        registered.clear();
        cudaGetLastError();
    }
    ~PlanCache() { unpin_all(); }
};
namespace {

void cluster_centres(const Rows& rows, int V, int cluster_size, Clustering& cl) {
    // greedy region growing: a seed takes its nearest still-unassigned centres (k-NN adjacency rows are ordered
    // by distance in the reference, the centre itself first; radius rows are in vertex order and scanned whole)
    const int n_centres = rows.n;
    cl.centre_of.assign(V, -1);
    for (int i = n_centres - 1; i >= 0; --i) {
        const int v = rows.anchor(i);
        if (v >= 0 && v < V) cl.centre_of[v] = i;
    }
    std::vector<char> assigned(n_centres, 0);
    cl.cl_off.assign(1, 0);
    cl.members.clear();
    cl.members.reserve(n_centres);
    for (int s = 0; s < n_centres; ++s) {
        if (assigned[s]) continue;
        assigned[s] = 1;
        cl.members.push_back(s);
        int count = 1;
        const int32_t* row = rows.row(s);
        const int reach = rows.off ? rows.len(s) : std::min(rows.k, 3 * cluster_size);
        for (int j = 0; j < reach && count < cluster_size; ++j) {
            const int v = row[j];
            if (v < 0 || v >= V) continue;
            const int c = cl.centre_of[v];
            if (c >= 0 && !assigned[c]) {
                assigned[c] = 1;
                cl.members.push_back(c);
                ++count;
            }
        }
        cl.cl_off.push_back((int)cl.members.size());
    }
}

// The plan of clusters [c_begin, c_end) of a clustering, with cluster indices local to that range: sorted unions and
// prefix-sharing groups.  This is the expensive host step (every neighbour list is walked twice); run_searchlight
// builds the plan of the next range while the GPU works on the current one.
void build_plan(const Rows& rows, int V, const Clustering& cl, int c_begin, int c_end, int group_size, int n_threads,
                Plan& plan) {
    const int n_centres = rows.n;
    const std::vector<int>& centre_of = cl.centre_of;
    const int m0 = cl.cl_off[c_begin];
    plan.cl_off.resize(c_end - c_begin + 1);
    for (int c = c_begin; c <= c_end; ++c) plan.cl_off[c - c_begin] = cl.cl_off[c] - m0;
    plan.members.assign(cl.members.begin() + m0, cl.members.begin() + cl.cl_off[c_end]);
    const int n_clusters = c_end - c_begin;
    // unions and prefix-sharing groups, in parallel over clusters
    struct ClusterGroups {
        std::vector<int> order;       // the cluster's members, groups contiguous
        std::vector<int> size, sp;    // per group
        std::vector<int> core;        // concatenated cores (sp entries each)
    };
    std::vector<std::vector<int>> unions(n_clusters);
    std::vector<ClusterGroups> cgroups(n_clusters);
    const bool grouping = !rows.off && group_size >= 2 && rows.k > MIN_SHARED_PREFIX;
    n_threads = std::max(1, std::min(n_threads, n_clusters));
    std::atomic<int> next(0);
    auto worker = [&]() {
        std::vector<int> stamp(V, -1);
        std::vector<int> member_of, grouped, vgen, cur, nxt;
        if (grouping) {
            member_of.assign(n_centres, -1);
            grouped.assign(n_centres, 0);
            vgen.assign(V, -1);
        }
        int gen = 0;
        for (;;) {
            const int c = next.fetch_add(1);
            if (c >= n_clusters) break;
            if (grouping) {
                // Greedy groups: a seed takes its nearest not yet grouped centres of the same cluster; the core is the
                // intersection of the members' neighbourhoods, cut to a multiple of 64 (any subset may be shared).
                ClusterGroups& cg = cgroups[c];
                for (int m = plan.cl_off[c]; m < plan.cl_off[c + 1]; ++m) member_of[plan.members[m]] = c;
                const int kp = round_up(rows.k, NB);
                for (int m = plan.cl_off[c]; m < plan.cl_off[c + 1]; ++m) {
                    const int s0 = plan.members[m];
                    if (grouped[s0]) continue;
                    grouped[s0] = 1;
                    const size_t first = cg.order.size();
                    cg.order.push_back(s0);
                    const int32_t* row = rows.row(s0);
                    const int scan = std::min(rows.k, 4 * group_size);
                    // The running intersection starts as the seed's neighbourhood; a candidate joins only if the
                    // per-centre factor/solve cost of the (larger group, smaller core) pair is lower.
                    cur.clear();
                    for (int j = 0; j < rows.k; ++j)
                        if (row[j] >= 0 && row[j] < V) cur.push_back(row[j]);
                    auto prefix_of = [&](size_t n_core) {
                        const int sp = std::min((int)(n_core / NB) * NB, kp - NB);
                        return sp < MIN_SHARED_PREFIX ? 0 : sp;
                    };
                    auto cost = [&](int g, int sp) {       // flops per centre up to the (unshared) backward solve
                        const double k3 = (double)kp * kp * kp, s3 = (double)sp * sp * sp;
                        const double k2 = (double)kp * kp, s2 = (double)sp * sp;
                        return (k3 - s3) / 3.0 + s3 / (3.0 * g) + EVAL_PAD * (k2 - s2) + EVAL_PAD * s2 / g;
                    };
                    int g = 1, sp = 0;
                    for (int j = 1; j < scan && g < group_size; ++j) {
                        const int v = row[j];
                        if (v < 0 || v >= V) continue;
                        const int cc = centre_of[v];
                        if (!(cc >= 0 && member_of[cc] == c && !grouped[cc] && rows.anchor(cc) == v)) continue;
                        ++gen;
                        const int32_t* r = rows.row(cc);
                        for (int q = 0; q < rows.k; ++q)
                            if (r[q] >= 0 && r[q] < V) vgen[r[q]] = gen;
                        nxt.clear();
                        for (int x : cur)
                            if (vgen[x] == gen) nxt.push_back(x);
                        const int sp_new = prefix_of(nxt.size());
                        if (sp_new == 0 || (g > 1 && cost(g + 1, sp_new) >= cost(g, sp))) continue;
                        grouped[cc] = 1;
                        cg.order.push_back(cc);
                        cur.swap(nxt);
                        sp = sp_new;
                        ++g;
                    }
                    if (g >= 2) {
                        std::sort(cur.begin(), cur.end());
                        cg.core.insert(cg.core.end(), cur.begin(), cur.begin() + sp);
                    } else {
                        sp = 0;
                    }
                    cg.size.push_back(g);
                    cg.sp.push_back(sp);
                }
            }
            std::vector<int>& u = unions[c];
            for (int m = plan.cl_off[c]; m < plan.cl_off[c + 1]; ++m) {
                const int32_t* row = rows.row(plan.members[m]);
                const int k = rows.len(plan.members[m]);
                for (int j = 0; j < k; ++j) {
                    const int v = row[j];
                    if (v < 0 || v >= V) continue;       // reported by the device-side check
                    if (stamp[v] != c) {
                        stamp[v] = c;
                        u.push_back(v);
                    }
                }
            }
            std::sort(u.begin(), u.end());
        }
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < n_threads; ++t) pool.emplace_back(worker);
    worker();
    for (auto& t : pool) t.join();
    plan.un_off.assign(1, 0);
    size_t total = 0;
    for (auto& u : unions) total += u.size();
    plan.un_vert.clear();
    plan.un_vert.reserve(total);
    for (auto& u : unions) {
        plan.un_vert.insert(plan.un_vert.end(), u.begin(), u.end());
        plan.un_off.push_back((long)plan.un_vert.size());
    }
    // groups: reorder the members of every cluster so that groups are contiguous
    plan.cl_grp_off.assign(1, 0);
    plan.grp_off.assign(1, 0);
    plan.grp_sp.clear();
    plan.core_off.clear();
    plan.core_vert.clear();
    for (int c = 0; c < n_clusters; ++c) {
        if (grouping) {
            const ClusterGroups& cg = cgroups[c];
            std::copy(cg.order.begin(), cg.order.end(), plan.members.begin() + plan.cl_off[c]);
            size_t core_pos = 0;
            for (size_t g = 0; g < cg.size.size(); ++g) {
                plan.grp_off.push_back(plan.grp_off.back() + cg.size[g]);
                plan.grp_sp.push_back(cg.sp[g]);
                plan.core_off.push_back((long)plan.core_vert.size());
                plan.core_vert.insert(plan.core_vert.end(), cg.core.begin() + core_pos, cg.core.begin() + core_pos + cg.sp[g]);
                core_pos += cg.sp[g];
            }
        } else {
            plan.grp_off.push_back(plan.cl_off[c + 1]);          // one group without a shared prefix
            plan.grp_sp.push_back(0);
            plan.core_off.push_back((long)plan.core_vert.size());
        }
        plan.cl_grp_off.push_back((int)plan.grp_sp.size());
    }
}

constexpr int MAX_CALL_BATCHES = 4096;
constexpr double I8_EPS_UNIT = 7.2 / 4398046511104.0;      // 7.2 * 2^-42, see score.cu

// X^T Y on tcgen05, exactly: the float16 betas and latents cut into six 7-bit digits, the digit-pair products accumulated in
// int32 (the rectangular mode of syrk_i8.cu).  40 TOP instead of 1.35 TFLOP on the float64 pipe.
int run_xty_i8(mmd_handle* h) {
    const int M = h->n_active ? h->n_active : h->V, N = h->Dp;
    const int upA = round_up(M, 128), upB = round_up(N, 128);
    const long plane_rows = 6L * upA + 6L * upB;
    CU(h->syrk_planes.ensure((size_t)plane_rows * h->ldt));
    h->syrk_plane_rows = h->syrk_planes.bytes / h->ldt;
    CU(h->syrk_scales.ensure((size_t)(upA + upB) * sizeof(double)));
    CU(h->syrk_flags.ensure(2 * sizeof(int)));
    if (!h->syrk_status.p) {
        CU(h->syrk_status.ensure(sizeof(int)));
        CU(cudaMemsetAsync(h->syrk_status.p, 0, sizeof(int), h->stream));
    }
    SyrkCluster cl[2] = {};
    cl[0].row_base = 0;
    cl[0].u = M;
    cl[0].up = upA;
    cl[0].rows = h->n_active ? h->active.as<int>() : nullptr;
    cl[0].scales = h->syrk_scales.as<double>();
    cl[0].inexact = h->syrk_flags.as<int>();
    cl[0].C = h->XtY.as<double>();
    cl[0].ldc = h->Dp;
    cl[0].u_b = N;
    cl[0].scales_b = h->syrk_scales.as<double>() + upA;
    cl[0].out_rows = cl[0].rows;                     // a vertex subset lands at its vertex rows
    cl[1] = cl[0];                                   // the latents: rows of Yt
    cl[1].row_base = 6L * upA;
    cl[1].u = N;
    cl[1].up = upB;
    cl[1].rows = nullptr;
    cl[1].scales = h->syrk_scales.as<double>() + upA;
    cl[1].inexact = h->syrk_flags.as<int>() + 1;
    std::vector<SyrkItem> items;
    items.reserve((size_t)(upA / 128) * (upB / 128));
    for (int m0 = 0; m0 < upA; m0 += 128)
        for (int n0 = 0; n0 < upB; n0 += 128)
            items.push_back(SyrkItem{m0, (int)(6L * upA) + n0, upA, 0, m0, n0, upB});
    const size_t cl_bytes = 2 * sizeof(SyrkCluster), it_bytes = items.size() * sizeof(SyrkItem);
    CU(h->tmp.ensure(cl_bytes + it_bytes));
    CU(cudaMemcpyAsync(h->tmp.p, cl, cl_bytes, cudaMemcpyHostToDevice, h->stream));
    CU(cudaMemcpyAsync(static_cast<char*>(h->tmp.p) + cl_bytes, items.data(), it_bytes, cudaMemcpyHostToDevice, h->stream));
    CU(cudaStreamSynchronize(h->stream));            // `cl` and `items` are pageable locals
    const SyrkCluster* cl_dev = h->tmp.as<SyrkCluster>();
    span_begin(h, ST_XTY);
    CU(cudaMemsetAsync(h->syrk_flags.p, 0, 2 * sizeof(int), h->stream));
    CU(launch_slice_union(h->Xt.as<__half>(), h->ldt, cl_dev, 1, M, h->syrk_planes.as<int8_t>(), h->stream));
    CU(launch_slice_union(h->Yt.as<__half>(), h->ldt, cl_dev + 1, 1, N, h->syrk_planes.as<int8_t>(), h->stream));
    CU(launch_syrk_i8(h->syrk_planes.as<int8_t>(), plane_rows, h->ldt,
                      reinterpret_cast<const SyrkItem*>(static_cast<char*>(h->tmp.p) + cl_bytes), (int)items.size(), cl_dev, h->n_sms,
                      h->syrk_status.as<int>(), h->stream));
    span_end(h, 3);
    return 0;
}

int run_xty(mmd_handle* h) {
    if (h->syrk_mode == 1 && h->ldt % 64 == 0 && (long)h->V * 6 < 0x7fffff00L / 2) {
        if (run_xty_i8(h)) return 1;
    } else {
        GemmProblem p{};
        p.A = h->Xt.as<__half>();
        p.lda = h->ldt;
        p.rowsA = h->n_active ? h->active.as<int>() : nullptr;      // a vertex subset: rows land at their vertex index
        p.scatter_c = h->n_active ? 1 : 0;
        p.M = h->n_active ? h->n_active : h->V;
        p.B = h->Yt.as<__half>();
        p.ldb = h->ldt;
        p.rowsB = nullptr;
        p.N = h->Dp;
        p.K = h->ldt;
        p.C = h->XtY.as<double>();
        p.ldc = h->Dp;
        p.lower_only = 0;
        if (to_device(h, h->probs, &p, 1)) return 1;
        span_begin(h, ST_XTY);
        CU(launch_gemm_gather(h->probs.as<GemmProblem>(), 1, p.M, p.N, h->stream));
        span_end(h);
    }
    if (h->predict_mode == 1) {
        // digit planes of the cross-covariance for the integer prediction GEMM (once per upload)
        CU(h->xplanes.ensure((size_t)I8_DIGITS * h->V * h->Dp));
        CU(h->sX.ensure((size_t)h->Dp * sizeof(double)));
        CU(h->colmax.ensure((size_t)h->Dp * sizeof(unsigned long long)));
        span_begin(h, ST_XTY);
        CU(launch_slice_xty(h->XtY.as<double>(), h->V, h->Dp, h->xplanes.as<int8_t>(), h->sX.as<double>(),
                            h->colmax.as<unsigned long long>(), h->n_active ? h->active.as<int>() : nullptr, h->n_active,
                            h->stream));
        span_end(h, 3);
    }
    return 0;
}

// Cross-validation part of a batch (after prep and the full union Grams): fold Grams, then for every fold, alpha and
// chunk of validation trials: factor + solve without the fold, raw predictions, pairwise-accuracy counts.
int run_batch_cv(mmd_handle* h, const SolveParams& base, const std::vector<GemmProblem>& probs, int n_tasks, int n_core, int kp,
                 int sp_max, int max_u) {
    CvState& cv = *h->cv;
    const int n_cl = (int)probs.size(), n_alphas = (int)cv.alphas.size();
    const double* pool0 = h->gram_pool.as<double>();
    // fold Grams of every cluster union, same layout as the full Gram pool
    CU(cv.probs.ensure((size_t)cv.n_folds * n_cl * sizeof(GemmProblem)));
    std::vector<GemmProblem> fp((size_t)cv.n_folds * n_cl);
    for (int f = 0; f < cv.n_folds; ++f)
        for (int c = 0; c < n_cl; ++c) {
            GemmProblem p = probs[c];
            p.only_if = nullptr;                             // fold Grams always run on the float64 kernel
            p.A = p.B = cv.xfold[f].as<__half>();
            p.lda = p.ldb = cv.ld[f];
            p.K = cv.ld[f];
            p.C = cv.pool.as<double>() + (size_t)f * cv.pool_stride + (probs[c].C - pool0);
            fp[(size_t)f * n_cl + c] = p;
        }
    CU(cudaMemcpyAsync(cv.probs.p, fp.data(), fp.size() * sizeof(GemmProblem), cudaMemcpyDefault, h->stream));
    span_begin(h, ST_SYRK);
    for (int f = 0; f < cv.n_folds; ++f)
        CU(launch_gemm_gather(cv.probs.as<GemmProblem>() + (size_t)f * n_cl, n_cl, max_u, max_u, h->stream));
    span_end(h, cv.n_folds);

    PredictParams pp{};
    pp.tasks = base.tasks;
    pp.n_tasks = n_tasks;
    pp.kp_max = kp;
    pp.work = h->work.as<double>();
    pp.Dp = h->Dp;
    pp.yhat = h->yhat.as<double>();
    pp.standardize = 0;
    CvScoreParams cs{};
    cs.tasks = base.tasks;
    cs.n_tasks = n_tasks;
    cs.yhat = h->yhat.as<double>();
    cs.Dp = h->Dp;
    cs.counts = cv.counts.as<unsigned long long>();
    cs.count_stride = (long)n_alphas * cv.n_folds;
    for (int f = 0; f < cv.n_folds; ++f) {
        const int n_val = cv.hi[f] - cv.lo[f];
        pp.XtY = cv.xty_tr[f].as<double>();
        cs.cand = cv.cand[f].as<double>();
        cs.cand_sq = cv.cand_sq[f].as<double>();
        cs.n_cand = n_val;
        for (int a = 0; a < n_alphas; ++a) {
            cs.count_off = a * cv.n_folds + f;
            for (size_t c = 0; c < cv.xe[f].size(); ++c) {
                SolveParams sp = base;
                sp.alpha = cv.alphas[a];
                sp.Xe = cv.xe[f][c].as<__half>();
                sp.gram_sub_off = (long)((cv.pool.as<double>() + (size_t)f * cv.pool_stride) - pool0);
                span_begin(h, ST_SOLVE);
                if (n_core > 0) {
                    CU(cudaMemsetAsync(h->group_flags.p, 0, (size_t)n_core * sizeof(int), h->stream));
                    CU(cudaMemsetAsync(h->core_done.p, 0, (size_t)(n_core + 1) * sizeof(int), h->stream));
                }
                CU(launch_factor_solve(sp, h->stream));            // cores and centres in one launch
                span_end(h, 1);
                span_begin(h, ST_PREDICT);
                CU(launch_predict(pp, h->stream));
                span_end(h);
                cs.row0 = (int)c * EVAL_PAD;
                cs.n_rows = std::min(EVAL_PAD, n_val - cs.row0);
                span_begin(h, ST_SCORE);
                CU(launch_cv_score(cs, h->stream));
                span_end(h);
            }
        }
    }
    return 0;
}

// One batch = clusters [c0, c1).  `slot0` = first output slot when yhat_only (diagnostic) is set.
int run_batch(mmd_handle* h, const Rows& rows, const Plan& plan, int c0, int c1, int k_max, int kp, int sp_max,
              double alpha, bool with_null, bool raw_yhat) {
    const int n_cl = c1 - c0;
    const int t0 = plan.cl_off[c0], n_tasks = plan.cl_off[c1] - t0;
    // core tasks (one per group with a shared prefix) follow the centre tasks in every per-task array
    int n_core = 0;
    for (int g = plan.cl_grp_off[c0]; g < plan.cl_grp_off[c1]; ++g) n_core += plan.grp_sp[g] > 0;
    const int n_all = n_tasks + n_core;
    const long adj_total = rows.total();                    // the cores sit behind the adjacency rows on the device
    const size_t core_stride = (size_t)(sp_max + EVAL_PAD) * sp_max;
    std::vector<int> task_k(n_all), task_un_n(n_all), task_core_n(n_all, 0);
    std::vector<long> task_row_off(n_all), task_un_off(n_all), task_core_off(n_all, 0);
    std::vector<CentreTask> tasks(n_all);
    std::vector<GemmProblem> probs(n_cl);
    size_t pool_doubles = 0;
    int max_u = 0, q = 0;
    for (int c = c0; c < c1; ++c) {
        const int u = (int)(plan.un_off[c + 1] - plan.un_off[c]);
        const long ldu = round_up(u, 8);
        GemmProblem& p = probs[c - c0];
        p.A = p.B = h->Xt.as<__half>();
        p.lda = p.ldb = h->ldt;
        p.rowsA = p.rowsB = h->union_vertices.as<int>() + plan.un_off[c];
        p.M = p.N = u;
        p.K = h->ldt;
        p.C = h->gram_pool.as<double>() + pool_doubles;     // pool base fixed before this call
        p.ldc = ldu;
        p.lower_only = 1;
        for (int g = plan.cl_grp_off[c]; g < plan.cl_grp_off[c + 1]; ++g) {
            const int sp = plan.grp_sp[g];
            const double* core_work = nullptr;
            const int* core_flag = nullptr;
            const int* core_done = nullptr;
            if (sp > 0) {
                const int t = n_tasks + q;
                task_row_off[t] = adj_total + plan.core_off[g];
                task_k[t] = sp;
                task_un_off[t] = plan.un_off[c];
                task_un_n[t] = u;
                CentreTask& ct = tasks[t];
                ct.gram = p.C;
                ct.gram_ld = ldu;
                ct.loc = h->loc.as<int>() + (long)t * kp;
                ct.nbr = h->nbr_sorted.as<int>() + (long)t * kp;
                ct.centre = q;
                ct.k = ct.kp = sp;
                ct.s_p = 0;
                ct.core_work = nullptr;
                ct.core_flag = nullptr;
                core_work = h->core_work.as<double>() + (size_t)q * core_stride;
                core_flag = h->group_flags.as<int>() + q;
                core_done = h->core_done.as<int>() + q;
                h->solve_flops += (double)sp * sp * sp / 3.0 + (double)EVAL_PAD * sp * sp;
                ++q;
            }
            for (int m = plan.grp_off[g]; m < plan.grp_off[g + 1]; ++m) {
                const int t = m - t0;
                task_row_off[t] = rows.row_off(plan.members[m]);
                task_k[t] = rows.len(plan.members[m]);
                task_un_off[t] = plan.un_off[c];
                task_un_n[t] = u;
                task_core_off[t] = adj_total + plan.core_off[g];
                task_core_n[t] = sp;
                CentreTask& ct = tasks[t];
                ct.gram = p.C;
                ct.gram_ld = ldu;
                ct.loc = h->loc.as<int>() + (long)t * kp;
                ct.nbr = h->nbr_sorted.as<int>() + (long)t * kp;
                ct.centre = plan.members[m];
                ct.k = task_k[t];
                ct.kp = round_up(task_k[t], NB);
                ct.s_p = sp;
                ct.core_work = core_work;
                ct.core_flag = core_flag;
                ct.core_done = core_done;
                const double kpd = ct.kp, spd = sp;
                h->solve_flops += (kpd * kpd * kpd - spd * spd * spd) / 3.0 + EVAL_PAD * (kpd * kpd - spd * spd) +
                                  EVAL_PAD * kpd * kpd;
            }
        }
        pool_doubles += (size_t)u * ldu;
        max_u = std::max(max_u, u);
        // executed flops of the lower-triangular 128 x 64 tiles
        const long tm = (u + 127) / 128;
        long tiles = 0;
        for (long i = 0; i < tm; ++i) tiles += std::min<long>((i * 128 + 127) / 64 + 1, (u + 63) / 64);
        h->syrk_flops += 2.0 * tiles * 128.0 * 64.0 * h->ldt;
    }
    h->n_batches += 1;
    h->n_core_tasks += n_core;
    if (pool_doubles * sizeof(double) > h->gram_pool.bytes) return fail(h, "internal: gram pool too small");
    // one packed block per batch: pinned staging (double-buffered) -> one asynchronous copy
    size_t blk = 0;
    auto place = [&](size_t bytes) {
        const size_t o = blk;
        blk = (blk + bytes + 255) & ~(size_t)255;
        return o;
    };
    const size_t o_row = place((size_t)n_all * sizeof(long)), o_k = place((size_t)n_all * sizeof(int)),
                 o_unoff = place((size_t)n_all * sizeof(long)), o_unn = place((size_t)n_all * sizeof(int)),
                 o_coff = place((size_t)n_all * sizeof(long)), o_cn = place((size_t)n_all * sizeof(int)),
                 o_tasks = place((size_t)n_all * sizeof(CentreTask)), o_probs = place((size_t)n_cl * sizeof(GemmProblem));
    // exact integer SYRK (syrk_i8.cu): per cluster six digit planes of `up` rows, 128 x 128 lower-triangle tiles
    const bool syrk_i8 = h->syrk_mode == 1 && h->ldt % 64 == 0;
    std::vector<SyrkCluster> sy_cl;
    std::vector<SyrkItem> sy_items;
    long plane_rows = 0;
    if (syrk_i8) {
        long scale_off = 0;
        sy_cl.resize(n_cl);
        for (int c = 0; c < n_cl; ++c) {
            const int u = probs[c].M, up = round_up(u, 128);
            SyrkCluster& sc = sy_cl[c];
            sc.row_base = plane_rows;
            sc.u = u;
            sc.up = up;
            sc.rows = probs[c].rowsA;
            sc.scales = h->syrk_scales.as<double>() + scale_off;
            sc.inexact = h->syrk_flags.as<int>() + c;
            sc.C = probs[c].C;
            sc.ldc = probs[c].ldc;
            sc.u_b = 0;
            sc.scales_b = nullptr;
            sc.out_rows = nullptr;
            probs[c].only_if = sc.inexact;                 // the float64 kernel redoes a cluster that was not exact
            for (int m0 = 0; m0 < up; m0 += 128)
                for (int n0 = 0; n0 <= m0; n0 += 128)
                    sy_items.push_back(SyrkItem{(int)(plane_rows + m0), (int)(plane_rows + n0), up, c, m0, n0, up});
            plane_rows += 6L * up;
            scale_off += up;
        }
        if ((size_t)plane_rows > h->syrk_plane_rows) return fail(h, "internal: digit-plane buffer too small");
    }
    const size_t o_sycl = place(sy_cl.size() * sizeof(SyrkCluster)), o_syit = place(sy_items.size() * sizeof(SyrkItem));
    const int half = h->batch_half;
    h->batch_half ^= 1;
    if (!h->batch_copied[half]) CU(cudaEventCreateWithFlags(&h->batch_copied[half], cudaEventDisableTiming));
    else CU(cudaEventSynchronize(h->batch_copied[half]));                // the copy that last used this half has landed
    CU(h->batch_host[half].ensure(blk));
    CU(h->batch_dev.ensure(blk));
    {
        char* hp = static_cast<char*>(h->batch_host[half].p);
        memcpy(hp + o_row, task_row_off.data(), (size_t)n_all * sizeof(long));
        memcpy(hp + o_k, task_k.data(), (size_t)n_all * sizeof(int));
        memcpy(hp + o_unoff, task_un_off.data(), (size_t)n_all * sizeof(long));
        memcpy(hp + o_unn, task_un_n.data(), (size_t)n_all * sizeof(int));
        memcpy(hp + o_coff, task_core_off.data(), (size_t)n_all * sizeof(long));
        memcpy(hp + o_cn, task_core_n.data(), (size_t)n_all * sizeof(int));
        memcpy(hp + o_tasks, tasks.data(), (size_t)n_all * sizeof(CentreTask));
        memcpy(hp + o_probs, probs.data(), (size_t)n_cl * sizeof(GemmProblem));
        if (!sy_cl.empty()) memcpy(hp + o_sycl, sy_cl.data(), sy_cl.size() * sizeof(SyrkCluster));
        if (!sy_items.empty()) memcpy(hp + o_syit, sy_items.data(), sy_items.size() * sizeof(SyrkItem));
    }
    CU(cudaMemcpyAsync(h->batch_dev.p, h->batch_host[half].p, blk, cudaMemcpyHostToDevice, h->stream));
    CU(cudaEventRecord(h->batch_copied[half], h->stream));
    char* dp = static_cast<char*>(h->batch_dev.p);
    const CentreTask* tasks_dev = reinterpret_cast<const CentreTask*>(dp + o_tasks);
    const GemmProblem* probs_dev = reinterpret_cast<const GemmProblem*>(dp + o_probs);

    span_begin(h, ST_PREP);
    CU(launch_prep_tasks(h->adj.as<int>(), reinterpret_cast<const long*>(dp + o_row), reinterpret_cast<const int*>(dp + o_k), k_max, kp,
                         reinterpret_cast<const long*>(dp + o_unoff), reinterpret_cast<const int*>(dp + o_unn),
                         h->union_vertices.as<int>(), reinterpret_cast<const long*>(dp + o_coff),
                         reinterpret_cast<const int*>(dp + o_cn), h->V,
                         h->vertex_bad.as<unsigned char>(), n_all, h->nbr_sorted.as<int>(), h->loc.as<int>(), h->misc_flag.as<int>(), h->stream));
    span_end(h);

    span_begin(h, ST_SYRK);
    if (syrk_i8) {
        const SyrkCluster* sycl_dev = reinterpret_cast<const SyrkCluster*>(dp + o_sycl);
        CU(cudaMemsetAsync(h->syrk_flags.p, 0, (size_t)n_cl * sizeof(int), h->stream));
        CU(launch_slice_union(h->Xt.as<__half>(), h->ldt, sycl_dev, n_cl, max_u, h->syrk_planes.as<int8_t>(), h->stream));
        CU(launch_syrk_i8(h->syrk_planes.as<int8_t>(), plane_rows, h->ldt, reinterpret_cast<const SyrkItem*>(dp + o_syit),
                          (int)sy_items.size(), sycl_dev, h->n_sms, h->syrk_status.as<int>(), h->stream));
    }
    CU(launch_gemm_gather(probs_dev, n_cl, max_u, max_u, h->stream));      // with syrk_i8: only the clusters flagged inexact
    span_end(h, syrk_i8 ? 3 : 1);

    SolveParams sp{};
    sp.tasks = tasks_dev;
    sp.n_tasks = n_tasks;
    sp.kp_max = kp;
    sp.alpha = alpha;
    sp.Xe = h->Xe.as<__half>();
    sp.work = h->work.as<double>();
    sp.flags = h->flags.as<int>();
    sp.core_only = 0;
    sp.gram_sub_off = 0;
    sp.n_core = n_core;
    sp.core_tasks = sp.tasks + n_tasks;
    sp.core_work = h->core_work.as<double>();
    sp.core_kp_max = sp_max;
    sp.core_flags = h->group_flags.as<int>();
    sp.core_done = h->core_done.as<int>();
    sp.ticket = h->core_done.as<int>() + n_core;
    if (h->cv) return run_batch_cv(h, sp, probs, n_tasks, n_core, kp, sp_max, max_u);
    // one merged launch: the groups' shared cores (factorisation + forward solve) and the centres extending them
    sp.n_core = n_core;
    sp.core_tasks = sp.tasks + n_tasks;
    sp.core_work = h->core_work.as<double>();
    sp.core_kp_max = sp_max;
    sp.core_flags = h->group_flags.as<int>();
    sp.core_done = h->core_done.as<int>();
    sp.ticket = h->core_done.as<int>() + n_core;
    span_begin(h, ST_SOLVE);
    if (n_core > 0) {
        CU(cudaMemsetAsync(h->group_flags.p, 0, (size_t)n_core * sizeof(int), h->stream));
        CU(cudaMemsetAsync(h->core_done.p, 0, (size_t)(n_core + 1) * sizeof(int), h->stream));
    }
    CU(launch_factor_solve(sp, h->stream));
    span_end(h, 1);

    PredictParams pp{};
    pp.tasks = sp.tasks;
    pp.n_tasks = n_tasks;
    pp.kp_max = kp;
    pp.work = h->work.as<double>();
    pp.XtY = h->XtY.as<double>();
    pp.Dp = h->Dp;
    pp.yhat = h->yhat.as<double>();
    pp.n_cap = h->n_cap;
    pp.n_img = h->n_img;
    pp.n_im = h->n_im;
    pp.standardize = raw_yhat ? 0 : 1;
    pp.task_list = nullptr;
    pp.n_list = nullptr;

    ScoreParams sc{};
    sc.tasks = sp.tasks;
    sc.n_tasks = n_tasks;
    sc.yhat = h->yhat.as<double>();
    sc.Dp = h->Dp;
    sc.cand = h->cand64.as<double>();
    sc.cand_sq = h->cand_sq.as<double>();
    sc.n_cap = h->n_cap;
    sc.n_img = h->n_img;
    sc.n_im = h->n_im;
    sc.scores = h->scores.as<double>();
    sc.dist_out = with_null ? h->dist.as<double>() : nullptr;
    sc.margin_out = h->want_margins ? h->margins.as<double>() : nullptr;

    const bool i8 = h->predict_mode == 1 && !raw_yhat;
    if (!i8) {
        span_begin(h, ST_PREDICT);
        CU(launch_predict(pp, h->stream));
        span_end(h);
        span_begin(h, ST_SCORE);
        CU(launch_score(sc, h->stream));
        span_end(h);
    } else {
        // integer GEMM on the tcgen05 tensor cores; the score kernel lists the centres whose comparisons are too close to
        // the GEMM's error bound, and those are redone by the float64 kernels (device-side list, no host round trip)
        if (h->call_batches >= MAX_CALL_BATCHES) return fail(h, "internal: more than %d batches in one call", MAX_CALL_BATCHES);
        int* count = h->reroute_counts.as<int>() + h->call_batches++;
        PredictI8Params ip{};
        ip.tasks = sp.tasks;
        ip.n_tasks = n_tasks;
        ip.kp_max = kp;
        ip.V = h->V;
        ip.Dp = h->Dp;
        ip.sX = h->sX.as<double>();
        ip.sZ = h->sZ.as<double>();
        ip.yhat = h->yhat.as<double>();
        ip.amp2 = h->amp2.as<double>();
        ip.n_cap = h->n_cap;
        ip.n_img = h->n_img;
        ip.n_im = h->n_im;
        ip.standardize = 1;
        ip.status = h->i8_status.as<int>();
        span_begin(h, ST_PREDICT);
        CU(cudaMemsetAsync(h->amp2.p, 0, (size_t)n_tasks * 4 * sizeof(double), h->stream));
        CU(launch_slice_zt(sp.tasks, n_tasks, kp, h->work.as<double>(), h->zplanes.as<int8_t>(), h->sZ.as<double>(), h->stream));
        CU(launch_predict_i8(ip, h->xplanes.as<int8_t>(), h->zplanes.as<int8_t>(), h->n_sms, h->stream));
        span_end(h, 2);
        sc.amp2 = h->amp2.as<double>();
        sc.sZ = h->sZ.as<double>();
        sc.eps_unit = I8_EPS_UNIT;
        sc.reroute_list = h->reroute_list.as<int>();
        sc.reroute_count = count;
        span_begin(h, ST_SCORE);
        CU(launch_score(sc, h->stream));
        span_end(h);
        // second pass, float64, over the listed centres only (CTAs beyond the count leave at once)
        pp.task_list = h->reroute_list.as<int>();
        pp.n_list = count;
        sc.task_list = pp.task_list;
        sc.n_list = count;
        sc.amp2 = nullptr;
        sc.sZ = nullptr;
        sc.reroute_list = nullptr;
        sc.reroute_count = nullptr;
        span_begin(h, ST_REROUTE);
        CU(launch_predict(pp, h->stream));
        CU(launch_score(sc, h->stream));
        span_end(h, 2);
    }
    if (with_null) {
        NullParams np{};
        np.tasks = sp.tasks;
        np.n_tasks = n_tasks;
        np.dist = h->dist.as<double>();
        np.perm_test = h->perm_test.as<int>();
        np.perm_imagery = h->perm_im.as<int>();
        np.perm_imagery_centre = h->use_perm_im_centre ? h->perm_im_centre.as<int>() : nullptr;
        np.n_perm = h->n_perm;
        np.n_cap = h->n_cap;
        np.n_img = h->n_img;
        np.n_im = h->n_im;
        np.null_out = h->null_out.as<double>();
        span_begin(h, ST_NULL);
        CU(launch_null_score(np, h->stream));
        span_end(h);
    }
    return 0;
}

// One chunk of clusters: upload its unions and shared cores, size the workspaces, run its batches.
int run_chunk(mmd_handle* h, const Rows& rows, const Plan& plan, size_t total, int k_max, int kp, double alpha,
              bool with_null, bool raw_yhat, int batch_cap) {
    const int n_clusters = (int)plan.cl_off.size() - 1;
    if (to_device(h, h->union_vertices, plan.un_vert.data(), plan.un_vert.size())) return 1;
    if (plan.core_vert.size() > total / 2) return fail(h, "internal: shared cores exceed their bound");
    if (!plan.core_vert.empty())
        CU(cudaMemcpyAsync(h->adj.as<int>() + total, plan.core_vert.data(), plan.core_vert.size() * sizeof(int),
                           cudaMemcpyDefault, h->stream));
    int sp_max = 0;
    for (size_t g = 0; g < plan.grp_sp.size(); ++g) {
        sp_max = std::max(sp_max, plan.grp_sp[g]);
        if (plan.grp_sp[g] > 0) {
            h->n_grouped += plan.grp_off[g + 1] - plan.grp_off[g];
            h->sum_sp += (int64_t)plan.grp_sp[g] * (plan.grp_off[g + 1] - plan.grp_off[g]);
        }
    }

    // batches: whole clusters, about batch_cap centres each (a batch may overshoot by a cluster or two; the cap keeps
    // half of the device memory free), the chunk split evenly so that no small remainder batch runs on an empty GPU
    const int chunk_tasks = plan.cl_off[n_clusters];
    const int n_batches = std::max(1, (chunk_tasks + batch_cap - 1) / batch_cap);
    const int batch_target = (chunk_tasks + n_batches - 1) / n_batches;
    std::vector<int> batch_start(1, 0);
    size_t max_pool = 0, cur_pool = 0;
    int cur_tasks = 0, max_tasks = 0, cur_core = 0, max_core = 0;
    for (int c = 0; c < n_clusters; ++c) {
        const int m = plan.cl_off[c + 1] - plan.cl_off[c];
        int n_core_c = 0;
        for (int g = plan.cl_grp_off[c]; g < plan.cl_grp_off[c + 1]; ++g) n_core_c += plan.grp_sp[g] > 0;
        const size_t u = (size_t)(plan.un_off[c + 1] - plan.un_off[c]);
        const size_t g = u * round_up((int)u, 8);
        // close a batch at the cluster boundary nearest to the even split; the last batch takes whatever is left
        if (cur_tasks > 0 && (int)batch_start.size() < n_batches && cur_tasks + m / 2 > batch_target) {
            batch_start.push_back(c);
            cur_tasks = 0;
            cur_pool = 0;
            cur_core = 0;
        }
        cur_tasks += m;
        cur_pool += g;
        cur_core += n_core_c;
        max_core = std::max(max_core, cur_core);
        max_tasks = std::max(max_tasks, cur_tasks);
        max_pool = std::max(max_pool, cur_pool);
    }
    batch_start.push_back(n_clusters);

    // workspaces only grow; a growing buffer is freed first, which waits for the kernels still using it
    CU(h->gram_pool.ensure(max_pool * sizeof(double)));
    CU(h->work.ensure((size_t)max_tasks * (kp + EVAL_PAD) * kp * sizeof(double)));
    CU(h->yhat.ensure((size_t)max_tasks * YHAT_ROWS * h->Dp * sizeof(double)));
    CU(h->nbr_sorted.ensure((size_t)(max_tasks + max_core) * kp * sizeof(int)));
    CU(h->loc.ensure((size_t)(max_tasks + max_core) * kp * sizeof(int)));
    CU(h->core_work.ensure(std::max<size_t>((size_t)max_core * (sp_max + EVAL_PAD) * sp_max, 1) * sizeof(double)));
    CU(h->group_flags.ensure(std::max<size_t>(max_core, 1) * sizeof(int)));
    CU(h->core_done.ensure((size_t)(max_core + 1) * sizeof(int)));
    if (with_null) CU(h->dist.ensure((size_t)max_tasks * NULL_DIST_ROWS * EVAL_PAD * sizeof(double)));
    if (h->syrk_mode == 1 && h->ldt % 64 == 0) {
        size_t max_rows = 0, max_scales = 0, max_cl = 0;
        for (size_t b = 0; b + 1 < batch_start.size(); ++b) {
            size_t rows_b = 0, sc_b = 0;
            for (int c = batch_start[b]; c < batch_start[b + 1]; ++c) {
                const size_t up = (size_t)round_up((int)(plan.un_off[c + 1] - plan.un_off[c]), 128);
                rows_b += 6 * up;
                sc_b += up;
            }
            max_rows = std::max(max_rows, rows_b);
            max_scales = std::max(max_scales, sc_b);
            max_cl = std::max<size_t>(max_cl, batch_start[b + 1] - batch_start[b]);
        }
        CU(h->syrk_planes.ensure(max_rows * h->ldt));
        h->syrk_plane_rows = h->syrk_planes.bytes / h->ldt;
        CU(h->syrk_scales.ensure(std::max<size_t>(max_scales, 1) * sizeof(double)));
        CU(h->syrk_flags.ensure(std::max<size_t>(max_cl, 1) * sizeof(int)));
        if (!h->syrk_status.p) {
            CU(h->syrk_status.ensure(sizeof(int)));
            CU(cudaMemsetAsync(h->syrk_status.p, 0, sizeof(int), h->stream));
        }
    }
    if (h->cv) {
        h->cv->pool_stride = std::max(h->cv->pool_stride, max_pool);
        CU(h->cv->pool.ensure((size_t)h->cv->n_folds * h->cv->pool_stride * sizeof(double)));
    }
    if (h->predict_mode == 1 && !raw_yhat) {
        CU(h->zplanes.ensure((size_t)max_tasks * I8_DIGITS * EVAL_PAD * kp));
        CU(h->sZ.ensure((size_t)max_tasks * EVAL_PAD * sizeof(double)));
        CU(h->amp2.ensure((size_t)max_tasks * 4 * sizeof(double)));
        CU(h->reroute_list.ensure((size_t)max_tasks * sizeof(int)));
    }

    for (size_t b = 0; b + 1 < batch_start.size(); ++b)
        if (run_batch(h, rows, plan, batch_start[b], batch_start[b + 1], k_max, kp, sp_max, alpha, with_null, raw_yhat)) return 1;
    return 0;
}

int run_searchlight(mmd_handle* h, const Rows& rows_in, double alpha, bool with_null = false, bool raw_yhat = false,
                    const int64_t* data64 = nullptr) {
    if (!h->uploaded) return fail(h, "mmd_upload must be called before mmd_searchlight");
    const int n_centres = rows_in.n;
    int n_threads = h->host_threads > 0 ? h->host_threads : (int)std::thread::hardware_concurrency();
    n_threads = std::max(1, std::min(n_threads, 32));
    int k_max = 0;
    for (int i = 0; i < n_centres; ++i) {
        const int k = rows_in.len(i);
        if (k < 1 || k > h->V) return fail(h, "k=%d out of range (1..%d) for centre %d", k, h->V, i);
        if (k > h->n_train)
            return fail(h, "k=%d exceeds n_train=%d: the primal normal equations are rank deficient only through "
                           "alpha; the dual form is not implemented on this path", k, h->n_train);
        if (k > 2000) return fail(h, "k=%d exceeds the supported maximum of 2000", k);
        k_max = std::max(k_max, k);
        if (!rows_in.off) break;              // rectangular: one length
    }
    // Stage the adjacency in pinned memory as int32 (sklearn hands out int64 indices): the H2D copy then runs
    // asynchronously while the host groups the centres.
    const auto t_stage0 = std::chrono::steady_clock::now();
    const size_t total = (size_t)rows_in.total();
    CU(h->adj_host.ensure(std::max<size_t>(total, 1) * sizeof(int32_t)));
    uint64_t hash1 = 0xcbf29ce484222325ull, hash2 = 0x84222325cbf29ce4ull;
    {
        int32_t* dst = reinterpret_cast<int32_t*>(h->adj_host.p);
        const int32_t* src32 = rows_in.data;
        const int nt = (int)std::max<size_t>(1, std::min<size_t>(n_threads, total / (1 << 20) + 1));
        std::vector<uint64_t> part1(nt), part2(nt);
        auto conv = [&](int t) {
            const size_t lo = total * t / nt, hi = total * (t + 1) / nt;
            if (data64) {
                for (size_t i = lo; i < hi; ++i) {
                    const int64_t v = data64[i];
                    dst[i] = (v < 0 || v > 0x7fffffff) ? -1 : (int32_t)v;     // out of range: reported by the device check
                }
            } else {
                memcpy(dst + lo, src32 + lo, (hi - lo) * sizeof(int32_t));
            }
            // two independent 64-bit hashes of the segment; four interleaved lanes each so the multiplies pipeline
            uint64_t a[4], b[4];
            for (int j = 0; j < 4; ++j) {
                a[j] = 0x9e3779b97f4a7c15ull ^ (lo + j);
                b[j] = 0xc2b2ae3d27d4eb4full + 31 * (lo + j);
            }
            size_t i = lo;
            for (; i + 8 <= hi; i += 8)
                for (int j = 0; j < 4; ++j) {
                    const uint64_t v = (uint64_t)(uint32_t)dst[i + 2 * j] | ((uint64_t)(uint32_t)dst[i + 2 * j + 1] << 32);
                    a[j] = (a[j] ^ v) * 0x100000001b3ull;
                    b[j] = (b[j] + v) * 0xff51afd7ed558ccdull;
                    b[j] ^= b[j] >> 29;
                }
            for (; i < hi; ++i) {
                const uint64_t v = (uint32_t)dst[i];
                a[0] = (a[0] ^ v) * 0x100000001b3ull;
                b[0] = (b[0] + v) * 0xff51afd7ed558ccdull;
                b[0] ^= b[0] >> 29;
            }
            uint64_t ha = 0, hb = 0;
            for (int j = 0; j < 4; ++j) {
                ha = (ha ^ a[j]) * 0x100000001b3ull;
                hb = (hb + b[j]) * 0xff51afd7ed558ccdull;
            }
            part1[t] = ha;
            part2[t] = hb;
        };
        std::vector<std::thread> pool;
        for (int t = 1; t < nt; ++t) pool.emplace_back(conv, t);
        conv(0);
        for (auto& t : pool) t.join();
        for (int t = 0; t < nt; ++t) {
            hash1 = (hash1 ^ part1[t]) * 0x100000001b3ull;
            hash2 = (hash2 + part2[t]) * 0xff51afd7ed558ccdull;
        }
        hash1 ^= (uint64_t)nt << 56;                       // the segmentation is part of the value
        if (rows_in.off)
            for (int i = 0; i <= n_centres; ++i) hash2 = (hash2 ^ (uint64_t)rows_in.off[i]) * 0x100000001b3ull;
        if (rows_in.anchors)
            for (int i = 0; i < n_centres; ++i) hash1 = (hash1 + (uint32_t)rows_in.anchors[i]) * 0xff51afd7ed558ccdull;
    }
    Rows rows = rows_in;
    rows.data = reinterpret_cast<const int32_t*>(h->adj_host.p);
    // adjacency rows, then (after grouping) the shared cores: at most half as many entries again
    const void* adj_before = h->adj.p;
    CU(h->adj.ensure((total + total / 2 + 2) * sizeof(int)));
    const int kp = round_up(k_max, NB);

    // Centres per batch: the configured size, capped so that the per-centre workspaces (factor + solve panel, shared
    // cores, predictions, index lists, a share of the cluster Grams) take at most half of the device memory that is
    // free, counting what the handle already holds.
    int batch_cap = h->batch_centres;
    {
        size_t free_b = 0, total_b = 0;
        if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) {
            const size_t held = h->work.bytes + h->yhat.bytes + h->core_work.bytes + h->nbr_sorted.bytes + h->loc.bytes +
                                h->dist.bytes + h->gram_pool.bytes + h->zplanes.bytes;
            const double per_centre = 1.25 * (double)(kp + EVAL_PAD) * kp * sizeof(double) +
                                      (double)YHAT_ROWS * h->Dp * sizeof(double) + 2.5 * kp * sizeof(int) +
                                      0.1 * (double)kp * kp * sizeof(double) * (h->cv ? 1 + h->cv->n_folds : 1) +
                                      (with_null ? (double)NULL_DIST_ROWS * EVAL_PAD * sizeof(double) : 0.0) +
                                      (h->predict_mode == 1 ? (double)I8_DIGITS * EVAL_PAD * kp : 0.0);
            const double fit = 0.5 * ((double)free_b + (double)held) / per_centre;
            batch_cap = (int)std::max(256.0, std::min((double)batch_cap, fit));
        }
    }

    CU(h->scores.ensure((size_t)n_centres * MMD_N_SCORES * sizeof(double)));
    CU(h->flags.ensure((size_t)n_centres * sizeof(int)));
    CU(h->misc_flag.ensure(sizeof(int)));
    CU(cudaMemsetAsync(h->misc_flag.p, 0, sizeof(int), h->stream));
    CU(cudaMemsetAsync(h->flags.p, 0, (size_t)n_centres * sizeof(int), h->stream));
    if (with_null) CU(h->null_out.ensure((size_t)n_centres * h->n_perm * 6 * sizeof(double)));
    h->call_batches = 0;
    if (h->predict_mode == 1 && !raw_yhat) {
        CU(h->reroute_counts.ensure((size_t)MAX_CALL_BATCHES * sizeof(int)));
        CU(h->i8_status.ensure(sizeof(int)));
        CU(cudaMemsetAsync(h->reroute_counts.p, 0, (size_t)MAX_CALL_BATCHES * sizeof(int), h->stream));
        CU(cudaMemsetAsync(h->i8_status.p, 0, sizeof(int), h->stream));
    }
    if (h->want_margins) CU(h->margins.ensure((size_t)n_centres * sizeof(double)));

    // Clusters for the whole call, then plans for ranges of clusters ("chunks" of two or more batches): the plan of chunk
    // i+1 is built by host threads while the GPU executes chunk i.  The plan of the last adjacency is kept (PlanCache).
    if (!h->plan_cache) h->plan_cache = new PlanCache();
    PlanCache& pc = *h->plan_cache;
    const bool hit = pc.valid && pc.h1 == hash1 && pc.h2 == hash2 && pc.total == (long)total && pc.n == n_centres &&
                     pc.k == k_max && pc.V == h->V && pc.cluster_size == h->cluster_size && pc.group_size == h->group_size &&
                     pc.batch_cap == batch_cap && pc.ragged == (rows.off ? 1 : 0);
    if (!(hit && pc.adj_on_device && adj_before == h->adj.p))
        CU(cudaMemcpyAsync(h->adj.p, rows.data, total * sizeof(int), cudaMemcpyHostToDevice, h->stream));
    h->adj_h2d_us += std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now() - t_stage0).count();
    auto t_host0 = std::chrono::steady_clock::now();
    pc.valid = false;                                          // set again when the call has gone through
    pc.adj_on_device = false;
    if (hit) {
        ++h->plan_hits;
    } else {
        pc.h1 = hash1;
        pc.h2 = hash2;
        pc.total = (long)total;
        pc.n = n_centres;
        pc.k = k_max;
        pc.V = h->V;
        pc.cluster_size = h->cluster_size;
        pc.group_size = h->group_size;
        pc.batch_cap = batch_cap;
        pc.ragged = rows.off ? 1 : 0;
        cluster_centres(rows, h->V, h->cluster_size, pc.cl);
        const int n_cl = (int)pc.cl.cl_off.size() - 1;
        pc.chunk_start.assign(1, 0);
        const int chunk_target = std::max(2 * batch_cap, 12288);      // centres per chunk: whole batches, never more
        for (int c = 0, acc = 0; c < n_cl; ++c) {
            const int m = pc.cl.cl_off[c + 1] - pc.cl.cl_off[c];
            if (acc > 0 && acc + m > chunk_target) {
                pc.chunk_start.push_back(c);
                acc = 0;
            }
            acc += m;
        }
        pc.chunk_start.push_back(n_cl);
        pc.unpin_all();
        pc.plans.assign(pc.chunk_start.size() - 1, Plan());
    }
    const Clustering& cl = pc.cl;
    const std::vector<int>& chunk_start = pc.chunk_start;
    const int n_clusters = (int)cl.cl_off.size() - 1;
    h->n_clusters += n_clusters;
    const int n_chunks = (int)chunk_start.size() - 1;
    const int group_size = h->group_size, V = h->V;
    std::mutex pin_mutex;
    const int device_id = h->device;
    auto pin_plan = [&](int i) {
        std::lock_guard<std::mutex> lock(pin_mutex);
        cudaSetDevice(device_id);                          // planner threads start on device 0
        Plan& pl = pc.plans[i];
        pc.pin(pl.un_vert.data(), pl.un_vert.size() * sizeof(int));
        pc.pin(pl.core_vert.data(), pl.core_vert.size() * sizeof(int));
    };
    auto plan_chunk = [&, group_size, V, n_threads, hit](int i) {
        if (hit) return;
        build_plan(rows, V, cl, chunk_start[i], chunk_start[i + 1], group_size, n_threads, pc.plans[i]);
        pin_plan(i);
    };
    std::future<void> pending = std::async(std::launch::async, plan_chunk, 0);
    int rc = 0;
    for (int i = 0; i < n_chunks && rc == 0; ++i) {
        pending.get();                                     // exposed host time: only what the GPU did not hide
        h->plan_us += std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now() - t_host0).count();
        if (i + 1 < n_chunks) pending = std::async(std::launch::async, plan_chunk, i + 1);
        rc = run_chunk(h, rows, pc.plans[i], total, k_max, kp, alpha, with_null, raw_yhat, batch_cap);
        t_host0 = std::chrono::steady_clock::now();
    }
    if (rc == 0) pc.valid = pc.adj_on_device = true;
    if (pending.valid()) pending.wait();                   // an error path must not leave the planner running
    return rc;
}

// The device-side checks of a searchlight call (prep_tasks_kernel's flag word, the latents' flag of mmd_upload).
int report_input_flags(mmd_handle* h, int bad, int y_bad, const char* who) {
    if (bad & 2) return fail(h, "%s: adjacency holds a vertex index outside [0, %d)", who, h->V);
    // sklearn's Ridge.fit / predict validate their inputs (check_array) and raise ValueError
    if (bad & 4) return fail(h, "Input X contains NaN or infinity (%s: a neighbourhood uses a vertex whose betas are not finite)", who);
    if (bad & 8) return fail(h, "%s: a neighbourhood uses a vertex outside the active set given to mmd_upload_active", who);
    if (y_bad) return fail(h, "Input y contains NaN or infinity (%s: training latents must be finite)", who);
    if (bad & 1) return fail(h, "internal: neighbour missing from its cluster union");
    return 0;
}

// every device buffer the handle owns
std::vector<DevBuf*> handle_buffers(mmd_handle* h) {
    return {&h->Xt, &h->Xe, &h->Yt, &h->cand, &h->cand64, &h->cand_sq, &h->XtY, &h->adj, &h->union_vertices,
                      &h->core_work, &h->group_flags, &h->nbr_sorted, &h->loc,
                      &h->probs, &h->gram_pool, &h->work, &h->yhat, &h->scores, &h->flags, &h->misc_flag, &h->tmp,
                      &h->st_in, &h->st_edges, &h->st_label, &h->st_count, &h->st_out, &h->dist, &h->null_out, &h->perm_test, &h->perm_im, &h->perm_im_centre, &h->pa_pred, &h->pa_lat, &h->pa_dist,
                      &h->pa_sq, &h->pa_count, &h->vertex_bad, &h->y_bad, &h->up_x, &h->up_y, &h->up_train, &h->up_eval,
                      &h->xplanes, &h->sX, &h->colmax, &h->zplanes, &h->sZ, &h->amp2, &h->i8_status, &h->reroute_list,
                      &h->reroute_counts, &h->margins, &h->active, &h->batch_dev, &h->syrk_planes, &h->syrk_scales, &h->syrk_flags,
                      &h->syrk_status, &h->core_done};
}

// the stream must have passed mmd_upload's kernels (any later synchronisation of h->stream will do)
void release_upload_staging(mmd_handle* h) {
    h->up_x.release();
    h->up_y.release();
    h->up_train.release();
    h->up_eval.release();
}

// after the stream is synchronised: how many centres went back to the float64 kernels; protocol errors of the integer GEMM
int collect_i8_stats(mmd_handle* h) {
    if (h->syrk_status.p) {
        int st = 0;
        CU(cudaMemcpy(&st, h->syrk_status.p, sizeof(int), cudaMemcpyDeviceToHost));
        if (st) {
            cudaMemset(h->syrk_status.p, 0, sizeof(int));
            return fail(h, "syrk_i8_kernel: a bounded wait ran out (role %d); results are not valid", st);
        }
    }
    if (h->call_batches == 0) return 0;
    std::vector<int> counts(h->call_batches);
    int status = 0;
    CU(cudaMemcpy(counts.data(), h->reroute_counts.p, counts.size() * sizeof(int), cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(&status, h->i8_status.p, sizeof(int), cudaMemcpyDeviceToHost));
    for (int c : counts) h->n_rerouted += c;
    h->call_batches = 0;
    if (status) return fail(h, "predict_i8_kernel: a bounded wait ran out (role %d); results are not valid", status);
    return 0;
}

int searchlight_and_fetch(mmd_handle* h, const Rows& rows, double alpha, double* scores_out, int32_t* flags_out,
                          const char* who, const int64_t* data64 = nullptr) {
    const int n_centres = rows.n;
    if (run_searchlight(h, rows, alpha, false, false, data64)) {
        cudaStreamSynchronize(h->stream);
        h->spans.clear();
        h->events_used = 0;
        return 1;
    }
    int bad = 0, y_bad = 0;
    CU(cudaMemcpyAsync(&bad, h->misc_flag.p, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaMemcpyAsync(&y_bad, h->y_bad.p, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaMemcpyAsync(scores_out, h->scores.p, (size_t)n_centres * MMD_N_SCORES * sizeof(double), cudaMemcpyDefault,
                       h->stream));
    if (flags_out)
        CU(cudaMemcpyAsync(flags_out, h->flags.p, (size_t)n_centres * sizeof(int), cudaMemcpyDefault, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    resolve_spans(h);
    release_upload_staging(h);
    if (collect_i8_stats(h)) return 1;
    return report_input_flags(h, bad, y_bad, who);
}

}  // namespace

// ------------------------------------------------------------------------------------------
// exported functions
// ------------------------------------------------------------------------------------------
extern "C" {

const char* mmd_version(void) { return "mmd_b200 0.1 (sm_100a)"; }
const char* mmd_score_name(int i) { return (i >= 0 && i < MMD_N_SCORES) ? kScoreNames[i] : ""; }
const char* mmd_stage_name(int i) { return (i >= 0 && i < MMD_N_STAGES) ? kStageNames[i] : ""; }

int mmd_create(int device, mmd_handle** out) {
    if (!out) return 1;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || device < 0 || device >= n) return 2;
    if (cudaSetDevice(device) != cudaSuccess) return 3;
    mmd_handle* h = new mmd_handle();
    h->device = device;
    if (cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete h;
        return 4;
    }
    cudaDeviceGetAttribute(&h->n_sms, cudaDevAttrMultiProcessorCount, device);
    if (const char* e = getenv("MMD_B200_PREDICT")) h->predict_mode = strcmp(e, "f64") == 0 ? 0 : 1;
    if (const char* e = getenv("MMD_B200_SYRK")) h->syrk_mode = strcmp(e, "f64") == 0 ? 0 : 1;
    *out = h;
    return 0;
}

void mmd_destroy(mmd_handle* h) {
    if (!h) return;
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    std::vector<DevBuf*> bufs = handle_buffers(h);
    for (DevBuf* b : bufs) b->release();
    release_upload_staging(h);
    if (h->up_copied) cudaEventDestroy(h->up_copied);
    h->adj_host.release();
    for (int i = 0; i < 2; ++i) {
        h->batch_host[i].release();
        if (h->batch_copied[i]) cudaEventDestroy(h->batch_copied[i]);
    }
    delete h->plan_cache;
    if (h->ridge) {
        for (DevBuf* b : h->ridge->all()) b->release();
        delete h->ridge;
    }
    for (cudaEvent_t e : h->event_pool) cudaEventDestroy(e);
    cudaStreamDestroy(h->stream);
    delete h;
}

int mmd_check_guards(mmd_handle* h) {
    if (!h) return -1;
    if (!guard_mode()) return 0;
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    for (DevBuf* b : handle_buffers(h)) b->check();
    const int n = g_guard_failures.exchange(0);
    if (n) fail(h, "mmd_check_guards: %d device buffer(s) were written outside their bounds (details on stderr)", n);
    return n;
}

int mmd_guard_selftest(mmd_handle* h) {
    if (!h) return -1;
    if (!guard_mode()) return -1;
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    DevBuf probe;
    if (probe.ensure(100) != cudaSuccess) return 2;
    const int before = g_guard_failures.load();
    // one byte behind the buffer, then one byte in front of it: both must be reported
    cudaMemset(static_cast<char*>(probe.p) + probe.bytes, 0, 1);
    const int behind = probe.check();
    cudaMemset(static_cast<char*>(probe.p) + probe.bytes, 0xA5, 1);
    cudaMemset(static_cast<char*>(probe.p) - 1, 0, 1);
    const int front = probe.check();
    cudaMemset(static_cast<char*>(probe.p) - 1, 0xA5, 1);
    const int clean = probe.check();
    probe.release();
    g_guard_failures.store(before);
    return (behind == 2 && front == 1 && clean == 0) ? 0 : 1;
}

void* mmd_stream(mmd_handle* h) { return h ? (void*)h->stream : nullptr; }

const char* mmd_last_error(const mmd_handle* h) { return h ? h->err.c_str() : "null handle"; }

int mmd_configure(mmd_handle* h, int cluster_size, int batch_centres, int host_threads) {
    if (!h) return 1;
    if (cluster_size > 0) h->cluster_size = cluster_size;
    if (batch_centres > 0) h->batch_centres = batch_centres;
    if (host_threads > 0) h->host_threads = host_threads;
    return 0;
}

void mmd_reset_stats(mmd_handle* h) {
    if (!h) return;
    if (!h->spans.empty()) {                 // mmd_upload leaves its kernels in flight
        cudaStreamSynchronize(h->stream);
        resolve_spans(h);
    }
    memset(h->stage_ms, 0, sizeof(h->stage_ms));
    h->launches = 0;
    h->n_clusters = 0;
    h->syrk_flops = 0.0;
    h->solve_flops = 0.0;
    h->n_batches = h->n_core_tasks = h->n_grouped = h->sum_sp = h->plan_us = h->adj_h2d_us = h->n_rerouted = h->plan_hits = 0;
}

int mmd_get_counters(mmd_handle* h, int64_t counters[MMD_N_COUNTERS], double* solve_flops) {
    if (!h) return 1;
    if (counters) {
        counters[0] = h->n_batches;
        counters[1] = h->n_core_tasks;
        counters[2] = h->n_grouped;
        counters[3] = h->sum_sp;
        counters[4] = h->plan_us;
        counters[5] = h->adj_h2d_us;
        counters[6] = h->n_rerouted;
        counters[7] = h->plan_hits;
    }
    if (solve_flops) *solve_flops = h->solve_flops;
    return 0;
}

int mmd_set_predict_mode(mmd_handle* h, int mode) {
    if (!h) return 1;
    if (mode != 0 && mode != 1) return fail(h, "mmd_set_predict_mode: mode must be 0 (float64) or 1 (tcgen05 int8 digits)");
    if (mode == 1 && h->uploaded && !h->xplanes.p)
        return fail(h, "mmd_set_predict_mode: switch to the integer GEMM before mmd_upload (its digit planes are cut there)");
    h->predict_mode = mode;
    return 0;
}

int mmd_set_syrk_mode(mmd_handle* h, int mode) {
    if (!h) return 1;
    if (mode != 0 && mode != 1) return fail(h, "mmd_set_syrk_mode: mode must be 0 (float64 DMMA) or 1 (exact integer SYRK on tcgen05)");
    h->syrk_mode = mode;
    if (h->plan_cache) h->plan_cache->valid = false;
    return 0;
}

int mmd_enable_margins(mmd_handle* h, int on) {
    if (!h) return 1;
    h->want_margins = on != 0;
    return 0;
}

int mmd_get_margins(mmd_handle* h, double* margins_out, int n_centres) {
    if (!h) return 1;
    if (!margins_out || n_centres < 0) return fail(h, "mmd_get_margins: bad arguments");
    if ((size_t)n_centres * sizeof(double) > h->margins.bytes) return fail(h, "mmd_get_margins: no margins recorded for %d centres", n_centres);
    CU(cudaSetDevice(h->device));
    CU(cudaMemcpyAsync(margins_out, h->margins.p, (size_t)n_centres * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return 0;
}

int mmd_set_group_size(mmd_handle* h, int group_size) {
    if (!h) return 1;
    if (group_size < 1 || group_size > 64) return fail(h, "mmd_set_group_size: group size must be 1..64");
    h->group_size = group_size;
    return 0;
}

int mmd_get_stats(mmd_handle* h, float stage_ms[MMD_N_STAGES], int64_t* n_launches, int64_t* n_clusters,
                  double* syrk_flops) {
    if (!h) return 1;
    if (!h->spans.empty()) {                 // mmd_upload leaves its kernels in flight
        CU(cudaStreamSynchronize(h->stream));
        resolve_spans(h);
    }
    if (stage_ms) memcpy(stage_ms, h->stage_ms, sizeof(h->stage_ms));
    if (n_launches) *n_launches = h->launches;
    if (n_clusters) *n_clusters = h->n_clusters;
    if (syrk_flops) *syrk_flops = h->syrk_flops;
    return 0;
}

int mmd_knn(mmd_handle* h, const double* coords, int n_vertices, int k, int q_begin, int q_end, int64_t* idx_out,
            double* dist_out) {
    if (!h) return 1;
    if (!coords || !idx_out || !dist_out) return fail(h, "mmd_knn: null pointer");
    if (q_begin < 0 || q_end > n_vertices || q_begin > q_end) return fail(h, "mmd_knn: bad query range");
    if (k < 1 || k > n_vertices || k > 2000) return fail(h, "mmd_knn: k=%d out of range", k);
    CU(cudaSetDevice(h->device));
    const size_t nq = (size_t)(q_end - q_begin);
    DevBuf dcoords, didx, ddist;
    const double* c_dev = coords;
    int rc = 0;
    do {
        if (!is_device_ptr(coords)) {
            if (dcoords.ensure((size_t)n_vertices * 3 * sizeof(double)) != cudaSuccess) { rc = fail(h, "mmd_knn: out of memory"); break; }
            cudaMemcpyAsync(dcoords.p, coords, (size_t)n_vertices * 3 * sizeof(double), cudaMemcpyHostToDevice, h->stream);
            c_dev = dcoords.as<double>();
        }
        int64_t* i_dev = idx_out;
        double* d_dev = dist_out;
        const bool idx_host = !is_device_ptr(idx_out), dist_host = !is_device_ptr(dist_out);
        if (idx_host) {
            if (didx.ensure(std::max<size_t>(nq * k, 1) * sizeof(int64_t)) != cudaSuccess) { rc = fail(h, "mmd_knn: out of memory"); break; }
            i_dev = didx.as<int64_t>();
        }
        if (dist_host) {
            if (ddist.ensure(std::max<size_t>(nq * k, 1) * sizeof(double)) != cudaSuccess) { rc = fail(h, "mmd_knn: out of memory"); break; }
            d_dev = ddist.as<double>();
        }
        if (h->misc_flag.ensure(sizeof(int)) != cudaSuccess) { rc = fail(h, "mmd_knn: out of memory"); break; }
        cudaMemsetAsync(h->misc_flag.p, 0, sizeof(int), h->stream);
        span_begin(h, ST_KNN);
        cudaError_t e = launch_knn(c_dev, n_vertices, k, q_begin, q_end, i_dev, d_dev, h->misc_flag.as<int>(), h->stream);
        span_end(h);
        if (e != cudaSuccess) { rc = fail(h, "knn launch failed: %s", cudaGetErrorString(e)); break; }
        int flag = 0;
        cudaMemcpyAsync(&flag, h->misc_flag.p, sizeof(int), cudaMemcpyDeviceToHost, h->stream);
        if (idx_host) cudaMemcpyAsync(idx_out, i_dev, nq * k * sizeof(int64_t), cudaMemcpyDeviceToHost, h->stream);
        if (dist_host) cudaMemcpyAsync(dist_out, d_dev, nq * k * sizeof(double), cudaMemcpyDeviceToHost, h->stream);
        e = cudaStreamSynchronize(h->stream);
        resolve_spans(h);
        if (e != cudaSuccess) { rc = fail(h, "mmd_knn failed: %s", cudaGetErrorString(e)); break; }
        if (flag) { rc = fail(h, "mmd_knn: more than 2048 candidates tie at the k-th distance (degenerate coordinates)"); break; }
    } while (0);
    dcoords.release();
    didx.release();
    ddist.release();
    return rc;
}

int mmd_radius_neighbors(mmd_handle* h, const double* coords, int n_vertices, double radius, int q_begin, int q_end,
                         const int64_t* offsets, int32_t* counts_out, int32_t* indices_out) {
    if (!h) return 1;
    if (!coords) return fail(h, "mmd_radius_neighbors: null pointer");
    if (q_begin < 0 || q_end > n_vertices || q_begin > q_end) return fail(h, "mmd_radius_neighbors: bad query range");
    if (!(radius >= 0.0)) return fail(h, "mmd_radius_neighbors: radius must be >= 0");
    const bool fill = indices_out != nullptr;
    if (fill ? !offsets : !counts_out) return fail(h, "mmd_radius_neighbors: the count pass needs counts_out, the fill pass offsets");
    CU(cudaSetDevice(h->device));
    const size_t nq = (size_t)(q_end - q_begin);
    if (nq == 0) return 0;
    DevBuf dcoords, dout, doff;
    const double* c_dev = coords;
    int rc = 0;
    do {
        if (!is_device_ptr(coords)) {
            if ((rc = to_device(h, dcoords, coords, (size_t)n_vertices * 3))) break;
            c_dev = dcoords.as<double>();
        }
        span_begin(h, ST_KNN);
        cudaError_t e;
        size_t n_out;
        if (!fill) {
            n_out = nq;
            if (dout.ensure(nq * sizeof(int)) != cudaSuccess) { rc = fail(h, "mmd_radius_neighbors: out of memory"); break; }
            e = launch_radius_count(c_dev, n_vertices, radius, q_begin, q_end, dout.as<int>(), h->stream);
        } else {
            if (offsets[0] != 0) { rc = fail(h, "mmd_radius_neighbors: offsets[0] must be 0"); break; }
            n_out = (size_t)offsets[nq];
            static_assert(sizeof(long) == sizeof(int64_t), "offsets are passed to the kernel as long");
            if ((rc = to_device(h, doff, reinterpret_cast<const long*>(offsets), nq + 1))) break;
            if (dout.ensure(std::max<size_t>(n_out, 1) * sizeof(int)) != cudaSuccess) { rc = fail(h, "mmd_radius_neighbors: out of memory"); break; }
            e = launch_radius_fill(c_dev, n_vertices, radius, q_begin, q_end, doff.as<long>(), dout.as<int>(), h->stream);
        }
        span_end(h);
        if (e != cudaSuccess) { rc = fail(h, "radius launch failed: %s", cudaGetErrorString(e)); break; }
        cudaMemcpyAsync(fill ? indices_out : counts_out, dout.p, n_out * sizeof(int), cudaMemcpyDeviceToHost, h->stream);
        e = cudaStreamSynchronize(h->stream);
        resolve_spans(h);
        if (e != cudaSuccess) { rc = fail(h, "mmd_radius_neighbors failed: %s", cudaGetErrorString(e)); break; }
    } while (0);
    dcoords.release();
    dout.release();
    doff.release();
    return rc;
}

// Host helper of the vertex-sharded path: mask[v] = 1 for every vertex id in `indices` (n entries, int64 when is_i64 else
// int32), by host threads.  What np.unique(A[lo:hi]) computes, without the sort (custom_search_light under sharding).
int mmd_mark_vertices(const void* indices, int is_i64, int64_t n, int n_vertices, uint8_t* mask) {
    if (!indices || !mask || n < 0 || n_vertices < 1) return 1;
    memset(mask, 0, (size_t)n_vertices);
    int nt = (int)std::max<int64_t>(1, std::min<int64_t>(std::min<unsigned>(std::thread::hardware_concurrency(), 32u), n / (1 << 20) + 1));
    std::atomic<int> bad(0);
    auto work = [&](int t) {
        const int64_t lo = n * t / nt, hi = n * (t + 1) / nt;
        for (int64_t i = lo; i < hi; ++i) {
            const int64_t v = is_i64 ? reinterpret_cast<const int64_t*>(indices)[i] : reinterpret_cast<const int32_t*>(indices)[i];
            if (v < 0 || v >= n_vertices) bad.store(1);
            else mask[v] = 1;                               // racing writers store the same value
        }
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < nt; ++t) pool.emplace_back(work, t);
    work(0);
    for (auto& t : pool) t.join();
    return bad.load() ? 2 : 0;
}

int mmd_upload(mmd_handle* h, const void* X, int n_total, int n_vertices, const void* Y, int n_dims,
               const int32_t* train_ids, int n_train, const int32_t* test_ids, int n_test,
               const uint8_t* test_is_caption, const int32_t* imagery_ids, int n_imagery) {
    return mmd_upload_active(h, X, n_total, n_vertices, Y, n_dims, train_ids, n_train, test_ids, n_test, test_is_caption,
                             imagery_ids, n_imagery, nullptr, 0);
}

int mmd_upload_active(mmd_handle* h, const void* X, int n_total, int n_vertices, const void* Y, int n_dims,
                      const int32_t* train_ids, int n_train, const int32_t* test_ids, int n_test,
                      const uint8_t* test_is_caption, const int32_t* imagery_ids, int n_imagery,
                      const int32_t* active_vertices, int n_active) {
    if (!h) return 1;
    if (n_active < 0 || (n_active > 0 && !active_vertices)) return fail(h, "mmd_upload_active: bad active-vertex list");
    for (int i = 0; i < n_active; ++i)
        if (active_vertices[i] < 0 || active_vertices[i] >= n_vertices || (i > 0 && active_vertices[i] <= active_vertices[i - 1]))
            return fail(h, "mmd_upload_active: active vertices must be ascending indices in [0, %d)", n_vertices);
    if (n_active == n_vertices) n_active = 0;
    if (!X || !Y || !train_ids || (n_test > 0 && (!test_ids || !test_is_caption)) || (n_imagery > 0 && !imagery_ids))
        return fail(h, "mmd_upload: null pointer");
    if (n_total < 1 || n_vertices < 1 || n_dims < 1 || n_train < 1) return fail(h, "mmd_upload: empty input");
    if (n_test + n_imagery > MMD_EVAL_PAD - 1)
        return fail(h, "mmd_upload: %d test + %d imagery trials exceed the supported %d", n_test, n_imagery,
                    MMD_EVAL_PAD - 1);
    if (n_imagery > 8) return fail(h, "mmd_upload: at most 8 imagery trials are supported");
    auto in_range = [&](const int32_t* ids, int n) {
        for (int i = 0; i < n; ++i)
            if (ids[i] < 0 || ids[i] >= n_total) return false;
        return true;
    };
    if (!in_range(train_ids, n_train) || !in_range(test_ids, n_test) || !in_range(imagery_ids, n_imagery))
        return fail(h, "mmd_upload: trial index out of range");
    CU(cudaSetDevice(h->device));
    h->uploaded = false;
    h->V = n_vertices;
    h->n_train = n_train;
    h->ldt = round_up(n_train, 64);
    h->n_test = n_test;
    h->n_im = n_imagery;
    h->D = n_dims;
    h->Dp = round_up(n_dims, 128);         // 128-feature tiles of the integer prediction GEMM
    // evaluation rows in the internal order: captions, images, imagery
    std::vector<int> eval_ids;
    h->n_cap = 0;
    h->int_of_test.assign(n_test, 0);
    for (int i = 0; i < n_test; ++i)
        if (test_is_caption[i]) { h->int_of_test[i] = (int)eval_ids.size(); eval_ids.push_back(test_ids[i]); ++h->n_cap; }
    for (int i = 0; i < n_test; ++i)
        if (!test_is_caption[i]) { h->int_of_test[i] = (int)eval_ids.size(); eval_ids.push_back(test_ids[i]); }
    h->n_img = n_test - h->n_cap;
    for (int i = 0; i < n_imagery; ++i) eval_ids.push_back(imagery_ids[i]);
    const int n_eval = (int)eval_ids.size();

    DevBuf &xtmp = h->up_x, &ytmp = h->up_y, &d_train = h->up_train, &d_eval = h->up_eval;
    if (!h->up_copied) CU(cudaEventCreateWithFlags(&h->up_copied, cudaEventDisableTiming));
    int rc = 0;
    do {
        const __half* Xd = reinterpret_cast<const __half*>(X);
        const __half* Yd = reinterpret_cast<const __half*>(Y);
        if (!is_device_ptr(X)) {
            if ((rc = to_device(h, xtmp, reinterpret_cast<const __half*>(X), (size_t)n_total * n_vertices))) break;
            Xd = xtmp.as<__half>();
        }
        if (!is_device_ptr(Y)) {
            if ((rc = to_device(h, ytmp, reinterpret_cast<const __half*>(Y), (size_t)n_total * n_dims))) break;
            Yd = ytmp.as<__half>();
        }
        if ((rc = to_device(h, d_train, train_ids, n_train))) break;
        if ((rc = to_device(h, d_eval, eval_ids.data(), n_eval))) break;
#define CUB(call)                                                                                  \
    {                                                                                              \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess) {                                                                  \
            rc = fail(h, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
            break;                                                                                 \
        }                                                                                          \
    }
        CUB(h->Xt.ensure((size_t)n_vertices * h->ldt * sizeof(__half)));
        CUB(h->Xe.ensure((size_t)n_vertices * EVAL_PAD * sizeof(__half)));
        CUB(h->Yt.ensure((size_t)h->Dp * h->ldt * sizeof(__half)));
        CUB(h->cand.ensure((size_t)EVAL_PAD * h->Dp * sizeof(__half)));
        CUB(h->cand64.ensure((size_t)EVAL_PAD * h->Dp * sizeof(double)));
        CUB(h->cand_sq.ensure((size_t)EVAL_PAD * sizeof(double)));
        CUB(h->XtY.ensure((size_t)n_vertices * h->Dp * sizeof(double)));
        CUB(cudaMemsetAsync(h->Yt.p, 0, (size_t)h->Dp * h->ldt * sizeof(__half), h->stream));
        CUB(h->vertex_bad.ensure((size_t)n_vertices));
        CUB(h->y_bad.ensure(sizeof(int)));
        h->n_active = n_active;
        if (n_active) {
            if ((rc = to_device(h, h->active, active_vertices, n_active))) break;
            CUB(cudaMemsetAsync(h->vertex_bad.p, 2, (size_t)n_vertices, h->stream));
            CUB(launch_clear_bytes(h->vertex_bad.as<unsigned char>(), h->active.as<int>(), n_active, h->stream));
        } else {
            CUB(cudaMemsetAsync(h->vertex_bad.p, 0, (size_t)n_vertices, h->stream));
        }
        CUB(cudaMemsetAsync(h->y_bad.p, 0, sizeof(int), h->stream));
        span_begin(h, ST_PACK);
        launch_pack_transpose(Xd, n_vertices, d_train.as<int>(), n_train, n_vertices, h->Xt.as<__half>(), h->ldt,
                              h->vertex_bad.as<unsigned char>(), nullptr, h->stream);
        launch_pack_transpose(Xd, n_vertices, d_eval.as<int>(), n_eval, n_vertices, h->Xe.as<__half>(), EVAL_PAD,
                              h->vertex_bad.as<unsigned char>(), nullptr, h->stream);
        launch_pack_transpose(Yd, n_dims, d_train.as<int>(), n_train, n_dims, h->Yt.as<__half>(), h->ldt, nullptr,
                              h->y_bad.as<int>(), h->stream);
        gather_rows_f16_kernel<<<EVAL_PAD, 256, 0, h->stream>>>(Yd, n_dims, d_eval.as<int>(), n_eval, n_dims,
                                                               h->cand.as<__half>(), EVAL_PAD, h->Dp);
        // squared norms of the candidate rows (identity row list = first EVAL_PAD entries of an iota)
        {
            std::vector<int> iota(EVAL_PAD);
            for (int i = 0; i < EVAL_PAD; ++i) iota[i] = i;
            if ((rc = to_device(h, h->tmp, iota.data(), EVAL_PAD))) break;
            CUB(cudaEventRecord(h->up_copied, h->stream));      // last read of caller memory
            row_sqnorm_kernel<<<EVAL_PAD, 128, 0, h->stream>>>(h->cand.as<__half>(), h->Dp, h->tmp.as<int>(), EVAL_PAD,
                                                              h->Dp, h->cand_sq.as<double>());
        }
        half_to_double_kernel<<<148, 256, 0, h->stream>>>(h->cand.as<__half>(), h->cand64.as<double>(), (long)EVAL_PAD * h->Dp);
        span_end(h, 6);
        CUB(cudaGetLastError());
        if ((rc = run_xty(h))) break;
        // The caller's buffers are free again once the copies have landed; the XtY contraction (and the staging
        // copies it reads) stays in flight on the stream, in front of whatever the next call enqueues.
        CUB(cudaEventSynchronize(h->up_copied));
        h->uploaded = true;
    } while (0);
    if (rc) {
        cudaStreamSynchronize(h->stream);
        resolve_spans(h);
        release_upload_staging(h);
    }
    return rc;
}

int mmd_searchlight(mmd_handle* h, const int32_t* adjacency, int n_centres, int k, double alpha, double* scores_out,
                    int32_t* flags_out) {
    if (!h) return 1;
    if (!adjacency || !scores_out) return fail(h, "mmd_searchlight: null pointer");
    if (n_centres < 0) return fail(h, "mmd_searchlight: negative centre count");
    if (n_centres == 0) return 0;
    if (is_device_ptr(adjacency)) return fail(h, "mmd_searchlight: adjacency must be a host pointer");
    CU(cudaSetDevice(h->device));
    const Rows rows{adjacency, nullptr, k, nullptr, n_centres};
    return searchlight_and_fetch(h, rows, alpha, scores_out, flags_out, "mmd_searchlight");
}

int mmd_searchlight_i64(mmd_handle* h, const int64_t* adjacency, int n_centres, int k, double alpha, double* scores_out,
                        int32_t* flags_out) {
    if (!h) return 1;
    if (!adjacency || !scores_out) return fail(h, "mmd_searchlight_i64: null pointer");
    if (n_centres < 0) return fail(h, "mmd_searchlight_i64: negative centre count");
    if (n_centres == 0) return 0;
    if (is_device_ptr(adjacency)) return fail(h, "mmd_searchlight_i64: adjacency must be a host pointer");
    CU(cudaSetDevice(h->device));
    const Rows rows{nullptr, nullptr, k, nullptr, n_centres};
    return searchlight_and_fetch(h, rows, alpha, scores_out, flags_out, "mmd_searchlight_i64", adjacency);
}

int mmd_searchlight_ragged(mmd_handle* h, const int32_t* indices, const int64_t* offsets, const int32_t* anchors,
                           int n_centres, double alpha, double* scores_out, int32_t* flags_out) {
    if (!h) return 1;
    if (!indices || !offsets || !scores_out) return fail(h, "mmd_searchlight_ragged: null pointer");
    if (n_centres < 0) return fail(h, "mmd_searchlight_ragged: negative centre count");
    if (n_centres == 0) return 0;
    if (is_device_ptr(indices) || is_device_ptr(offsets))
        return fail(h, "mmd_searchlight_ragged: indices and offsets must be host pointers");
    if (offsets[0] != 0) return fail(h, "mmd_searchlight_ragged: offsets[0] must be 0");
    for (int i = 0; i < n_centres; ++i) {
        if (offsets[i + 1] < offsets[i]) return fail(h, "mmd_searchlight_ragged: offsets must be non-decreasing");
        if (offsets[i + 1] == offsets[i])
            return fail(h, "mmd_searchlight_ragged: centre %d has an empty neighbourhood (sklearn's Ridge.fit rejects "
                           "an array with 0 features)", i);
    }
    CU(cudaSetDevice(h->device));
    const Rows rows{indices, offsets, 0, anchors, n_centres};
    return searchlight_and_fetch(h, rows, alpha, scores_out, flags_out, "mmd_searchlight_ragged");
}

int mmd_predict_centre(mmd_handle* h, const int32_t* neighbours, int k, double alpha, double* yhat_out) {
    if (!h) return 1;
    if (!neighbours || !yhat_out) return fail(h, "mmd_predict_centre: null pointer");
    CU(cudaSetDevice(h->device));
    const Rows one{neighbours, nullptr, k, nullptr, 1};
    if (run_searchlight(h, one, alpha, false, true)) {
        cudaStreamSynchronize(h->stream);
        h->spans.clear();
        h->events_used = 0;
        return 1;
    }
    const int n_eval = h->n_test + h->n_im;
    CU(cudaMemcpy2DAsync(yhat_out, (size_t)h->D * sizeof(double), h->yhat.p, (size_t)h->Dp * sizeof(double),
                         (size_t)h->D * sizeof(double), n_eval, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    resolve_spans(h);
    release_upload_staging(h);
    return 0;
}

int mmd_searchlight_null(mmd_handle* h, const int32_t* adjacency, int n_centres, int k, double alpha,
                         const int32_t* perm_test, const int32_t* perm_imagery, int n_perm, double* scores_out,
                         int32_t* flags_out, double* null_out) {
    return mmd_searchlight_null_pc(h, adjacency, n_centres, k, alpha, perm_test, perm_imagery, nullptr, n_perm, scores_out,
                                   flags_out, null_out);
}

static int searchlight_null_rows(mmd_handle* h, const Rows& rows, double alpha, const int32_t* perm_test,
                                 const int32_t* perm_imagery, const int32_t* perm_imagery_centre, int n_perm, double* scores_out,
                                 int32_t* flags_out, double* null_out) {
    const int n_centres = rows.n;
    if (perm_imagery_centre && !perm_imagery) perm_imagery = perm_imagery_centre;      // any valid table for the shared slot
    if (!rows.data || !scores_out || !perm_test || !null_out || (h->n_im > 0 && !perm_imagery))
        return fail(h, "mmd_searchlight_null: null pointer");
    if (!h->uploaded) return fail(h, "mmd_upload must be called before mmd_searchlight_null");
    if (n_centres < 0 || n_perm < 1 || n_perm > 1700) return fail(h, "mmd_searchlight_null: n_perm must be 1..1700 per call (split larger sets)");
    if (n_centres == 0) return 0;
    const int n_test = h->n_test;
    for (long i = 0; i < (long)n_perm * n_test; ++i)
        if (perm_test[i] < 0 || perm_test[i] >= n_test) return fail(h, "mmd_searchlight_null: test permutation index out of range");
    for (long i = 0; i < (long)n_perm * h->n_im; ++i)
        if (perm_imagery[i] < 0 || perm_imagery[i] >= h->n_im) return fail(h, "mmd_searchlight_null: imagery permutation index out of range");
    CU(cudaSetDevice(h->device));
    h->n_perm = n_perm;
    {   // test-order positions -> internal evaluation rows, for rows and values alike
        std::vector<int> internal((size_t)n_perm * n_test);
        for (int p = 0; p < n_perm; ++p)
            for (int pos = 0; pos < n_test; ++pos)
                internal[(size_t)p * n_test + h->int_of_test[pos]] = h->int_of_test[perm_test[(size_t)p * n_test + pos]];
        if (to_device(h, h->perm_test, internal.data(), internal.size())) return 1;
        CU(cudaStreamSynchronize(h->stream));          // `internal` is pageable and goes out of scope
    }
    if (to_device(h, h->perm_im, perm_imagery, (size_t)n_perm * h->n_im)) return 1;
    h->use_perm_im_centre = false;
    if (perm_imagery_centre && h->n_im > 0) {
        const size_t n_pc = (size_t)n_centres * n_perm * h->n_im;
        for (size_t i = 0; i < n_pc; ++i)
            if (perm_imagery_centre[i] < 0 || perm_imagery_centre[i] >= h->n_im)
                return fail(h, "mmd_searchlight_null: imagery permutation index out of range");
        if (to_device(h, h->perm_im_centre, perm_imagery_centre, n_pc)) return 1;
        h->use_perm_im_centre = true;
    }
    const int rc_null = run_searchlight(h, rows, alpha, true);
    h->use_perm_im_centre = false;
    if (rc_null) {
        cudaStreamSynchronize(h->stream);
        h->spans.clear();
        h->events_used = 0;
        return 1;
    }
    int bad = 0, y_bad = 0;
    CU(cudaMemcpyAsync(&bad, h->misc_flag.p, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaMemcpyAsync(&y_bad, h->y_bad.p, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaMemcpyAsync(scores_out, h->scores.p, (size_t)n_centres * MMD_N_SCORES * sizeof(double), cudaMemcpyDefault, h->stream));
    CU(cudaMemcpyAsync(null_out, h->null_out.p, (size_t)n_centres * n_perm * 6 * sizeof(double), cudaMemcpyDefault, h->stream));
    if (flags_out) CU(cudaMemcpyAsync(flags_out, h->flags.p, (size_t)n_centres * sizeof(int), cudaMemcpyDefault, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    resolve_spans(h);
    release_upload_staging(h);
    if (collect_i8_stats(h)) return 1;
    return report_input_flags(h, bad, y_bad, "mmd_searchlight_null");
}

int mmd_searchlight_null_pc(mmd_handle* h, const int32_t* adjacency, int n_centres, int k, double alpha,
                            const int32_t* perm_test, const int32_t* perm_imagery, const int32_t* perm_imagery_centre, int n_perm,
                            double* scores_out, int32_t* flags_out, double* null_out) {
    if (!h) return 1;
    const Rows rows{adjacency, nullptr, k, nullptr, n_centres};
    return searchlight_null_rows(h, rows, alpha, perm_test, perm_imagery, perm_imagery_centre, n_perm, scores_out, flags_out, null_out);
}

int mmd_searchlight_null_ragged(mmd_handle* h, const int32_t* indices, const int64_t* offsets, const int32_t* anchors, int n_centres,
                                double alpha, const int32_t* perm_test, const int32_t* perm_imagery,
                                const int32_t* perm_imagery_centre, int n_perm, double* scores_out, int32_t* flags_out,
                                double* null_out) {
    if (!h) return 1;
    if (!indices || !offsets) return fail(h, "mmd_searchlight_null_ragged: null pointer");
    if (n_centres < 0) return fail(h, "mmd_searchlight_null_ragged: negative centre count");
    if (n_centres > 0) {
        if (offsets[0] != 0) return fail(h, "mmd_searchlight_null_ragged: offsets[0] must be 0");
        for (int i = 0; i < n_centres; ++i)
            if (offsets[i + 1] <= offsets[i])
                return fail(h, "mmd_searchlight_null_ragged: centre %d has an empty neighbourhood or offsets decrease", i);
    }
    const Rows rows{indices, offsets, 0, anchors, n_centres};
    return searchlight_null_rows(h, rows, alpha, perm_test, perm_imagery, perm_imagery_centre, n_perm, scores_out, flags_out, null_out);
}

// GridSearchCV semantics per neighbourhood (ridge_regression_decoding.py:111-117; KFold without shuffling:
// fold f = training positions fold_bounds[f] .. fold_bounds[f+1]-1 in train_ids order).  fold_scores_out is
// (n_centres x n_alphas x n_folds): pairwise_accuracy(latents_val, predictions_val) of eval.py:115-123.
int mmd_searchlight_cv(mmd_handle* h, const int32_t* adjacency, int n_centres, int k, const double* alphas, int n_alphas,
                       const int32_t* fold_bounds, int n_folds, double* fold_scores_out, int32_t* flags_out) {
    if (!h) return 1;
    if (!adjacency || !alphas || !fold_bounds || !fold_scores_out) return fail(h, "mmd_searchlight_cv: null pointer");
    if (!h->uploaded) return fail(h, "mmd_upload must be called before mmd_searchlight_cv");
    if (n_centres < 0 || n_alphas < 1 || n_alphas > 64) return fail(h, "mmd_searchlight_cv: need 1..64 alphas and a non-negative centre count");
    if (n_folds < 2 || n_folds > h->n_train)
        return fail(h, "mmd_searchlight_cv: k-fold cross-validation requires at least one train/test split by setting n_splits=2 "
                       "or more and at most n_samples=%d, got n_splits=%d", h->n_train, n_folds);
    if (fold_bounds[0] != 0 || fold_bounds[n_folds] != h->n_train) return fail(h, "mmd_searchlight_cv: fold bounds must run from 0 to n_train");
    for (int f = 0; f < n_folds; ++f)
        if (fold_bounds[f + 1] <= fold_bounds[f]) return fail(h, "mmd_searchlight_cv: empty fold %d", f);
    for (int a = 0; a < n_alphas; ++a)
        if (!(alphas[a] > 0.0)) return fail(h, "mmd_searchlight_cv: alpha must be positive (got %g)", alphas[a]);
    if (n_centres == 0) return 0;
    if (is_device_ptr(adjacency)) return fail(h, "mmd_searchlight_cv: adjacency must be a host pointer");
    CU(cudaSetDevice(h->device));

    CvState cv;
    cv.n_folds = n_folds;
    cv.alphas.assign(alphas, alphas + n_alphas);
    cv.xfold.resize(n_folds);
    cv.xty_tr.resize(n_folds);
    cv.cand.resize(n_folds);
    cv.cand_sq.resize(n_folds);
    cv.xe.resize(n_folds);
    int rc = 0;
    auto prepare = [&]() -> int {
        int ld_max = 0;
        for (int f = 0; f < n_folds; ++f) {
            cv.lo.push_back(fold_bounds[f]);
            cv.hi.push_back(fold_bounds[f + 1]);
            cv.ld.push_back(round_up(fold_bounds[f + 1] - fold_bounds[f], 64));
            ld_max = std::max(ld_max, cv.ld.back());
        }
        CU(cv.yfold.ensure((size_t)h->Dp * ld_max * sizeof(__half)));
        CU(cv.counts.ensure((size_t)n_centres * n_alphas * n_folds * sizeof(unsigned long long)));
        CU(cudaMemsetAsync(cv.counts.p, 0, cv.counts.bytes, h->stream));
        CU(cv.probs.ensure(sizeof(GemmProblem)));
        span_begin(h, ST_XTY);
        int launches = 0;
        for (int f = 0; f < n_folds; ++f) {
            const int lo = cv.lo[f], n_val = cv.hi[f] - lo, ld = cv.ld[f];
            const int n_chunks = (n_val + EVAL_PAD - 1) / EVAL_PAD;
            CU(cv.xfold[f].ensure((size_t)h->V * ld * sizeof(__half)));
            CU(cv.xty_tr[f].ensure((size_t)h->V * h->Dp * sizeof(double)));
            CU(cv.cand[f].ensure((size_t)n_chunks * EVAL_PAD * h->Dp * sizeof(double)));
            CU(cv.cand_sq[f].ensure((size_t)n_chunks * EVAL_PAD * sizeof(double)));
            CU(launch_copy_cols_f16(h->Xt.as<__half>(), h->ldt, lo, n_val, cv.xfold[f].as<__half>(), ld, h->V, h->stream));
            CU(launch_copy_cols_f16(h->Yt.as<__half>(), h->ldt, lo, n_val, cv.yfold.as<__half>(), ld, h->Dp, h->stream));
            // XtY without the fold: XtY - Xfold Yfold^T
            GemmProblem p{};
            p.A = cv.xfold[f].as<__half>();
            p.lda = ld;
            p.M = h->V;
            p.B = cv.yfold.as<__half>();
            p.ldb = ld;
            p.N = h->Dp;
            p.K = ld;
            p.C = cv.xty_tr[f].as<double>();
            p.ldc = h->Dp;
            CU(cudaMemcpyAsync(cv.probs.p, &p, sizeof(p), cudaMemcpyDefault, h->stream));
            CU(launch_gemm_gather(cv.probs.as<GemmProblem>(), 1, p.M, p.N, h->stream));
            CU(launch_rsub(h->XtY.as<double>(), cv.xty_tr[f].as<double>(), (long)h->V * h->Dp, h->stream));
            CU(launch_fold_latents(h->Yt.as<__half>(), h->ldt, lo, n_val, n_chunks * EVAL_PAD, h->Dp, cv.cand[f].as<double>(),
                                   cv.cand_sq[f].as<double>(), h->stream));
            cv.xe[f].resize(n_chunks);
            for (int c = 0; c < n_chunks; ++c) {
                CU(cv.xe[f][c].ensure((size_t)h->V * EVAL_PAD * sizeof(__half)));
                CU(launch_copy_cols_f16(h->Xt.as<__half>(), h->ldt, lo + c * EVAL_PAD, std::min(EVAL_PAD, n_val - c * EVAL_PAD),
                                        cv.xe[f][c].as<__half>(), EVAL_PAD, h->V, h->stream));
            }
            launches += 5 + n_chunks;
            CU(cudaStreamSynchronize(h->stream));          // `p` is a stack object; the fold panels are ready behind this point
        }
        span_end(h, launches);
        return 0;
    };
    rc = prepare();
    if (!rc) {
        h->cv = &cv;
        const Rows rows{adjacency, nullptr, k, nullptr, n_centres};
        rc = run_searchlight(h, rows, cv.alphas[0], false, true);
        h->cv = nullptr;
    }
    std::vector<unsigned long long> counts;
    int bad = 0, y_bad = 0;
    if (!rc) {
        counts.resize((size_t)n_centres * n_alphas * n_folds);
        auto fetch = [&]() -> int {
            CU(cudaMemcpyAsync(&bad, h->misc_flag.p, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
            CU(cudaMemcpyAsync(&y_bad, h->y_bad.p, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
            CU(cudaMemcpyAsync(counts.data(), cv.counts.p, counts.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost, h->stream));
            if (flags_out) CU(cudaMemcpyAsync(flags_out, h->flags.p, (size_t)n_centres * sizeof(int), cudaMemcpyDefault, h->stream));
            return 0;
        };
        rc = fetch();
    }
    cudaStreamSynchronize(h->stream);
    if (rc) {
        h->spans.clear();
        h->events_used = 0;
    } else {
        resolve_spans(h);
        release_upload_staging(h);
    }
    cv.release();
    if (rc) return rc;
    if ((rc = report_input_flags(h, bad, y_bad, "mmd_searchlight_cv"))) return rc;
    for (int i = 0; i < n_centres; ++i)
        for (int a = 0; a < n_alphas; ++a)
            for (int f = 0; f < n_folds; ++f) {
                const long n_val = fold_bounds[f + 1] - fold_bounds[f];
                const long den = n_val * n_val - n_val;                 // dist_mat.size - diag.size (eval.py:111)
                const size_t o = ((size_t)i * n_alphas + a) * n_folds + f;
                fold_scores_out[o] = den > 0 ? (double)counts[o] / (double)den : std::nan("");
            }
    return 0;
}


// ------------------------------------------------------------------------------------------
// whole-brain ridge: load / cross-validate / fit / predict (ridge_regression_decoding.py:111-135)
// ------------------------------------------------------------------------------------------
namespace {

// One ridge solve on the loaded problem without the trials `val` (empty = all trials), regulariser alpha.
// Leaves the TRANSPOSED solution in r.rhs: dual form (dp x mp) dual coefficients over the remaining trials `tr`,
// primal form (dp x vp) weights.  mp_out = padded order of the factored matrix.
int ridge_solve(mmd_handle* h, RidgeState& r, const std::vector<int>& tr, const std::vector<int>& val, double alpha, int* mp_out) {
    const int n_tr = (int)tr.size(), n_val = (int)val.size();
    if (to_device(h, r.idx_tr, tr.data(), tr.size())) return 1;
    if (to_device(h, r.idx_val, val.data(), val.size())) return 1;
    CU(cudaStreamSynchronize(h->stream));                       // the index vectors are pageable
    const int mp = r.dual ? round_up(n_tr, NB) : r.vp;
    *mp_out = mp;
    CU(r.A.ensure((size_t)mp * mp * sizeof(double)));
    CU(r.Lt.ensure((size_t)mp * mp * sizeof(double)));
    CU(r.linv.ensure((size_t)mp * NB * sizeof(double)));
    CU(r.linvT.ensure((size_t)mp * NB * sizeof(double)));
    CU(r.rhs.ensure((size_t)r.dp * mp * sizeof(double)));
    CU(r.flag.ensure(sizeof(int)));
    if (r.dual) {
        // K[tr, tr] + alpha I ; right-hand sides Y[tr]^T
        CU(launch_gather_block(r.G.as<double>(), r.np, r.idx_tr.as<int>(), n_tr, r.idx_tr.as<int>(), n_tr, r.A.as<double>(), mp, mp,
                               alpha, 1.0, h->stream));
        CU(launch_gather_block(r.Yt.as<double>(), r.np, nullptr, r.dp, r.idx_tr.as<int>(), n_tr, r.rhs.as<double>(), r.dp, mp, 0.0,
                               0.0, h->stream));
        h->launches += 2;
    } else {
        // X^T X - Xf^T Xf + alpha I ; right-hand sides (X^T Y - Xf^T Yf)^T
        CU(launch_gather_block(r.G.as<double>(), r.vp, nullptr, r.v, nullptr, r.v, r.A.as<double>(), mp, mp, alpha, 1.0, h->stream));
        CU(cudaMemcpyAsync(r.rhs.p, r.bt.p, (size_t)r.dp * mp * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
        h->launches += 1;
        if (n_val > 0) {
            const int kv = round_up(n_val, 16);
            CU(r.xft.ensure((size_t)r.vp * kv * sizeof(double)));
            CU(r.yft.ensure((size_t)r.dp * kv * sizeof(double)));
            CU(launch_gather_block(r.Xt.as<double>(), r.np, nullptr, r.vp, r.idx_val.as<int>(), n_val, r.xft.as<double>(), r.vp, kv,
                                   0.0, 0.0, h->stream));
            CU(launch_gather_block(r.Yt.as<double>(), r.np, nullptr, r.dp, r.idx_val.as<int>(), n_val, r.yft.as<double>(), r.dp, kv,
                                   0.0, 0.0, h->stream));
            DenseGemm g{};
            g.A = g.B = r.xft.as<double>(); g.lda = g.ldb = kv; g.C = r.A.as<double>(); g.ldc = mp;
            g.M = g.N = mp; g.K = kv; g.mode = 1; g.lower_only = 1;
            CU(launch_dense_gemm(g, h->stream));
            DenseGemm b{};
            b.A = r.yft.as<double>(); b.lda = kv; b.B = r.xft.as<double>(); b.ldb = kv; b.C = r.rhs.as<double>(); b.ldc = mp;
            b.M = r.dp; b.N = mp; b.K = kv; b.mode = 1;
            CU(launch_dense_gemm(b, h->stream));
            h->launches += 4;
        }
    }
    CU(cudaMemsetAsync(r.flag.p, 0, sizeof(int), h->stream));
    CU(dense_potrf(r.A.as<double>(), mp, r.linv.as<double>(), r.linvT.as<double>(), r.flag.as<int>(), h->stream));
    CU(launch_transpose_f64(r.A.as<double>(), mp, mp, mp, r.Lt.as<double>(), mp, mp, h->stream));
    CU(dense_potrs_t(r.A.as<double>(), r.Lt.as<double>(), r.linv.as<double>(), r.linvT.as<double>(), mp, r.rhs.as<double>(), r.dp,
                     h->stream));
    h->launches += 7 * (mp / NB) + 1;
    return 0;
}

// predictions of the solution in r.rhs for the trials `val` -> r.pred (n_val x dp)
int ridge_predict_val(mmd_handle* h, RidgeState& r, int n_tr, int n_val, int mp) {
    const int nvp = round_up(n_val, NB);
    CU(r.pred_t.ensure((size_t)r.dp * nvp * sizeof(double)));
    CU(r.pred.ensure((size_t)nvp * r.dp * sizeof(double)));
    CU(r.kval.ensure((size_t)nvp * mp * sizeof(double)));
    if (r.dual)      // K[val, tr]
        CU(launch_gather_block(r.G.as<double>(), r.np, r.idx_val.as<int>(), n_val, r.idx_tr.as<int>(), n_tr, r.kval.as<double>(), nvp,
                               mp, 0.0, 0.0, h->stream));
    else             // X[val]
        CU(launch_gather_block(r.X.as<double>(), r.vp, r.idx_val.as<int>(), n_val, nullptr, r.vp, r.kval.as<double>(), nvp, mp, 0.0,
                               0.0, h->stream));
    DenseGemm g{};
    g.A = r.rhs.as<double>(); g.lda = mp; g.B = r.kval.as<double>(); g.ldb = mp; g.C = r.pred_t.as<double>(); g.ldc = nvp;
    g.M = r.dp; g.N = n_val; g.K = mp;
    CU(launch_dense_gemm(g, h->stream));
    CU(launch_transpose_f64(r.pred_t.as<double>(), nvp, r.dp, n_val, r.pred.as<double>(), r.dp, nvp, h->stream));
    h->launches += 3;
    return 0;
}

}  // namespace

int mmd_ridge_load(mmd_handle* h, const void* X, int x_is_f32, int n, int v, const void* Y, int y_is_f32, int d) {
    if (!h) return 1;
    if (!X || !Y || n < 1 || v < 1 || d < 1) return fail(h, "mmd_ridge_load: bad arguments");
    CU(cudaSetDevice(h->device));
    if (!h->ridge) h->ridge = new RidgeState();
    RidgeState& r = *h->ridge;
    r.loaded = r.fitted = false;
    r.n = n; r.v = v; r.d = d;
    r.np = round_up(n, NB); r.vp = round_up(v, NB); r.dp = round_up(d, NB);
    r.dual = v > n;                                     // sklearn's cholesky solver: kernel form when n_features > n_samples
    CU(r.X.ensure((size_t)r.np * r.vp * sizeof(double)));
    CU(r.Xt.ensure((size_t)r.vp * r.np * sizeof(double)));
    CU(r.Y.ensure((size_t)r.np * r.dp * sizeof(double)));
    CU(r.Yt.ensure((size_t)r.dp * r.np * sizeof(double)));
    const size_t xb = (size_t)n * v * (x_is_f32 ? 4 : 8), yb = (size_t)n * d * (y_is_f32 ? 4 : 8);
    CU(r.stage.ensure(std::max(xb, yb)));
    span_begin(h, ST_XTY);
    CU(cudaMemcpyAsync(r.stage.p, X, xb, cudaMemcpyDefault, h->stream));
    CU(launch_widen_pad(r.stage.p, x_is_f32, v, n, v, r.X.as<double>(), r.np, r.vp, h->stream));
    CU(cudaMemcpyAsync(r.stage.p, Y, yb, cudaMemcpyDefault, h->stream));
    CU(launch_widen_pad(r.stage.p, y_is_f32, d, n, d, r.Y.as<double>(), r.np, r.dp, h->stream));
    CU(launch_transpose_f64(r.X.as<double>(), r.vp, r.np, r.vp, r.Xt.as<double>(), r.np, r.vp, h->stream));
    CU(launch_transpose_f64(r.Y.as<double>(), r.dp, r.np, r.dp, r.Yt.as<double>(), r.np, r.dp, h->stream));
    DenseGemm g{};
    if (r.dual) {                                       // K = X X^T
        CU(r.G.ensure((size_t)r.np * r.np * sizeof(double)));
        g.A = g.B = r.X.as<double>(); g.lda = g.ldb = r.vp; g.C = r.G.as<double>(); g.ldc = r.np; g.M = g.N = r.np; g.K = r.vp;
        CU(launch_dense_gemm(g, h->stream));
    } else {                                            // X^T X and (X^T Y)^T
        CU(r.G.ensure((size_t)r.vp * r.vp * sizeof(double)));
        CU(r.bt.ensure((size_t)r.dp * r.vp * sizeof(double)));
        g.A = g.B = r.Xt.as<double>(); g.lda = g.ldb = r.np; g.C = r.G.as<double>(); g.ldc = r.vp; g.M = g.N = r.vp; g.K = r.np;
        CU(launch_dense_gemm(g, h->stream));
        DenseGemm b{};
        b.A = r.Yt.as<double>(); b.lda = r.np; b.B = r.Xt.as<double>(); b.ldb = r.np; b.C = r.bt.as<double>(); b.ldc = r.vp;
        b.M = r.dp; b.N = r.vp; b.K = r.np;
        CU(launch_dense_gemm(b, h->stream));
    }
    span_end(h, 6);
    CU(cudaStreamSynchronize(h->stream));
    resolve_spans(h);
    r.loaded = true;
    return 0;
}

int mmd_ridge_cv(mmd_handle* h, const double* alphas, int n_alphas, const int32_t* fold_bounds, int n_folds, double* fold_scores_out) {
    if (!h) return 1;
    if (!h->ridge || !h->ridge->loaded) return fail(h, "mmd_ridge_load must be called before mmd_ridge_cv");
    RidgeState& r = *h->ridge;
    if (!alphas || !fold_bounds || !fold_scores_out || n_alphas < 1) return fail(h, "mmd_ridge_cv: bad arguments");
    if (n_folds < 2 || n_folds > r.n)
        return fail(h, "mmd_ridge_cv: k-fold cross-validation requires at least one train/test split by setting n_splits=2 or more "
                       "and at most n_samples=%d, got n_splits=%d", r.n, n_folds);
    if (fold_bounds[0] != 0 || fold_bounds[n_folds] != r.n) return fail(h, "mmd_ridge_cv: fold bounds must run from 0 to n");
    for (int f = 0; f < n_folds; ++f)
        if (fold_bounds[f + 1] <= fold_bounds[f]) return fail(h, "mmd_ridge_cv: empty fold %d", f);
    CU(cudaSetDevice(h->device));
    CU(r.count.ensure(sizeof(unsigned long long)));
    int any_flag = 0;
    for (int f = 0; f < n_folds; ++f) {
        const int lo = fold_bounds[f], hi = fold_bounds[f + 1], n_val = hi - lo;
        std::vector<int> tr, val;
        for (int i = 0; i < r.n; ++i) (i >= lo && i < hi ? val : tr).push_back(i);
        const int nvp = round_up(n_val, NB);
        CU(r.yval.ensure((size_t)nvp * r.dp * sizeof(double)));
        CU(r.dist.ensure((size_t)n_val * n_val * sizeof(double)));
        CU(r.sq.ensure((size_t)2 * n_val * sizeof(double)));
        for (int a = 0; a < n_alphas; ++a) {
            int mp = 0;
            span_begin(h, ST_SOLVE);
            if (ridge_solve(h, r, tr, val, alphas[a], &mp)) return 1;
            span_end(h, 0);
            span_begin(h, ST_PREDICT);
            if (ridge_predict_val(h, r, (int)tr.size(), n_val, mp)) return 1;
            span_end(h, 0);
            span_begin(h, ST_SCORE);
            CU(launch_gather_block(r.Y.as<double>(), r.dp, r.idx_val.as<int>(), n_val, nullptr, r.dp, r.yval.as<double>(), nvp, r.dp,
                                   0.0, 0.0, h->stream));
            CU(launch_cosine_dist(r.pred.as<double>(), n_val, r.yval.as<double>(), n_val, r.d, r.dp, r.sq.as<double>(),
                                  r.sq.as<double>() + n_val, r.dist.as<double>(), h->stream));
            CU(launch_pairwise_count(r.dist.as<double>(), n_val, n_val, r.count.as<unsigned long long>(), h->stream));
            span_end(h, 5);
            unsigned long long count = 0;
            int flag = 0;
            CU(cudaMemcpyAsync(&count, r.count.p, sizeof(count), cudaMemcpyDeviceToHost, h->stream));
            CU(cudaMemcpyAsync(&flag, r.flag.p, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
            CU(cudaStreamSynchronize(h->stream));
            any_flag |= flag;
            const long den = (long)n_val * n_val - n_val;
            fold_scores_out[(size_t)a * n_folds + f] = den > 0 ? (double)count / (double)den : std::nan("");
        }
    }
    resolve_spans(h);
    if (any_flag) return fail(h, "mmd_ridge_cv: a regularised Gram matrix was not positive definite");
    return 0;
}

int mmd_ridge_fit(mmd_handle* h, double alpha, double* coef_out) {
    if (!h) return 1;
    if (!h->ridge || !h->ridge->loaded) return fail(h, "mmd_ridge_load must be called before mmd_ridge_fit");
    RidgeState& r = *h->ridge;
    CU(cudaSetDevice(h->device));
    std::vector<int> tr(r.n), val;
    for (int i = 0; i < r.n; ++i) tr[i] = i;
    int mp = 0;
    span_begin(h, ST_SOLVE);
    if (ridge_solve(h, r, tr, val, alpha, &mp)) return 1;
    CU(r.coef_t.ensure((size_t)r.dp * r.vp * sizeof(double)));
    if (r.dual) {                                       // W^T = C^T X  (dp x vp)
        DenseGemm g{};
        g.A = r.rhs.as<double>(); g.lda = mp; g.B = r.Xt.as<double>(); g.ldb = r.np; g.C = r.coef_t.as<double>(); g.ldc = r.vp;
        g.M = r.dp; g.N = r.vp; g.K = mp;               // mp == np: all trials
        CU(launch_dense_gemm(g, h->stream));
    } else {
        CU(cudaMemcpyAsync(r.coef_t.p, r.rhs.p, (size_t)r.dp * r.vp * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    }
    span_end(h, 1);
    int flag = 0;
    CU(cudaMemcpyAsync(&flag, r.flag.p, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    if (coef_out)                                       // sklearn's coef_ layout: (n_targets, n_features)
        CU(cudaMemcpy2DAsync(coef_out, (size_t)r.v * sizeof(double), r.coef_t.p, (size_t)r.vp * sizeof(double),
                             (size_t)r.v * sizeof(double), r.d, cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    resolve_spans(h);
    if (flag) return fail(h, "mmd_ridge_fit: the regularised Gram matrix is not positive definite");
    r.fitted = true;
    return 0;
}

int mmd_ridge_predict(mmd_handle* h, const void* X, int x_is_f32, int m, double* out) {
    if (!h) return 1;
    if (!h->ridge || !h->ridge->fitted) return fail(h, "mmd_ridge_fit must be called before mmd_ridge_predict");
    if (!X || !out || m < 1) return fail(h, "mmd_ridge_predict: bad arguments");
    RidgeState& r = *h->ridge;
    CU(cudaSetDevice(h->device));
    const int mpad = round_up(m, NB);
    const size_t xb = (size_t)m * r.v * (x_is_f32 ? 4 : 8);
    CU(r.stage.ensure(xb));
    CU(r.kval.ensure((size_t)mpad * r.vp * sizeof(double)));
    CU(r.pred_t.ensure((size_t)r.dp * mpad * sizeof(double)));
    CU(r.pred.ensure((size_t)mpad * r.dp * sizeof(double)));
    span_begin(h, ST_PREDICT);
    CU(cudaMemcpyAsync(r.stage.p, X, xb, cudaMemcpyDefault, h->stream));
    CU(launch_widen_pad(r.stage.p, x_is_f32, r.v, m, r.v, r.kval.as<double>(), mpad, r.vp, h->stream));
    DenseGemm g{};
    g.A = r.coef_t.as<double>(); g.lda = r.vp; g.B = r.kval.as<double>(); g.ldb = r.vp; g.C = r.pred_t.as<double>(); g.ldc = mpad;
    g.M = r.dp; g.N = m; g.K = r.vp;
    CU(launch_dense_gemm(g, h->stream));
    CU(launch_transpose_f64(r.pred_t.as<double>(), mpad, r.dp, m, r.pred.as<double>(), r.dp, mpad, h->stream));
    span_end(h, 3);
    CU(cudaMemcpy2DAsync(out, (size_t)r.d * sizeof(double), r.pred.p, (size_t)r.dp * sizeof(double), (size_t)r.d * sizeof(double), m,
                         cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    resolve_spans(h);
    return 0;
}

// ---- eval.py primitives on arbitrary shapes (host float64 buffers) ----
static int stage_eval_inputs(mmd_handle* h, const double* predictions, int n_pred, const double* latents, int n_lat,
                             int n_dims, int std_pred, int std_lat) {
    if (to_device(h, h->pa_pred, predictions, (size_t)n_pred * n_dims)) return 1;
    if (to_device(h, h->pa_lat, latents, (size_t)n_lat * n_dims)) return 1;
    CU(h->pa_dist.ensure((size_t)n_pred * n_lat * sizeof(double)));
    CU(h->pa_sq.ensure((size_t)(n_pred + n_lat) * sizeof(double)));
    CU(h->pa_count.ensure(sizeof(unsigned long long)));
    if (std_pred) CU(launch_standardize_columns(h->pa_pred.as<double>(), n_pred, n_dims, n_dims, h->stream));
    if (std_lat) CU(launch_standardize_columns(h->pa_lat.as<double>(), n_lat, n_dims, n_dims, h->stream));
    CU(launch_cosine_dist(h->pa_pred.as<double>(), n_pred, h->pa_lat.as<double>(), n_lat, n_dims, n_dims,
                          h->pa_sq.as<double>(), h->pa_sq.as<double>() + n_pred, h->pa_dist.as<double>(), h->stream));
    h->launches += 3 + (std_pred ? 1 : 0) + (std_lat ? 1 : 0);
    return 0;
}

int mmd_cosine_distances(mmd_handle* h, const double* predictions, int n_pred, const double* originals, int n_orig,
                         int n_dims, double* dist_out) {
    if (!h) return 1;
    if (!predictions || !originals || !dist_out) return fail(h, "mmd_cosine_distances: null pointer");
    if (n_pred < 1 || n_orig < 1 || n_dims < 1) return fail(h, "mmd_cosine_distances: empty input");
    CU(cudaSetDevice(h->device));
    if (stage_eval_inputs(h, predictions, n_pred, originals, n_orig, n_dims, 0, 0)) return 1;
    CU(cudaMemcpyAsync(dist_out, h->pa_dist.p, (size_t)n_pred * n_orig * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return 0;
}

int mmd_pairwise_accuracy(mmd_handle* h, const double* latents, int n_lat, const double* predictions, int n_pred,
                          int n_dims, int standardize_predictions, int standardize_latents, double* acc_out) {
    if (!h) return 1;
    if (!latents || !predictions || !acc_out) return fail(h, "mmd_pairwise_accuracy: null pointer");
    if (n_pred < 1 || n_lat < 1 || n_dims < 1) return fail(h, "mmd_pairwise_accuracy: empty input");
    if (n_pred > n_lat && n_lat != 1)
        return fail(h, "mmd_pairwise_accuracy: %d predictions but %d candidate latents: the diagonal (eval.py:108) "
                       "cannot be broadcast against the distance matrix", n_pred, n_lat);
    CU(cudaSetDevice(h->device));
    if (stage_eval_inputs(h, predictions, n_pred, latents, n_lat, n_dims, standardize_predictions, standardize_latents)) return 1;
    CU(launch_pairwise_count(h->pa_dist.as<double>(), n_pred, n_lat, h->pa_count.as<unsigned long long>(), h->stream));
    h->launches += 1;
    unsigned long long count = 0;
    CU(cudaMemcpyAsync(&count, h->pa_count.p, sizeof(count), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    const long rows = n_pred < n_lat ? n_pred : n_lat;
    *acc_out = (double)count / (double)((long)n_pred * n_lat - rows);
    return 0;
}

int mmd_dist_mat_to_pairwise_acc(mmd_handle* h, const double* dist_mat, int n_rows, int n_cols, double* acc_out) {
    if (!h) return 1;
    if (!dist_mat || !acc_out) return fail(h, "mmd_dist_mat_to_pairwise_acc: null pointer");
    if (n_rows < 1 || n_cols < 1) return fail(h, "mmd_dist_mat_to_pairwise_acc: empty matrix");
    if (n_rows > n_cols && n_cols != 1) return fail(h, "mmd_dist_mat_to_pairwise_acc: more rows than columns (eval.py:108-109 cannot broadcast)");
    CU(cudaSetDevice(h->device));
    if (to_device(h, h->pa_dist, dist_mat, (size_t)n_rows * n_cols)) return 1;
    CU(h->pa_count.ensure(sizeof(unsigned long long)));
    CU(launch_pairwise_count(h->pa_dist.as<double>(), n_rows, n_cols, h->pa_count.as<unsigned long long>(), h->stream));
    h->launches += 1;
    unsigned long long count = 0;
    CU(cudaMemcpyAsync(&count, h->pa_count.p, sizeof(count), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    const long rows = n_rows < n_cols ? n_rows : n_cols;
    *acc_out = (double)count / (double)((long)n_rows * n_cols - rows);
    return 0;
}

// ---- group-level statistics on score maps (HOST buffers) ----
int mmd_ttest_1samp(mmd_handle* h, const double* data, int n_subjects, int n_vertices, double popmean, double sigma,
                    int flags, double* t_out) {
    if (!h) return 1;
    if (!data || !t_out) return fail(h, "mmd_ttest_1samp: null pointer");
    if (n_subjects < 2 || n_subjects > 128) return fail(h, "mmd_ttest_1samp: 2..128 subjects are supported (got %d)", n_subjects);
    if (n_vertices < 1) return fail(h, "mmd_ttest_1samp: empty map");
    CU(cudaSetDevice(h->device));
    if (to_device(h, h->st_in, data, (size_t)n_subjects * n_vertices)) return 1;
    CU(h->st_out.ensure((size_t)n_vertices * sizeof(double)));
    CU(launch_ttest(h->st_in.as<double>(), n_subjects, n_vertices, popmean, sigma, flags & 1, (flags >> 1) & 1, h->st_out.as<double>(),
                    h->stream));
    h->launches += 1;
    CU(cudaMemcpyAsync(t_out, h->st_out.p, (size_t)n_vertices * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return 0;
}

static int check_edges(mmd_handle* h, const int32_t* edges, int n_edges, int n_vertices, const char* who) {
    if (n_edges < 0 || (n_edges > 0 && !edges)) return fail(h, "%s: bad edge list", who);
    for (long i = 0; i < 2L * n_edges; ++i)
        if (edges[i] < 0 || edges[i] >= n_vertices) return fail(h, "%s: edge endpoint %d outside [0, %d)", who, edges[i], n_vertices);
    return 0;
}

int mmd_tfce(mmd_handle* h, const double* values, int n_maps, int n_vertices, const int32_t* edges, int n_edges,
             double h_exponent, double e_exponent, double dh, float* tfce_out) {
    if (!h) return 1;
    if (!values || !tfce_out) return fail(h, "mmd_tfce: null pointer");
    if (n_maps < 1 || n_vertices < 1) return fail(h, "mmd_tfce: empty input");
    if (check_edges(h, edges, n_edges, n_vertices, "mmd_tfce")) return 1;
    CU(cudaSetDevice(h->device));
    if (to_device(h, h->st_edges, edges, (size_t)2 * std::max(n_edges, 1))) return 1;
    const int chunk = std::max(1, std::min(n_maps, (int)std::min<size_t>(4096, ((size_t)2 << 30) / ((size_t)n_vertices * 8))));
    CU(h->st_label.ensure((size_t)chunk * n_vertices * sizeof(int)));
    CU(h->st_count.ensure((size_t)chunk * n_vertices * sizeof(int)));
    CU(h->st_out.ensure((size_t)chunk * n_vertices * sizeof(float)));
    for (int m0 = 0; m0 < n_maps; m0 += chunk) {
        const int n = std::min(chunk, n_maps - m0);
        if (to_device(h, h->st_in, values + (size_t)m0 * n_vertices, (size_t)n * n_vertices)) return 1;
        CU(launch_tfce(h->st_in.as<double>(), n, n_vertices, h->st_edges.as<int>(), n_edges, h_exponent, e_exponent, dh,
                       h->st_label.as<int>(), h->st_count.as<int>(), h->st_out.as<float>(), h->stream));
        h->launches += 1;
        CU(cudaMemcpyAsync(tfce_out + (size_t)m0 * n_vertices, h->st_out.p, (size_t)n * n_vertices * sizeof(float),
                           cudaMemcpyDeviceToHost, h->stream));
        CU(cudaStreamSynchronize(h->stream));
    }
    return 0;
}

int mmd_cluster_labels(mmd_handle* h, const double* values, int n_vertices, const int32_t* edges, int n_edges,
                       double threshold, int32_t* labels_out) {
    if (!h) return 1;
    if (!values || !labels_out) return fail(h, "mmd_cluster_labels: null pointer");
    if (n_vertices < 1) return fail(h, "mmd_cluster_labels: empty map");
    if (check_edges(h, edges, n_edges, n_vertices, "mmd_cluster_labels")) return 1;
    CU(cudaSetDevice(h->device));
    if (to_device(h, h->st_edges, edges, (size_t)2 * std::max(n_edges, 1))) return 1;
    if (to_device(h, h->st_in, values, (size_t)n_vertices)) return 1;
    CU(h->st_label.ensure((size_t)n_vertices * sizeof(int)));
    CU(h->st_count.ensure((size_t)n_vertices * sizeof(int)));
    CU(h->st_out.ensure((size_t)n_vertices * sizeof(int)));
    CU(launch_cluster_labels(h->st_in.as<double>(), n_vertices, h->st_edges.as<int>(), n_edges, threshold,
                             h->st_label.as<int>(), h->st_count.as<int>(), h->st_out.as<int>(), h->stream));
    h->launches += 1;
    CU(cudaMemcpyAsync(labels_out, h->st_out.p, (size_t)n_vertices * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return 0;
}

}  // extern "C"
