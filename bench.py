"""Benchmark of the searchlight ridge-decoding hot path (BASELINE.json metric: searchlight centres/sec).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (BASELINE.json configs[1]): one synthetic fsaverage7-sized hemisphere (icosphere-7, 163,842 vertices),
k = 750 nearest neighbours, N_train = 4000, 140 test + 3 imagery trials, D = 1024 (ImageBind), alpha = 1,
modality-agnostic.  A *step* is one pass of the hot path over a whole hemisphere (all 163,842 centres; with
`--centres-per-step` a range of that many consecutive centres).  `value` is measured with the trial arrays already
staged in HBM (CUDA events on the engine's stream); `e2e` calls the public drop-in `custom_search_light` with HOST
buffers every step (pinned X, y and the int64 adjacency in; the object array of score dicts out).

N > 1 (torchrun): the headline is the north star's split -- ONE hemisphere, its centres in Morton order, rank r scores the
r-th contiguous range, the score blocks are all-gathered once per step ("scaling": "strong") -- and the replica layout
(every rank runs one whole hemisphere job of its own, no data-path collective) is measured in the same run and reported in
`replicas`.  `--scaling weak` makes the replica layout the headline.

At N = 1 the line also carries `cpu_baseline` (the UNMODIFIED reference from baseline/_ref -- custom_search_light + sklearn
Ridge -- on a contiguous sample of the same workload, all host cores) and `parity_sample` (the same centres scored on the GPU,
compared).  `--impl reference` times that reference alone, step by step, and prints the own arm's `config`.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "searchlight centers/sec (750-nbr ridge+pairwise acc)"
WORKLOAD = ("searchlight imagebind --n-neighbors 750 modality-agnostic, one synthetic fsaverage7-sized hemisphere "
            "(BASELINE.json configs[1])")
UNIT = "centers/s"


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=3)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--mesh-level", type=int, default=7, help="icosphere level (7 = 163,842 vertices)")
    p.add_argument("--n-neighbors", type=int, default=750)
    p.add_argument("--n-train", type=int, default=4000)
    p.add_argument("--n-dims", type=int, default=1024)
    p.add_argument("--centres-per-step", type=int, default=0,
                   help="centres per step and GPU; 0 = the whole hemisphere (weak) or the rank's whole shard (strong)")
    p.add_argument("--scaling", default="strong", choices=["weak", "strong"],
                   help="N > 1 only.  strong (default, north_star (5)): ONE hemisphere, centres sharded across the GPUs, the "
                        "replica layout measured beside it; weak: every GPU runs one whole hemisphere job per step")
    p.add_argument("--no-replicas", action="store_true", help="strong scaling: skip the replica-layout measurement")
    p.add_argument("--cpu-centres", type=int, default=768, help="centres of the CPU-baseline sample")
    p.add_argument("--batch-centres", type=int, default=0, help="engine knob: centres per kernel batch (0 = default)")
    p.add_argument("--cluster-size", type=int, default=0, help="engine knob: centres per union-Gram cluster (0 = default)")
    p.add_argument("--group-size", type=int, default=0, help="engine knob: centres per prefix-sharing group (1 = off, 0 = default)")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-e2e", action="store_true")
    return p.parse_args()


# ----------------------------------------------------------------------------------------------------------
# synthetic workload
# ----------------------------------------------------------------------------------------------------------
def make_workload(args, device, seed=0):
    """Mesh on the host (numpy), trials on the GPU (torch as plumbing): low-rank stimulus signal with smooth spatial
    loadings plus noise, z-scored with the TRAIN statistics and cast to float16 exactly like searchlight.py:199-248."""
    import torch
    from multimodal_decoding_b200 import synthetic as syn
    mesh = syn.make_mesh(level=args.mesh_level, seed=7)
    coords = mesh.coords
    V = len(coords)
    n_tr, n_te, n_im, D, r = args.n_train, 140, 3, args.n_dims, 32
    g = torch.Generator(device=device).manual_seed(seed)
    z_tr = torch.randn(n_tr, r, device=device, generator=g)
    z_half = torch.randn(70, r, device=device, generator=g)
    z = torch.cat([z_tr, z_half, z_half, torch.randn(n_im, r, device=device, generator=g)])
    mix = torch.randn(r, D, device=device, generator=g)
    Y = z @ mix + 0.5 * torch.randn(len(z), D, device=device, generator=g)
    Y[n_tr + 70:n_tr + 140] = Y[n_tr:n_tr + 70]                     # caption i and image i share a latent (`avg`)
    c = torch.from_numpy(coords).to(device)
    bumps = c[torch.randint(0, V, (r,), device=device, generator=g)]
    load = torch.exp(-0.5 * torch.cdist(c, bumps) ** 2 / 20.0 ** 2)  # (V, r) smooth loadings, sigma 20 mm
    X = torch.empty(len(z), V, device=device, dtype=torch.float32)
    for lo in range(0, len(z), 512):                                # row blocks keep the temporaries small
        hi = min(lo + 512, len(z))
        X[lo:hi] = 0.35 * (z[lo:hi] @ load.T) + torch.randn(hi - lo, V, device=device, generator=g)

    def zscore(a):
        mu, sd = a[:n_tr].mean(0), a[:n_tr].std(0, unbiased=False)
        sd[sd == 0] = 1
        return ((a - mu) / sd).to(torch.float16)

    X16, Y16 = zscore(X), zscore(Y)
    del X
    ids = (list(range(n_tr)), list(range(n_tr, n_tr + n_te)), list(range(n_tr + n_te, n_tr + n_te + n_im)))
    return coords, X16, Y16, ids


# ----------------------------------------------------------------------------------------------------------
# clocks during the timed region (B200_PROFILING.md recipe)
# ----------------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *exc):
        if self.proc:
            self.proc.terminate()
            self.thread.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------------------------------------
# CPU baseline: the oracle's restatement of the reference loop (float64, BLAS on all host threads)
# ----------------------------------------------------------------------------------------------------------
def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def _cpu_chunk(X, Y, A_rows, ids, centres):
    """Worker of the CPU arm: the oracle's per-centre loop on a chunk of centres, one BLAS thread (the reference
    runs `n_jobs` loky processes over chunks of centres, searchlight.py:130-148)."""
    from threadpoolctl import threadpool_limits
    from oracle import searchlight_oracle as orc          # bench.py's cpu_baseline leg may run the oracle
    with threadpool_limits(limits=1):
        return orc.search_light(X, Y, A_rows, ids[0], ids[1], ids[2], 1.0, orc.default_is_caption(len(ids[1])),
                                centres=list(centres))


class CpuArm:
    """The reference's CPU path for this workload: joblib processes over chunks of centres on all host cores."""

    def __init__(self, X, Y, A_rows, ids):
        from joblib import Parallel
        self.X, self.Y, self.A, self.ids = X, Y, A_rows, ids
        self.cores = host_cores()
        self.pool = Parallel(n_jobs=self.cores)
        self.pool.__enter__()                              # keep the worker processes between steps

    def close(self):
        self.pool.__exit__(None, None, None)

    def step(self, centres):
        from joblib import delayed
        chunks = [c for c in np.array_split(np.asarray(centres), self.cores) if len(c)]
        out = self.pool(delayed(_cpu_chunk)(self.X, self.Y, self.A, self.ids, c) for c in chunks)
        return np.concatenate(out)


def cpu_centres_per_sec(X, Y, A_rows, ids, n_centres, budget_s=30.0):
    arm = CpuArm(X, Y, A_rows, ids)
    per_step = 2 * arm.cores
    arm.step(range(min(per_step, len(A_rows))))          # start the workers, page the arrays in
    t0, done = time.perf_counter(), 0
    while done < min(n_centres, len(A_rows)):
        hi = min(done + per_step, n_centres, len(A_rows))
        arm.step(range(done, hi))
        done = hi
        if time.perf_counter() - t0 > budget_s:
            break
    dt = time.perf_counter() - t0
    arm.close()
    return done / dt, done, dt, arm.cores


SCORE_KEYS = ("pairwise_acc_captions", "pairwise_acc_images", "pairwise_acc_cross_captions_to_images",
              "pairwise_acc_cross_images_to_captions", "pairwise_acc_imagery", "pairwise_acc_imagery_whole_test_set",
              "pairwise_acc_imagery_no_std", "pairwise_acc_imagery_whole_test_set_no_std")


def reference_available():
    from baseline import make_ref
    return make_ref.available()


class RefArm:
    """The UNMODIFIED reference (baseline/_ref, copied from /root/reference by baseline/make_ref.py at build time):
    custom_search_light(X16, Y16, Ridge(alpha=1, fit_intercept=False), A[sample], ...) -> joblib over chunks of centres ->
    sklearn Ridge.fit / predict + eval.py scoring, on the full-width X of the workload (the strided column gather of
    searchlight.py:110 is paid).  X is handed over as a read-only np.memmap so that joblib passes it to its workers by
    reference instead of dumping 1.4 GB per call (a per-call cost the real 164k-centre run amortises)."""

    def __init__(self, X, Y, ids):
        import tempfile
        from baseline import make_ref
        self.sl = make_ref.activate()
        base = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else None
        self.dir = tempfile.mkdtemp(prefix="mmd_ref_", dir=base)
        path = os.path.join(self.dir, "X.npy")
        np.save(path, np.asarray(X))
        self.X = np.load(path, mmap_mode="r")
        self.Y = np.ascontiguousarray(Y)
        self.ids = [list(map(int, i)) for i in ids]
        self.cores = host_cores()
        import sklearn
        self.versions = f"sklearn {sklearn.__version__}, numpy {np.__version__}"
        try:
            from threadpoolctl import threadpool_info
            self.blas = ", ".join(sorted({f"{i.get('internal_api')} {i.get('version')}" for i in threadpool_info()
                                          if i.get("user_api") == "blas"})) or "unknown"
        except Exception:
            self.blas = "unknown"

    def close(self):
        import shutil
        del self.X
        shutil.rmtree(self.dir, ignore_errors=True)

    def run(self, A_rows, n_jobs):
        from sklearn.linear_model import Ridge
        out = self.sl.custom_search_light(self.X, self.Y, estimator=Ridge(alpha=1.0, fit_intercept=False), A=np.asarray(A_rows),
                                          train_ids=self.ids[0], test_ids=self.ids[1], imagery_ids=self.ids[2], n_jobs=n_jobs)
        return np.array([[d[k] for k in SCORE_KEYS] for d in out], dtype=np.float64)


def reference_centres_per_sec(arm, A_rows, n_centres):
    """The reference both ways BASELINE.md section 4 names: n_jobs = cores (its default layout) and n_jobs = 1 with the BLAS
    threads; returns the better rate, its description and the scores of the joblib run (the parity sample)."""
    n = min(n_centres, len(A_rows))
    arm.run(A_rows[:2 * arm.cores], arm.cores)                    # start the loky workers (they import torch via data.py)
    t0 = time.perf_counter()
    scores = arm.run(A_rows[:n], arm.cores)
    dt_jobs = time.perf_counter() - t0
    n1 = min(n, max(8, int(n / arm.cores)))
    t0 = time.perf_counter()
    arm.run(A_rows[:n1], 1)
    dt_one = time.perf_counter() - t0
    rate_jobs, rate_one = n / dt_jobs, n1 / dt_one
    desc = (f"unmodified custom_search_light + sklearn Ridge(alpha=1) on {n} consecutive centres of the same workload, "
            f"full-width X: n_jobs={arm.cores} -> {rate_jobs:.1f} centres/s ({dt_jobs:.1f} s); n_jobs=1 with BLAS threads on "
            f"{n1} centres -> {rate_one:.1f} centres/s; the better one is reported ({arm.versions}, BLAS {arm.blas})")
    return max(rate_jobs, rate_one), desc, scores


# ----------------------------------------------------------------------------------------------------------
# roofline bookkeeping
# ----------------------------------------------------------------------------------------------------------
# DRAM bytes per centre of the per-centre kernels: dram__bytes_read.sum + dram__bytes_write.sum of one `ncu --set full` capture
# of a full batch, divided by the batch's centres -- extracted from the .ncu-rep files by tools/ncu_traffic.py into
# profiles/r2_traffic.json (capture file names and sha256 inside).  factor_solve = the per-centre launch + the shared-core launch.
TRAFFIC_FILE = os.path.join(ROOT, "profiles", "r2_traffic.json")
FP64_TENSOR_PEAK_TFLOPS = 37.1      # DMMA m8n8k4, measured on this pool with tools/peak_probe.cu (profiles/r1_measured_peaks_fp64.log)


def load_traffic():
    if not os.path.exists(TRAFFIC_FILE):
        return {}, None
    with open(TRAFFIC_FILE) as f:
        d = json.load(f)
    k = d["kernels"]
    per = {name: v["dram_bytes_per_centre"] for name, v in k.items()}
    if "factor_solve_core" in per:
        per["factor_solve"] = per.get("factor_solve", 0.0) + per.pop("factor_solve_core")
    src = "profiles/r2_traffic.json <- " + ", ".join(f"{v['capture']} (sha256 {v['sha256'][:12]})" for v in k.values())
    return per, src


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            d = json.load(f)
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"], "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "source": "fallback"}


def stage_flops_per_centre(stage, k, n_train, n_eval_pad, D):
    """EXECUTED float64 flops per centre of each per-centre kernel.  The Gram stage is
    accounted from the library's own executed-flop counter because it depends on the cluster unions."""
    kp = (k + 63) // 64 * 64
    ldt = (n_train + 63) // 64 * 64
    Dp = (D + 63) // 64 * 64
    return {
        "factor_solve": kp ** 3 / 3.0 + 2.0 * kp * kp * n_eval_pad,     # Cholesky + forward and backward solves, 144 RHS
        "predict": 2.0 * n_eval_pad * kp * Dp,
        "syrk_algorithmic": 2.0 * n_train * k * k,                      # what the reference executes per centre
        "ldt": ldt,
    }[stage]


def roofline_entry(stats, n_centres, args, peaks):
    ms = stats["stage_ms"]
    per_centre = ["prep", "syrk", "factor_solve", "predict", "score"]
    top = max(per_centre, key=lambda s: ms[s])
    k, ntr, D = args.n_neighbors, args.n_train, args.n_dims
    t = ms[top] * 1e-3
    shares = {s: round(ms[s] / max(sum(ms[x] for x in per_centre), 1e-9), 4) for s in per_centre}
    n_batches = max(stats["batches"], 1)                             # every batch launches each per-centre kernel once
    config2 = (k, ntr, D) == (750, 4000, 1024)
    per_centre_bytes, traffic_src = load_traffic()
    traffic = round(per_centre_bytes[top] * n_centres / n_batches) if config2 and top in per_centre_bytes else None
    out = {"kernel": top, "stage_ms": {s: round(ms[s], 3) for s in per_centre}, "stage_share": shares,
           "launches_of_kernel": n_batches, "avg_launch_ms": round(ms[top] / n_batches, 3),
           "traffic_source": (f"ncu --set full capture of one full batch: DRAM bytes per centre x centres per launch; {traffic_src}"
                              if traffic else None)}
    if top in ("syrk", "factor_solve", "predict"):
        executed = (stats["syrk_flops"] if top == "syrk" else stats["solve_flops"] if top == "factor_solve"
                    else stage_flops_per_centre(top, k, ntr, 144, D) * n_centres)
        achieved = executed / t / 1e12
        out.update({"bound": "tensor", "achieved": round(achieved, 3), "peak": FP64_TENSOR_PEAK_TFLOPS,
                    "unit": "TFLOP/s", "frac": round(achieved / FP64_TENSOR_PEAK_TFLOPS, 4), "traffic": traffic,
                    "peak_source": "fp64 tensor pipe (DMMA) measured with tools/peak_probe.cu on this pool; "
                                   "MEASURED_PEAKS.json holds no fp64 figure",
                    "frac_of_bf16_peak": round(achieved / peaks["bf16_tflops"], 5),
                    "flops": "executed float64 flops (<= the reference's algorithmic count)"})
        if top == "factor_solve":
            full = stage_flops_per_centre(top, k, ntr, 144, D) * n_centres
            out["prefix_sharing"] = {
                "grouped_centres_frac": round(stats["grouped_centres"] / max(n_centres, 1), 4),
                "mean_shared_prefix": round(stats["sum_shared_prefix"] / max(stats["grouped_centres"], 1), 1),
                "shared_core_factorisations": int(stats["core_tasks"]),
                "executed_over_unshared_flops": round(executed / full, 4)}
        if top == "syrk":
            out["algorithmic_equiv_tflops"] = round(2.0 * ntr * k * k * n_centres / t / 1e12, 2)
    else:
        n_eval = 143
        bytes_alg = n_centres * (152.0 * ((D + 63) // 64 * 64) * 8 + 64) if top == "score" else n_centres * k * 8.0
        achieved = bytes_alg / t / 1e9
        out.update({"bound": "hbm", "achieved": round(achieved, 2), "peak": peaks["hbm_gbs"], "unit": "GB/s",
                    "frac": round(achieved / peaks["hbm_gbs"], 4), "traffic": traffic,
                    "peak_source": peaks["source"]})
    return out


def workload_config(args, V, world, weak, cps):
    """The `config` of the JSON line: the workload (BASELINE.json configs[1]) and how it is laid out over the GPUs.  Both arms
    print the same dict (the reference arm times a bounded sample of this workload, stated in its cpu_baseline.sample)."""
    return {"workload": WORKLOAD,
            "n_vertices": int(V), "n_neighbors": args.n_neighbors, "n_train": args.n_train, "n_test": 140, "n_imagery": 3,
            "n_dims": args.n_dims, "alpha": 1.0, "centres_per_step_per_gpu": int(cps),
            "sharding": (f"{world} hemisphere job(s), one per GPU" if weak else
                         f"one hemisphere, Morton-ordered vertex ranges x{world}, score maps all-gathered"),
            "l2": "inputs larger than L2 (Xt 1.3 GB, per-step workspaces > 126 MB); no flush needed"}


# ----------------------------------------------------------------------------------------------------------
# arms
# ----------------------------------------------------------------------------------------------------------
def dist_setup(args):
    import torch
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return rank, world, local


def knn_rows(eng, coords, k, lo, hi, chunk=32768):
    """Adjacency rows lo..hi (the k-NN branch of get_adjacency_matrix) from the GPU kernel, as int64 like sklearn."""
    out = np.empty((hi - lo, k), dtype=np.int64)
    for s in range(lo, hi, chunk):
        e = min(s + chunk, hi)
        _, idx = eng.knn(coords, k, s, e)
        out[s - lo:e - lo] = idx
    return out


def measure(args, mode, rank, world, local, eng, with_cpu):
    """One measurement of the hot path in `mode`:
      "strong"  ONE hemisphere, centres in Morton order, rank r takes the r-th contiguous range (vertex-sharded,
                north_star (5)); the score blocks are all-gathered once per step
      "weak"    every rank runs one whole hemisphere job of its own (replicas; BASELINE configs[4] distributes whole
                searchlights), no collective on the data path
    Returns the pieces of the JSON line."""
    import torch
    import torch.distributed as dist
    from multimodal_decoding_b200.analyses.decoding.searchlight import searchlight as sl
    from multimodal_decoding_b200.sharding import gather_scores, spatial_order, split_bounds
    from sklearn.linear_model import Ridge

    device = torch.device("cuda", local)
    stream = torch.cuda.ExternalStream(eng.stream_ptr, device=device)
    K, W, k = args.steps, args.warmup, args.n_neighbors
    weak = mode == "weak" or world == 1

    t_setup = time.perf_counter()
    coords, X16, Y16, ids = make_workload(args, device, seed=rank if weak else 0)
    V = len(coords)
    torch.cuda.synchronize()
    t_knn = time.perf_counter()
    A_all = knn_rows(eng, coords, k, 0, V)                      # k-NN branch of get_adjacency_matrix, all vertices
    knn_s = time.perf_counter() - t_knn
    # weak: this rank's job is one whole hemisphere (its own subject/hemisphere, seed = rank), in the
    # reference's vertex order.  strong: one hemisphere; the rows are put in Morton order first so that the
    # contiguous range a rank receives (the reference's array_split chunking) is a compact patch of cortex.
    order = np.arange(V) if weak and args.centres_per_step == 0 else spatial_order(coords)
    A_all = A_all[order]
    lo, hi = (0, V) if weak else split_bounds(V, world)[rank]
    cps = hi - lo if args.centres_per_step == 0 else min(args.centres_per_step, hi - lo)
    if not weak:                                                # every rank times the same number of centres
        cps = min(cps, min(b - a for a, b in split_bounds(V, world)))
    n_rows = min(hi - lo, (W + K) * cps)
    A = A_all[lo:lo + n_rows]
    # strong scaling through the public call: every rank holds the whole (Morton-ordered) adjacency and takes its own range
    A_full = A_all if (not weak and args.centres_per_step == 0) else None
    del A_all
    is_cap = np.arange(140) < 70
    estimator = Ridge(alpha=1.0, fit_intercept=False)

    def rows_of(step):
        s = (step * cps) % max(n_rows - cps + 1, 1)
        return A[s:s + cps]

    A32 = A.astype(np.int32)                                    # the C ABI's own index type (mmd_searchlight)

    def rows32_of(step):
        s = (step * cps) % max(n_rows - cps + 1, 1)
        return A32[s:s + cps]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-resident arm: X / y staged once, adjacency rows in, score block out -------------------------
    t_up = time.perf_counter()
    eng.upload(X16, Y16, ids[0], ids[1], ids[2], is_cap)
    upload_s = time.perf_counter() - t_up
    upload_stats = eng.stats()
    setup_s = time.perf_counter() - t_setup
    for s in range(W):
        eng.searchlight(rows32_of(s), 1.0)
    eng.reset_stats()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    with ClockSampler(local) as clocks:
        ev0.record(stream)
        t0 = time.perf_counter()
        for s in range(W, W + K):
            scores, flags = eng.searchlight(rows32_of(s), 1.0)
            if world > 1 and not weak:
                gather_scores(scores, cps * world)           # the path's one collective: (V_shard x 8) blocks
        ev1.record(stream)
        barrier()
        wall = time.perf_counter() - t0
    dev_s = max_over_ranks(ev0.elapsed_time(ev1) * 1e-3)
    wall = max_over_ranks(wall)
    stats = eng.stats()
    value = world * cps * K / dev_s

    # ---- end to end through the public drop-in, host buffers every step -----------------------------------
    e2e = None
    if not args.no_e2e:
        Xh = torch.empty(X16.shape, dtype=torch.float16, pin_memory=True).copy_(X16).numpy()
        Yh = torch.empty(Y16.shape, dtype=torch.float16, pin_memory=True).copy_(Y16).numpy()
        del X16, Y16

        def e2e_step(step):
            rows = rows_of(step)
            if weak:                            # n_jobs=1: the whole job on this process's GPU, no collective
                return sl.custom_search_light(Xh, Yh, estimator, rows, ids[0], ids[1], ids[2], n_jobs=1)
            if A_full is not None:              # custom_search_light takes the r-th range of the hemisphere itself
                return sl.custom_search_light(Xh, Yh, estimator, A_full, ids[0], ids[1], ids[2], n_jobs=world)
            full = np.zeros((world * cps, k), dtype=np.int64)
            full[rank * cps:(rank + 1) * cps] = rows
            return sl.custom_search_light(Xh, Yh, estimator, full, ids[0], ids[1], ids[2], n_jobs=world)

        for s in range(min(W, 3)):
            e2e_step(s)
        eng.reset_stats()
        barrier()
        t0 = time.perf_counter()
        step_ms, call_ms = [], []
        for s in range(W, W + K):
            t_step = time.perf_counter()
            out = e2e_step(s)
            step_ms.append(round((time.perf_counter() - t_step) * 1e3, 1))      # includes freeing the previous result
            call_ms.append(round(sl.LAST_TIMING.get("total", 0.0), 1))
        barrier()
        e2e_s = max_over_ranks(time.perf_counter() - t0)
        e2e_stats = eng.stats()
        assert len(out) == (cps if weak else (V if A_full is not None else world * cps)) and len(out[0]) == 8
        h2d = Xh.nbytes + Yh.nbytes + cps * k * 4
        n_e2e = V if A_full is not None else world * cps
        e2e = {"value": round(n_e2e * K / e2e_s, 2), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(cps * (8 * 8 + 4)), "ms_per_step": round(e2e_s / K * 1e3, 2),
               "api": "custom_search_light(X_host_f16, y_host_f16, Ridge(alpha=1), A_rows, ...)",
               "host_ms_last_call_rank0": {k_: round(v_, 1) for k_, v_ in sl.LAST_TIMING.items()},
               "gpu_stage_ms_per_step_rank0": {k_: round(v_ / K, 1) for k_, v_ in e2e_stats["stage_ms"].items() if v_ > 0},
               "step_ms_rank0": step_ms, "call_ms_rank0": call_ms,
               "plan_cache_hits": int(e2e_stats["plan_cache_hits"]),
               "host_plan_ms_per_step": round(e2e_stats["host_plan_ms"] / K, 1),
               "host_adjacency_ms_per_step": round(e2e_stats["host_adjacency_h2d_ms"] / K, 1)}
    else:
        Xh, Yh = X16.cpu().numpy(), Y16.cpu().numpy()
        del X16, Y16

    cpu = parity = None
    if rank == 0 and with_cpu:
        n_cpu = min(args.cpu_centres, len(A))
        if reference_available():
            arm = RefArm(Xh, Yh, ids)
            v, desc, ref_scores = reference_centres_per_sec(arm, A, n_cpu)
            arm.close()
            cpu = {"value": round(v, 3), "unit": UNIT, "cores": arm.cores, "kind": "reference", "sample": desc}
            # parity of the same centres: GPU scores against the reference's
            eng.upload(Xh, Yh, ids[0], ids[1], ids[2], is_cap)
            eng.enable_margins(True)
            got, fl = eng.searchlight(A[:len(ref_scores)], 1.0)
            margins = eng.margins(len(ref_scores))
            eng.enable_margins(False)
            diff = np.abs(got - ref_scores)
            parity = {"n": int(len(ref_scores)), "max_abs_diff": float(diff.max()),
                      "n_centres_differing": int((diff.max(axis=1) > 0).sum()), "min_margin": float(margins.min()),
                      "flags": int(fl.sum()), "against": "unmodified reference (baseline/_ref), same X, same centres"}
        else:
            # only the columns the sample touches travel to the worker processes
            used = np.unique(A[:n_cpu])
            A_cpu = np.searchsorted(used, A[:n_cpu])
            v, n, dt, cores = cpu_centres_per_sec(np.ascontiguousarray(Xh[:, used]), Yh, A_cpu, ids, n_cpu)
            cpu = {"value": round(v, 3), "unit": UNIT, "cores": cores, "kind": "port",
                   "sample": f"{n} consecutive centres of the same workload in {dt:.1f} s (baseline/_ref absent: float64 "
                             f"numpy/scipy oracle of the reference loop, {cores} joblib processes over chunks of centres like "
                             f"searchlight.py:130-148, one BLAS thread each)"}
    del Xh, Yh
    return {"value": value, "dev_s": dev_s, "wall": wall, "stats": stats, "e2e": e2e, "cpu": cpu, "parity": parity,
            "clocks": clocks.summary(), "cps": cps, "V": V, "weak": weak, "n_rows": n_rows, "knn_s": knn_s,
            "upload_s": upload_s, "upload_stats": upload_stats, "setup_s": setup_s}


def run_ours(args):
    import torch
    import torch.distributed as dist
    from multimodal_decoding_b200.analyses.decoding.searchlight import searchlight as sl

    rank, world, local = dist_setup(args)
    torch.cuda.set_device(local)
    eng = sl.get_engine(local)
    if args.batch_centres or args.cluster_size or args.group_size:
        eng.configure(cluster_size=args.cluster_size, batch_centres=args.batch_centres, group_size=args.group_size)
    K, W, k = args.steps, args.warmup, args.n_neighbors
    # N > 1: the headline is the north star's split -- ONE hemisphere, vertex-sharded ("strong"); the replica layout (one
    # whole hemisphere job per GPU) is measured in the same run and reported beside it
    mode = "weak" if world == 1 else args.scaling
    # the CPU leg (the unmodified reference on the host cores + the live parity sample) runs at N = 1 only
    m = measure(args, mode, rank, world, local, eng, with_cpu=not args.no_cpu_baseline and world == 1)
    replicas = None
    if world > 1 and mode == "strong" and not args.no_replicas:
        torch.cuda.empty_cache()
        r = measure(args, "weak", rank, world, local, eng, with_cpu=False)
        replicas = {"value": round(r["value"], 2), "unit": UNIT, "scaling": "weak",
                    "ms_per_step": round(r["dev_s"] / K * 1e3, 3), "e2e": r["e2e"],
                    "layout": f"{world} whole-hemisphere jobs, one per GPU, no data-path collective"}
    if rank == 0:
        peaks = load_peaks()
        stats, cps, V, weak = m["stats"], m["cps"], m["V"], m["weak"]
        line = {
            "metric": METRIC, "value": round(m["value"], 2), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": round(m["dev_s"] / K * 1e3, 3), "higher_is_better": True,
            "scaling": "weak" if (weak and world > 1) or args.scaling == "weak" else "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args, V, world, weak, cps),
            "wall_ms_per_step": round(m["wall"] / K * 1e3, 3),
            "e2e": m["e2e"], "gpu_launches": int(stats["launches"]),
            "roofline": roofline_entry(stats, cps * K, args, peaks),
            "cpu_baseline": m["cpu"], "parity_sample": m["parity"], "replicas": replicas, "clocks": m["clocks"],
            "setup": {"knn_rows": int(m["n_rows"]), "knn_s": round(m["knn_s"], 2), "upload_s": round(m["upload_s"], 3),
                      "upload_stage_ms": {s: round(v, 2) for s, v in m["upload_stats"]["stage_ms"].items() if v > 0},
                      "total_s": round(m["setup_s"], 1)},
            "clusters_per_step": stats["clusters"] / K,
            "rerouted_centres_per_step": stats["rerouted_centres"] / K,
            "host_ms_per_step": {"plan": round(stats["host_plan_ms"] / K, 1),
                                 "adjacency_h2d": round(stats["host_adjacency_h2d_ms"] / K, 1),
                                 "plan_cache_hits": int(stats["plan_cache_hits"])},
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_reference(args):
    """The reference's own CPU implementation of the path on this box's host cores: the unmodified `custom_search_light`
    (baseline/_ref) on a bounded sample of the own arm's workload -- same generator, same seed, full-width X; falls back
    to the oracle port only when baseline/_ref is absent (and says so)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    k, n_tr, D = args.n_neighbors, args.n_train, args.n_dims
    cores = host_cores()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    n_c = 8 * cores                                             # centres per step (bounded sample)
    n_rows = (args.warmup + args.steps) * n_c
    device = torch.device("cuda", 0) if torch.cuda.is_available() else torch.device("cpu")
    coords, X16, Y16, ids = make_workload(args, device, seed=0)  # the own arm's generator and seed
    Xh, Yh = X16.cpu().numpy(), Y16.cpu().numpy()
    del X16, Y16
    if device.type == "cuda":
        torch.cuda.empty_cache()
    from sklearn.neighbors import NearestNeighbors              # the k-NN branch of get_adjacency_matrix (searchlight.py:173)
    n_rows = min(n_rows, len(coords))
    A = NearestNeighbors().fit(coords).kneighbors(coords[:n_rows], n_neighbors=k, return_distance=False)
    use_ref = reference_available()
    if use_ref:
        arm = RefArm(Xh, Yh, ids)
        step = lambda s: arm.run(A[s * n_c:(s + 1) * n_c], cores)
        kind = "reference"
        how = (f"unmodified custom_search_light + sklearn Ridge(alpha=1) from baseline/_ref, n_jobs={cores} "
               f"(searchlight.py:130-148), full-width X ({arm.versions}, BLAS {arm.blas})")
    else:
        used = np.unique(A)
        arm = CpuArm(np.ascontiguousarray(Xh[:, used]), Yh, np.searchsorted(used, A), ids)
        step = lambda s: arm.step(range(s * n_c, (s + 1) * n_c))
        kind = "port"
        how = (f"baseline/_ref absent: float64 numpy/scipy oracle of the reference loop, {cores} joblib processes over chunks "
               f"of centres, one BLAS thread each")
    n_steps_avail = max(n_rows // n_c, 1)
    for s in range(max(args.warmup, 1)):
        step(s % n_steps_avail)
    t0 = time.perf_counter()
    for s in range(args.warmup, args.warmup + args.steps):
        step(s % n_steps_avail)
    dt = time.perf_counter() - t0
    arm.close()
    v = n_c * args.steps / dt
    # the own arm's layout for this N (run_ours / measure): strong scaling when N > 1
    from multimodal_decoding_b200.sharding import split_bounds
    weak_own = world == 1 or args.scaling == "weak"
    V_all = len(coords)
    cps_own = V_all if weak_own else min(b - a for a, b in split_bounds(V_all, world))
    if args.centres_per_step:
        cps_own = min(args.centres_per_step, cps_own)
    sample = f"{n_c} consecutive centres per step x {args.steps} steps, k={k}, N_train={n_tr}, D={D}; {how}"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": round(v, 3), "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 2),
        "higher_is_better": True, "scaling": "weak" if (weak_own and world > 1) or args.scaling == "weak" else "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        # the own arm's workload and layout; each step here is a bounded sample of its centres (cpu_baseline.sample)
        "config": workload_config(args, len(coords), world, weak_own, cps_own),
        "cpu_baseline": {"value": round(v, 3), "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": round(v, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "the reference's CPU path (Python + sklearn/scipy, float64) on all host cores; inputs generated like the own "
                "arm's (same seed)"}), flush=True)


if __name__ == "__main__":
    a = parse_args()
    run_reference(a) if a.impl == "reference" else run_ours(a)
