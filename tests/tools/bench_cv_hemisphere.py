"""The north star's target run: a full-hemisphere searchlight (163,842 vertices, k = 750, N_train = 4000, D = 1024) with the whole
alpha grid and 5-fold cross-validation per vertex, vertex-sharded over the GPUs of one box through the public drop-in

    custom_search_light(X, y, GridSearchCV(Ridge(fit_intercept=False), dict(alpha=DEFAULT_ALPHAS), cv=5), A, ...)

Launch: python -m torch.distributed.run --nproc-per-node N tests/tools/bench_cv_hemisphere.py  (N_CENTRES=... for a prefix of the
Morton-ordered hemisphere).  Rank 0 prints one JSON line; it also re-scores a sample of vertices with the fixed-alpha searchlight
at the alpha the cross-validation selected and checks the refit scores are identical."""
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from multimodal_decoding_b200.analyses.decoding import ridge_regression_decoding as rr  # noqa: E402
from multimodal_decoding_b200.analyses.decoding.searchlight import searchlight as sl  # noqa: E402
from multimodal_decoding_b200.sharding import spatial_order  # noqa: E402
from sklearn.linear_model import Ridge  # noqa: E402
from sklearn.model_selection import GridSearchCV  # noqa: E402

sys.argv = [sys.argv[0]]
args = bench.parse_args()
rank, world, local = bench.dist_setup(args)
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
eng = sl.get_engine(local)
coords, X16, Y16, ids = bench.make_workload(args, dev, seed=0)
V = len(coords)
A = bench.knn_rows(eng, coords, args.n_neighbors, 0, V)[spatial_order(coords)]
n = int(os.environ.get("N_CENTRES", "0")) or V
A = A[:n]
Xh, Yh = X16.cpu().numpy(), Y16.cpu().numpy()
del X16, Y16
torch.cuda.empty_cache()
alphas = [float(a) for a in rr.DEFAULT_ALPHAS]
model = GridSearchCV(estimator=Ridge(fit_intercept=False), param_grid=dict(alpha=alphas), cv=rr.NUM_CV_SPLITS)
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
t0 = time.perf_counter()
scores = sl.custom_search_light(Xh, Yh, model, A, ids[0], ids[1], ids[2], n_jobs=world)
if world > 1:
    dist.barrier()
dt = time.perf_counter() - t0
best = sl.LAST_CV["best_alpha"]                       # this rank's shard
lo, hi = (0, n) if world == 1 else __import__("multimodal_decoding_b200.sharding", fromlist=["x"]).shard_range(n)
check = None
if rank == 0:
    # refit check on a sample of this rank's vertices: fixed-alpha searchlight at the selected alpha gives the same 8 scores
    pick = np.arange(lo, hi, max(1, (hi - lo) // 64))[:64]
    ok = True
    for a in np.unique(best[pick - lo]):
        sel = pick[best[pick - lo] == a]
        eng.upload(Xh, Yh, ids[0], ids[1], ids[2], np.arange(140) < 70)
        s, f = eng.searchlight(A[sel], float(a))
        keys = list(scores[sel[0]].keys())
        ok &= bool(np.array_equal(s, np.array([[scores[i][k] for k in keys] for i in sel])))
    check = {"vertices": int(len(pick)), "refit_scores_identical": ok}
    print(json.dumps({
        "metric": "full-hemisphere cross-validated searchlight (6 alphas x 5 folds per vertex + refit), vertex-sharded",
        "n_gpus": world, "n_centres": int(n), "n_neighbors": args.n_neighbors, "n_train": args.n_train, "n_dims": args.n_dims,
        "alphas": alphas, "seconds": round(dt, 1), "centres_per_s": round(n / dt, 1),
        "selected_alpha_counts_rank0": {str(a): int((best == a).sum()) for a in alphas}, "refit_check": check,
        "api": "custom_search_light(X_host_f16, y_host_f16, GridSearchCV(Ridge(fit_intercept=False), dict(alpha=DEFAULT_ALPHAS), cv=5), A, ...)"}),
        flush=True)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
