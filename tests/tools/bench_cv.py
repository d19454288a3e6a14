"""Throughput of the cross-validated searchlight (mmd_searchlight_cv: DEFAULT_ALPHAS x 5 folds per centre, then the refit of
the selected alpha) next to the fixed-alpha searchlight, at config-2 sizes (k = 750, N_train = 4000, D = 1024).  Prints one
JSON line.  The reference's counterpart is a per-centre GridSearchCV: 30 sklearn fits + 30 scorings of 800 x 800 distances."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from multimodal_decoding_b200.analyses.decoding import ridge_regression_decoding as rr  # noqa: E402
from multimodal_decoding_b200.analyses.decoding.searchlight import searchlight as sl  # noqa: E402

n_centres = int(os.environ.get("N_CENTRES", "2048"))
sys.argv = [sys.argv[0]]
args = bench.parse_args()
dev = torch.device("cuda", 0)
eng = sl.get_engine(0)
coords, X16, Y16, ids = bench.make_workload(args, dev)
A = bench.knn_rows(eng, coords, args.n_neighbors, 0, n_centres)
eng.upload(X16, Y16, ids[0], ids[1], ids[2], np.arange(140) < 70)
alphas = [float(a) for a in rr.DEFAULT_ALPHAS]
out = {}
eng.searchlight(A, 1.0)
eng.reset_stats()
t0 = time.perf_counter()
eng.searchlight(A, 1.0)
dt = time.perf_counter() - t0
out["fixed_alpha"] = {"centres_per_s": round(n_centres / dt, 1), "s": round(dt, 3)}
sl.search_light_cv(eng, A[:256], alphas, 5, len(ids[0]))                   # warm-up (workspaces)
eng.reset_stats()
t0 = time.perf_counter()
scores, flags, best_alpha, mean = sl.search_light_cv(eng, A, alphas, 5, len(ids[0]))
dt = time.perf_counter() - t0
st = eng.stats()["stage_ms"]
out["cross_validated"] = {"centres_per_s": round(n_centres / dt, 1), "s": round(dt, 3),
                          "stage_ms": {k: round(v, 1) for k, v in st.items() if v > 0},
                          "selected_alpha_counts": {str(a): int((best_alpha == a).sum()) for a in alphas},
                          "flags": int(flags.sum())}
print(json.dumps({"metric": "searchlight centres/s with a per-centre alpha grid x 5-fold CV (GridSearchCV semantics) + refit",
                  "n_centres": n_centres, "alphas": alphas, "n_folds": 5, "fits_per_centre": 31, **out}))
