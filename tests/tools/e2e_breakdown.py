"""Where the end-to-end time of custom_search_light goes on a whole hemisphere (development aid)."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from multimodal_decoding_b200.analyses.decoding.searchlight import searchlight as sl  # noqa: E402

args = bench.parse_args()
dev = torch.device("cuda", 0)
eng = sl.get_engine(0)
coords, X16, Y16, ids = bench.make_workload(args, dev)
A = bench.knn_rows(eng, coords, args.n_neighbors, 0, len(coords))
Xh = torch.empty(X16.shape, dtype=torch.float16, pin_memory=True).copy_(X16).numpy()
Yh = torch.empty(Y16.shape, dtype=torch.float16, pin_memory=True).copy_(Y16).numpy()
is_cap = np.arange(140) < 70
for it in range(3):
    t0 = time.perf_counter()
    eng.upload(Xh, Yh, ids[0], ids[1], ids[2], is_cap)
    t1 = time.perf_counter()
    scores, flags = eng.searchlight(A, 1.0)
    t2 = time.perf_counter()
    out = sl._scores_to_dicts(scores)
    t3 = time.perf_counter()
    print(f"iter {it}: upload {1e3*(t1-t0):.1f} ms, searchlight {1e3*(t2-t1):.1f} ms, dicts {1e3*(t3-t2):.1f} ms, "
          f"stats {eng.stats()['stage_ms']}", flush=True)
    eng.reset_stats()
