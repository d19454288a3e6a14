"""Throughput of the searchlight with the null-distribution branch on (searchlight.py:61-79, `--create-null-distr`,
100 permutations per centre) next to the plain searchlight, at config-2 sizes.  Prints one JSON line."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from multimodal_decoding_b200 import data as d  # noqa: E402
from multimodal_decoding_b200.analyses.decoding.searchlight import searchlight as sl  # noqa: E402
from multimodal_decoding_b200.analyses.decoding.searchlight.null_distribution import null_scores  # noqa: E402

sys.argv = [sys.argv[0]]
args = bench.parse_args()
n_centres, n_perm = 32768, 100
dev = torch.device("cuda", 0)
eng = sl.get_engine(0)
coords, X16, Y16, ids = bench.make_workload(args, dev)
A = bench.knn_rows(eng, coords, args.n_neighbors, 0, n_centres)
eng.upload(X16, Y16, ids[0], ids[1], ids[2], np.arange(140) < 70)
perms = d.create_null_distr_shuffled_indices(n_perm)
np.random.seed(0)
out = {}
for name, fn in (("plain", lambda: eng.searchlight(A, 1.0)),
                 ("with_null", lambda: null_scores(eng, A, 1.0, perms))):
    fn()
    eng.reset_stats()
    t0 = time.perf_counter()
    res = fn()
    dt = time.perf_counter() - t0
    st = eng.stats()["stage_ms"]
    out[name] = {"centres_per_s": round(n_centres / dt, 1), "s": round(dt, 3),
                 "stage_ms": {k: round(v, 1) for k, v in st.items() if v > 0}}
null = res[2]
print(json.dumps({"metric": "searchlight centres/s with the null-distribution branch (100 permutations per centre)",
                  "n_centres": n_centres, "n_permutations": n_perm, "null_block_shape": list(null.shape), **out,
                  "cpu_reference_note": "the reference re-scores on the CPU: ~3 s per centre for 100 permutations "
                                        "(SURVEY.md §8 a11)"}))
