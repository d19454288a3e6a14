"""Throughput of the group-level statistics kernels at fsaverage7 size: TFCE maps/s on the GPU
next to the numpy/scipy oracle on the host, and the t-map kernel.  Prints one JSON line.

    python tests/tools/bench_group_stats.py [--maps 296] [--cpu-maps 1]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from multimodal_decoding_b200 import synthetic as syn  # noqa: E402
from multimodal_decoding_b200.engine import Engine  # noqa: E402
from oracle import stats_oracle as so  # noqa: E402  (checker / CPU baseline only)

p = argparse.ArgumentParser()
p.add_argument("--level", type=int, default=7)
p.add_argument("--maps", type=int, default=296)
p.add_argument("--cpu-maps", type=int, default=1)
p.add_argument("--subjects", type=int, default=6)
a = p.parse_args()

mesh = syn.make_mesh(level=a.level, seed=7)
edges = np.unique(np.sort(syn.mesh_edges(mesh.faces), axis=1), axis=0).astype(np.int32)
V = len(mesh.coords)
rng = np.random.default_rng(0)
# null-like t-maps: smooth noise (a few diffusion steps over the mesh) scaled to t-values with maxima around 4-6
import scipy.sparse as sp
adj = sp.coo_matrix((np.ones(len(edges)), (edges[:, 0], edges[:, 1])), shape=(V, V))
adj = (adj + adj.T).tocsr()
deg = np.asarray(adj.sum(1)).ravel()
op = sp.diags(0.5 / deg) @ adj + 0.5 * sp.identity(V)
maps = rng.standard_normal((a.maps, V))
for _ in range(6):
    maps = (op @ maps.T).T
maps = maps / maps.std() * 1.3

eng = Engine(0)
eng.tfce(maps[:2], edges)                                     # warm-up
t0 = time.perf_counter()
got = eng.tfce(maps, edges)
gpu_s = time.perf_counter() - t0
t0 = time.perf_counter()
want = np.stack([so.tfce(maps[i], edges) for i in range(a.cpu_maps)])
cpu_s = time.perf_counter() - t0
exact = bool(np.array_equal(got[:a.cpu_maps], want))

data = 0.5 + 0.03 * rng.standard_normal((a.subjects, V))
eng.ttest_1samp(data, 0.5)
t0 = time.perf_counter()
for _ in range(10):
    t = eng.ttest_1samp(data, 0.5)
tt_s = (time.perf_counter() - t0) / 10
t0 = time.perf_counter()
t_ref = so.calc_image_t_values(data, 0.5)
tt_cpu_s = time.perf_counter() - t0
print(json.dumps({
    "metric": "TFCE maps/s (fsaverage7-sized hemisphere, h=2 e=1 dh=0.1)", "n_vertices": V, "n_edges": int(len(edges)),
    "maps": a.maps, "thresholds_per_map_mean": float(np.mean(np.ceil(maps.max(1) / 0.1))),
    "gpu_maps_per_s": round(a.maps / gpu_s, 2), "gpu_s": round(gpu_s, 3),
    "cpu_oracle_maps_per_s": round(a.cpu_maps / cpu_s, 4), "cpu_s_per_map": round(cpu_s / a.cpu_maps, 2),
    "bit_exact_vs_oracle": exact,
    "ttest": {"subjects": a.subjects, "gpu_ms": round(tt_s * 1e3, 3), "cpu_oracle_ms": round(tt_cpu_s * 1e3, 1),
              "bit_exact": bool(np.array_equal(t, t_ref, equal_nan=True))},
    "note": "host buffers in and out (H2D/D2H inside the timing); the CPU column is the vectorised numpy/scipy oracle, "
            "not the reference's pure-Python cluster loop (cluster_analysis.py:15-75), which is far slower"}))
