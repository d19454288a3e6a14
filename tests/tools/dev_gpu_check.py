"""Development check on a GPU box: every stage against the oracle / golden fixtures, with diagnostics."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from multimodal_decoding_b200.engine import Engine  # noqa: E402
from oracle import searchlight_oracle as orc  # noqa: E402

G = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "golden")
eng = Engine(0)

g = np.load(os.path.join(G, "knn_small.npz"))
for c, a, d in (("coords", "adjacency", "distances"), ("coords2", "adjacency2", "distances2")):
    dist, idx = eng.knn(g[c], g[a].shape[1])
    print("knn", c, "idx equal:", np.array_equal(idx, g[a]), "dist equal:", np.array_equal(dist, g[d]),
          "mismatch rows:", int((idx != g[a]).any(1).sum()))

for name in ("searchlight_small.npz", "searchlight_smooth.npz", "searchlight_captions.npz"):
    g = np.load(os.path.join(G, name))
    ntr, nte, nim = int(g["n_train"]), int(g["n_test"]), int(g["n_imagery"])
    tr, te, im = np.arange(ntr), np.arange(ntr, ntr + nte), np.arange(ntr + nte, ntr + nte + nim)
    is_cap = orc.default_is_caption(nte)
    t0 = time.time()
    eng.upload(g["X"], g["latents"], tr, te, im, is_cap)
    A = g["adjacency"]
    # predictions of a few centres
    for c in (0, 7, len(A) - 1):
        yh = eng.predict_centre(A[c], float(g["alpha"]))
        _, pt, pi = orc.train_and_test(g["X"][:, A[c]], g["latents"], tr, te, im, float(g["alpha"]), is_cap,
                                       return_preds=True)
        ref = np.concatenate([pt[is_cap], pt[~is_cap], pi])
        print(name, "centre", c, "pred max rel err", np.abs(yh - ref).max() / np.abs(ref).max())
    scores, flags = eng.searchlight(A, float(g["alpha"]))
    dt = time.time() - t0
    diff = np.abs(scores - g["scores"])
    print(name, "scores equal:", np.array_equal(scores, g["scores"]), "max diff", diff.max(), "rows differing",
          int((diff > 0).any(1).sum()), "of", len(A), "flags", int(flags.sum()), f"{dt:.2f}s")
    if diff.max() > 0:
        r = int(np.argmax(diff.max(1)))
        print("  worst row", r, "\n  got ", scores[r], "\n  want", g["scores"][r])
    print("  stats", eng.stats())
