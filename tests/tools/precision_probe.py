"""Design probe (CPU): which stages of the per-centre ridge path tolerate fp32?

Compares the float64 oracle with fp32-Gram / fp32-Cholesky variants on synthetic smooth betas
and prints, per variant, how many of the 8 per-centre scores differ and by how much.
Result summary lives in DESIGN.md §"Precision".  Run: python tests/tools/precision_probe.py [level] [k] [n_centres]
"""
import sys
import time

import numpy as np
import scipy.linalg

sys.path.insert(0, ".")
from multimodal_decoding_b200 import synthetic as syn  # noqa: E402
from oracle import searchlight_oracle as orc  # noqa: E402


def variant_scores(S16, Y16, ntr, nte, alpha, gram_dt, chol_dt, pred_dt, is_cap):
    s_tr = S16[:ntr]
    if gram_dt == np.float32:
        g = (s_tr.astype(np.float32).T @ s_tr.astype(np.float32)).astype(np.float64)
        b = (s_tr.astype(np.float32).T @ Y16[:ntr].astype(np.float32)).astype(np.float64)
    else:
        g = s_tr.astype(np.float64).T @ s_tr.astype(np.float64)
        b = s_tr.astype(np.float64).T @ Y16[:ntr].astype(np.float64)
    a = g.astype(chol_dt)
    a[np.diag_indices_from(a)] += chol_dt(alpha)
    e = S16[ntr:].astype(chol_dt).T                       # k x 143
    c, low = scipy.linalg.cho_factor(a, lower=True)
    z = scipy.linalg.cho_solve((c, low), e)               # k x 143
    pred = (z.astype(pred_dt).T @ b.astype(pred_dt)).astype(np.float64)
    y64 = Y16.astype(np.float64)
    sc = orc.score_centre(y64[ntr:ntr + nte], pred[:nte], y64[ntr + nte:], pred[nte:], is_cap)
    return np.array([sc[k] for k in orc.SCORE_KEYS]), pred


def main():
    level = int(sys.argv[1]) if len(sys.argv) > 1 else 6
    k = int(sys.argv[2]) if len(sys.argv) > 2 else 750
    n_centres = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    smooth = int(sys.argv[4]) if len(sys.argv) > 4 else 4
    t0 = time.time()
    mesh = syn.make_mesh(level=level)
    trials = syn.make_trials(mesh, n_train=4000, n_dims=1024, smooth_steps=smooth)
    inp = syn.make_searchlight_inputs(mesh, trials)
    print(f"data {time.time() - t0:.1f}s  X {inp.X.shape}")
    rng = np.random.default_rng(1)
    centres = rng.choice(len(inp.coords), n_centres, replace=False)
    c64 = inp.coords.astype(np.float64)
    ntr, nte = len(inp.train_ids), len(inp.test_ids)
    is_cap = orc.default_is_caption(nte)
    variants = {
        "g32_c32_p32": (np.float32, np.float32, np.float32),
        "g32_c64_p64": (np.float32, np.float64, np.float64),
        "g64_c32_p64": (np.float64, np.float32, np.float64),
        "g64_c64_p32": (np.float64, np.float64, np.float32),
    }
    stats = {v: [] for v in variants}
    conds = []
    for c in centres:
        d2 = ((c64 - c64[c]) ** 2).sum(1)
        nb = np.argsort(d2, kind="stable")[:k]
        S = inp.X[:, nb]
        ref, pref = variant_scores(S, inp.latents, ntr, nte, 1.0, np.float64, np.float64, np.float64, is_cap)
        ev = np.linalg.eigvalsh(S[:ntr].astype(np.float64).T @ S[:ntr].astype(np.float64))
        conds.append((ev[-1], ev[0], np.sum(ev > 1.0)))
        for name, (g, ch, p) in variants.items():
            sc, pr = variant_scores(S, inp.latents, ntr, nte, 1.0, g, ch, p, is_cap)
            rel = np.abs(pr - pref).max() / np.abs(pref).max()
            stats[name].append((np.abs(sc - ref), rel))
    conds = np.array(conds)
    print(f"eig max {conds[:, 0].mean():.3g}  eig min {conds[:, 1].mean():.3g}  #eig>1: {conds[:, 2].mean():.0f} of {k}")
    print("ref scores (last centre):", np.round(ref, 4))
    for name, lst in stats.items():
        diffs = np.array([d for d, _ in lst])
        rels = np.array([r for _, r in lst])
        print(f"{name}: max|dscore| per metric {np.round(diffs.max(0), 5)}  centres with any diff "
              f"{(diffs.max(1) > 0).sum()}/{len(lst)}  pred rel err max {rels.max():.2e} median {np.median(rels):.2e}")


if __name__ == "__main__":
    main()
