"""Design probe (CPU) for the next round's tcgen05 route: can the prediction GEMM  yhat = Zt^T (144 x k) . XtY[nbr] (k x D)
run as exact INT8 slice GEMMs (tcgen05 kind::i8, int32 accumulation) and still reproduce every strict-`<` comparison
of the float64 path?

Both operands are float64.  Each row of Zt^T and each column of XtY[nbr] is scaled by a power of two and cut into
`s` signed digits (6 bits, then 7 bits per further digit; every digit fits int8), so that a slice pair's GEMM is an
exact integer GEMM (|sum| <= k * 64 * 64 < 2^31).  The triangular scheme keeps the slice pairs with t + u < s
(s (s + 1) / 2 GEMMs); the full scheme keeps all s^2.  The probe prints, per scheme, the relative error of yhat and
how many of the 8 scores differ from the float64 path on ill-conditioned synthetic betas, next to the cost of the
slice GEMMs at the int8 dense peak against the fp64 GEMM at the measured DMMA peak.

Run: python tests/tools/int8_slice_probe.py [level] [k] [n_centres]      (log: profiles/r1_int8_slice_probe.log)
"""
import sys
import time

import numpy as np
import scipy.linalg

sys.path.insert(0, ".")
from multimodal_decoding_b200 import synthetic as syn  # noqa: E402
from oracle import searchlight_oracle as orc  # noqa: E402

INT8_PEAK = 4.5e15       # dense int8 op/s (nominal B200)
DMMA_PEAK = 37.1e12      # measured fp64 tensor pipe


def digits(a, axis, s):
    """Power-of-two scale along `axis` and `s` signed digits with a = scale * sum_t q_t 2^-(6 + 7 t) + residual."""
    amax = np.abs(a).max(axis=axis, keepdims=True)
    scale = np.exp2(np.ceil(np.log2(np.where(amax > 0, amax, 1.0))))
    r = a / scale
    out = []
    for t in range(s):
        r = r * (64.0 if t == 0 else 128.0)
        q = np.rint(r)
        assert np.abs(q).max() <= 64
        out.append(q)
        r = r - q
    return scale, out


def sliced_gemm(a, b, s, triangular):
    sa, qa = digits(a, 1, s)          # rows of a
    sb, qb = digits(b, 0, s)          # columns of b
    acc = np.zeros((a.shape[0], b.shape[1]))
    n = 0
    # smallest terms first
    for w in sorted({t + u for t in range(s) for u in range(s)}, reverse=True):
        if triangular and w >= s:
            continue
        part = np.zeros_like(acc)
        for t in range(s):
            u = w - t
            if 0 <= u < s:
                part += qa[t] @ qb[u]              # exact: integer-valued float64 below 2^53
                n += 1
        acc += part * 2.0 ** (-(12 + 7 * w))
    return acc * sa * sb, n


def main():
    level = int(sys.argv[1]) if len(sys.argv) > 1 else 5
    k = int(sys.argv[2]) if len(sys.argv) > 2 else 750
    n_centres = int(sys.argv[3]) if len(sys.argv) > 3 else 16
    t0 = time.time()
    mesh = syn.make_mesh(level=level)
    trials = syn.make_trials(mesh, n_train=4000, n_dims=1024, smooth_steps=4)
    inp = syn.make_searchlight_inputs(mesh, trials)
    print(f"data {time.time() - t0:.1f}s  X {inp.X.shape}  k={k}")
    rng = np.random.default_rng(1)
    centres = rng.choice(len(inp.coords), n_centres, replace=False)
    c64 = inp.coords.astype(np.float64)
    ntr, nte = len(inp.train_ids), len(inp.test_ids)
    is_cap = orc.default_is_caption(nte)
    y64 = inp.latents.astype(np.float64)
    schemes = [(s, tri) for s in (4, 5, 6, 7, 8) for tri in (True, False) if not (s >= 7 and not tri)]
    res = {sc: [] for sc in schemes}
    margins = []
    for c in centres:
        nb = np.argsort(((c64 - c64[c]) ** 2).sum(1), kind="stable")[:k]
        S = inp.X[:, nb].astype(np.float64)
        g = S[:ntr].T @ S[:ntr]
        g[np.diag_indices_from(g)] += 1.0
        b = S[:ntr].T @ y64[:ntr]                                    # XtY[nbr]  (k x D)
        z = scipy.linalg.cho_solve(scipy.linalg.cho_factor(g, lower=True), S[ntr:].T)   # Zt (k x 143)
        pred = z.T @ b
        ref = orc.score_centre(y64[ntr:ntr + nte], pred[:nte], y64[ntr + nte:], pred[nte:], is_cap)
        ref = np.array([ref[key] for key in orc.SCORE_KEYS])
        margins.append(orc.score_margins(y64[ntr:ntr + nte], pred[:nte], y64[ntr + nte:], pred[nte:], is_cap).min())
        for sc in schemes:
            got, n = sliced_gemm(z.T, b, *sc)
            s8 = orc.score_centre(y64[ntr:ntr + nte], got[:nte], y64[ntr + nte:], got[nte:], is_cap)
            s8 = np.array([s8[key] for key in orc.SCORE_KEYS])
            res[sc].append((np.abs(got - pred).max() / np.abs(pred).max(), int((s8 != ref).sum()), n))
    print(f"smallest comparison margin of the float64 path over the centres: {min(margins):.2e} "
          f"(median {np.median(margins):.2e})")
    flops = 2.0 * 144 * k * 1024
    print(f"fp64 DMMA GEMM at peak: {flops / DMMA_PEAK * 1e6:.2f} us per centre")
    for (s, tri), rows in res.items():
        rel = np.array([r[0] for r in rows])
        flips = np.array([r[1] for r in rows])
        n = rows[0][2]
        print(f"s={s} {'triangular' if tri else 'full      '}: {n:2d} int8 GEMMs = {n * flops / INT8_PEAK * 1e6:.2f} us at peak; "
              f"yhat rel err max {rel.max():.2e} median {np.median(rel):.2e}; centres with a changed score "
              f"{(flips > 0).sum()}/{len(rows)}")


if __name__ == "__main__":
    main()
