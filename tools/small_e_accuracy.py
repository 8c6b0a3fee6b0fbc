"""Accuracy of the two arithmetic modes against a long-double Cholesky as the uncertainties e shrink (the a-priori
column bound |W_ik| <= 1 / e_k of the emulated mode gets looser as e -> 0)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np
from oracle import gp_oracle as orc
from gaussian_process_b200 import engine, pso


def longdouble_selfpred(x, y, e, ls):
    """mu - y and sd at the training points in 80-bit arithmetic (Cholesky by hand)."""
    n = x.shape[1]
    K = orc.rbf_kernel(x, x, ls).astype(np.longdouble)
    d = (e.astype(np.longdouble)) ** 2
    A = K + np.diag(d)
    L = np.zeros_like(A)
    for j in range(n):
        s = A[j, j] - np.dot(L[j, :j], L[j, :j])
        L[j, j] = np.sqrt(s)
        L[j + 1:, j] = (A[j + 1:, j] - L[j + 1:, :j] @ L[j, :j]) / L[j, j]
    W = np.zeros_like(A)
    for i in range(n):
        b = np.zeros(n, dtype=np.longdouble); b[i] = 1
        for r in range(i, n):
            W[r, i] = (b[r] - np.dot(L[r, i:r], W[i:r, i])) / L[r, r]
    alpha = W.T @ (W @ y.astype(np.longdouble))
    kinv_diag = (W * W).sum(0)
    resid = -d * alpha
    var = d - d * d * kinv_diag
    return resid.astype(np.float64), np.sqrt(np.maximum(var, 1e-12)).astype(np.float64)


s, ex = pso.sigma_grid()
n, dd = (int(sys.argv[1]) if len(sys.argv) > 1 else 384), 2
MODES = ("fp64", "emu", "emu28")
for emin, emax in ((0.05, 0.2), (1e-2, 1e-1), (1e-3, 1e-1), (1e-4, 1e-2), (1e-5, 1e-3)):
    x, y, _ = orc.synthetic_problem(n, dd, 3)
    rng = np.random.default_rng(1)
    e = np.exp(rng.uniform(np.log(emin), np.log(emax), n))
    ls = np.array([0.3, 0.4])
    r_ref, sd_ref = longdouble_selfpred(x, y, e, ls)
    row = []
    for mode in MODES:
        engine.set_precision(mode)
        c, info, r, sd = engine.LossContext(x, y, e, s).counts(ls[None, :], want_detail=True)
        row.append((np.max(np.abs(r[0] - r_ref) / np.abs(r_ref).max()), np.max(np.abs(sd[0] - sd_ref) / sd_ref), int(info[0]), c[0].copy()))
    print(f"N={n} e in [{emin:g}, {emax:g}]: " + " | ".join(f"{m} resid {v[0]:.2e} sd {v[1]:.2e}" for m, v in zip(MODES, row))
          + f" | counts equal to fp64: {[bool(np.array_equal(v[3], row[0][3])) for v in row]}", flush=True)
