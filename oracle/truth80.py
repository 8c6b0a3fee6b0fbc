"""80-bit (x87 long double, 64-bit mantissa) evaluation of the path's outputs -- TEST INFRASTRUCTURE.

Used where the FP64 results of the CUDA path and of the reference (oracle/gp_oracle.py, NumPy/LAPACK in FP64) differ by
more than the 1e-9 parity bar: on ill-conditioned problems both carry a forward error ~ cond(K) * 2^-53, and only a
more accurate evaluation can say which side is off.  The factorisation is done by hand (LAPACK has no long double); the
kernel matrix itself is formed from the reference's FP64 expression (GP_func.py:49-65: its rounding is part of the
reference's definition of K), everything after it runs with eps = 1.08e-19.  Cost ~ n^3 long-double operations in NumPy's
generic loops: seconds at n = 1000, so this is for selected cases only."""
import numpy as np

from . import gp_oracle as orc

LD = np.longdouble


def cholesky_ld(A):
    """Lower Cholesky factor of a symmetric positive definite long-double matrix (left-looking, one column per step)."""
    n = A.shape[0]
    L = np.zeros_like(A)
    for j in range(n):
        s = A[j, j] - np.dot(L[j, :j], L[j, :j])
        if not s > 0:
            raise np.linalg.LinAlgError("Matrix is not positive definite")
        L[j, j] = np.sqrt(s)
        if j + 1 < n:
            L[j + 1:, j] = (A[j + 1:, j] - L[j + 1:, :j] @ L[j, :j]) / L[j, j]
    return L


def forward_solve_ld(L, B):
    """L^-1 B by forward substitution, all right-hand sides at once (B is (n, m))."""
    n = L.shape[0]
    V = np.array(B, dtype=LD, copy=True)
    for r in range(n):
        if r:
            V[r] -= L[r, :r] @ V[:r]
        V[r] /= L[r, r]
    return V


def fit_ld(x, y, e, ls):
    """(L, alpha, K + diag(e^2)) in long double; K from the reference's FP64 kernel expression."""
    A = orc.rbf_kernel(x, x, ls).astype(LD)
    A[np.diag_indices_from(A)] += e.astype(LD) ** 2
    L = cholesky_ld(A)
    z = forward_solve_ld(L, y.astype(LD)[:, None])[:, 0]
    # back substitution L^T alpha = z
    n = L.shape[0]
    alpha = z.copy()
    for r in range(n - 1, -1, -1):
        if r + 1 < n:
            alpha[r] -= L[r + 1:, r] @ alpha[r + 1:]
        alpha[r] /= L[r, r]
    return L, alpha, A


def selfpredict_ld(x, y, e, ls):
    """(mu - y, sd) at the training points (find_len_scales.py:159 via GP_func.py:32-40), long double, returned as FP64.
    Uses the reference's own formulae: mu = K_s^T alpha, v = L^-1 K_s, var = clip(1 - sum v^2, 1e-12)."""
    L, alpha, A = fit_ld(x, y, e, ls)
    Ks = A.copy()
    Ks[np.diag_indices_from(Ks)] -= e.astype(LD) ** 2
    mu = Ks.T @ alpha
    V = forward_solve_ld(L, Ks)
    var = LD(1) - (V * V).sum(axis=0)
    return (mu - y.astype(LD)).astype(np.float64), np.sqrt(np.maximum(var, LD(1e-12))).astype(np.float64)


def predict_ld(x, y, e, grid, ls):
    """GP() at grid points (d, M): (mu, sd) in long double, returned as FP64.  GP_func.py:21-45."""
    L, alpha, _ = fit_ld(x, y, e, ls)
    Ks = orc.rbf_kernel(x, grid, ls).astype(LD)
    mu = Ks.T @ alpha
    V = forward_solve_ld(L, Ks)
    var = LD(1) - (V * V).sum(axis=0)
    return mu.astype(np.float64), np.sqrt(np.maximum(var, LD(1e-12))).astype(np.float64)


def rel_err(a, ref):
    """max |a - ref| / max |ref| (the scale a 1e-9 relative tolerance on a vector refers to)."""
    return float(np.max(np.abs(a - ref)) / max(float(np.max(np.abs(ref))), 1e-300))


def rel_err_pointwise(a, ref):
    return float(np.max(np.abs(a / ref - 1)))


BAR = 1e-9        # north_star: predictive mean / variance within 1e-9 relative of the reference's NumPy path


def check_selfpredict(got_resid, got_sd, ora_resid, ora_sd, x, y, e, ls, label=""):
    """The parity rule for (mu - y, sd) at the training points, with the tolerance written as a function of conditioning:

    * within BAR of the FP64 oracle (max-norm relative for mu - y, pointwise relative for sd): pass;
    * otherwise two FP64 evaluations of an ill-conditioned problem are being compared (each carries ~ cond(K) 2^-53) and an
      80-bit evaluation decides which side is off: the CUDA result must be within BAR of it, or at least as close to it as the
      reference's own arithmetic is (factor 2).

    Returns (err_resid, err_sd, None) or (err_resid, err_sd, dict with the errors of both sides against the truth)."""
    er, es = rel_err(got_resid, ora_resid), rel_err_pointwise(got_sd, ora_sd)
    if er <= BAR and es <= BAR:
        return er, es, None
    r_t, sd_t = selfpredict_ld(x, y, e, ls)
    d = dict(gpu_resid=rel_err(got_resid, r_t), gpu_sd=rel_err_pointwise(got_sd, sd_t),
             ref_resid=rel_err(ora_resid, r_t), ref_sd=rel_err_pointwise(ora_sd, sd_t))
    assert d["gpu_resid"] <= max(BAR, 2 * d["ref_resid"]) and d["gpu_sd"] <= max(BAR, 2 * d["ref_sd"]), (label, er, es, d)
    return er, es, d
