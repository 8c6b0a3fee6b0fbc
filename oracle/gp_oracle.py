"""CPU oracle for the GP-fitting hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

A plain NumPy/SciPy restatement of the reference algorithm (rferguson22/Gaussian-Process):
covariance, GP fit/predict, the calibration ("Wasserstein") loss, the particle-swarm
optimiser and the probability-surface reduction.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference`` legs may
import this module, and only as the checker or the timed CPU baseline.  The product path
(``gaussian-process_b200``) never imports it and has no CPU fallback.

Parity status: PINNED.  ``oracle/make_golden.py`` imports the *unmodified* Python reference
from ``/root/reference`` in the build container and stores its outputs under
``tests/golden/``; ``tests/test_oracle_golden.py`` checks every function here against those
vectors (bit-for-bit in the container that generated them, 1e-11 relative elsewhere because
a different host CPU may pick different OpenBLAS kernels).

Third-party piece that is absent from the reference tree: ``smt.sampling_methods.LHS``
(requirements.txt:3, unpinned; call site find_len_scales.py:69-70).  It only supplies the
initial swarm.  ``center_lhs`` below restates the published pyDOE/SMT "center" criterion
(one point at the centre of each of P equal cells per dimension, independently permuted);
parity at that single call is *unpinned* (the reference itself is unseeded there), so the
tests hand the same seeded initial swarm to both sides.

Each function cites the reference lines it follows.
"""
import os
import numpy as np
from numpy.linalg import cholesky, solve
from scipy.stats import norm

N_SIGMA = 1000          # find_len_scales.py:66
PENALTY_OOB = 1e13      # find_len_scales.py:156-157,182


# --------------------------------------------------------------------------- covariance
def rbf_kernel(x1, x2, ls):
    """ARD-RBF covariance, expanded squared distance with clamp.  GP_func.py:49-65.

    x1 (d,N1), x2 (d,N2), ls (d,) -> (N1,N2)."""
    a = x1 / ls[:, None]                                   # :56
    b = x2 / ls[:, None]                                   # :57
    a_sq = np.sum(a ** 2, axis=0).reshape(-1, 1)           # :59
    b_sq = np.sum(b ** 2, axis=0).reshape(1, -1)           # :60
    sq = a_sq + b_sq - 2 * np.dot(a.T, b)                  # :62
    sq = np.maximum(sq, 0)                                 # :63
    return np.exp(-0.5 * sq)                               # :65


# --------------------------------------------------------------------------- fit / predict
def gp_fit(x_known, y_known, e_known, ls):
    """K = k(X,X)+diag(e^2); L = chol(K); alpha = L^-T L^-1 y.  GP_func.py:21-24.
    The reference calls the general solver on the triangular factor; so does the oracle."""
    K = rbf_kernel(x_known, x_known, ls) + np.diag(e_known ** 2)
    L = cholesky(K)
    alpha = solve(L.T, solve(L, y_known))
    return L, alpha


def gp_predict(x_known, L, alpha, x_fit, ls, batch_size=10000):
    """Batched predictive mean and clipped std-dev.  GP_func.py:26-45."""
    mu_parts, sd_parts = [], []
    m = x_fit.shape[1]
    for lo in range(0, m, batch_size):
        xb = x_fit[:, lo:min(lo + batch_size, m)]
        Ks = rbf_kernel(x_known, xb, ls)                   # :32
        prior = np.ones(xb.shape[1])                       # :34
        mu = Ks.T @ alpha                                  # :36
        v = solve(L, Ks)                                   # :38
        var = np.clip(prior - np.sum(v ** 2, axis=0), 1e-12, None)   # :39
        mu_parts.append(mu)
        sd_parts.append(np.sqrt(var))                      # :40
    return np.concatenate(mu_parts), np.concatenate(sd_parts)


def gp(x_known, y_known, e_known, x_fit, ls, batch_size=10000):
    """GP(x_known, y_known, e_known, x_fit, lengths, batch_size).  GP_func.py:12-45."""
    L, alpha = gp_fit(x_known, y_known, e_known, ls)
    return gp_predict(x_known, L, alpha, x_fit, ls, batch_size)


# --------------------------------------------------------------------------- loss
def sigma_to_percent(s):
    """Phi(s)-Phi(-s).  find_len_scales.py:192-201."""
    return norm.cdf(s) - norm.cdf(-s)


def sigma_grid():
    """find_len_scales.py:66-67."""
    s = np.linspace(0.001, 3, N_SIGMA)
    return s, sigma_to_percent(s)


def out_of_bounds(ls, lower, upper):
    """On-bound counts as outside.  find_len_scales.py:156."""
    return bool(np.any(ls <= lower) or np.any(ls >= upper))


def calibration_counts(ls, x, y, e, sigma_vals):
    """Integer counts #{i : |pull_i(s)| <= 1} for every s, plus the self-prediction
    residual and std-dev they come from.  find_len_scales.py:159-163 (before the mean)."""
    n = x.shape[1]
    mu, sd = gp(x, y, e, x, ls, batch_size=n)              # :159
    scaled = sd[:, None] * sigma_vals[None, :]             # :161
    pulls = (mu[:, None] - y[:, None]) / np.maximum(scaled, 1e-12)   # :162
    inside = np.abs(pulls) <= 1                            # :163
    counts = inside.sum(axis=0).astype(np.int64)
    return counts, mu - y, sd, pulls


def finish_loss(counts, n, ls, sigma_vals, expected, lower, upper):
    """From counts to the PSO score (positive, minimised).  find_len_scales.py:163-177,182.
    ``counts/n`` is exactly what np.mean over booleans returns."""
    measured = counts.astype(np.float64) / n               # :163
    diff = np.abs(measured - expected)                     # :165
    wasserstein = np.trapz(diff, sigma_vals)               # :166
    span = upper - lower                                   # :168
    d_lo = (ls - lower) / span                             # :169
    d_hi = (upper - ls) / span                             # :170
    d_min = min(np.minimum(d_lo, d_hi))                    # :172-173
    proximity = np.clip(1 - 2 * d_min, 0.0, 1.0)           # :175
    return -(-wasserstein - (0.01 * proximity))            # :177 then the sign flip of :182


def evaluate_loss(ls, x, y, e, sigma_vals, expected, lower, upper):
    """evaluate_loss == -wass_loss.  find_len_scales.py:154-182."""
    if out_of_bounds(ls, lower, upper):
        return PENALTY_OOB                                 # -(-1e13)
    n = x.shape[1]
    mu, sd = gp(x, y, e, x, ls, batch_size=n)
    scaled = sd[:, None] * sigma_vals[None, :]
    pulls = (mu[:, None] - y[:, None]) / np.maximum(scaled, 1e-12)
    measured = np.mean(np.abs(pulls) <= 1, axis=0)
    diff = np.abs(measured - expected)
    wasserstein = np.trapz(diff, sigma_vals)
    span = upper - lower
    d_min = min(np.minimum((ls - lower) / span, (upper - ls) / span))
    proximity = np.clip(1 - 2 * d_min, 0.0, 1.0)
    return -(-wasserstein - (0.01 * proximity))


def _eval_tuple(t):
    return evaluate_loss(*t)


# --------------------------------------------------------------------------- PSO
def center_lhs(bounds, n, rng):
    """'center' Latin hypercube (pyDOE `_lhscentered`, the routine SMT's LHS wraps):
    cell centres (k+0.5)/n, an independent permutation per dimension, scaled to bounds.
    `rng` is a numpy RandomState; draw order = one permutation per dimension."""
    bounds = np.asarray(bounds, dtype=np.float64)
    d = bounds.shape[0]
    cut = np.linspace(0, 1, n + 1)
    centres = (cut[:n] + cut[1:n + 1]) / 2
    h = np.zeros((n, d))
    for j in range(d):
        h[:, j] = rng.permutation(centres)
    return bounds[:, 0] + h * (bounds[:, 1] - bounds[:, 0])


def kmeans_subsample(x, y, e, max_points=100):
    """find_len_scales.py:25-47 (sklearn KMeans, random_state=0, closest member per centre)."""
    from sklearn.cluster import KMeans
    X = x.T
    km = KMeans(n_clusters=max_points, n_init='auto', random_state=0)
    labels = km.fit_predict(X)
    centres = km.cluster_centers_
    keep = []
    for k in range(max_points):
        members = np.where(labels == k)[0]
        if len(members) == 0:
            continue
        d2 = np.sum((X[members] - centres[k]) ** 2, axis=1)
        keep.append(members[np.argmin(d2)])
    keep = np.array(keep)
    return x[:, keep], y[keep], e[keep]


def pso_bounds(x):
    """lower = smallest positive gap of the sorted unique coordinates, upper = range.
    find_len_scales.py:56-64."""
    ranges = np.max(x, axis=1) - np.min(x, axis=1)
    gaps = [np.diff(np.unique(row)) for row in x]
    lower = np.array([np.min(g[g > 0]) if np.any(g > 0) else 0 for g in gaps])
    return lower, ranges


def len_scale_opt(x, y, e, progress=False, *, init_positions=None, lhs_rng=None,
                  num_particles=40, max_iter=500, patience=100, max_points=100,
                  trace=None, map_fn=None, printer=print):
    """The PSO of find_len_scales.py:22-150, evaluated serially (or through `map_fn`).

    Random draws come from the global legacy ``np.random`` stream in the reference's order
    (:91-92 two rand(P,d) per iteration, :128 one uniform per restart); the caller seeds it.
    `trace`, if a list, receives (positions.copy(), scores.copy()) for every swarm evaluation."""
    if x.shape[1] > max_points:
        printer(f"Dataset too large ({x.shape[1]} points). Subsampling to {max_points} for hyperparameter optimisation.")
        x, y, e = kmeans_subsample(x, y, e, max_points)
    ndim = len(x)
    lower, upper = pso_bounds(x)
    v_max = 1.0 * (upper - lower)
    sigma_vals, expected = sigma_grid()
    if init_positions is None:
        rng = lhs_rng if lhs_rng is not None else np.random.RandomState()
        init_positions = center_lhs(np.column_stack((lower, upper)), num_particles, rng)
    pos = np.array(init_positions, dtype=np.float64)
    vel = np.zeros_like(pos)
    mapper = map_fn if map_fn is not None else (lambda f, it: [f(t) for t in it])

    def swarm_scores(p):
        args = [(q, x, y, e, sigma_vals, expected, lower, upper) for q in p]
        s = np.array(list(mapper(_eval_tuple, args)))
        if trace is not None:
            trace.append((p.copy(), s.copy()))
        return s

    pbest_s = swarm_scores(pos)
    pbest = pos.copy()
    g = np.argmin(pbest_s)
    gbest, gbest_s = pbest[g].copy(), pbest_s[g]
    stall, restarts = 0, 0
    for i in range(max_iter):
        w = max(0.4, 0.9 - i * 0.002)
        c1 = c2 = 1.4
        r1 = np.random.rand(num_particles, ndim)
        r2 = np.random.rand(num_particles, ndim)
        vel = (w * vel + c1 * r1 * (pbest - pos) + c2 * r2 * (gbest - pos))
        vel = np.clip(vel, -v_max, v_max)
        pos += vel
        pos = np.clip(pos, lower, upper)
        s = swarm_scores(pos)
        better = s < pbest_s
        pbest[better] = pos[better]
        pbest_s[better] = s[better]
        c = np.argmin(pbest_s)
        if pbest_s[c] < gbest_s:
            gbest_s, gbest, stall = pbest_s[c], pbest[c].copy(), 0
        else:
            stall += 1
        if progress and i % 20 == 0:
            printer(f"Iter {i}: Best Score = {gbest_s:.6f}, No Improve = {stall}")
        if stall >= patience:
            if progress:
                printer(f"Stagnation at iter {i}, soft-restarting swarm...")
            noise = 0.1 * (upper - lower)
            pbest += np.random.uniform(-noise, noise, pbest.shape)
            pbest = np.clip(pbest, lower, upper)
            pos = pbest.copy()
            vel = np.zeros_like(pos)
            pbest_s = swarm_scores(pbest)
            g = np.argmin(pbest_s)
            gbest, gbest_s = pbest[g].copy(), pbest_s[g]
            stall = 0
            restarts += 1
    printer("PSO completed.")
    printer("Optimal length scale found:")
    printer(gbest)
    printer("Value of loss function:")
    printer(gbest_s)
    printer(f"Total soft restarts: {restarts}")
    return gbest


# --------------------------------------------------------------------------- surface
def sum_gaussians(ys, gaus):
    """Mean over experiments of the Gaussian probability in a bin of width dy around each y.
    dy = |max-min| / len(ys)  (100, not 99).  calc_prob_surf.py:15-30."""
    z = np.zeros(len(ys))
    dy = abs(max(ys) - min(ys)) / len(ys)
    n_exp = int(len(gaus) / 2)
    for i in range(n_exp):
        z += norm.cdf(ys + dy / 2, loc=gaus[2 * i], scale=gaus[2 * i + 1])
        z -= norm.cdf(ys - dy / 2, loc=gaus[2 * i], scale=gaus[2 * i + 1])
    return z / n_exp


def prob_surface_rows(table, points=100):
    """Arithmetic of generate_prob_surf's row loop (calc_prob_surf.py:67-81) on a plain
    (R, 2E) array of interleaved (mean, std) with non-finite gaps.  Returns
    (row_index[Rv], y[Rv,points], p[Rv,points]) for the rows that pass the :71 test."""
    idx, ys, ps = [], [], []
    for r in range(table.shape[0]):
        row = table[r]
        g = row[np.isfinite(row)]                          # :69
        if len(g) >= 2 and len(g) % 2 == 0:                # :71
            means, stds = g[::2], g[1::2]                  # :72-73
            ymin = min(means - 3 * stds)
            ymax = max(means + 3 * stds)
            y = np.linspace(ymin, ymax, points)            # :77
            idx.append(r)
            ys.append(y)
            ps.append(sum_gaussians(y, g))                 # :78
    if not idx:
        return np.zeros(0, dtype=np.int64), np.zeros((0, points)), np.zeros((0, points))
    return np.array(idx, dtype=np.int64), np.array(ys), np.array(ps)


# --------------------------------------------------------------------------- synthetic data
def synthetic_problem(n, d, seed=0):
    """X uniform on the unit box plus its 2^d corners, y = sum sin(2 pi x) + 0.1 N(0,1),
    e ~ U(0.05, 0.15).  Returns x (d,n), y (n,), e (n,)."""
    rng = np.random.default_rng(seed)
    corners = np.array(np.meshgrid(*[[0.0, 1.0]] * d, indexing="ij")).reshape(d, -1)
    k = corners.shape[1]
    x = np.concatenate([rng.random((d, n - k)), corners], axis=1)
    y = np.sin(2 * np.pi * x).sum(axis=0) + 0.1 * rng.standard_normal(n)
    e = rng.uniform(0.05, 0.15, n)
    return x, y, e


def synthetic_particles(x, p, seed=0):
    """Seeded centre-LHS strictly inside the PSO bounds (no particle short-circuits)."""
    lower, upper = pso_bounds(x)
    pos = center_lhs(np.column_stack((lower, upper)), p, np.random.RandomState(seed))
    return pos, lower, upper


def synthetic_surface_table(r, n_exp, seed=0, p_gap=0.25):
    """(R, 2E) merged-table payload: mu ~ N(0.5,0.2), sd ~ U(0.05,0.2), pairs inf-masked."""
    rng = np.random.default_rng(seed)
    mu = rng.normal(0.5, 0.2, (r, n_exp))
    sd = rng.uniform(0.05, 0.2, (r, n_exp))
    gap = rng.random((r, n_exp)) < p_gap
    mu[gap] = np.inf
    sd[gap] = np.inf
    t = np.empty((r, 2 * n_exp))
    t[:, 0::2] = mu
    t[:, 1::2] = sd
    return t


def adversarial_problem(n, d, seed=0, e_lo=1e-6, e_hi=1.0, n_isolated=8):
    """Stress case for the emulated-FP64 row scales: e log-uniform over six decades, and a few isolated points
    (far from every other point, so their row of K is ~ e_i^2 on the diagonal only) carrying the smallest e.
    Returns x (d,n), y (n,), e (n,)."""
    rng = np.random.default_rng(seed)
    x = rng.random((d, n))
    x[:, :n_isolated] = 3.0 + 2.0 * np.arange(n_isolated)[None, :] + rng.random((d, n_isolated))
    y = np.sin(2 * np.pi * x).sum(axis=0) + 0.1 * rng.standard_normal(n)
    e = np.exp(rng.uniform(np.log(e_lo), np.log(e_hi), n))
    e[:n_isolated] = e_lo * (1.0 + rng.random(n_isolated))
    return x, y, e
