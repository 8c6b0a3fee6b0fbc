"""Length-scale optimiser: host side of find_len_scales.py on top of the batched GPU loss.

The particle-swarm state update stays in NumPy with the reference's expression order and RNG call
order (find_len_scales.py:88-100,123-141) so that a seeded run reproduces the reference trajectory;
what the reference fans out over a fork pool (`executor.map(evaluate_loss_helper, ...)`, :77,:103,:134)
is one `gpk_loss_batched` launch per swarm evaluation.  The GPU returns the integer count matrix and
the host finishes :163-177 with the same NumPy expressions, so the score is bit-identical whenever
the counts are."""
import numpy as np
from scipy.stats import norm

from . import engine, lhs as _lhs, sharding

N_SIGMA = 1000
OOB_SCORE = 1e13


def sigma_to_percent(x):
    """Fraction of a normal distribution within x standard deviations.  find_len_scales.py:192-201."""
    return norm.cdf(x) - norm.cdf(-x)


def sigma_grid():
    sigma_vals = np.linspace(0.001, 3, N_SIGMA)           # find_len_scales.py:66
    return sigma_vals, sigma_to_percent(sigma_vals)        # :67


def out_of_bounds_mask(positions, lower, upper):
    """True where a particle touches or leaves the box (find_len_scales.py:156: <= / >=)."""
    positions = np.atleast_2d(positions)
    return np.any(positions <= lower, axis=1) | np.any(positions >= upper, axis=1)


def finish_losses(counts, n_points, positions, sigma_vals, expected, lower, upper, oob, work=None, wasserstein=None):
    """counts (P,S) -> PSO scores (P,), the arithmetic of find_len_scales.py:163-177 and the sign
    flip of :182, vectorised over particles (row-wise identical to the scalar expressions).
    `work` = (d_sigma (1,S-1), buffer (P,S), buffer (P,S-1)), optional, for the PSO loop: np.trapz recomputes
    diff(sigma_vals) and allocates five (P,S) temporaries per call; with `work` the integral is formed by trapz's own
    expression `(d * (y[1:] + y[:-1]) / 2.0).sum(-1)`, same operations in the same order, in preallocated storage."""
    positions = np.atleast_2d(positions)
    if wasserstein is not None:
        pass                                       # formed by libgpk's one-pass host routine (SwarmLoss._integral)
    elif work is None:
        measured = counts.astype(np.float64) / n_points                   # np.mean(bool, axis=0)
        diff = np.abs(measured - expected[None, :])
        wasserstein = np.trapz(diff, sigma_vals, axis=1)
    else:
        d_sigma, a, b = work
        np.divide(counts, float(n_points), out=a)                         # int32 -> float64, then the division
        np.subtract(a, expected[None, :], out=a)
        np.abs(a, out=a)
        np.add(a[:, 1:], a[:, :-1], out=b)
        np.multiply(d_sigma, b, out=b)
        np.divide(b, 2.0, out=b)
        wasserstein = b.sum(-1)
    span = upper - lower
    with np.errstate(invalid="ignore", divide="ignore"):     # degenerate (zero-range) dims are always out of bounds
        d_lo = (positions - lower) / span
        d_hi = (upper - positions) / span
        d_min = np.min(np.minimum(d_lo, d_hi), axis=1)
        proximity = np.clip(1 - 2 * d_min, 0.0, 1.0)
    scores = -(-wasserstein - (0.01 * proximity))
    return np.where(oob, OOB_SCORE, scores)


_FAST_FINISH = {}      # sigma-grid length -> True / False: libgpk's host finish checked against NumPy once per process


class SwarmLoss:
    """evaluate_loss for a whole swarm.  Holds the device-resident training set."""

    def __init__(self, x_known, y_known, e_known, sigma_vals, expected_percents, lower_bounds, upper_bounds):
        self.ctx = engine.LossContext(x_known, y_known, e_known, sigma_vals)
        self.n = np.asarray(x_known).shape[1]
        self.sigma_vals = np.ascontiguousarray(sigma_vals, dtype=np.float64)
        self.expected = np.ascontiguousarray(expected_percents, dtype=np.float64)
        self.lower = np.asarray(lower_bounds, dtype=np.float64)
        self.upper = np.asarray(upper_bounds, dtype=np.float64)
        self.evaluations = 0
        self._d_sigma = np.diff(self.sigma_vals).reshape(1, -1)
        self._work = None
        self._table = np.arange(self.n + 1, dtype=np.int32).astype(np.float64) / self.n      # counts / n, every value
        self._scratch = np.empty(2 * self.sigma_vals.shape[0])
        self._fin = None

    def _integral(self, counts):
        """np.trapz(|counts / n - expected|, sigma_vals, axis=1) by libgpk's host routine (gpk_finish_wasserstein: NumPy's
        operation order and pairwise summation in one pass).  Verified against the NumPy expressions on the first swarm
        evaluation of the process (per grid length); any difference switches it off for good."""
        key = counts.shape[1]
        if _FAST_FINISH.get(key) is False or counts.dtype != np.int32 or not counts.flags.c_contiguous:
            return None
        fin = self._fin
        if fin is None or fin[0] != counts.shape[0]:
            import ctypes as C
            out = np.empty(counts.shape[0])
            vp = C.c_void_p
            fin = self._fin = (counts.shape[0], engine._lib.load().gpk_finish_wasserstein, out, vp(self._table.ctypes.data),
                               vp(self.expected.ctypes.data), vp(self._d_sigma.ctypes.data), vp(self._scratch.ctypes.data),
                               vp(out.ctypes.data))
        _p, fn, out, p_table, p_exp, p_dsig, p_scr, p_out = fin
        st = fn(counts.ctypes.data, counts.shape[0], counts.shape[1], p_table, self.n, p_exp, p_dsig, p_scr, p_out)
        out = out.copy()
        if st != 0:
            _FAST_FINISH[key] = False
            return None
        if key not in _FAST_FINISH:
            ref = np.trapz(np.abs(counts.astype(np.float64) / self.n - self.expected[None, :]), self.sigma_vals, axis=1)
            _FAST_FINISH[key] = bool(np.array_equal(ref, out))
            if not _FAST_FINISH[key]:
                return None
        return out

    def __call__(self, positions):
        positions = np.atleast_2d(np.asarray(positions, dtype=np.float64))
        oob = out_of_bounds_mask(positions, self.lower, self.upper)
        if np.all(oob):
            return np.full(positions.shape[0], OOB_SCORE)
        counts, info = self.ctx.counts_view(positions, active=~oob)         # views: consumed before the next evaluation
        if np.any(info != 0):
            raise np.linalg.LinAlgError("Matrix is not positive definite")   # GP_func.py:22 behaviour
        self.evaluations += positions.shape[0]
        p, s = counts.shape
        if self._work is None or self._work[1].shape[0] != p:
            self._work = (self._d_sigma, np.empty((p, s)), np.empty((p, s - 1)))
        return finish_losses(counts, self.n, positions, self.sigma_vals, self.expected, self.lower, self.upper, oob,
                             work=self._work, wasserstein=self._integral(counts))


def evaluate_loss_batch(positions, x_known, y_known, e_known, sigma_vals, expected_percents, lower_bounds,
                        upper_bounds):
    """Batched evaluate_loss: positions (P,d) -> (P,) float64 scores."""
    return SwarmLoss(x_known, y_known, e_known, sigma_vals, expected_percents, lower_bounds, upper_bounds)(positions)


def kmeans_subsample(x_known, y_known, e_known, max_points, backend=None):
    """KMeans cap of find_len_scales.py:25-47: KMeans(n_clusters=max_points, n_init='auto', random_state=0) on the points,
    then the member closest to each centroid.  backend "device" (default; env GPK_KMEANS): the Lloyd iterations run in
    libgpk (kmeans.fit_predict: sklearn's labels and centres); "sklearn": the library itself on the host."""
    import os
    backend = backend or os.environ.get("GPK_KMEANS", "device")
    pts = x_known.T
    if backend == "device":
        from . import kmeans
        labels, centers, _n_iter = kmeans.fit_predict(pts, max_points, random_state=0)
    elif backend == "sklearn":
        from sklearn.cluster import KMeans
        km = KMeans(n_clusters=max_points, n_init='auto', random_state=0)
        labels = km.fit_predict(pts)
        centers = km.cluster_centers_
    else:
        raise ValueError(f"GPK_KMEANS / backend must be 'device' or 'sklearn' (got {backend!r})")
    chosen = []
    for k in range(max_points):
        members = np.where(labels == k)[0]
        if len(members) == 0:
            continue
        dist2 = np.sum((pts[members] - centers[k]) ** 2, axis=1)
        chosen.append(members[np.argmin(dist2)])
    chosen = np.array(chosen)
    return x_known[:, chosen], y_known[chosen], e_known[chosen]


def swarm_bounds(x_known):
    """find_len_scales.py:56-64: lower = smallest positive coordinate gap, upper = coordinate range."""
    upper = np.max(x_known, axis=1) - np.min(x_known, axis=1)
    gaps = [np.diff(np.unique(row)) for row in x_known]
    lower = np.array([np.min(g[g > 0]) if np.any(g > 0) else 0 for g in gaps])
    return lower, upper


PSO_HDR_WORDS = 32            # include/gpk.h gpk_pso_layout: int32 header of the state buffer
_HDR = {"P": 0, "d": 1, "S": 2, "n_points": 3, "max_iter": 4, "patience": 5, "max_evals": 6, "iter": 7, "pending": 8,
        "stall": 9, "restarts": 10, "n_evals": 11, "error": 12, "rng_cursor": 13, "fallbacks": 14, "rng_len": 15, "logged": 16}
_OFFS = ("scalars", "lower", "upper", "vmax", "pos", "vel", "pbest_pos", "pbest_sc", "gbest_pos", "scores", "expected",
         "d_sigma", "measured_of", "log_best", "log_stall", "log_restart", "trace_pos", "trace_sc", "rng")


def device_loop_enabled():
    import os
    return os.environ.get("GPK_PSO_DEVICE", "1") != "0"


def _device_swarm_loop(x_known, y_known, e_known, positions, lower, upper, v_max, sigma_vals, expected, num_particles,
                       max_iter, patience, inertia_decay):
    """The swarm loop of find_len_scales.py:76-141 without a host round trip per iteration: max_evals pairs of
    (gpk_loss_batched, gpk_pso_step) are enqueued on the current stream and the state is read back once.  The random
    numbers the reference would draw from np.random during the loop are generated up front with the same generator and
    consumed on the device in the same order; afterwards the global np.random state is left exactly where the host loop
    would have left it.  Returns a dict with the final state, the per-iteration log and the trace of every evaluation."""
    import ctypes as C
    import torch
    lib = engine._lib.load()
    P, d = positions.shape
    n = x_known.shape[1]
    S = sigma_vals.shape[0]
    max_restarts = (max_iter // patience + 1) if patience >= 1 else max_iter
    max_evals = 1 + max_iter + max_restarts
    rng_len = (2 * max_iter + max_restarts) * P * d
    offs = (C.c_ulonglong * len(_OFFS))()
    total = lib.gpk_pso_layout(P, d, S, n, max_iter, max_evals, rng_len, offs)
    if total == 0:
        raise ValueError(f"swarm of {P} particles in {d} dimensions with {S} sigma scales is outside the kernels' limits")
    off = dict(zip(_OFFS, (int(v) for v in offs)))
    ctx = engine.LossContext(x_known, y_known, e_known, sigma_vals)
    if ctx._sig_unsort is not None:
        return None                                  # unsorted sigma grid: the host loop handles the permutation
    # ---- initial state (everything up to the logs), one upload
    rng_state = np.random.get_state()
    stream_vals = np.random.random_sample(rng_len)
    host = np.zeros(total, dtype=np.uint8)
    hdr = host[:PSO_HDR_WORDS * 4].view(np.int32)
    for k, v in (("P", P), ("d", d), ("S", S), ("n_points", n), ("max_iter", max_iter), ("patience", patience),
                 ("max_evals", max_evals), ("rng_len", rng_len)):
        hdr[_HDR[k]] = v

    def f64(name, count):
        return host[off[name]:off[name] + 8 * count].view(np.float64)

    f64("scalars", 8)[1] = inertia_decay
    f64("lower", d)[:] = lower
    f64("upper", d)[:] = upper
    f64("vmax", d)[:] = v_max
    f64("pos", P * d)[:] = positions.reshape(-1)
    f64("expected", S)[:] = expected
    f64("d_sigma", S)[:S - 1] = np.diff(sigma_vals)
    f64("measured_of", n + 1)[:] = np.arange(n + 1, dtype=np.int32).astype(np.float64) / n      # counts / n as NumPy divides
    f64("rng", rng_len)[:] = stream_vals
    dev = engine.device()
    state = torch.from_numpy(host).to(dev)
    d_pos = torch.from_numpy(np.ascontiguousarray(positions)).to(dev)
    oob0 = out_of_bounds_mask(positions, lower, upper)
    d_act = torch.from_numpy((~oob0).astype(np.uint8)).to(dev)
    counts = torch.zeros((P, S), dtype=torch.int32, device=dev)
    info = torch.zeros(P, dtype=torch.int32, device=dev)
    fb = torch.zeros(P, dtype=torch.uint8, device=dev)
    ws = ctx._ws(P)
    st = torch.cuda.current_stream(dev.index).cuda_stream
    tol = engine.get_check_tol() if ctx.mode != 0 else 0.0
    loss_args = (ctx.x.data_ptr(), ctx.y.data_ptr(), ctx.e.data_ptr(), d_pos.data_ptr(), d_act.data_ptr(), ctx.sig.data_ptr(),
                 P, n, d, S, ctx.mode, ctx._form(), tol, counts.data_ptr(), info.data_ptr(), None, None, None, fb.data_ptr(),
                 ws.data_ptr(), ws.numel(), st)
    step_args = (state.data_ptr(), total, P, d, S, n, max_iter, max_evals, rng_len, counts.data_ptr(), info.data_ptr(),
                 fb.data_ptr(), d_pos.data_ptr(), d_act.data_ptr(), st)
    loss, step = lib.gpk_loss_batched, lib.gpk_pso_step
    for _ in range(max_evals):
        rc = loss(*loss_args)
        if rc == 0:
            rc = step(*step_args)
        if rc != 0:
            engine._lib.check(rc, "device swarm loop")
    out = state.cpu().numpy()
    hdr = out[:PSO_HDR_WORDS * 4].view(np.int32)

    def g64(name, count):
        return out[off[name]:off[name] + 8 * count].view(np.float64)

    n_evals = int(hdr[_HDR["n_evals"]])
    consumed = int(hdr[_HDR["rng_cursor"]])
    np.random.set_state(rng_state)                   # leave the global stream where the host loop would have left it
    if consumed:
        np.random.random_sample(consumed)
    return {
        "error": int(hdr[_HDR["error"]]), "restarts": int(hdr[_HDR["restarts"]]), "n_evals": n_evals,
        "fallbacks": int(hdr[_HDR["fallbacks"]]), "iters_logged": min(int(hdr[_HDR["logged"]]), max_iter),
        "gbest_pos": g64("gbest_pos", d).copy(), "gbest_score": float(g64("scalars", 8)[0]),
        "log_best": g64("log_best", max_iter).copy(),
        "log_stall": out[off["log_stall"]:off["log_stall"] + 4 * max_iter].view(np.int32).copy(),
        "log_restart": out[off["log_restart"]:off["log_restart"] + 4 * max_iter].view(np.int32).copy(),
        "trace_pos": g64("trace_pos", max_evals * P * d).reshape(max_evals, P, d)[:n_evals].copy(),
        "trace_sc": g64("trace_sc", max_evals * P).reshape(max_evals, P)[:n_evals].copy(),
    }


def len_scale_opt(x_known, y_known, e_known, PSO_progress, *, num_particles=40, max_iter=500, patience=100,
                  max_points=100, inertia_decay=0.002, lhs_random_state=None, swarm_fn=None, trace=None,
                  distributed=None, timings=None):
    """Drop-in for find_len_scales.len_scale_opt (:22-150); keyword defaults are the reference's
    hard-coded constants (:24,:49-54).  `trace` (a list) receives (positions, scores) of every swarm
    evaluation; `swarm_fn` overrides the evaluator (tests).

    Under an initialised torch.distributed group (world_size > 1) every rank must follow the same trajectory.  The
    reference is unseeded (SMT's LHS draws from a private RandomState, r1 / r2 / the restart noise from the
    global np.random stream), so nothing may be assumed identical across ranks: rank 0's initial swarm and rank 0's
    np.random state are broadcast once, before the first evaluation (every other rank's global np.random state is
    replaced by it: the run then equals the single-process run of rank 0, seeded or not), and the result is broadcast
    at the end.  `distributed`: True shards the swarm evaluation over the ranks (each evaluates its slice, one
    all-gather of the losses per iteration), False keeps it local (every rank evaluates the whole swarm redundantly,
    no communication per iteration); None (default) shards only when one swarm evaluation is worth more than the
    exchange (sharding.worth_sharding).  `timings` (a dict) accumulates the seconds of the KMeans subsample and of the
    swarm loop under "kmeans" / "pso"."""
    import time as _time
    x_known = np.asarray(x_known, dtype=np.float64)
    y_known = np.asarray(y_known, dtype=np.float64)
    e_known = np.asarray(e_known, dtype=np.float64)
    _t0 = _time.perf_counter()
    if x_known.shape[1] > max_points:
        print(f"Dataset too large ({x_known.shape[1]} points). Subsampling to {max_points} for hyperparameter optimisation.")
        x_known, y_known, e_known = kmeans_subsample(x_known, y_known, e_known, max_points)
    _t1 = _time.perf_counter()

    ndim = len(x_known)
    restart_count = 0
    lower_bounds, upper_bounds = swarm_bounds(x_known)
    v_max = 1.0 * (upper_bounds - lower_bounds)
    sigma_vals, expected_percents = sigma_grid()

    lhs_cls = _lhs.get_lhs_class()
    if lhs_random_state is None:
        sampler = lhs_cls(xlimits=np.column_stack((lower_bounds, upper_bounds)), criterion="center")
    else:
        sampler = lhs_cls(xlimits=np.column_stack((lower_bounds, upper_bounds)), criterion="center",
                          random_state=lhs_random_state)
    positions = np.array(sampler(num_particles), dtype=np.float64)
    velocities = np.zeros_like(positions)

    in_group = sharding.world()[1] > 1
    if in_group:
        # one exchange per optimisation: the start of the trajectory (initial swarm + RNG stream) is rank 0's
        positions, rng_state = sharding.broadcast_object((positions, np.random.get_state()))
        positions = np.array(positions, dtype=np.float64)
        np.random.set_state(rng_state)
    if distributed is None:
        distributed = in_group and sharding.worth_sharding(num_particles, x_known.shape[1])
    sharded = bool(distributed and in_group)
    dev_run = None
    if swarm_fn is None and not sharded and device_loop_enabled():
        # the whole loop on the device (one kernel pair per swarm evaluation, no host round trip): same trajectory, bit
        # for bit, as the host loop below, which stays for sharded swarms, `swarm_fn` and GPK_PSO_DEVICE=0
        dev_run = _device_swarm_loop(x_known, y_known, e_known, positions, lower_bounds, upper_bounds, v_max, sigma_vals,
                                     expected_percents, num_particles, max_iter, patience, inertia_decay)
    if dev_run is not None:
        if trace is not None:
            for k in range(dev_run["n_evals"]):
                trace.append((dev_run["trace_pos"][k].copy(), dev_run["trace_sc"][k].copy()))
        for i in range(dev_run["iters_logged"]):
            if PSO_progress and i % 20 == 0:
                print(f"Iter {i}: Best Score = {dev_run['log_best'][i]:.6f}, No Improve = {int(dev_run['log_stall'][i])}")
            if dev_run["log_restart"][i] and PSO_progress:
                print(f"Stagnation at iter {i}, soft-restarting swarm...")
        if dev_run["error"] == 1:
            raise np.linalg.LinAlgError("Matrix is not positive definite")   # GP_func.py:22 behaviour
        if dev_run["error"]:
            raise RuntimeError("device swarm loop: random stream exhausted (more restarts than max_iter // patience + 1)")
        gbest_pos, gbest_score, restart_count = dev_run["gbest_pos"], np.float64(dev_run["gbest_score"]), dev_run["restarts"]
        evaluate = None
    else:
        evaluate = swarm_fn if swarm_fn is not None else SwarmLoss(
            x_known, y_known, e_known, sigma_vals, expected_percents, lower_bounds, upper_bounds)
        if sharded:
            evaluate = sharding.ShardedSwarm(evaluate)

    def scored(p):
        s = np.asarray(evaluate(p), dtype=np.float64)
        if trace is not None:
            trace.append((p.copy(), s.copy()))
        return s

    if dev_run is None:
        pbest_scores = scored(positions)
        pbest_pos = positions.copy()
        g = np.argmin(pbest_scores)
        gbest_pos = pbest_pos[g].copy()
        gbest_score = pbest_scores[g]
    stall = 0

    for i in range(max_iter if dev_run is None else 0):          # host loop (the device loop above has done all of it)
        w = max(0.4, 0.9 - i * inertia_decay)
        c1 = c2 = 1.4
        r1 = np.random.rand(num_particles, ndim)
        r2 = np.random.rand(num_particles, ndim)
        velocities = (w * velocities + c1 * r1 * (pbest_pos - positions) + c2 * r2 * (gbest_pos - positions))
        velocities = np.clip(velocities, -v_max, v_max)
        positions += velocities
        positions = np.clip(positions, lower_bounds, upper_bounds)

        scores = scored(positions)
        improved = scores < pbest_scores
        pbest_pos[improved] = positions[improved]
        pbest_scores[improved] = scores[improved]
        cur = np.argmin(pbest_scores)
        if pbest_scores[cur] < gbest_score:
            gbest_score = pbest_scores[cur]
            gbest_pos = pbest_pos[cur].copy()
            stall = 0
        else:
            stall += 1

        if PSO_progress and i % 20 == 0:
            print(f"Iter {i}: Best Score = {gbest_score:.6f}, No Improve = {stall}")

        if stall >= patience:
            if PSO_progress:
                print(f"Stagnation at iter {i}, soft-restarting swarm...")
            noise = 0.1 * (upper_bounds - lower_bounds)
            pbest_pos += np.random.uniform(-noise, noise, pbest_pos.shape)
            pbest_pos = np.clip(pbest_pos, lower_bounds, upper_bounds)
            positions = pbest_pos.copy()
            velocities = np.zeros_like(positions)
            pbest_scores = scored(pbest_pos)
            g = np.argmin(pbest_scores)
            gbest_pos = pbest_pos[g].copy()
            gbest_score = pbest_scores[g]
            stall = 0
            restart_count += 1

    if in_group:
        gbest_pos, gbest_score = sharding.broadcast_object((gbest_pos, gbest_score))
    if timings is not None:
        timings["kmeans"] = timings.get("kmeans", 0.0) + (_t1 - _t0)
        timings["pso"] = timings.get("pso", 0.0) + (_time.perf_counter() - _t1)
    print("PSO completed.")
    print("Optimal length scale found:")
    print(gbest_pos)
    print("Value of loss function:")
    print(gbest_score)
    print(f"Total soft restarts: {restart_count}")
    return gbest_pos
