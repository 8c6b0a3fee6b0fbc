"""Randomised parity sweep of the loss path against the CPU checker (oracle/), both arithmetic modes, beyond the sizes
the test suite pins: ragged N around every 128-block boundary, 1 <= d <= 6, odd swarm sizes, inactive particles, wide
uncertainty ranges, clustered points, short and long length scales.  Prints one line per case and a summary; exit code 1
when any count differs outside a rounding tie or a residual / sd exceeds the tolerance.
    python tools/parity_sweep.py [seconds]"""
import os
import sys
import time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from oracle import gp_oracle as orc, truth80
from gaussian_process_b200 import engine

SIZES = [3, 64, 127, 128, 129, 130, 200, 255, 256, 257, 383, 385, 511, 513, 640, 777, 1000, 1023, 1025, 1300, 1537, 2047]


def run(seconds=150.0, max_cases=None, seed=2026, sizes=SIZES, verbose=True):
    """Returns (cases, failures, worst residual error, worst sd error)."""
    s, ex = orc.sigma_grid()
    rng = np.random.default_rng(seed)
    t_start = time.time()
    bad = 0
    worst = {"resid": 0.0, "sd": 0.0}
    n_cases = 0
    while time.time() - t_start < seconds and (max_cases is None or n_cases < max_cases):
        bad += _case(rng, sizes, s, worst)
        if n_cases % 3 == 0:
            bad += _predict_case(rng, sizes, worst)
        n_cases += 1
        if verbose and n_cases % 50 == 0:
            print(f"{n_cases} cases, {bad} failures, worst resid {worst['resid']:.2e} sd {worst['sd']:.2e}, {time.time() - t_start:.0f} s", flush=True)
    engine.set_precision("auto")
    if verbose:
        print(f"prediction cases: worst mean error {worst.get('pred_mu', 0.0):.2e}, worst sd error {worst.get('pred_sd', 0.0):.2e}")
    return n_cases, bad, worst["resid"], worst["sd"]


def _predict_case(rng, sizes, worst):
    """GP() fit + predict at random grid sizes (single point, ragged chunk tails) against the checker, both modes."""
    bad = 0
    n = int(rng.choice(sizes))
    d = int(rng.integers(1, 7))
    m = int(rng.choice([1, 7, 100, 1000, 8193, 20001]))
    x = rng.random((d, n))
    y = np.sin(3 * x).sum(axis=0) + 0.1 * rng.standard_normal(n)
    e = rng.uniform(0.05, 0.2, n)
    ls = rng.uniform(0.1, 1.2, d)
    grid = rng.random((d, m)) * 1.2 - 0.1                      # a little outside the hull too
    mu_r, sd_r = orc.gp(x, y, e, grid, ls)
    for mode in ("fp64", "emu"):
        engine.set_precision(mode)
        mu, sd = engine.gp(x, y, e, grid, ls)
        em = np.abs(mu - mu_r).max() / max(np.abs(mu_r).max(), 1e-300)
        es = np.abs(sd / sd_r - 1).max()
        worst["pred_mu"] = max(worst.get("pred_mu", 0.0), em)
        worst["pred_sd"] = max(worst.get("pred_sd", 0.0), es)
        if mu.shape == (m,) and not (em < 1e-9 and es < 1e-9) and n <= 2100:
            # two FP64 evaluations differ by more than the bar: an 80-bit evaluation decides (on a sample of the grid)
            pick = np.unique(np.concatenate([np.argsort(-np.abs(mu - mu_r))[:20], np.argsort(-np.abs(sd / sd_r - 1))[:20]]))
            mu_t, sd_t = truth80.predict_ld(x, y, e, grid[:, pick], ls)
            sc = max(np.abs(mu_r).max(), 1e-300)
            g_mu, r_mu = np.abs(mu[pick] - mu_t).max() / sc, np.abs(mu_r[pick] - mu_t).max() / sc
            g_sd, r_sd = np.abs(sd[pick] / sd_t - 1).max(), np.abs(sd_r[pick] / sd_t - 1).max()
            print(f"  80-bit truth decides predict n={n} d={d} m={m} {mode}: vs checker {em:.2e} / {es:.2e}; vs truth gpu "
                  f"{g_mu:.2e} / {g_sd:.2e}, checker {r_mu:.2e} / {r_sd:.2e}", flush=True)
            if g_mu <= max(1e-9, 2 * r_mu) and g_sd <= max(1e-9, 2 * r_sd):
                em, es = min(em, 1e-10), min(es, 1e-10)
        if not (mu.shape == (m,) and em < 1e-9 and es < 1e-9):
            bad += 1
            print(f"FAIL predict n={n} d={d} m={m} {mode}: mean err {em:.2e} sd err {es:.2e}", flush=True)
    return bad


def _case(rng, sizes, s, worst):
    bad = 0
    if True:
        n = int(rng.choice(sizes))
        d = int(rng.integers(1, 7))
        p = int(rng.choice([1, 2, 3, 5, 7]))
        kind = rng.choice(["uniform", "clustered", "wide_e", "short_ls", "long_ls"])
        x = rng.random((d, n))
        if kind == "clustered":
            centers = rng.random((d, 8))
            x = centers[:, rng.integers(0, 8, n)] + 0.02 * rng.standard_normal((d, n))
        y = np.sin(3 * x).sum(axis=0) + 0.1 * rng.standard_normal(n)
        e = rng.uniform(0.05, 0.2, n)
        if kind == "wide_e":
            e = 10.0 ** rng.uniform(-2.5, 0.3, n)
        lo_ls, hi_ls = (0.02, 0.1) if kind == "short_ls" else ((1.5, 4.0) if kind == "long_ls" else (0.05, 1.5))
        pos = rng.uniform(lo_ls, hi_ls, (p, d))
        active = np.ones(p, dtype=bool)
        if p > 2 and rng.random() < 0.3:
            active[rng.integers(0, p)] = False
        for mode in ("fp64", "emu"):
            engine.set_precision(mode)
            try:
                counts, info, resid, sd = engine.LossContext(x, y, e, s).counts(pos, active=active, want_detail=True)
            except Exception as exc:                                   # noqa: BLE001
                print(f"FAIL n={n} d={d} p={p} {kind} {mode}: {type(exc).__name__}: {exc}", flush=True)
                bad += 1
                continue
            for k in range(p):
                if not active[k]:
                    continue
                if info[k] != 0:
                    # the checker must fail too (Cholesky breakdown)
                    try:
                        orc.calibration_counts(pos[k], x, y, e, s)
                        print(f"FAIL n={n} d={d} {kind} {mode} particle {k}: info={info[k]} but the checker factorises", flush=True)
                        bad += 1
                    except np.linalg.LinAlgError:
                        pass
                    continue
                c_ref, r_ref, sd_ref, pulls = orc.calibration_counts(pos[k], x, y, e, s)
                try:
                    er, es, detail = truth80.check_selfpredict(resid[k], sd[k], r_ref, sd_ref, x, y, e, pos[k])
                    within = True
                except AssertionError as exc:
                    er, es, detail = truth80.rel_err(resid[k], r_ref), truth80.rel_err_pointwise(sd[k], sd_ref), str(exc)
                    within = False
                if detail is not None and within:
                    worst["decided_by_truth"] = worst.get("decided_by_truth", 0) + 1
                    print(f"  80-bit truth decides n={n} d={d} {kind} {mode} particle {k}: vs checker {er:.2e} / {es:.2e}; vs truth "
                          f"gpu {detail['gpu_resid']:.2e} / {detail['gpu_sd']:.2e}, checker {detail['ref_resid']:.2e} / {detail['ref_sd']:.2e}",
                          flush=True)
                worst["resid"] = max(worst["resid"], er)
                worst["sd"] = max(worst["sd"], es)
                tie = np.min(np.abs(np.abs(pulls) - 1))
                dc = int(np.abs(counts[k] - c_ref).max())
                ok = within and (dc == 0 or (tie < 1e-7 and dc <= 2))
                if not ok:
                    bad += 1
                    print(f"FAIL n={n} d={d} p={p} {kind} {mode} particle {k}: resid err {er:.2e} sd err {es:.2e} count diff {dc} (closest pull to 1: {tie:.1e})", flush=True)
    return bad


if __name__ == "__main__":
    n_cases, bad, wr, ws = run(float(sys.argv[1]) if len(sys.argv) > 1 else 150.0)
    print(f"parity sweep: {n_cases} cases x 2 modes, failures {bad}, worst relative residual error {wr:.2e}, worst sd error {ws:.2e}")
    sys.exit(1 if bad else 0)
