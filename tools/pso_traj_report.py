"""Seeded PSO trajectories against the reference's recorded ones (tests/golden/pso_trace_*.npz), for both algebraic forms
of the one-CTA loss (engine.set_loss_form): how many swarm evaluations stay bit-identical, the final length scales, and how
far the downstream GP() prediction moves when the trajectories split.
    python tools/pso_traj_report.py > profiles/r02_pso_trajectories.log"""
import contextlib
import io
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from oracle import gp_oracle as orc, ref_shim
from gaussian_process_b200 import engine, pso

ref_shim.install_smt_stub()
G = os.path.join(ROOT, "tests", "golden")
gp_cases = np.load(os.path.join(G, "gp_cases.npz"))
s, ex = orc.sigma_grid()
print("| trace | form | identical swarm evaluations | differing scores before the split | final length scales (ours) | "
      "reference's | loss ours / reference's (oracle) | max abs d(mu), d(sd) of GP() on the grid |")
print("|---|---|---|---|---|---|---|---|")
for name in ("Test_file1_exp1", "Test_file3_exp2", "syn100_d2"):
    g = np.load(os.path.join(G, f"pso_trace_{name}.npz"))
    x, y, e, seed = g["x"], g["y"], g["e"], int(g["seed"])
    lo, hi = orc.pso_bounds(x)
    if f"{name}__grid" in gp_cases.files:
        grid = gp_cases[f"{name}__grid"].T
    else:
        grid = np.random.default_rng(0).random((x.shape[0], 2000))
    mu_ref, sd_ref = orc.gp(x, y, e, grid, g["best"])
    for form in ("reduced", "reference"):
        engine.set_loss_form(form)
        np.random.seed(seed)
        trace = []
        with contextlib.redirect_stdout(io.StringIO()):
            best = pso.len_scale_opt(x, y, e, True, lhs_random_state=seed, trace=trace)
        pos = np.array([t[0] for t in trace])
        sc = np.array([t[1] for t in trace])
        n_common = min(len(pos), len(g["positions"]))
        same = [np.array_equal(pos[i], g["positions"][i]) for i in range(n_common)]
        first = same.index(False) if False in same else n_common
        ndiff = int(np.sum(sc[:first] != g["scores"][:first]))
        mu, sd = orc.gp(x, y, e, grid, best)
        mine = orc.evaluate_loss(best, x, y, e, s, ex, lo, hi)
        theirs = orc.evaluate_loss(g["best"], x, y, e, s, ex, lo, hi)
        print(f"| {name} | {form} | {first} / {len(g['positions'])} | {ndiff} of {first * sc.shape[1]} | {np.array2string(best, precision=6)} | "
              f"{np.array2string(g['best'], precision=6)} | {mine:.6f} / {theirs:.6f} | {np.abs(mu - mu_ref).max():.3e}, {np.abs(sd - sd_ref).max():.3e} |",
              flush=True)
engine.set_loss_form("reduced")
