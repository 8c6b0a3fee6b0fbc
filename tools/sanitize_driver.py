"""Small end-to-end pass over every kernel family for compute-sanitizer (tools/sanitize.sh): the one-CTA loss (both forms),
the blocked path in native FP64 and in the emulated mode (tcgen05 / TMEM / mbarrier pipeline) with the residual probe and
a forced fallback, fit + predict in both modes, the surface reduction."""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from gaussian_process_b200 import engine, pso, synthetic

which = sys.argv[1] if len(sys.argv) > 1 else "all"
s, ex = pso.sigma_grid()
if which in ("all", "small"):
    x, y, e = synthetic.problem(60, 2, 0)
    pos, lo, hi = synthetic.particles(x, 5, 0)
    for form in ("reduced", "reference"):
        engine.set_loss_form(form)
        print("small", form, pso.evaluate_loss_batch(pos, x, y, e, s, ex, lo, hi)[:2])
    engine.set_loss_form("reduced")
if which in ("all", "fp64"):
    x, y, e = synthetic.problem(300, 3, 1)
    pos, lo, hi = synthetic.particles(x, 3, 1)
    engine.set_precision("fp64")
    print("blocked fp64", pso.evaluate_loss_batch(pos, x, y, e, s, ex, lo, hi))
    mu, sd = engine.gp(x, y, e, np.random.default_rng(0).random((3, 300)), pos[0])
    print("predict fp64", mu[:2], sd[:2])
if which in ("all", "emu"):
    x, y, e = synthetic.problem(384, 2, 2)
    pos, lo, hi = synthetic.particles(x, 3, 2)
    for mode in ("emu", "emu21"):
        engine.set_precision(mode)
        print("blocked", mode, pso.evaluate_loss_batch(pos, x, y, e, s, ex, lo, hi))
    engine.set_precision("emu")
    engine.set_check_tol(1e-30)                       # every particle fails the probe: the FP64 fallback pass runs
    ctx = engine.LossContext(x, y, e, s)
    c, i = ctx.counts(pos)
    print("forced fallback", ctx.fallbacks, c[:, 500])
    engine.set_check_tol(1e-10)
    mu, sd = engine.gp(x, y, e, np.random.default_rng(0).random((2, 300)), pos[0])
    print("predict emu", mu[:2], sd[:2])
if which in ("all", "surface"):
    t = synthetic.surface_table(500, 6, seed=1)
    idx, yy, pp = engine.prob_surface(t)
    print("surface", idx.shape, pp.sum())
engine.set_precision("auto")
