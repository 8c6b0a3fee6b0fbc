"""Where one PSO iteration of the default workflow goes (N -> 100 points, 40 particles, 500 iterations): cProfile of
pso.len_scale_opt plus explicit timers around the swarm evaluation's stages.
    python tools/pso_host_profile.py > profiles/r02_pso_host_profile.log"""
import contextlib
import cProfile
import io
import os
import pstats
import sys
import time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from gaussian_process_b200 import engine, pso, synthetic

x, y, e = synthetic.problem(100, 2, 0)
with contextlib.redirect_stdout(io.StringIO()):
    np.random.seed(0)
    pso.len_scale_opt(x, y, e, False, lhs_random_state=0, max_iter=30)           # warm-up
    np.random.seed(0)
    t0 = time.perf_counter()
    pso.len_scale_opt(x, y, e, False, lhs_random_state=0)
    total = time.perf_counter() - t0
    pr = cProfile.Profile()
    np.random.seed(0)
    pr.enable()
    pso.len_scale_opt(x, y, e, False, lhs_random_state=0)
    pr.disable()
print(f"len_scale_opt, N=100, 40 particles, 500 iterations: {total * 1e3:.1f} ms = {total / 502 * 1e6:.0f} us per swarm evaluation")
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(22)
print(s.getvalue()[:6000])

# the kernel alone, device timed
sv, ex = pso.sigma_grid()
pos, lo, hi = synthetic.particles(x, 40, 0)
ctx = engine.LossContext(x, y, e, sv)
dpos = engine.to_dev(pos)
dact = torch.ones(40, dtype=torch.uint8, device="cuda")
for _ in range(5):
    ctx.counts_device(dpos, dact, 40)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(200):
    ctx.counts_device(dpos, dact, 40)
e1.record()
torch.cuda.synchronize()
print(f"k_small_loss back to back (device events, 200 launches): {e0.elapsed_time(e1) / 200 * 1e3:.1f} us per swarm")
t0 = time.perf_counter()
for _ in range(200):
    ctx.counts_view(pos)
print(f"counts_view (H2D, kernel, D2H, sync) from the host: {(time.perf_counter() - t0) / 200 * 1e6:.1f} us per swarm")
